"""The op-level oracle (oracle/onnx_numpy.py: numpy restatement of GraphExecutor.executeNode, GraphExecutor.swift:591-2663)
against the model-level oracle (oracle/vits_oracle.py) on the same exported file: two independent restatements - one
of the 50 ONNX ops evaluated node by node over torch's own export, one of the VITS network - must give the same
waveform, durations and frame count.  Neither is a reference-held vector (parity stays unpinned),
but a mistake in either restatement would have to be made twice, in different code, to pass."""
import numpy as np
import pytest

from common import S, oracle, snr_db, fixture_ids
from oracle import onnx_numpy, refrng
from oracle.onnx_reader import load_model
from test_real_export_cpu import exported


@pytest.mark.parametrize("name,factor", [("tiny", 1), ("tiny", 4), ("medium", 2)])
def test_op_interpreter_agrees_with_the_model_oracle(name, factor):
    m = load_model(exported(name))
    ids = fixture_ids(factor, S.SPECS[name].n_vocab)
    ns, ls, nw = S.FIXTURE_SCALES
    audio, v = onnx_numpy.synthesize(m, ids, ns, ls, nw, keep_all=True)
    o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
    assert audio.shape == o["audio"].shape
    assert snr_db(o["audio"], audio) >= 100.0
    assert np.abs(audio - o["audio"]).max() <= 2e-5
    # the integer path: the one Ceil node of the graph holds ceil(w) per phoneme id
    ceil_out = [n.outputs[0] for n in m.nodes if n.op_type == "Ceil"]
    assert len(ceil_out) == 1
    assert np.array_equal(v[ceil_out[0]].reshape(-1).astype(np.int64), np.asarray(o["w_ceil"]).reshape(-1).astype(np.int64))


def test_op_semantics_small_cases():
    """Known answers of the ops whose ONNX semantics are easy to get wrong (negative-step Slice, integer Div, Cast
    truncation, GatherND / ScatterND, NonZero order, ConvTranspose padding)."""
    from oracle.onnx_reader import Node
    g = onnx_numpy.NumpyGraph.__new__(onnx_numpy.NumpyGraph)
    g.seed = 1234

    def run(op, ins, **attrs):
        n = Node()
        n.op_type, n.attrs, n.outputs = op, attrs, ["y"]
        return g._op(n, ins)
    x = np.arange(10, dtype=np.float32).reshape(1, 10)
    i64 = lambda *v: np.array(v, np.int64)
    assert np.array_equal(run("Slice", [x, i64(-1), i64(-(2 ** 63) + 1), i64(1), i64(-1)]), x[:, ::-1])
    assert np.array_equal(run("Slice", [x, i64(2), i64(100), i64(1), i64(3)]), x[:, 2::3])
    assert np.array_equal(run("Div", [i64(-7, 7), i64(2, 2)]), i64(-3, 3))
    assert np.array_equal(run("Cast", [np.array([-1.7, 1.7], np.float32)], to=7), i64(-1, 1))
    d = np.arange(12, dtype=np.float32).reshape(3, 4)
    assert np.array_equal(run("GatherND", [d, i64(0, 1, 2, 3).reshape(2, 2)]), np.array([1, 11], np.float32))
    assert np.array_equal(run("ScatterND", [d, i64(1).reshape(1, 1), np.zeros((1, 4), np.float32)])[1], np.zeros(4, np.float32))
    assert np.array_equal(run("NonZero", [np.array([[0, 1], [1, 1]], bool)]), i64(0, 1, 1, 1, 0, 1).reshape(2, 3))
    import torch
    import torch.nn.functional as F
    rng = np.random.default_rng(3)
    xx, ww, bb = rng.standard_normal((1, 4, 9), np.float32), rng.standard_normal((4, 3, 8), np.float32), rng.standard_normal(3, np.float32)
    ref = F.conv_transpose1d(torch.from_numpy(xx), torch.from_numpy(ww), torch.from_numpy(bb), stride=4, padding=2).numpy()
    assert np.allclose(run("ConvTranspose", [xx, ww, bb], strides=[4], pads=[2, 2]), ref, atol=1e-5)
    wc = rng.standard_normal((6, 2, 3), np.float32)
    refc = F.conv1d(torch.from_numpy(xx), torch.from_numpy(wc), None, dilation=2, padding=2, groups=2).numpy()
    assert np.allclose(run("Conv", [xx, wc], dilations=[2], pads=[2, 2], group=2), refc, atol=1e-5)
