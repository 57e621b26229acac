"""the generic ONNX-op executor (csrc/graph_exec.cu) through the C ABI, against the op-level numpy oracle
(oracle/onnx_numpy.py) node by node, against the model-level oracle, and against the fused pipeline.
Tolerances: integer / bool values bit-exact; float values max-abs <= 1e-4 * max(1, max|ref|) per node (fp32 summation
order and libm differences only); waveform max-abs <= 2e-4 like the fp32 parity mode."""
import os

import numpy as np
import pytest

from common import S, voice_files, oracle, snr_db, fixture_ids, VOICE_DIR
from oracle import onnx_numpy, refrng
from oracle.onnx_reader import load_model
from piper_b200.runtime import PiperCUDARuntime, Options, PiperB200Error
from test_real_export_cpu import exported

pytestmark = pytest.mark.gpu


def test_every_node_of_the_exported_graph_against_the_numpy_interpreter():
    name = "tiny"
    path = exported(name)
    m = load_model(path)
    ids = fixture_ids(2, S.SPECS[name].n_vocab)
    ns, ls, nw = S.FIXTURE_SCALES
    audio_ref, v = onnx_numpy.synthesize(m, ids, ns, ls, nw, keep_all=True)
    rt = PiperCUDARuntime(path, voice_files(name)[1], Options(precision="fp32"))
    out = rt.run_graph(ids, ns, ls, nw)
    assert out.reshape(-1).shape == audio_ref.shape
    assert np.abs(out.reshape(-1) - audio_ref).max() <= 2e-4
    checked, by_op = 0, {}
    for node in m.nodes:
        if node.op_type == "Constant":
            continue
        by_op.setdefault(node.op_type, []).append(node)
    for op, nodes in sorted(by_op.items()):
        for node in nodes[:3] + nodes[-2:]:          # the first three and the last two nodes of every op type
            for o in node.outputs:
                if not o:
                    continue
                ref = np.asarray(v[o])
                got = rt.run_graph(ids, ns, ls, nw, value_name=o)
                assert got.shape == ref.shape, (op, node.name, got.shape, ref.shape)
                if ref.dtype.kind == "f":
                    assert got.dtype == np.float32
                    tol = 1e-4 * max(1.0, float(np.abs(ref).max()) if ref.size else 1.0)
                    assert ref.size == 0 or float(np.abs(got - ref).max()) <= tol, (op, node.name)
                else:
                    assert np.array_equal(got.astype(np.int64), ref.astype(np.int64)), (op, node.name)
                checked += 1
    assert len(by_op) == len(S.REFERENCE_GRAPH_OPS) and checked >= 150
    rt.close()


@pytest.mark.parametrize("name", ["tiny", "medium"])
def test_generic_synthesize_equals_fused_pipeline_and_oracle(name):
    path = exported(name)
    ids_a, ids_b = fixture_ids(2, S.SPECS[name].n_vocab), fixture_ids(1, S.SPECS[name].n_vocab)
    ns, ls, nw = S.FIXTURE_SCALES
    rt = PiperCUDARuntime(path, voice_files(name)[1], Options(precision="fp32"))
    fused = rt.synthesize_batch([ids_a, ids_b], ns, ls, nw)
    gen = rt.synthesize_batch([ids_a, ids_b], ns, ls, nw, generic=True)
    for ids, a, b in zip((ids_a, ids_b), fused.audio, gen.audio):
        o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
        assert a.shape == b.shape == o["audio"].shape
        assert np.abs(a - b).max() <= 2e-4 and np.abs(b - o["audio"]).max() <= 2e-4
        assert snr_db(o["audio"], b) >= 80.0
    assert rt.last_run_timings["kernel_launches"] > 1000       # node by node: thousands of small launches
    rt.close()


def _patched(src: str, old: bytes, new: bytes, tag: str) -> str:
    assert len(old) == len(new)
    data = open(src, "rb").read()
    assert old in data
    dst = os.path.join(VOICE_DIR, f"patched-{tag}-" + os.path.basename(src))
    with open(dst, "wb") as f:
        f.write(data.replace(old, new))
    return dst


def test_unrecognised_graph_falls_back_to_the_generic_executor():
    """A graph the fused pipeline cannot bind (a LayerNorm scale renamed everywhere, same length so the protobuf stays
    valid; convolution weights would still be found through their nodes) loads anyway and synthesizes through the generic executor, like the reference interprets any graph of its ops."""
    name = "tiny"
    path = _patched(exported(name), b"norm_layers_1.0.gamma", b"norm_layers_1.0.gammx", "renamed")
    ids = fixture_ids(2, S.SPECS[name].n_vocab)
    ns, ls, nw = S.FIXTURE_SCALES
    rt = PiperCUDARuntime(path, voice_files(name)[1], Options(precision="fp32"))
    assert int(rt.info.hidden) == 0 and int(rt.info.sample_rate) == S.SPECS[name].sample_rate     # only the config is known
    a = rt.synthesize(ids, ns, ls, nw)
    o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
    assert a.shape == o["audio"].shape and np.abs(a - o["audio"]).max() <= 2e-4
    with pytest.raises(PiperB200Error) as e:            # modes the generic path does not have fail loudly
        rt.synthesize_batch([ids], ns, ls, nw, pcm16=True)
    assert e.value.code == 4
    rt.close()


def test_unknown_op_is_reported_as_unsupported():
    name = "tiny"
    path = _patched(exported(name), b"\x03Erf", b"\x03Erg", "badop")       # op_type field (length 3) of the Erf nodes
    ids = fixture_ids(1, S.SPECS[name].n_vocab)
    rt = PiperCUDARuntime(path, voice_files(name)[1], Options(precision="fp32"))
    with pytest.raises(PiperB200Error) as e:
        rt.run_graph(ids, *S.FIXTURE_SCALES)
    assert e.value.code == 4 and "Erg" in str(e.value)
    rt.close()
