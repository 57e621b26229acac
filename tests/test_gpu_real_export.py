"""The C++ loader (csrc/onnx_reader.cpp + voice.cpp) on a voice file this repository's writer did NOT produce:
the upstream module tree exported by torch's own ONNX exporter (tests/test_real_export_cpu.py describes the file:
anonymous `onnx::Conv_NNNN` flow weights, embedding table called `sid`, folded ElementwiseAffine scale, no
dp.flows.1, Constant nodes, 6 k nodes).  Loaded through piper_b200_load_voice it must synthesize what the oracle
computes from the plainly named file: durations bit-exact, audio within the fp32 / f16 tolerances."""
import numpy as np
import pytest

from common import S, voice_files, oracle, snr_db, fixture_ids
from oracle import refrng
from piper_b200.runtime import PiperCUDARuntime, Options
from test_real_export_cpu import exported

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,precision,snr_min,abs_max", [("tiny", "fp32", 80.0, 2e-4), ("medium", "fp32", 80.0, 2e-4),
                                                            ("medium", "f16", 58.0, 4e-3)])
def test_real_export_loads_and_matches_the_oracle(name, precision, snr_min, abs_max):
    spec = S.SPECS[name]
    ids = fixture_ids(2, spec.n_vocab)
    ns, ls, nw = S.FIXTURE_SCALES
    o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
    rt = PiperCUDARuntime(exported(name), voice_files(name)[1], Options(precision=precision))
    r = rt.synthesize_batch([ids], ns, ls, nw)
    assert np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))
    assert int(r.frame_lengths[0]) == o["y_len"]
    assert r.audio[0].shape == o["audio"].shape
    assert np.abs(r.audio[0] - o["audio"]).max() <= abs_max
    assert snr_db(o["audio"], r.audio[0]) >= snr_min
    # same result as the plainly named file through the same loader, up to the g * v / ||v|| rounding of the flow weights
    plain = PiperCUDARuntime(*voice_files(name), Options(precision=precision)).synthesize_batch([ids], ns, ls, nw)
    assert np.array_equal(plain.durations[0], r.durations[0])
    assert np.abs(plain.audio[0] - r.audio[0]).max() <= abs_max
    assert int(rt.info.num_symbols) == spec.n_vocab and int(rt.info.hidden) == spec.hidden
    rt.close()


@pytest.mark.parametrize("which", ["plain", "torch-export"])
def test_route_b_equals_route_a(which):
    """init(model:config:options:) (PiperMetalRuntime.swift:54-59) through the C ABI's route B - set_weight per
    initializer, add_node per graph node, set_config, finalize - must give the same bits as route A
    (piper_b200_load_voice on the same file).  The parser on the host side here is the oracle's pure-Python ONNX reader,
    standing in for the reference's PiperONNX."""
    from oracle.onnx_reader import load_model
    name = "medium"
    path = voice_files(name)[0] if which == "plain" else exported(name)
    m = load_model(path)
    ids = fixture_ids(2)
    ns, ls, nw = S.FIXTURE_SCALES
    a = PiperCUDARuntime(path, voice_files(name)[1], Options(precision="f16"))
    b = PiperCUDARuntime.from_parsed(m.initializers, m.nodes, sample_rate=S.SPECS[name].sample_rate,
                                     options=Options(precision="f16"))
    ra, rb = a.synthesize_batch([ids], ns, ls, nw), b.synthesize_batch([ids], ns, ls, nw)
    assert np.array_equal(ra.durations[0], rb.durations[0])
    assert np.array_equal(ra.audio[0], rb.audio[0])
    # n_params counts what the host pushed as weights: route B also pushes the float values of Constant nodes (under
    # their output names), route A keeps those out of the count
    n_const = sum(int(np.asarray(n.attrs["value"]).size) for n in m.nodes
                  if n.op_type == "Constant" and isinstance(n.attrs.get("value"), np.ndarray)
                  and n.attrs["value"].dtype == np.float32 and n.outputs and n.outputs[0] not in m.initializers)
    assert int(b.info.n_params) == int(a.info.n_params) + n_const
    assert int(b.info.num_symbols) == S.SPECS[name].n_vocab
    assert np.array_equal(b.synthesize(ids), a.synthesize(ids))       # config defaults arrived (set_config)
    a.close()
    b.close()
