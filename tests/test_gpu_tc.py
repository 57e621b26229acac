"""GPU tests of the tensor-core (tcgen05 / TMEM / TMA) back end through the C ABI.

Kernel level : piper_b200_selftest_conv - the tcgen05 implicit-GEMM convolution against the fp32
               CUDA-core kernel on identical, exactly representable operands.  Tolerance: 2e-3 * max|ref|
               for the fp32 outputs (observed: bit-exact), 2e-2 for the 16-bit gate output.
Graph level  : precision modes "f16" / "bf16" against the fp32 oracle.  The text encoder, duration
               predictor and length regulator stay in fp32, so durations / y_len must be bit-exact;
               waveform tolerance (stated, measured on B200, profiles/r01_tc_parity_vs_oracle.log):
                 f16  : SNR >= 58 dB, max-abs <= 4e-3      (observed 63-65 dB, 1.2e-3)
                 bf16 : SNR >= 40 dB, max-abs <= 3e-2      (observed 45-47 dB, 9.5e-3)
"""
import numpy as np
import pytest

from common import S, voice_files, oracle, snr_db, fixture_ids
from oracle import refrng
from piper_b200.runtime import PiperCUDARuntime, Options, PiperB200Error
from piper_b200.synth_voice import synthetic_utterances

pytestmark = pytest.mark.gpu

_rt = {}


def rt(name, precision):
    key = (name, precision)
    if key not in _rt:
        mp, cp = voice_files(name)
        _rt[key] = PiperCUDARuntime(mp, cp, Options(precision=precision))
    return _rt[key]


CASES = [  # mode, cin, cout, k, dil, stride, rows
    (0, 64, 64, 1, 1, 1, 300), (0, 64, 64, 3, 1, 1, 1000), (0, 32, 32, 7, 3, 1, 5000), (0, 32, 32, 3, 1, 1, 129),
    (0, 96, 192, 1, 1, 1, 700), (0, 192, 96, 1, 1, 1, 700), (0, 192, 256, 7, 1, 1, 515), (0, 128, 128, 5, 6, 1, 2000),
    (0, 192, 512, 7, 1, 1, 400), (0, 256, 256, 11, 5, 1, 900), (0, 64, 64, 7, 12, 1, 4000), (0, 32, 32, 11, 5, 1, 1),
    (1, 192, 384, 5, 1, 1, 1000), (2, 192, 384, 1, 1, 1, 1000),
    (3, 256, 128, 16, 1, 8, 300), (3, 128, 64, 16, 1, 8, 1000), (3, 64, 32, 8, 1, 4, 3000), (3, 128, 64, 4, 1, 2, 500),
]


@pytest.mark.parametrize("prec", ["bf16", "f16"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "m%d_%dto%d_k%d_d%d_s%d_r%d" % c)
def test_tcgen05_conv_matches_fp32_kernel(case, prec):
    r = rt("tiny", "fp32").selftest_conv(*case, precision=prec)
    tol = (2e-2 if case[0] == 1 else 2e-3) * max(1.0, r["max_abs_ref"])
    assert r["max_abs_ref"] > 0.5
    assert r["max_abs_diff"] <= tol, r


KAT = [  # transposed, cin, cout, k, dil, stride, rows  - the node shapes of the medium / high generator, flow, encoder
    (0, 192, 256, 7, 1, 1, 515), (0, 192, 384, 5, 1, 1, 700), (0, 192, 192, 1, 1, 1, 300), (0, 192, 768, 3, 1, 1, 260),
    (0, 128, 128, 3, 1, 1, 1000), (0, 64, 64, 5, 6, 1, 2100), (0, 32, 32, 7, 12, 1, 4000), (0, 256, 256, 11, 5, 1, 900),
    (0, 32, 32, 3, 1, 1, 1), (0, 64, 64, 7, 3, 1, 129),
    (1, 256, 128, 16, 1, 8, 300), (1, 128, 64, 16, 1, 8, 777), (1, 64, 32, 8, 1, 4, 1500), (1, 512, 256, 16, 1, 8, 100),
    (1, 128, 64, 4, 1, 2, 500),
]


@pytest.mark.parametrize("prec", ["fp32", "f16", "bf16"])
@pytest.mark.parametrize("case", KAT, ids=lambda c: "%s_%dto%d_k%d_d%d_s%d_r%d" % (("ct" if c[0] else "conv",) + c[1:]))
def test_conv_kernels_against_torch(case, prec):
    """Per-op known-answer test: the product's convolution kernels - conv_f32_*_kernel in fp32,
    conv_tc_persist_kernel (tcgen05) in f16 / bf16 - through piper_b200_conv1d on seeded random operands against
    torch.nn.functional.conv1d / conv_transpose1d on the CPU.  In the 16-bit modes torch sees the same rounded operands
    in float64, so what is left is fp32 accumulation order: tolerance 2e-5 * sqrt(cin*k) (observed ~1e-6)."""
    import torch
    import torch.nn.functional as F
    tr, cin, cout, k, dil, stride, rows = case
    g = torch.Generator().manual_seed(1000 + cin + 7 * cout + 31 * k + rows)
    x = torch.randn(rows, cin, generator=g)
    w = torch.randn((cin, cout, k) if tr else (cout, cin, k), generator=g) / float(np.sqrt(cin * k / (stride if tr else 1)))
    b = torch.randn(cout, generator=g) * 0.1
    if prec != "fp32":
        dt = torch.float16 if prec == "f16" else torch.bfloat16
        x, w = x.to(dt).float(), w.to(dt).float()
    xt = x.t().unsqueeze(0).double()
    if tr:
        ref = F.conv_transpose1d(xt, w.double(), b.double(), stride=stride, padding=(k - stride) // 2)
    else:
        ref = F.conv1d(xt, w.double(), b.double(), dilation=dil, padding=dil * (k - 1) // 2)
    ref = ref[0].t().numpy()
    out = rt("tiny", "fp32").conv1d(x.numpy(), w.numpy(), b.numpy(), dilation=dil, stride=stride, transposed=bool(tr),
                                    precision=prec)
    assert out.shape == ref.shape
    tol = 2e-5 * float(np.sqrt(cin * k)) * max(1.0, float(np.abs(ref).max()))
    assert float(np.abs(out - ref).max()) <= tol, (float(np.abs(out - ref).max()), tol)


def test_selftest_rejects_bad_shapes():
    with pytest.raises(PiperB200Error):
        rt("tiny", "fp32").selftest_conv(0, 24, 64, 3, 1, 1, 100, precision="bf16")    # cin not a multiple of 32
    with pytest.raises(PiperB200Error):
        rt("tiny", "fp32").selftest_conv(0, 64, 64, 3, 1, 1, 100, precision="fp32")


TOL = {"f16": (58.0, 4e-3), "bf16": (40.0, 3e-2)}


@pytest.mark.parametrize("prec", ["f16", "bf16"])
@pytest.mark.parametrize("name,factor,sid", [("medium", 1, None), ("medium", 8, None), ("high", 2, None),
                                             ("medium-multispeaker", 2, 7), ("x_low", 4, None)])
def test_tc_modes_against_oracle(name, factor, sid, prec):
    ids = fixture_ids(factor)
    ns, ls, nw = S.FIXTURE_SCALES
    o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main, sid=sid)
    r = rt(name, prec).synthesize_batch([ids], ns, ls, nw, speaker_ids=None if sid is None else [sid], keep_stages=True)
    assert np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))       # front end is fp32: bit-exact
    assert int(r.frame_lengths[0]) == o["y_len"]
    zp = rt(name, prec).get_stage("z_p", 0)
    assert np.abs(zp - o["z_p"].reshape(zp.shape)).max() < 5e-5             # still fp32 up to here
    snr_min, abs_max = TOL[prec]
    a = r.audio[0]
    assert a.shape == o["audio"].shape
    assert snr_db(o["audio"], a) >= snr_min
    assert np.abs(a - o["audio"]).max() <= abs_max
    t = rt(name, prec).last_run_timings
    assert t["kernel_launches"] > 50


def test_tc_mode_unsupported_voice_fails_loudly():
    """tiny has 16/8-channel convolutions that do not fit the tcgen05 tile: error, no silent fallback."""
    with pytest.raises(PiperB200Error) as e:
        rt("tiny", "f16").synthesize(S.FIXTURE_IDS)
    assert e.value.code == 4


def test_tc_batch_invariance_and_masks():
    """Ragged batch in f16 mode: every utterance equals its B=1 result (gap rows stay zero)."""
    utts = [fixture_ids(1), fixture_ids(3), [1, 5, 2], fixture_ids(2)[:-3]]
    r = rt("medium", "f16")
    batch = r.synthesize_batch(utts, *S.FIXTURE_SCALES)
    for u, a, d in zip(utts, batch.audio, batch.durations):
        s = r.synthesize_batch([u], *S.FIXTURE_SCALES)
        assert np.array_equal(d, s.durations[0])
        assert np.array_equal(a, s.audio[0])


def test_tc_full_size_batch_properties():
    """BASELINE config 3 shape in f16 mode: size-independent properties + oracle spot check."""
    utts = synthetic_utterances(256, 256)
    r = rt("medium", "f16")
    out = r.synthesize_batch(utts, *S.FIXTURE_SCALES)
    ref = rt("medium", "fp32").synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert np.array_equal(out.frame_lengths, ref.frame_lengths)
    assert all(np.array_equal(x, y) for x, y in zip(out.durations, ref.durations))
    allv = np.concatenate(out.audio)
    assert np.isfinite(allv).all() and np.abs(allv).max() <= 1.0
    assert snr_db(np.concatenate(ref.audio), allv) >= 58.0
    again = r.synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert all(np.array_equal(x, y) for x, y in zip(out.audio, again.audio))
    # a batch this size leaves for the host in row ranges while the last generator stage is still running; a PCM16
    # request takes the single-launch, single-copy path: the float audio of the two must be the same bits
    whole = r.synthesize_batch(utts, *S.FIXTURE_SCALES, pcm16=True)
    assert all(np.array_equal(x, y) for x, y in zip(out.audio, whole.audio))


def test_tc_ragged_batch_leaves_in_row_ranges():
    """Host-bound audio of a ragged batch (lengths 20..400 ids, multi-speaker voice) goes out in row ranges under the
    last generator stage; the PCM16 request keeps the single launch + single copy: same float bits, same durations,
    and PCM16 follows the WavFileWriter rule on those floats."""
    utts = synthetic_utterances(48, 256, 20, 400, seed=5)
    sids = [b % 4 for b in range(len(utts))]
    r = rt("medium-multispeaker", "f16")
    ranged = r.synthesize_batch(utts, *S.FIXTURE_SCALES, speaker_ids=sids)
    whole = r.synthesize_batch(utts, *S.FIXTURE_SCALES, speaker_ids=sids, pcm16=True)
    assert sum(len(a) for a in ranged.audio) >= 1 << 22           # large enough for the ranged path
    assert np.array_equal(ranged.frame_lengths, whole.frame_lengths)
    for a, b, d0, d1, pc in zip(ranged.audio, whole.audio, ranged.durations, whole.durations, whole.pcm16):
        assert np.array_equal(d0, d1)
        assert np.array_equal(a, b)
        assert np.array_equal(pc, (np.clip(b, -1.0, 1.0) * np.float32(32767.0)).astype(np.int32).astype(np.int16))


@pytest.mark.gpu
@pytest.mark.timeout(120)
@pytest.mark.parametrize("prec,min_snr", [("f16", 58.0), ("bf16", 40.0)])
def test_tc_long_form_high_voice(prec, min_snr):
    """BASELINE config 4 shape in the tensor-core modes: one long utterance of the ResBlock1 voice, so that every
    fused ResBlock1 unit launch walks several windows per CTA (the 28-id parity cases fit in one); against the fp32
    CUDA path (durations exact, audio SNR), and twice (idempotence).
    This is also the regression case of the round-1 hang (tools/hang_repro.py: `high` voice, one
    600-id utterance, seed 11): the ResBlock1 unit launches run in the streamed weight-ring order with ring = k = 11,
    the geometry whose parity-aliased mbarrier wait hung before the fix - twice per precision, under the timeout."""
    ids = synthetic_utterances(1, 256, 600, 600, seed=11)[0]
    ref = rt("high", "fp32").synthesize_batch([ids], *S.FIXTURE_SCALES)
    r = rt("high", prec)
    out = r.synthesize_batch([ids], *S.FIXTURE_SCALES)
    assert np.array_equal(out.durations[0], ref.durations[0])
    assert np.isfinite(out.audio[0]).all() and out.audio[0].shape == ref.audio[0].shape
    assert snr_db(ref.audio[0], out.audio[0]) >= min_snr
    again = r.synthesize_batch([ids], *S.FIXTURE_SCALES)
    assert np.array_equal(out.audio[0], again.audio[0])


@pytest.mark.timeout(240)
def test_flow_window_operand_path_against_oracle(tmp_path):
    """PIPER_B200_FLOW_WINDOW=1 (off by default): the WN in_layer convolutions take their operand as a row window of the
    group-planar residual copy (conv_tc.h, TcConvArgs::in_planar).  The switch is read once per process, so the run is a
    child process; its audio is checked against the oracle with the f16 tolerance and against the default path."""
    import os
    import subprocess
    import sys
    ids = fixture_ids(8)
    ns, ls, nw = S.FIXTURE_SCALES
    out = tmp_path / "win.npy"
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    code = (
        "import sys, numpy as np\n"
        "sys.path[:0] = [%r, %r, %r]\n"
        "from common import S, voice_files, fixture_ids\n"
        "from piper_b200.runtime import PiperCUDARuntime, Options\n"
        "mp, cp = voice_files('medium')\n"
        "rt = PiperCUDARuntime(mp, cp, Options(precision='f16'))\n"
        "r = rt.synthesize_batch([fixture_ids(8), fixture_ids(2)], *S.FIXTURE_SCALES)\n"
        "np.save(%r, r.audio[0])\n"
        "rt.close()\n" % (root, os.path.join(root, "piper-swift_b200"), here, str(out)))
    env = dict(os.environ, PIPER_B200_FLOW_WINDOW="1")
    subprocess.run([sys.executable, "-c", code], check=True, env=env, timeout=200)
    a = np.load(out)
    o = oracle("medium").infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
    snr_min, abs_max = TOL["f16"]
    assert a.shape == o["audio"].shape
    assert snr_db(o["audio"], a) >= snr_min
    assert np.abs(a - o["audio"]).max() <= abs_max
    d = rt("medium", "f16").synthesize_batch([ids, fixture_ids(2)], ns, ls, nw).audio[0]
    assert np.abs(a - d).max() <= 2e-3     # same arithmetic, K blocks and taps accumulated in the other order
