"""GPU parity tests: the CUDA path, called through the C ABI (piper_b200.runtime -> ctypes ->
libpiper_b200.so), against the oracle on the same seeded inputs and against the committed
golden vectors.  Tolerances (fp32 mode):
  durations / y_len / frame index : bit-exact
  stage tensors                   : max-abs 5e-5
  audio                           : max-abs 2e-4 and SNR >= 80 dB
"""
import glob
import os

import numpy as np
import pytest

from common import S, ROOT, voice_files, oracle, snr_db, fixture_ids
from oracle import refrng
from oracle.vits_oracle import VitsOracle
from piper_b200 import runtime
from piper_b200.runtime import PiperCUDARuntime, Options, PiperB200Error
from piper_b200.synth_voice import synthetic_utterances

pytestmark = pytest.mark.gpu

_rt = {}


def rt(name, precision="fp32"):
    key = (name, precision)
    if key not in _rt:
        mp, cp = voice_files(name)
        _rt[key] = PiperCUDARuntime(mp, cp, Options(precision=precision))
    return _rt[key]


def oracle_run(name, ids, sid=None, scales=S.FIXTURE_SCALES, nd=None, nm=None):
    ns, ls, nw = scales
    nd = refrng.noise_dp(len(ids)) if nd is None else nd
    kw = dict(noise_main_fn=refrng.noise_main) if nm is None else dict(noise_main=nm)
    return oracle(name).infer(ids, ns, ls, nw, nd, sid=sid, **kw)


@pytest.mark.parametrize("name,factor,sid", [("tiny", 1, None), ("tiny", 4, None), ("medium", 1, None),
                                             ("medium", 2, None), ("medium", 8, None), ("high", 2, None),
                                             ("medium-multispeaker", 2, 7), ("x_low", 4, None)])
def test_stage_and_audio_parity(name, factor, sid):
    ids = fixture_ids(factor, S.SPECS[name].n_vocab)
    o = oracle_run(name, ids, sid)
    r = rt(name).synthesize_batch([ids], *S.FIXTURE_SCALES, speaker_ids=None if sid is None else [sid],
                                  keep_stages=True)
    assert np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))
    assert int(r.frame_lengths[0]) == o["y_len"]
    for st in ("x", "m_p", "logs_p", "logw", "z_p", "z"):
        g = rt(name).get_stage(st, 0)
        assert np.abs(g - o[st].reshape(g.shape)).max() < 5e-5, st
    a = r.audio[0]
    assert a.shape == o["audio"].shape
    assert np.abs(a - o["audio"]).max() < 2e-4
    assert snr_db(o["audio"], a) > 80.0


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))),
                         ids=lambda p: os.path.basename(p))
def test_against_committed_golden(path):
    g = np.load(path)
    name = os.path.basename(path).rsplit("_x", 1)[0]
    sid = int(g["sid"])
    ns, ls, nw = (float(v) for v in g["scales"])
    r = rt(name).synthesize_batch([g["ids"].tolist()], ns, ls, nw, speaker_ids=None if sid < 0 else [sid])
    assert np.array_equal(r.durations[0], g["w_ceil"])
    assert int(r.frame_lengths[0]) == int(g["y_len"])
    assert np.abs(r.audio[0][:2048] - g["audio_head"]).max() < 1e-3
    assert np.abs(r.audio[0] - g["audio"].astype(np.float32)).max() < 3e-3


def test_reference_api_surface():
    """synthesize / synthesizeStream / lastRunTimings of PiperMetalRuntime.swift:33,62,84."""
    r = rt("medium")
    a = r.synthesize(S.FIXTURE_IDS)                       # defaults come from the voice config (:63-65)
    b = r.synthesize(S.FIXTURE_IDS, noise_scale=0.667, length_scale=1.0, noise_w=0.8)
    assert a.dtype == np.float32 and a.ndim == 1 and a.size > 0 and np.isfinite(a).all()   # SmokeTests.swift:18-19
    assert np.array_equal(a, b)
    t = r.last_run_timings
    assert t["kernel_launches"] > 100 and t["gpu_ms"] > 0
    chunks = list(r.synthesize_stream(S.FIXTURE_IDS, chunk_size=2048))
    assert all(c.samples.size == 2048 for c in chunks[:-1]) and chunks[-1].is_final
    assert [c.start_sample_index for c in chunks] == list(range(0, a.size, 2048))
    assert np.array_equal(np.concatenate([c.samples for c in chunks]), a)
    assert chunks[0].format.sample_rate == 22050


@pytest.mark.parametrize("name,precision,chunk", [("medium", "fp32", 2048), ("medium", "fp32", 5000),
                                                  ("medium", "f16", 4096), ("medium", "bf16", 8192),
                                                  ("high", "fp32", 4096), ("high", "f16", 3000),
                                                  ("x_low", "f16", 2048), ("tiny", "fp32", 512)])
def test_incremental_stream_is_bit_identical(name, precision, chunk):
    """true streaming.  piper_b200_stream_next decodes `chunk` samples at a time (plus the
    generator's receptive-field halo); the concatenation must equal the one-shot waveform bit for bit, in every
    precision mode, and the chunk boundaries must follow synthesizeStream (PiperMetalRuntime.swift:84-121)."""
    r = rt(name, precision)
    ids = fixture_ids(4, S.SPECS[name].n_vocab)
    a = r.synthesize(ids)
    chunks = list(r.synthesize_stream(ids, chunk_size=chunk))
    assert len(chunks) == -(-a.size // chunk) and len(chunks) > 2
    assert all(c.samples.size == chunk and not c.is_final for c in chunks[:-1]) and chunks[-1].is_final
    assert [c.start_sample_index for c in chunks] == list(range(0, a.size, chunk))
    assert np.array_equal(np.concatenate([c.samples for c in chunks]), a)
    assert r.last_run_timings["kernel_launches"] > 0
    # the C-ABI state machine: no stream open -> ERR_STATE; a batch of two cannot stream
    import ctypes as C
    rs, fin = runtime.Result(), C.c_int32(0)
    assert r._lib.piper_b200_stream_next(r._ctx, 8, C.byref(rs), C.byref(fin)) == 6   # PIPER_B200_ERR_STATE
    with pytest.raises(PiperB200Error):
        r.synthesize_packed(np.asarray(ids + ids, np.int64), np.asarray([len(ids)] * 2, np.int64), stream=True)
    assert np.array_equal(r.synthesize(ids), a)


def test_ipa_front_door_and_profile_flag():
    """idsFromIPA -> synthesize (PiperCLI.swift:115-136 -> :173) through the context, and the per-class kernel
    timings the bench reads (PIPER_B200_FLAG_PROFILE)."""
    r = rt("medium")
    ids = r.ipa_to_ids("hello world.")
    assert ids[0] == 1 and ids[-1] == 2 and ids[2:-1:2] == [0] * 12 and len(ids) == 2 * 12 + 2
    assert np.array_equal(r.synthesize_ipa("hello world."), r.synthesize(ids))
    with pytest.raises(PiperB200Error):
        r.ipa_to_ids("h\u00e9llo")
    arr = np.asarray(ids, dtype=np.int64)
    out = r.synthesize_packed(arr, np.asarray([len(ids)], dtype=np.int64), profile=True)
    t = r.last_run_timings
    assert sum(t["class_launches"]) > 0 and sum(t["class_ms"]) > 0 and sum(t["class_ms"]) <= t["gpu_ms"] * 1.05
    assert out.total_samples > 0


def test_injected_noise_overrides():
    """The reference's `overrides` hook (GraphExecutor.swift:190-193): noise as inputs."""
    rng = np.random.default_rng(7)
    ids = fixture_ids(2)
    nd = rng.standard_normal((2, len(ids))).astype(np.float32)
    nm = rng.standard_normal((192, 400)).astype(np.float32)
    o = oracle_run("medium", ids, nd=nd, nm=nm)
    r = rt("medium").synthesize_batch([ids], *S.FIXTURE_SCALES, noise_dp=[nd], noise_main=[nm])
    assert np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))
    assert np.abs(r.audio[0] - o["audio"]).max() < 2e-4


def test_scales_change_the_result_like_the_oracle():
    ids = fixture_ids(1)
    for scales in ((0.3, 1.4, 0.5), (1.0, 0.7, 1.0)):
        o = oracle_run("medium", ids, scales=scales)
        r = rt("medium").synthesize_batch([ids], *scales)
        assert np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))
        assert np.abs(r.audio[0] - o["audio"]).max() < 2e-4


def test_ragged_batch_equals_single_utterances():
    """Batch invariance: packed rows with zero gaps give each utterance its B=1 result."""
    utts = [fixture_ids(1), fixture_ids(3), [1, 5, 2], fixture_ids(2)[:-3], [1, 0, 9, 0, 2]]
    r = rt("medium")
    batch = r.synthesize_batch(utts, *S.FIXTURE_SCALES)
    for u, a, d in zip(utts, batch.audio, batch.durations):
        s = r.synthesize_batch([u], *S.FIXTURE_SCALES)
        assert np.array_equal(d, s.durations[0])
        assert np.array_equal(a, s.audio[0])          # same kernels, same order of operations
    o = oracle_run("medium", utts[1])
    assert np.abs(batch.audio[1] - o["audio"]).max() < 2e-4


def test_multispeaker_batch_mixed_speakers():
    utts = [fixture_ids(1), fixture_ids(2), fixture_ids(1)]
    sids = [3, 900, 0]
    r = rt("medium-multispeaker").synthesize_batch(utts, *S.FIXTURE_SCALES, speaker_ids=sids)
    for u, sid, a, d in zip(utts, sids, r.audio, r.durations):
        o = oracle_run("medium-multispeaker", u, sid)
        assert np.array_equal(d, o["w_ceil"].astype(np.int32))
        assert np.abs(a - o["audio"]).max() < 2e-4
    with pytest.raises(PiperB200Error):
        rt("medium-multispeaker").synthesize_batch(utts, *S.FIXTURE_SCALES)   # sid required


def test_length_regulator_bit_exact():
    """Integer part alone on oracle-provided logw, incl. edge cases."""
    import torch
    rng = np.random.default_rng(3)
    logws = [rng.normal(0.9, 0.6, T).astype(np.float32) for T in (1, 2, 14, 257, 1000)]
    logws.append(np.full(5, -200.0, np.float32))              # all durations zero -> y_len clipped to 1
    logws.append(np.log(np.array([1.0, 2.0, 3.0, 4.0], np.float32)))   # exact integers
    for ls in (1.0, 0.37, 2.5):
        dur, fl, idx = rt("tiny").length_regulate(logws, ls)
        for lw, d, f, ix in zip(logws, dur, fl, idx):
            w_ceil, y_len, ref_idx = VitsOracle.length_regulate(torch.from_numpy(lw).view(1, 1, -1), ls)
            assert np.array_equal(d, w_ceil.numpy().astype(np.int32))
            assert int(f) == y_len
            assert np.array_equal(ix, ref_idx.numpy().astype(np.int32))


def test_random_normal_matches_reference_generator():
    g = rt("tiny").random_normal(4096, 1234)
    ref = refrng.random_normal_like(4096, 1234)
    assert np.abs(g - ref).max() < 2e-5
    assert not np.array_equal(g, rt("tiny").random_normal(4096, 99))


def test_pcm16_rule():
    r = rt("medium").synthesize_batch([fixture_ids(1)], *S.FIXTURE_SCALES, pcm16=True)
    a = r.audio[0]
    expect = (np.clip(a, -1, 1) * np.float32(32767)).astype(np.int32).astype(np.int16)   # WavFileWriter.swift:20-30
    assert np.array_equal(r.pcm16[0], expect)


def test_error_behaviour():
    r = rt("tiny")
    with pytest.raises(PiperB200Error) as e:
        r.synthesize_batch([], 0.667, 1.0, 0.8)
    assert e.value.code == 1
    with pytest.raises(PiperB200Error):
        r.synthesize_batch([[]], 0.667, 1.0, 0.8)
    with pytest.raises(PiperB200Error) as e:
        PiperCUDARuntime("/nonexistent.onnx", voice_files("tiny")[1])
    assert e.value.code == 2
    bad = os.path.join(os.path.dirname(voice_files("tiny")[0]), "truncated.onnx")
    with open(voice_files("tiny")[0], "rb") as f, open(bad, "wb") as g:
        g.write(f.read()[:1000])
    with pytest.raises(PiperB200Error) as e:
        PiperCUDARuntime(bad, voice_files("tiny")[1])
    assert e.value.code in (3, 4)
    # out-of-range ids embed as zero rows (gather.metal OOB rule) and still synthesize
    a = r.synthesize([1, 9999, 2])
    assert np.isfinite(a).all() and a.size > 0


def test_full_size_batch_properties():
    """BASELINE config 3 shape (256 utterances, ~100 ids each): size-independent properties."""
    utts = synthetic_utterances(256, 256)
    r = rt("medium")
    out = r.synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert out.total_frames == int(out.frame_lengths.sum())
    for u, a, d, f in zip(utts, out.audio, out.durations, out.frame_lengths):
        assert d.shape[0] == len(u) and d.min() >= 1
        assert int(d.sum()) == int(f) and a.size == int(f) * 256
    allv = np.concatenate(out.audio)
    assert np.isfinite(allv).all() and np.abs(allv).max() <= 1.0
    again = r.synthesize_batch(utts, *S.FIXTURE_SCALES)                      # idempotence
    assert all(np.array_equal(x, y) for x, y in zip(out.audio, again.audio))
    for i in (0, 97, 255):                                                   # batch invariance + oracle spot checks
        s = r.synthesize_batch([utts[i]], *S.FIXTURE_SCALES)
        assert np.array_equal(s.audio[0], out.audio[i])
    o = oracle_run("medium", utts[5])
    assert np.array_equal(out.durations[5], o["w_ceil"].astype(np.int32))
    assert np.abs(out.audio[5] - o["audio"]).max() < 2e-4


def test_long_form_high_voice():
    """BASELINE config 4 shape: high voice, one long utterance (kept to 600 ids for oracle time)."""
    ids = synthetic_utterances(1, 256, 600, 600, seed=11)[0]
    o = oracle_run("high", ids)
    r = rt("high").synthesize_batch([ids], *S.FIXTURE_SCALES)
    mism = int((r.durations[0] != o["w_ceil"].astype(np.int32)).sum())
    assert mism == 0, f"{mism} of {len(ids)} durations differ"
    assert snr_db(o["audio"], r.audio[0]) > 80.0
