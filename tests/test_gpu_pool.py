"""GPU tests of the multi-GPU product path and of parity at the benchmark's own scale.

  * PiperCUDAPool / caller-owned output arena / PCM16-only output on however many GPUs the box has (one is enough:
    two contexts on the same device exercise the same code); with >= 2 GPUs the sharded result must equal the
    one-GPU result bit for bit.
  * BASELINE config 3 at full size: ALL 256 utterances against the oracle (not a spot check), f16 and fp32 directly.
  * BASELINE config 4 as stated: the high voice on one 2000-id utterance against the oracle.
"""
import multiprocessing as mp
import os
from concurrent.futures import ProcessPoolExecutor

import numpy as np
import pytest

from common import S, voice_files, snr_db, VOICE_DIR
from piper_b200 import runtime
from piper_b200.runtime import PiperCUDARuntime, Options, PiperB200Error, HostArena
from piper_b200.sharding import PiperCUDAPool
from piper_b200.synth_voice import synthetic_utterances

pytestmark = pytest.mark.gpu

TOL = {"fp32": (80.0, 2e-4), "f16": (58.0, 4e-3), "bf16": (40.0, 3e-2)}

# ---- the oracle over a whole batch: one single-threaded worker process per host core ------------------------------------
_ORC = None


def _init(voice):
    global _ORC
    import torch
    torch.set_num_threads(1)
    from oracle.vits_oracle import VitsOracle
    _ORC = VitsOracle.from_file(voice_files(voice)[0])


def _run(job):
    from oracle import refrng
    ns, ls, nw = S.FIXTURE_SCALES
    out = {}
    for i, u, sid in job:
        o = _ORC.infer(u, ns, ls, nw, refrng.noise_dp(len(u)), noise_main_fn=refrng.noise_main, sid=sid)
        out[i] = (o["audio"].astype(np.float32), o["w_ceil"].astype(np.int32), int(o["y_len"]))
    return out


def oracle_batch(voice, utts, sids=None):
    voice_files(voice)
    cores = min(os.cpu_count() or 1, 32)
    jobs = [[] for _ in range(cores)]
    for i, u in enumerate(utts):
        jobs[i % cores].append((i, list(map(int, u)), None if sids is None else int(sids[i])))
    with ProcessPoolExecutor(max_workers=cores, mp_context=mp.get_context("spawn"), initializer=_init,
                             initargs=(voice,)) as ex:
        parts = list(ex.map(_run, [j for j in jobs if j]))
    out = {}
    for p in parts:
        out.update(p)
    return out


def check_against_oracle(res, ref, precision):
    snr_min, abs_max = TOL[precision]
    worst_snr, worst_abs = 1e9, 0.0
    for i, (audio, dur, y_len) in ref.items():
        assert np.array_equal(np.asarray(res.durations[i], np.int32), dur), f"durations of utterance {i}"
        assert int(res.frame_lengths[i]) == y_len
        assert res.audio[i].shape == audio.shape
        worst_snr = min(worst_snr, snr_db(audio, res.audio[i]))
        worst_abs = max(worst_abs, float(np.abs(res.audio[i] - audio).max()))
    assert worst_snr >= snr_min and worst_abs <= abs_max, (worst_snr, worst_abs)
    return worst_snr, worst_abs


@pytest.fixture(scope="module")
def config3():
    utts = synthetic_utterances(256, 256)
    return utts, oracle_batch("medium", utts)


@pytest.mark.parametrize("precision", ["f16", "fp32"])
def test_config3_all_256_utterances_against_the_oracle(config3, precision):
    utts, ref = config3
    rt = PiperCUDARuntime(*voice_files("medium"), Options(precision=precision))
    res = rt.synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert len(ref) == 256
    check_against_oracle(res, ref, precision)
    rt.close()


def test_pool_shards_and_gathers_into_one_arena(config3):
    """Two contexts (two GPUs when the box has them, else the same one twice): every utterance comes back in the original
    order, bit-identical to the single-context call, and the audio lies inside the pool's one pinned arena."""
    utts, ref = config3
    utts = utts[:96]
    n_dev = runtime.device_count()
    assert n_dev >= 1
    devices = [0, 1] if n_dev >= 2 else [0, 0]
    pool = PiperCUDAPool(*voice_files("medium"), devices=devices, options=Options(precision="f16"))
    res = pool.synthesize_batch(utts, *S.FIXTURE_SCALES)
    single = PiperCUDARuntime(*voice_files("medium"), Options(precision="f16"))
    one = single.synthesize_batch(utts, *S.FIXTURE_SCALES)
    lo, hi = pool._arena.ptr, pool._arena.ptr + pool._arena.nbytes
    for i in range(len(utts)):
        assert np.array_equal(res.durations[i], one.durations[i])
        assert np.array_equal(res.audio[i], one.audio[i])           # utterances are independent: sharding changes nothing
        assert lo <= res.audio[i].ctypes.data < hi
    assert sorted(i for s in pool.last_shards for i in s) == list(range(len(utts)))
    assert all(len(s) > 0 for s in pool.last_shards)
    check_against_oracle(res, {i: ref[i] for i in range(len(utts))}, "f16")
    again = pool.synthesize_batch(utts, *S.FIXTURE_SCALES)           # arena reused
    assert all(np.array_equal(a, b) for a, b in zip(again.audio, one.audio))
    a1 = pool.synthesize(S.FIXTURE_IDS)                              # the reference's one-utterance call still works
    assert np.array_equal(a1, single.synthesize(S.FIXTURE_IDS))
    single.close()
    pool.close()


def test_pool_grows_a_short_arena():
    utts = synthetic_utterances(24, 256, 30, 60, seed=3)
    pool = PiperCUDAPool(*voice_files("medium"), devices=[0], options=Options(precision="f16"), bytes_per_id=256 * 4)
    res = pool.synthesize_batch(utts, *S.FIXTURE_SCALES)             # 1 frame per id reserved, ~3 needed
    assert pool._bytes_per_id > 256 * 4
    ref = PiperCUDARuntime(*voice_files("medium"), Options(precision="f16")).synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert all(np.array_equal(a, b) for a, b in zip(res.audio, ref.audio))
    pool.close()


@pytest.mark.parametrize("n_utts,t_lo,t_hi", [(3, 10, 20), (64, 80, 120)])   # single-copy path / ranged copy-out path
def test_output_buffer_and_pcm16_only(n_utts, t_lo, t_hi):
    utts = synthetic_utterances(n_utts, 256, t_lo, t_hi, seed=9)
    rt = PiperCUDARuntime(*voice_files("medium"), Options(precision="f16"))
    ref = rt.synthesize_batch(utts, *S.FIXTURE_SCALES, pcm16=True)
    total = sum(a.size for a in ref.audio)
    arena = HostArena(total * 4 + 4096)
    rt.set_output_buffer(arena.ptr, arena.nbytes)
    own = rt.synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert own.audio[0].ctypes.data == arena.ptr                     # written in place, no second copy
    assert all(np.array_equal(a, b) for a, b in zip(own.audio, ref.audio))
    p16 = rt.synthesize_batch(utts, *S.FIXTURE_SCALES, pcm16_only=True)
    assert p16.audio == [] and p16.pcm16[0].ctypes.data == arena.ptr
    assert all(np.array_equal(a, b) for a, b in zip(p16.pcm16, ref.pcm16))
    assert rt.last_run_timings["d2h_bytes"] < total * 2 + 65536      # half the float bytes
    rt.set_output_buffer(arena.ptr, 1024)                            # too small: a clear capacity error, nothing written
    with pytest.raises(PiperB200Error) as e:
        rt.synthesize_batch(utts, *S.FIXTURE_SCALES)
    assert e.value.code == 7 and "output buffer" in str(e.value)
    rt.set_output_buffer(None)
    lib_owned = rt.synthesize_batch(utts, *S.FIXTURE_SCALES, pcm16_only=True)
    assert all(np.array_equal(a, b) for a, b in zip(lib_owned.pcm16, ref.pcm16))
    pageable = np.zeros(total, np.float32)
    with pytest.raises(PiperB200Error):                              # pageable memory is refused, not silently staged
        rt.set_output_buffer(pageable.ctypes.data, pageable.nbytes)
    rt.close()
    arena.close()


def test_streaming_pcm16_chunks():
    """PIPER_B200_FLAG_STREAM | PCM16: every chunk carries int16 samples following the WavFileWriter rule."""
    import ctypes as C
    rt = PiperCUDARuntime(*voice_files("medium"), Options(precision="f16"))
    ids = np.asarray(S.FIXTURE_IDS * 4, np.int64)
    whole = rt.synthesize_batch([ids.tolist()], *S.FIXTURE_SCALES, pcm16=True)
    rt.synthesize_packed(ids, np.array([len(ids)], np.int64), *S.FIXTURE_SCALES, stream=True, pcm16=True)
    got, final = [], C.c_int32(0)
    while not final.value:
        rs = runtime.Result()
        rt._check(rt._lib.piper_b200_stream_next(rt._ctx, 16, C.byref(rs), C.byref(final)))
        assert rs.pcm16
        got.append(np.ctypeslib.as_array(C.cast(rs.pcm16, C.POINTER(C.c_int16)), shape=(int(rs.total_samples),)).copy())
    assert np.array_equal(np.concatenate(got), whole.pcm16[0])
    rt.close()


NEAR_TIE = 5e-5      # |w - round(w)| in float64 below which a ceil() decision is a near-tie


@pytest.mark.timeout(600)
@pytest.mark.parametrize("precision", ["f16", "fp32"])
def test_config4_high_voice_2000_ids_against_the_oracle(precision):
    """BASELINE config 4 as stated (high voice, one 2000-id utterance) against the oracle: durations exact, audio within
    the mode's tolerance.  Data seed 2025: the float64 oracle puts no duration of this utterance closer than 3e-5 to a
    ceil() boundary (tools/diag_config4.py, profiles/r02_diag_config4.log), so every fp32-grade evaluation must agree.
    The oracle needs ~10-20 s once and is cached under VOICE_DIR."""
    ids = synthetic_utterances(1, 256, 2000, 2000, seed=2025)[0]
    cache = os.path.join(VOICE_DIR, "oracle_high_2000_seed2025.npz")
    if os.path.exists(cache):
        z = np.load(cache)
        ref = {0: (z["audio"], z["dur"], int(z["y_len"]))}
    else:
        ref = oracle_batch("high", [ids])
        np.savez(cache, audio=ref[0][0], dur=ref[0][1], y_len=ref[0][2])
    rt = PiperCUDARuntime(*voice_files("high"), Options(precision=precision))
    res = rt.synthesize_batch([ids], *S.FIXTURE_SCALES)
    check_against_oracle(res, ref, precision)
    rt.close()


@pytest.mark.timeout(600)
def test_duration_mismatches_are_float64_near_ties():
    """ceil() is a discrete decision on a float: two correct fp32 evaluations of logw (different summation orders) can
    disagree where w = exp(logw) * length_scale lies within rounding error of an integer.  Data seed 2024 at T = 2000 has
    such a phoneme (float64 w = 3.9999927).  Adjudicated by the float64 oracle: every duration of every precision mode
    either equals ceil(w64) or sits at a near-tie (|w64 - round(w64)| < NEAR_TIE); the fp32 CUDA path and the fp32
    oracle agree everywhere on this utterance."""
    import torch
    from common import oracle
    from oracle import refrng
    ids = synthetic_utterances(1, 256, 2000, 2000, seed=2024)[0]
    ns, ls, nw = S.FIXTURE_SCALES
    o64 = oracle("high", torch.float64)
    with torch.no_grad():
        x, _, _ = o64.text_encoder(torch.tensor(ids, dtype=torch.int64).view(1, -1))
        logw = o64.duration_predictor(x, torch.as_tensor(refrng.noise_dp(len(ids))).double().view(1, 2, -1), nw)
    w64 = np.exp(logw.view(-1).numpy()) * ls
    d64 = np.ceil(w64).astype(np.int32)
    tie = np.abs(w64 - np.round(w64)) < NEAR_TIE
    assert tie.sum() <= 4                                             # near-ties are rare: this is not a loophole
    for precision in ("fp32", "f16", "bf16"):
        rt = PiperCUDARuntime(*voice_files("high"), Options(precision=precision))
        d = rt.synthesize_batch([ids], ns, ls, nw).durations[0]
        rt.close()
        bad = np.nonzero(d != d64)[0]
        assert all(tie[i] and abs(int(d[i]) - int(d64[i])) == 1 for i in bad), (precision, bad, w64[bad])
        if precision == "fp32":
            assert len(bad) == 0
