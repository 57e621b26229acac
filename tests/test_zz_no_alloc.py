"""piper_b200_length_regulate_noalloc (no counterpart in the reference):
the length regulator out of a fixed `static __device__` scratch array, for callers that must not call cudaMalloc.
Collected last on purpose.  Checked through the C ABI against the existing entry point
piper_b200_length_regulate on the same inputs - integers, so the tolerance is the one the other length-regulator tests
use: bit-exact - on two calls in a row (the scratch array keeps the previous call's contents), and against the size
limit."""
import numpy as np
import pytest

from common import voice_files
from piper_b200 import runtime
from piper_b200.runtime import PiperCUDARuntime, Options, PiperB200Error

pytestmark = pytest.mark.gpu


def _logw(rng, lengths):
    out = [rng.normal(0.8, 0.9, n).astype(np.float32) for n in lengths]
    out[0][:3] = [np.log(2.0), -200.0, np.log(3.0) + 1e-7]     # exact integer, underflow to 0, just above an integer
    return out


def test_noalloc_matches_the_allocating_entry_point_twice():
    mp, cp = voice_files("tiny")
    rt = PiperCUDARuntime(mp, cp, Options(precision="fp32"))
    rng = np.random.default_rng(7)
    # second call is SMALLER than the first and uses other values: stale contents of the shared array must not leak
    for lengths, ls in (([300, 17, 1, 1025, 64], 1.0), ([5, 200, 33], 1.3)):
        lw = _logw(rng, lengths)
        dur_ref, fl_ref, _ = rt.length_regulate(lw, ls)
        dur, fl = rt.length_regulate_noalloc(lw, ls)
        assert np.array_equal(fl, fl_ref)
        for a, b in zip(dur, dur_ref):
            assert np.array_equal(a, b)
        assert all(int(f) == max(1, int(d.sum())) for f, d in zip(fl, dur))
    # all-zero durations clip to one frame (Clip min 1), like the allocating path
    z = [np.full(4, -200.0, np.float32)]                    # exp underflows to exactly 0: ceil 0
    assert rt.length_regulate_noalloc(z, 1.0)[1][0] == rt.length_regulate(z, 1.0)[1][0] == 1
    rt.close()


def test_noalloc_rejects_sizes_beyond_its_scratch():
    mp, cp = voice_files("tiny")
    rt = PiperCUDARuntime(mp, cp, Options(precision="fp32"))
    with pytest.raises(PiperB200Error) as e:
        rt.length_regulate_noalloc([np.zeros(runtime.NOALLOC_MAX_IDS + 1, np.float32)], 1.0)
    assert e.value.code == 7 and "at most" in str(e.value)
    with pytest.raises(PiperB200Error) as e:
        rt.length_regulate_noalloc([np.zeros(1, np.float32)] * (runtime.NOALLOC_MAX_BATCH + 1), 1.0)
    assert e.value.code == 7
    full = [np.zeros(runtime.NOALLOC_MAX_IDS, np.float32)]      # exactly the maximum still runs: exp(0) = 1 frame per id
    dur, fl = rt.length_regulate_noalloc(full, 1.0)
    assert fl[0] == runtime.NOALLOC_MAX_IDS and dur[0].min() == dur[0].max() == 1
    rt.close()
