"""Microbenchmark: tcgen05 conv with row-shifted A descriptors (window mode) for aligned vs misaligned tap shifts."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200")):
    sys.path.insert(0, p)
from piper_b200 import spec as S, synth_voice            # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options  # noqa: E402

if __name__ == "__main__":
    mp, cp = synth_voice.write_voice(S.TINY, "/tmp/piper_b200_voices")
    rt = PiperCUDARuntime(mp, cp, Options())
    for case in [(0, 128, 128, 5, 6, 1, 400000), (0, 128, 128, 5, 8, 1, 400000), (0, 128, 128, 5, 1, 1, 400000),
                 (0, 128, 128, 1, 1, 1, 400000),
                 (0, 32, 32, 7, 3, 1, 2000000), (0, 32, 32, 7, 8, 1, 2000000), (0, 64, 64, 7, 1, 1, 1000000),
                 (0, 64, 64, 7, 8, 1, 1000000)]:
        for rep in range(2):
            r = rt.selftest_conv(*case, precision="f16")
        fl = 2.0 * case[1] * case[2] * case[3] * case[6]
        print(os.environ.get("PIPER_B200_TC_WINDOW", "1"), case, "tc_ms %.3f  %.1f TFLOP/s  diff %.1e" % (r["tc_ms"], fl / r["tc_ms"] / 1e9, r["max_abs_diff"]), flush=True)
