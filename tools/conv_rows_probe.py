"""Launch time of the tcgen05 convolution kernel against the number of rows (fixed cost per launch vs per-tile cost).
Run on the GPU box: PIPER_B200_SELFTEST_REPEAT=20 [PIPER_B200_SELFTEST_NT=192] python tools/conv_rows_probe.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "piper-swift_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from common import voice_files
from piper_b200.runtime import PiperCUDARuntime, Options

os.environ.setdefault("PIPER_B200_SELFTEST_REPEAT", "20")
mp, cp = voice_files("tiny")
rt = PiperCUDARuntime(mp, cp, Options(precision="fp32"))
CASES = [(0, 192, 192, 1, 1, 1), (0, 192, 384, 5, 1, 1), (1, 192, 384, 5, 1, 1), (0, 768, 96, 1, 1, 1)]
for case in CASES:
    for tiles in (1, 2, 3, 5, 8):
        rows = 128 * 148 * tiles
        try:
            r = rt.selftest_conv(*case, rows=rows, precision="f16")
            print(case, "rows", rows, "tiles/CTA", tiles, "us %.1f" % (r["tc_ms"] * 1e3), "err %.2g" % r["max_abs_diff"], flush=True)
        except Exception as e:
            print(case, rows, "failed", e, flush=True)
            break
