"""Runs piper_b200_selftest_conv over the convolution shapes of the medium/high voices (GPU box)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200")):
    sys.path.insert(0, p)
from piper_b200 import spec as S, synth_voice            # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options  # noqa: E402

CASES = [  # mode, cin, cout, k, dil, stride, rows
    (0, 64, 64, 1, 1, 1, 300), (0, 64, 64, 3, 1, 1, 1000), (0, 32, 32, 7, 3, 1, 5000), (0, 32, 32, 3, 1, 1, 129),
    (0, 96, 192, 1, 1, 1, 700), (0, 192, 96, 1, 1, 1, 700), (0, 192, 256, 7, 1, 1, 515), (0, 128, 128, 5, 6, 1, 2000),
    (0, 192, 512, 7, 1, 1, 400), (0, 256, 256, 11, 5, 1, 900), (0, 64, 64, 7, 12, 1, 4000),
    (1, 192, 384, 5, 1, 1, 1000), (2, 192, 384, 1, 1, 1, 1000),
    (3, 256, 128, 16, 1, 8, 300), (3, 128, 64, 16, 1, 8, 1000), (3, 64, 32, 8, 1, 4, 3000), (3, 128, 64, 4, 1, 2, 500),
]

if __name__ == "__main__":
    mp, cp = synth_voice.write_voice(S.TINY, os.environ.get("PIPER_B200_VOICE_DIR", "/tmp/piper_b200_voices"))
    rt = PiperCUDARuntime(mp, cp, Options())
    bad = 0
    for prec in ("bf16", "f16"):
        for c in CASES:
            r = rt.selftest_conv(*c, precision=prec)
            tol = (2e-2 if c[0] == 1 else 2e-3) * max(1.0, r["max_abs_ref"])
            ok = r["max_abs_diff"] <= tol
            bad += not ok
            print(prec, c, {k: round(v, 5) for k, v in r.items()}, "ok" if ok else "FAIL", flush=True)
    print("failures:", bad)
    sys.exit(1 if bad else 0)
