"""Per-launch CUDA-event times of the tensor-core kernels of one config-3 step (medium voice, 256 utterances, f16):
PIPER_B200_PROF_DUMP=1 makes the library print one stderr line per profiled launch.  usage:
    PIPER_B200_PROF_DUMP=1 python tools/prof_launches.py 2> gpurun_out/prof.txt ; python tools/prof_launches.py --digest gpurun_out/prof.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

if len(sys.argv) > 2 and sys.argv[1] == "--digest":
    rows = [ln.split() for ln in open(sys.argv[2]) if ln.startswith("prof cls")]
    # the last step's launches: spans are printed per call; take the last block of equal length
    per_call = []
    cur = []
    for r in rows:
        cur.append((int(r[2]), float(r[4]), float(r[6])))
    n_calls = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    n = len(cur) // n_calls
    last = cur[-n:]
    mrf = [(ms, gf) for c, ms, gf in last if c == 1]
    ctc = [(ms, gf) for c, ms, gf in last if c == 0]
    print("mrf launches:", " ".join("%.3f ms (%.0f TF/s)" % (ms, gf / ms) for ms, gf in mrf), " total %.3f" % sum(ms for ms, _ in mrf))
    print("conv_tc total %.3f ms over %d launches" % (sum(ms for ms, _ in ctc), len(ctc)))
    print("conv_tc launches (ms):", " ".join("%.3f" % ms for ms, _ in ctc))
    sys.exit(0)

from common import S, voice_files                                   # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options           # noqa: E402
from piper_b200.synth_voice import synthetic_utterances            # noqa: E402

mp, cp = voice_files("medium")
rt = PiperCUDARuntime(mp, cp, Options(precision="f16"))
# PROF_UTTS / PROF_IDS: batch size and ids per utterance (default: config 3; PROF_UTTS=1 PROF_IDS=14 = config 1)
n_utts, n_ids = int(os.environ.get("PROF_UTTS", "256")), int(os.environ.get("PROF_IDS", "0"))
utts = synthetic_utterances(n_utts, 256) if not n_ids else synthetic_utterances(n_utts, 256, n_ids, n_ids)
import numpy as np
ids = np.concatenate([np.asarray(u, np.int64) for u in utts]); lengths = np.array([len(u) for u in utts], np.int64)
for i in range(3):
    rt.synthesize_packed(ids, lengths, *S.FIXTURE_SCALES, keep_on_device=True, profile=True)
    print("gpu_ms", rt.last_run_timings["gpu_ms"])
rt.close()
