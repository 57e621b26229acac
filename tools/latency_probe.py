"""Single-utterance latency through the host API (BASELINE configs 1-2: the reference's own ms_mean protocol,
bench/benchmark_onnxruntime.py:36-37 - warmup 2, iters 10): wall ms per call, device ms, launches.
    python tools/latency_probe.py [precision]"""
import json
import sys
import time

sys.path.insert(0, "piper-swift_b200")
sys.path.insert(0, "tests")
import numpy as np                         # noqa: E402
from common import voice_files, S          # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options   # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
mp, cp = voice_files("medium")
r = PiperCUDARuntime(mp, cp, Options(precision=prec))
ns, ls, nw = S.FIXTURE_SCALES
for factor in (1, 2, 4, 8):
    ids = list(S.FIXTURE_IDS) * factor
    for _ in range(2):
        r.synthesize(ids, ns, ls, nw)
    wall, dev = [], []
    for _ in range(10):
        t0 = time.perf_counter()
        a = r.synthesize(ids, ns, ls, nw)
        wall.append((time.perf_counter() - t0) * 1e3)
        dev.append(r.last_run_timings["gpu_ms"])
    t = r.last_run_timings
    print(json.dumps({"voice": "medium", "precision": prec, "phonemes": len(ids), "samples": int(a.size),
                      "ms_mean": float(np.mean(wall)), "ms_p50": float(np.median(wall)), "ms_max": float(np.max(wall)),
                      "device_ms_mean": float(np.mean(dev)), "launches": int(t["kernel_launches"]),
                      "audio_s_per_s": a.size / S.SPECS["medium"].sample_rate / (np.mean(wall) / 1e3)}))
