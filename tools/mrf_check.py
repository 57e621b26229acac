"""Fused MRF kernel vs the per-convolution tensor-core path (run twice with PIPER_B200_MRF=0/1)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from common import S, voice_files, oracle, snr_db, fixture_ids   # noqa: E402
from oracle import refrng                                          # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options           # noqa: E402
from piper_b200.synth_voice import synthetic_utterances            # noqa: E402

if __name__ == "__main__":
    tag = sys.argv[1]
    out = {}
    for prec in ("f16", "bf16"):
        mp, cp = voice_files("medium")
        rt = PiperCUDARuntime(mp, cp, Options(precision=prec))
        for factor in (1, 8):
            ids = fixture_ids(factor)
            ns, ls, nw = S.FIXTURE_SCALES
            a = rt.synthesize_batch([ids], ns, ls, nw).audio[0]
            o = oracle("medium").infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
            print(tag, prec, factor, "snr vs oracle %.1f dB" % snr_db(o["audio"], a), "maxabs %.2e" % np.abs(a - o["audio"]).max(),
                  "gpu_ms %.3f launches %d" % (rt.last_run_timings["gpu_ms"], rt.last_run_timings["kernel_launches"]), flush=True)
            out[f"{prec}_{factor}"] = a
        utts = synthetic_utterances(48, 256)
        r = rt.synthesize_batch(utts, *S.FIXTURE_SCALES)
        out[f"{prec}_batch"] = np.concatenate(r.audio)[:2000000]
        print(tag, prec, "batch48 gpu_ms %.3f" % rt.last_run_timings["gpu_ms"], flush=True)
        rt.close()
    np.savez(os.path.join(ROOT, "gpurun_out", f"mrf_check_{tag}.npz"), **out)
    other = os.path.join(ROOT, "gpurun_out", "mrf_check_%s.npz" % ("off" if tag == "on" else "on"))
    if os.path.exists(other):
        g = np.load(other)
        for k in out:
            print("fused vs unfused", k, "snr %.1f dB" % snr_db(g[k], out[k]), "maxabs %.2e" % np.abs(g[k] - out[k]).max())
