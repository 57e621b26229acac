"""Diagnostic: per-stage error of the CUDA path against the oracle (run on the GPU box)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from common import S, voice_files, oracle, snr_db, fixture_ids   # noqa: E402
from oracle import refrng                                      # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options        # noqa: E402


def run(name, factors, sid=None, precision="fp32"):
    mp, cp = voice_files(name)
    rt = PiperCUDARuntime(mp, cp, Options(precision=precision))
    orc = oracle(name)
    sp = S.SPECS[name]
    out = []
    for f in factors:
        ids = fixture_ids(f, sp.n_vocab)
        t0 = time.time()
        o = orc.infer(ids, 0.667, 1.0, 0.8, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main, sid=sid)
        t_or = time.time() - t0
        r = rt.synthesize_batch([ids], 0.667, 1.0, 0.8, speaker_ids=None if sid is None else [sid], keep_stages=True)
        rec = {"voice": name, "factor": f, "oracle_s": t_or, "y_len": int(o["y_len"]), "gpu_y_len": int(r.frame_lengths[0]),
               "dur_equal": bool(np.array_equal(o["w_ceil"].astype(np.int32), r.durations[0])),
               "timings": rt.last_run_timings}
        for st in ("x", "m_p", "logs_p", "logw"):
            g = rt.get_stage(st, 0)
            ref = o[st].reshape(g.shape)
            rec[st] = float(np.abs(g - ref).max())
        if rec["dur_equal"]:
            for st in ("z_p", "z"):
                g = rt.get_stage(st, 0)
                rec[st] = float(np.abs(g - o[st]).max())
            a = r.audio[0]
            rec["audio_maxabs"] = float(np.abs(a - o["audio"]).max())
            rec["audio_snr_db"] = snr_db(o["audio"], a)
        else:
            rec["dur_gpu"] = r.durations[0].tolist()[:32]
            rec["dur_ref"] = o["w_ceil"].astype(int).tolist()[:32]
        print(json.dumps(rec), flush=True)
        out.append(rec)
    rt.close()
    return out


if __name__ == "__main__":
    res = []
    res += run("tiny", [1, 4])
    res += run("medium", [1, 8])
    res += run("high", [2])
    res += run("medium-multispeaker", [2], sid=7)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/diag_parity.json", "w") as fh:
        json.dump(res, fh, indent=1)
