"""Prints max-abs / SNR of the 16-bit tensor-core modes against the fp32 oracle (GPU box)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from common import S, voice_files, oracle, snr_db, fixture_ids   # noqa: E402
from oracle import refrng                                          # noqa: E402
from piper_b200.runtime import PiperCUDARuntime, Options           # noqa: E402

if __name__ == "__main__":
    for name, factor, sid in (("medium", 1, None), ("medium", 8, None), ("high", 2, None), ("medium-multispeaker", 2, 7),
                              ("x_low", 4, None)):
        ids = fixture_ids(factor)
        ns, ls, nw = S.FIXTURE_SCALES
        o = oracle(name).infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main, sid=sid)
        for prec in ("fp32", "bf16", "f16"):
            mp, cp = voice_files(name)
            rt = PiperCUDARuntime(mp, cp, Options(precision=prec))
            r = rt.synthesize_batch([ids], ns, ls, nw, speaker_ids=None if sid is None else [sid], keep_stages=True)
            z = rt.get_stage("z", 0)
            a = r.audio[0]
            print(name, factor, prec, "dur_exact", bool(np.array_equal(r.durations[0], o["w_ceil"].astype(np.int32))),
                  "z_maxabs %.3e" % np.abs(z - o["z"].reshape(z.shape)).max(), "z_snr %.1f" % snr_db(o["z"].reshape(z.shape), z),
                  "audio_maxabs %.3e" % np.abs(a - o["audio"]).max(), "audio_snr %.1f dB" % snr_db(o["audio"], a),
                  "rms %.3f" % float(np.sqrt((o["audio"] ** 2).mean())), flush=True)
            rt.close()
