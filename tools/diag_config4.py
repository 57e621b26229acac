"""Where does the f16-mode front end disagree with the oracle on the 2000-id `high` utterance (BASELINE config 4)?
Prints every phoneme whose duration differs, with w = exp(logw) from: the CUDA f16 mode (f16x3 tensor-core front end),
the CUDA fp32 mode, the fp32 oracle and the float64 oracle.  A flip is a `ceil` near-tie when the float64 w lies within
a few fp32 ulps of an integer."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch
from common import S, voice_files, oracle
from oracle import refrng
from piper_b200.runtime import PiperCUDARuntime, Options
from piper_b200.synth_voice import synthetic_utterances

voice = sys.argv[1] if len(sys.argv) > 1 else "high"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
seeds = [int(x) for x in sys.argv[3:]] or [2024]
ns, ls, nw = S.FIXTURE_SCALES
for seed in seeds:
    ids = synthetic_utterances(1, 256, T, T, seed=seed)[0]
    nd = refrng.noise_dp(len(ids))
    lw = {}
    for name, dt in (("o32", torch.float32), ("o64", torch.float64)):
        o = oracle(voice, dt)
        idt = torch.tensor(ids, dtype=torch.int64).view(1, -1)
        with torch.no_grad():
            x, _, _ = o.text_encoder(idt)
            lw[name] = o.duration_predictor(x, torch.as_tensor(nd).to(dt).view(1, 2, -1), nw).view(-1).double().numpy()
    for prec in ("f16", "fp32"):
        rt = PiperCUDARuntime(*voice_files(voice), Options(precision=prec))
        r = rt.synthesize_batch([ids], ns, ls, nw, keep_stages=True)
        lw[prec] = rt.get_stage("logw", 0).reshape(-1).astype(np.float64)
        lw[prec + "_dur"] = r.durations[0]
        rt.close()
    w = {k: np.exp(v) * ls for k, v in lw.items() if not k.endswith("_dur")}
    d64 = np.ceil(w["o64"]).astype(np.int64)
    d32 = np.ceil(np.exp(lw["o32"].astype(np.float32)) * np.float32(ls)).astype(np.int64)
    print(f"seed {seed}: T={len(ids)}  max|logw - o64|: f16 {np.abs(lw['f16'] - lw['o64']).max():.2e}  "
          f"fp32 {np.abs(lw['fp32'] - lw['o64']).max():.2e}  o32 {np.abs(lw['o32'] - lw['o64']).max():.2e}")
    frac = np.abs(w["o64"] - np.round(w["o64"]))
    print(f"  closest float64 w to an integer: {frac.min():.3e} at id {int(frac.argmin())}")
    for prec in ("f16", "fp32"):
        bad = np.nonzero(lw[prec + "_dur"] != d32)[0]
        print(f"  {prec}: {len(bad)} durations differ from the fp32 oracle, {int((lw[prec + '_dur'] != d64).sum())} from float64")
        for i in bad[:10]:
            print(f"    id {i}: w f16 {w['f16'][i]:.9f} fp32 {w['fp32'][i]:.9f} o32 {w['o32'][i]:.9f} o64 {w['o64'][i]:.9f}"
                  f"  dur {prec} {lw[prec + '_dur'][i]} o32 {d32[i]} o64 {d64[i]}")
