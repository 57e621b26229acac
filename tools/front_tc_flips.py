"""Duration flips of the fp32-emulating tensor-core text encoder (PIPER_B200_FRONT_TC=1, bf16x6 split operands)
against the fp32 CUDA-core text encoder (PIPER_B200_FRONT_TC=0): same voice, same utterances, same seeds.
The env switch is read once per process, so each arm runs in its own child process.
    python tools/front_tc_flips.py [--voices medium,high] [--batches 4] [--utts 256]"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(args):
    sys.path.insert(0, os.path.join(ROOT, "piper-swift_b200"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import voice_files, S
    from piper_b200.runtime import PiperCUDARuntime, Options
    from piper_b200.synth_voice import synthetic_utterances
    out = {}
    for name in args.voices.split(","):
        mp, cp = voice_files(name)
        r = PiperCUDARuntime(mp, cp, Options(precision=args.precision))
        durs, logw = [], []
        for b in range(args.batches):
            utts = synthetic_utterances(args.utts, S.SPECS[name].n_vocab, seed=1000 + b)
            res = r.synthesize_batch(utts, seed=77 + b, keep_stages=(b == 0))
            durs.append(np.concatenate(res.durations))
            if b == 0:
                logw = np.concatenate([r.get_stage("logw", u).reshape(-1) for u in range(min(16, args.utts))])
                out[name + "_audio0"] = res.audio[0]
        out[name + "_dur"] = np.concatenate(durs)
        out[name + "_logw"] = logw
    np.savez(args.child, **out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--voices", default="medium,high")
    ap.add_argument("--batches", type=int, default=4)
    ap.add_argument("--utts", type=int, default=256)
    ap.add_argument("--precision", default="f16")
    ap.add_argument("--child", default=None)
    args = ap.parse_args()
    if args.child:
        return child(args)
    files = {}
    with tempfile.TemporaryDirectory() as td:
        for arm in ("0", "1"):
            f = os.path.join(td, f"arm{arm}.npz")
            env = dict(os.environ, PIPER_B200_FRONT_TC=arm)
            subprocess.run([sys.executable, __file__, "--voices", args.voices, "--batches", str(args.batches),
                            "--utts", str(args.utts), "--precision", args.precision, "--child", f], check=True, env=env)
            files[arm] = dict(np.load(f))
    for name in args.voices.split(","):
        d0, d1 = files["0"][name + "_dur"], files["1"][name + "_dur"]
        l0, l1 = files["0"][name + "_logw"], files["1"][name + "_logw"]
        a0, a1 = files["0"][name + "_audio0"], files["1"][name + "_audio0"]
        snr = None
        if a0.shape == a1.shape:
            snr = float(10 * np.log10((a0.astype(np.float64) ** 2).sum() / max(1e-30, ((a0 - a1).astype(np.float64) ** 2).sum())))
        print(json.dumps({"voice": name, "precision": args.precision, "phonemes": int(d0.size),
                          "duration_flips": int((d0 != d1).sum()), "frames": int(d0.sum()),
                          "logw_max_abs_diff": float(np.abs(l0 - l1).max()), "logw_rms_diff": float(np.sqrt(((l0 - l1) ** 2).mean())),
                          "audio0_snr_db_tc_front_vs_simt_front": snr}))


if __name__ == "__main__":
    main()
