#!/usr/bin/env python
"""bench.py - audio-seconds synthesized per second (BASELINE.json metric) on N B200s.

A step = one pass of the hot path (phoneme ids -> waveforms) over one batch of synthetic
utterances.  Workload at every N: BASELINE config 3 - the medium voice, a batch of 256
utterances of ~100 phoneme ids PER GPU (weak scaling: independent utterances sharded
across ranks, no data-path collective, final host gather of lengths only).
`value`  : inputs (ids) already in HBM, audio left in HBM, CUDA-event timed.
`e2e`    : the public API with HOST buffers - ids H2D and audio D2H inside the timed region.
`--impl reference` : the reference's CPU path (the torch-CPU oracle of the same graph; ORT
and the Swift reference are unavailable on this box) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np   # noqa: E402

from piper_b200 import spec as S            # noqa: E402
from piper_b200 import synth_voice          # noqa: E402

VOICE_DIR = os.environ.get("PIPER_B200_VOICE_DIR", "/tmp/piper_b200_voices")
METRIC = "audio_seconds_per_second"
UNIT = "audio-s/s"


def workload(args, rank):
    spec = S.SPECS[args.voice]
    utts = synth_voice.synthetic_utterances(args.batch, spec.n_vocab, args.t_lo, args.t_hi, seed=2024 + rank)
    sids = None
    if spec.n_speakers > 1:
        sids = np.random.default_rng(5 + rank).integers(0, spec.n_speakers, args.batch).astype(np.int64)
    return spec, utts, sids


def config_dict(args, n_gpus):
    return {"workload": f"BASELINE config 3: {args.voice} voice, batch of {args.batch} utterances "
                        f"(T~U[{args.t_lo},{args.t_hi}] phoneme ids) per GPU, scales 0.667/1.0/0.8",
            "voice": args.voice, "utterances_per_gpu": args.batch, "global_utterances": args.batch * n_gpus,
            "precision": args.precision, "sharding": f"utterances x{n_gpus}, no collective",
            "l2": "per-step activation working set > 1 GB >> 126 MB L2 (no explicit flush)"}


class ClockSampler(threading.Thread):
    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.stop_flag, self.sm, self.reasons, self.sm_max = device, False, [], set(), None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(f[0]))
                self.sm_max = float(f[1])
                for n, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.2)

    def result(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ---- CPU arm: the torch-CPU oracle of the same graph on all host cores ---------------------------
# One single-threaded worker process per host core, utterances dealt round-robin (the reference's CPU
# path is one utterance per call; B=1 graphs do not scale across intra-op threads, processes do).
_ORC = None


def _cpu_worker_init(voice, voice_dir):
    global _ORC
    import torch
    torch.set_num_threads(1)
    from oracle.vits_oracle import VitsOracle
    mp, _ = synth_voice.write_voice(S.SPECS[voice], voice_dir)
    _ORC = (VitsOracle.from_file(mp), S.SPECS[voice].sample_rate)


def _cpu_worker_run(job):
    from oracle import refrng
    orc, sr = _ORC
    ns, ls, nw = S.FIXTURE_SCALES
    total = 0.0
    for u, sid in job:
        o = orc.infer(u, ns, ls, nw, refrng.noise_dp(len(u)), noise_main_fn=refrng.noise_main, sid=sid)
        total += o["audio"].size / sr
    return total


class CpuOracle:
    def __init__(self, args):
        import multiprocessing as mp
        from concurrent.futures import ProcessPoolExecutor
        self.cores = os.cpu_count() or 1
        synth_voice.write_voice(S.SPECS[args.voice], VOICE_DIR)      # written once, before the workers read it
        self.pool = ProcessPoolExecutor(max_workers=self.cores, mp_context=mp.get_context("spawn"),
                                        initializer=_cpu_worker_init, initargs=(args.voice, VOICE_DIR))
        list(self.pool.map(_cpu_worker_run, [[([1, 5, 9, 2], None if S.SPECS[args.voice].n_speakers <= 1 else 0)]] * self.cores))

    def rate(self, utts, sids, n):
        """audio-s/s over the first n utterances of the batch, all cores busy."""
        import torch
        n = min(n, len(utts))
        jobs = [[] for _ in range(self.cores)]
        for i in range(n):
            jobs[i % self.cores].append((list(map(int, utts[i])), None if sids is None else int(sids[i])))
        t0 = time.perf_counter()
        audio_s = sum(self.pool.map(_cpu_worker_run, [j for j in jobs if j]))
        dt = time.perf_counter() - t0
        return {"value": audio_s / dt, "unit": UNIT, "cores": self.cores, "kind": "port",
                "sample": f"first {n} utterances of the batch, one per call, {self.cores} single-threaded worker "
                          f"processes, torch {torch.__version__} CPU fp32 oracle ({dt:.1f} s wall, "
                          f"{dt * self.cores:.0f} core-s)", "ms_per_utterance": 1000 * dt / n}

    def close(self):
        self.pool.shutdown()


def cpu_oracle_rate(args, utts, sids, n=128):
    orc = CpuOracle(args)
    try:
        orc.rate(utts, sids, min(n, 2 * orc.cores))     # warm-up pass
        return orc.rate(utts, sids, n)
    finally:
        orc.close()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, utts, sids = workload(args, 0)
    orc = CpuOracle(args)
    per_step, base = [], None
    t0 = time.perf_counter()
    try:
        for i in range(args.warmup + args.steps):
            base = orc.rate(utts, sids, 64)             # a step = a 64-utterance sample of the batch
            if i >= args.warmup:
                per_step.append(base)
    finally:
        orc.close()
    v = float(np.mean([b["value"] for b in per_step]))
    ms = float(np.mean([b["ms_per_utterance"] * 64 for b in per_step]))
    base = dict(per_step[-1])
    base["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args, args.gpus), "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--voice", default="medium")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--t-lo", type=int, default=80)
    ap.add_argument("--t-hi", type=int, default=120)
    ap.add_argument("--precision", default=os.environ.get("PIPER_B200_PRECISION", "f16"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from piper_b200.runtime import PiperCUDARuntime, Options

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    # stdout carries exactly one JSON line: everything else written to fd 1 (NCCL prints its version banner there from C,
    # whatever NCCL_DEBUG says) is sent to stderr, the line itself goes to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    spec, utts, sids = workload(args, rank)
    mp, cp = synth_voice.write_voice(spec, VOICE_DIR) if rank == 0 or world == 1 else (None, None)
    if world > 1:
        dist.barrier()
        mp, cp = synth_voice.write_voice(spec, VOICE_DIR)
    rt = PiperCUDARuntime(mp, cp, Options(device=local, precision=args.precision))
    stream = torch.cuda.current_stream()
    rt.set_stream(stream.cuda_stream)

    lengths = np.array([len(u) for u in utts], dtype=np.int64)
    ids = np.concatenate([np.asarray(u, np.int64) for u in utts])
    ids_pinned = torch.from_numpy(ids).pin_memory()
    ids_dev = ids_pinned.cuda(non_blocking=True)
    ns, ls, nw = S.FIXTURE_SCALES

    def step_resident():
        return rt.synthesize_packed(None, lengths, ns, ls, nw, sids, ids_device_ptr=ids_dev.data_ptr(),
                                    keep_on_device=True, profile=True)

    def step_e2e():
        return rt.synthesize_packed(ids_pinned.numpy(), lengths, ns, ls, nw, sids, copy=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        samples, launches, tm = 0, 0, None
        cls = {"ms": [0.0] * 3, "flops": [0.0] * 3, "launches": [0] * 3}
        for _ in range(steps):
            r = fn()
            samples += r.total_samples
            tm = rt.last_run_timings
            launches += tm["kernel_launches"]
            for i in range(3):
                cls["ms"][i] += tm["class_ms"][i]
                cls["flops"][i] += tm["class_flops"][i]
                cls["launches"][i] += tm["class_launches"][i]
        tm["cls"] = cls
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        s = torch.tensor([float(samples)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(s, op=dist.ReduceOp.SUM)
        return float(t.item()), float(s.item()), launches, tm

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    ms, samples, launches, tm = timed(step_resident, args.steps)
    sampler.stop_flag = True
    sampler.join(2)
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e, samples_e, _, tm_e = timed(step_e2e, args.steps)

    sr = spec.sample_rate
    value = samples / sr / (ms / 1000.0)
    e2e = samples_e / sr / (ms_e / 1000.0)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = peaks.get("bf16_tflops_sustained", 1400.0)
        back_tflops = (tm["dec_flops"] + tm["flow_flops"]) / (tm["decoder_ms"] / 1000.0) / 1e12
        # dominant tensor-core kernel class by device time inside the timed region (CUDA events around each launch
        # on the library's stream): 0 conv_tc_kernel, 1 mrf_tc_kernel, 2 conv_f32_kernel
        names = ["conv_tc_kernel (tcgen05 implicit-GEMM conv: flow, conv_pre, ups, C>=128 resblocks)",
                 "mrf_tc_kernel (fused HiFi-GAN MRF stage, tcgen05)",
                 "conv_f32_kernel (fp32 CUDA-core conv: text encoder + duration predictor)"]
        cls = tm["cls"]
        tc_classes = [i for i in (0, 1) if cls["launches"][i] > 0]
        top = max(tc_classes, key=lambda i: cls["ms"][i]) if tc_classes else 2
        top_tflops = cls["flops"][top] / max(cls["ms"][top], 1e-9) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json"))).get(str(top))
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"fp32": "f32", "bf16": "bf16", "f16": "f16"}[args.precision],
            "data": "synthetic", "config": config_dict(args, world),
            "ms_per_utterance": ms / args.steps / args.batch,
            "stage_ms": {k: tm[k] for k in ("front_ms", "flow_ms", "decoder_ms", "h2d_ms", "d2h_ms", "gpu_ms")},
            "roofline": {"bound": "tensor", "achieved": top_tflops, "peak": peak, "unit": "TFLOP/s",
                         "frac": top_tflops / peak, "traffic": traffic, "kernel": names[top],
                         "launches_per_step": cls["launches"][top] / args.steps,
                         "avg_launch_ms": cls["ms"][top] / max(cls["launches"][top], 1),
                         "flops_per_launch": cls["flops"][top] / max(cls["launches"][top], 1),
                         "share_of_step": cls["ms"][top] / ms,
                         "peak_source": ("MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks
                                         else "fallback 1.4 PF (of fallback)"),
                         "back_end_aggregate_tflops": back_tflops,
                         "by_class": {names[i].split(" ")[0]: {"ms_per_step": cls["ms"][i] / args.steps,
                                                                "tflops": cls["flops"][i] / max(cls["ms"][i], 1e-9) / 1e9,
                                                                "launches_per_step": cls["launches"][i] / args.steps}
                                      for i in range(3)}},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": tm_e["h2d_bytes"],
                    "d2h_bytes_per_step": tm_e["d2h_bytes"], "ms_per_step": ms_e / args.steps},
            "gpu_launches": launches, "clocks": sampler.result(),
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_oracle_rate(args, utts, sids)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    rt.close()


if __name__ == "__main__":
    main()
