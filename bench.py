#!/usr/bin/env python
"""bench.py - audio-seconds synthesized per second (BASELINE.json metric) on N B200s.

A step = one pass of the hot path (phoneme ids -> waveforms) over one batch of synthetic utterances.

Headline workload at every N: BASELINE config 3 AS STATED - the medium voice, ONE batch of 256 utterances of ~100
phoneme ids, sharded over the N GPUs (`piper_b200.sharding.shard_utterances`, 256 / 128 / 64 / 32 utterances per GPU):
strong scaling, no data-path collective.
  value : ids already in HBM, audio left in HBM, CUDA-event timed, max over ranks.
  e2e   : the product's public call with HOST buffers - ids H2D, every rank's audio D2H straight into one shared
          page-locked host arena, the frame-count all_gather and the per-utterance result views on rank 0 all inside
          the timed region (N = 1: PiperCUDARuntime.synthesize_packed; N > 1: ShardedSynthesizer.synthesize_batch).
Extra keys of the same JSON line (fewer steps each): `weak` (256 utterances per GPU, round 1's workload), `config5`
(multi-speaker medium voice, 1024 mixed-speaker utterances sharded over the N GPUs), `pcm16_e2e` (int16-only host
output), and at N = 1 `config1` / `config2` (single utterance, the reference's ms_mean protocol), `config4` (high voice,
one 2000-id utterance), `parity` (the timed batch against the CPU oracle).
`--impl reference` : the reference's CPU path (the torch-CPU oracle of the same graph; ORT and the Swift reference are
unavailable on this box) on a bounded sample.
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
from piper_b200.sharding import shard_utterances   # noqa: E402

VOICE_DIR = os.environ.get("PIPER_B200_VOICE_DIR", "/tmp/piper_b200_voices")
METRIC = "audio_seconds_per_second"
UNIT = "audio-s/s"


def workload(voice, n, t_lo, t_hi, seed=2024):
    spec = S.SPECS[voice]
    utts = synth_voice.synthetic_utterances(n, spec.n_vocab, t_lo, t_hi, seed=seed)
    sids = None
    if spec.n_speakers > 1:
        sids = np.random.default_rng(seed + 5).integers(0, spec.n_speakers, n).astype(np.int64)
    return spec, utts, sids


def config_dict(args, n_gpus):
    per = args.batch // n_gpus
    return {"workload": f"BASELINE config 3: {args.voice} voice, ONE batch of {args.batch} utterances "
                        f"(T~U[{args.t_lo},{args.t_hi}] phoneme ids, data seed 2024) sharded over {n_gpus} GPU(s) "
                        f"(~{per} utterances per GPU, strong scaling), scales 0.667/1.0/0.8",
            "voice": args.voice, "global_utterances": args.batch, "utterances_per_gpu": per,
            "precision": args.precision, "sharding": f"LPT-greedy utterance shards x{n_gpus}, no data-path collective; "
                                                     "e2e gathers into one shared pinned host arena",
            "l2": ("inputs larger than L2: per-step activation working set "
                   f"~{14.0 * per / 256:.1f} GB vs 126 MB L2 (no explicit flush)")}


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
    """job = (keep_outputs, [(index, ids, sid), ...]) -> (audio seconds, {index: (audio, w_ceil)})"""
    from oracle import refrng
    keep, items = job
    orc, sr = _ORC
    ns, ls, nw = S.FIXTURE_SCALES
    total, kept = 0.0, {}
    for i, u, sid in items:
        o = orc.infer(u, ns, ls, nw, refrng.noise_dp(len(u)), noise_main_fn=refrng.noise_main, sid=sid)
        total += o["audio"].size / sr
        if keep:
            kept[i] = (o["audio"].astype(np.float32), o["w_ceil"].astype(np.int32))
    return total, kept


class CpuOracle:
    def __init__(self, voice):
        import multiprocessing as mp
        from concurrent.futures import ProcessPoolExecutor
        self.cores = os.cpu_count() or 1
        synth_voice.write_voice(S.SPECS[voice], VOICE_DIR)      # written once, before the workers read it
        self.pool = ProcessPoolExecutor(max_workers=self.cores, mp_context=mp.get_context("spawn"),
                                        initializer=_cpu_worker_init, initargs=(voice, VOICE_DIR))
        sid0 = None if S.SPECS[voice].n_speakers <= 1 else 0
        list(self.pool.map(_cpu_worker_run, [(False, [(0, [1, 5, 9, 2], sid0)])] * self.cores))

    def rate(self, utts, sids, n, keep=False):
        """audio-s/s over the first n utterances of the batch, all cores busy; keep=True also returns the outputs."""
        import torch
        n = min(n, len(utts))
        jobs = [[] for _ in range(self.cores)]
        for i in range(n):
            jobs[i % self.cores].append((i, list(map(int, utts[i])), None if sids is None else int(sids[i])))
        t0 = time.perf_counter()
        parts = list(self.pool.map(_cpu_worker_run, [(keep, j) for j in jobs if j]))
        dt = time.perf_counter() - t0
        audio_s = sum(p[0] for p in parts)
        kept = {}
        for p in parts:
            kept.update(p[1])
        info = {"value": audio_s / dt, "unit": UNIT, "cores": self.cores, "kind": "port",
                "sample": f"first {n} utterances of the batch, one per call, {self.cores} single-threaded worker "
                          f"processes, torch {torch.__version__} CPU fp32 oracle ({dt:.1f} s wall, "
                          f"{dt * self.cores:.0f} core-s)", "ms_per_utterance": 1000 * dt / n}
        return (info, kept) if keep else info

    def close(self):
        self.pool.shutdown()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, utts, sids = workload(args.voice, args.batch, args.t_lo, args.t_hi)
    orc = CpuOracle(args.voice)
    per_step = []
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
    cfg = config_dict(args, args.gpus)
    cfg["precision"] = "fp32"                           # the CPU arm computes in fp32 whatever --precision says
    cfg["sample_per_step"] = "first 64 utterances of the batch on the host cores"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


def parity_report(rt_result, kept, precision):
    """The batch that was timed against the CPU oracle (the cpu_baseline leg's outputs): durations must match exactly,
    audio within the precision mode's stated tolerance."""
    tol = {"fp32": (80.0, 2e-4), "f16": (58.0, 4e-3), "bf16": (40.0, 3e-2)}[precision]
    mism, snr_min, max_abs, len_mism = 0, float("inf"), 0.0, 0
    for i, (ref_audio, ref_dur) in sorted(kept.items()):
        got = rt_result.audio[i]
        if not np.array_equal(np.asarray(rt_result.durations[i], np.int32), ref_dur):
            mism += 1
        if got.size != ref_audio.size:
            len_mism += 1
            continue
        err = got.astype(np.float64) - ref_audio.astype(np.float64)
        snr = 10 * np.log10((ref_audio.astype(np.float64) ** 2).sum() / max((err ** 2).sum(), 1e-300))
        snr_min = min(snr_min, float(snr))
        max_abs = max(max_abs, float(np.abs(err).max()))
    ok = mism == 0 and len_mism == 0 and snr_min >= tol[0] and max_abs <= tol[1]
    return {"utterances": len(kept), "duration_mismatches": mism, "length_mismatches": len_mism,
            "snr_db_min": snr_min, "max_abs": max_abs, "tolerance": {"snr_db_min": tol[0], "max_abs": tol[1]},
            "against": "oracle/vits_oracle.py (torch CPU fp32), same ids / scales / reference RNG noise", "ok": bool(ok)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--voice", default="medium")
    ap.add_argument("--batch", type=int, default=256, help="utterances in the ONE global batch (sharded over --gpus)")
    ap.add_argument("--t-lo", type=int, default=80)
    ap.add_argument("--t-hi", type=int, default=120)
    ap.add_argument("--precision", default=os.environ.get("PIPER_B200_PRECISION", "f16"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only (weak / config1,2,4,5 / pcm16 legs skipped)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from piper_b200.runtime import PiperCUDARuntime, Options
    from piper_b200.sharding import ShardedSynthesizer

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
    stream = torch.cuda.current_stream()
    ns, ls, nw = S.FIXTURE_SCALES
    runtimes = {}

    def runtime_for(voice):
        if voice not in runtimes:
            spec = S.SPECS[voice]
            if rank == 0 or world == 1:
                synth_voice.write_voice(spec, VOICE_DIR)
            if world > 1:
                dist.barrier()
            mp, cp = synth_voice.write_voice(spec, VOICE_DIR)
            rt = PiperCUDARuntime(mp, cp, Options(device=local, precision=args.precision))
            rt.set_stream(stream.cuda_stream)
            syn = None
            if world > 1:
                syn = ShardedSynthesizer(rt, int(rt.info.hop), tag=f"bench_{voice}_{os.environ.get('MASTER_PORT', '0')}")
            runtimes[voice] = (spec, rt, syn)
        return runtimes[voice]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """K steps between barrier + synchronize pairs, CUDA events on the launching stream, max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        acc = [fn() for _ in range(steps)]
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), acc

    def all_sum(x):
        t = torch.tensor([float(x)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def measure(voice, utts, sids, steps, warmup, shard=True, pcm16_only=False, want_resident=True, want_e2e=True):
        """One workload: `utts` is the GLOBAL batch when shard=True (dealt to the ranks), else this rank's own batch."""
        spec, rt, syn = runtime_for(voice)
        lens_all = [len(u) for u in utts]
        mine = shard_utterances(lens_all, world)[rank] if (shard and world > 1) else list(range(len(utts)))
        my_utts = [utts[i] for i in mine]
        my_sids = None if sids is None else np.asarray([sids[i] for i in mine], np.int64)
        lengths = np.array([len(u) for u in my_utts], dtype=np.int64)
        ids = np.concatenate([np.asarray(u, np.int64) for u in my_utts])
        ids_pinned = torch.from_numpy(ids).pin_memory()
        ids_dev = ids_pinned.cuda(non_blocking=True)
        out = {"utterances_this_rank": len(mine)}
        sr = spec.sample_rate
        if want_resident:
            def step_resident(profile=False):
                r = rt.synthesize_packed(None, lengths, ns, ls, nw, my_sids, ids_device_ptr=ids_dev.data_ptr(),
                                         keep_on_device=True, profile=profile)
                return r.total_samples, rt.last_run_timings
            for _ in range(warmup):
                step_resident()
            ms, acc = timed(step_resident, steps)
            samples = all_sum(sum(a[0] for a in acc))
            # The roofline's per-launch CUDA events come from a SECOND pass over the same K steps: an event recorded
            # between two kernels ends the programmatic dependent launch chain (kernels.h), so the pass that gives `value`
            # carries none.  Both passes are timed; the profiled one is reported as ms_profiled.
            ms_p, acc_p = timed(lambda: step_resident(True), steps)
            out.update(ms=ms, value=samples / sr / (ms / 1000.0), ms_per_step=ms / steps, timings=[a[1] for a in acc_p],
                       ms_profiled=ms_p)
        if want_e2e:
            rt.set_output_buffer(None)
            sharded = syn is not None and shard
            if sharded:
                # the caller holds its phoneme ids as int64 arrays (the reference API takes [Int] buffers): no per-call
                # list -> array conversion inside the timed region
                utts_a = [np.asarray(u, np.int64) for u in utts]

                def step_e2e():      # all ranks pass the same global batch; rank 0 receives every utterance's view
                    r = syn.synthesize_batch(utts_a, ns, ls, nw, speaker_ids=sids, pcm16_only=pcm16_only)
                    return (r.total_samples if rank == 0 else 0), rt.last_run_timings, r
            else:
                def step_e2e():
                    r = rt.synthesize_packed(ids_pinned.numpy(), lengths, ns, ls, nw, my_sids, copy=False,
                                             pcm16_only=pcm16_only)
                    return r.total_samples, rt.last_run_timings, r
            for _ in range(max(2, warmup // 2)):
                step_e2e()
            ms_e, acc = timed(step_e2e, steps)
            samples_e = all_sum(sum(a[0] for a in acc))
            tm_e = acc[-1][1]
            last = acc[-1][2]
            if getattr(last, "audio", None):     # views into the pinned result buffer, which later legs reuse: copy
                last.audio = [np.array(a) for a in last.audio]      # (after the timed region)
            out.update(e2e_ms=ms_e, e2e_value=samples_e / sr / (ms_e / 1000.0), e2e_ms_per_step=ms_e / steps,
                       h2d=all_sum(tm_e["h2d_bytes"]), d2h=all_sum(tm_e["d2h_bytes"]), last=last)
            if sharded:
                rt.set_output_buffer(None)
        return out

    def brief(m, n_utts):
        d = {"utterances": n_utts}
        if "value" in m:
            d.update(value=m["value"], ms_per_step=m["ms_per_step"])
        if "e2e_value" in m:
            d["e2e"] = {"value": m["e2e_value"], "unit": UNIT, "ms_per_step": m["e2e_ms_per_step"],
                        "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"]}
        return d

    # ---- headline: config 3, strong scaling ---------------------------------------------------------------------
    spec, utts, sids = workload(args.voice, args.batch, args.t_lo, args.t_hi)
    runtime_for(args.voice)
    sampler = ClockSampler(local)
    sampler.start()
    head = measure(args.voice, utts, sids, args.steps, args.warmup)
    sampler.stop_flag = True
    sampler.join(2)
    extras = {}
    x_steps = max(3, args.steps // 2)
    if not args.no_extras:
        if world > 1:      # round 1's workload: 256 utterances PER GPU (independent batches, seed 2024 + rank)
            _, wu, ws = workload(args.voice, args.batch, args.t_lo, args.t_hi, seed=2024 + rank)
            w = measure(args.voice, wu, ws, x_steps, 3, shard=False)
            extras["weak"] = dict(brief(w, args.batch * world), scaling="weak",
                                  workload=f"{args.batch} utterances per GPU, independent batches")
        # int16-only host output (WavFileWriter.swift:20-30 rule on the device): half the device->host bytes
        p16 = measure(args.voice, utts, sids, x_steps, 3, pcm16_only=True, want_resident=False)
        extras["pcm16_e2e"] = brief(p16, args.batch)["e2e"]
        # BASELINE config 5: multi-speaker medium voice, 1024 mixed-speaker utterances sharded over the N GPUs
        _, c5u, c5s = workload("medium-multispeaker", 1024, args.t_lo, args.t_hi)
        c5 = measure("medium-multispeaker", c5u, c5s, max(3, args.steps // 3), 2)
        extras["config5"] = dict(brief(c5, 1024), workload="BASELINE config 5: medium-multispeaker (904 speakers, "
                                 f"gin 512), 1024 mixed-speaker utterances sharded over {world} GPU(s)")
        if world == 1:
            # configs 1-2: one utterance, the reference's own protocol (benchmark_onnxruntime.py:36-37,154-165: warmup 2,
            # iters 10, wall clock around the public call, ms_mean / p50 / p95 / max), fixture tiled x1, 2, 4, 8
            _, rt, _ = runtime_for(args.voice)
            sweep = {}
            for factor in (1, 2, 4, 8):
                ids1 = [i % spec.n_vocab for i in S.FIXTURE_IDS * factor]
                for _ in range(2):
                    rt.synthesize(ids1, ns, ls, nw)
                ts, n_s = [], 0
                for _ in range(10):
                    t0 = time.perf_counter()
                    a = rt.synthesize(ids1, ns, ls, nw)
                    ts.append(1000 * (time.perf_counter() - t0))
                    n_s = a.size
                sweep[str(len(ids1))] = {"ms_mean": float(np.mean(ts)), "ms_p50": float(np.percentile(ts, 50)),
                                         "ms_p95": float(np.percentile(ts, 95)), "ms_max": float(np.max(ts)),
                                         "audio_s_per_s": n_s / spec.sample_rate / (np.mean(ts) / 1000.0),
                                         "launches": rt.last_run_timings["kernel_launches"]}
            extras["config1"] = dict(sweep["14"], workload="BASELINE config 1: medium voice, the 14-id fixture, one call")
            extras["config2"] = dict(sweep, workload="BASELINE config 2: fixture tiled x1,2,4,8 (14-112 ids), one call each")
            # config 4: high voice (512-channel generator, ResBlock1), one 2000-id utterance
            _, hu, _ = workload("high", 1, 2000, 2000)
            c4 = measure("high", hu, None, max(3, args.steps // 2), 2)
            extras["config4"] = dict(brief(c4, 1), workload="BASELINE config 4: high voice, one 2000-id utterance")

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = peaks.get("bf16_tflops_sustained", 1400.0)
        tms = head["timings"]
        cls = {"ms": [0.0] * 3, "flops": [0.0] * 3, "launches": [0] * 3}
        launches = 0
        for tm in tms:
            launches += tm["kernel_launches"]
            for i in range(3):
                cls["ms"][i] += tm["class_ms"][i]
                cls["flops"][i] += tm["class_flops"][i]
                cls["launches"][i] += tm["class_launches"][i]
        tm = tms[-1]
        back_tflops = (tm["dec_flops"] + tm["flow_flops"]) / (tm["decoder_ms"] / 1000.0) / 1e12
        # dominant tensor-core kernel class by device time inside the timed region (CUDA events around each launch
        # on the library's stream): 0 conv_tc_kernel, 1 mrf_tc_kernel, 2 conv_f32_kernel
        names = ["conv_tc_kernel (tcgen05 implicit-GEMM conv: flow, conv_pre, ups, text encoder / duration predictor)",
                 "mrf_tc_kernel (fused HiFi-GAN MRF stage, tcgen05)",
                 "conv_f32_kernel (fp32 CUDA-core conv: projection layers of the front end)"]
        tc_classes = [i for i in (0, 1) if cls["launches"][i] > 0]
        top = max(tc_classes, key=lambda i: cls["ms"][i]) if tc_classes else 2
        top_tflops = cls["flops"][top] / max(cls["ms"][top], 1e-9) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json"))).get(str(top))
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": {"fp32": "f32", "bf16": "bf16", "f16": "f16"}[args.precision],
            "data": "synthetic", "config": config_dict(args, world),
            "ms_per_utterance": head["ms_per_step"] / args.batch,
            "stage_ms": {k: tm[k] for k in ("front_ms", "flow_ms", "decoder_ms", "h2d_ms", "d2h_ms", "gpu_ms")},
            "roofline": {"bound": "tensor", "achieved": top_tflops, "peak": peak, "unit": "TFLOP/s",
                         "frac": top_tflops / peak, "traffic": traffic, "kernel": names[top],
                         "launches_per_step": cls["launches"][top] / args.steps,
                         "avg_launch_ms": cls["ms"][top] / max(cls["launches"][top], 1),
                         "flops_per_launch": cls["flops"][top] / max(cls["launches"][top], 1),
                         "flops_counted": ("SURVEY 8d rows of this class only - for mrf_tc_kernel the resblock rows "
                                           "(7.864 + 15.729 + 15.729 MFLOP/frame) - times the ROWS the launch covers, "
                                           "zero gap rows between utterances included (~1 % above valid frames)"),
                         "share_of_step": cls["ms"][top] / head["ms_profiled"],
                         "timing": ("CUDA events around every launch of the class, on the library's stream, over a second "
                                    "pass of the same %d steps (%.3f ms per step with the events, %.3f without: the events "
                                    "end the programmatic-dependent-launch chain, so the pass that gives `value` has none)"
                                    % (args.steps, head["ms_profiled"] / args.steps, head["ms_per_step"])),
                         "peak_source": ("MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks
                                         else "fallback 1.4 PF (of fallback)"),
                         "back_end_aggregate_tflops": back_tflops,
                         "by_class": {names[i].split(" ")[0]: {"ms_per_step": cls["ms"][i] / args.steps,
                                                                "tflops": cls["flops"][i] / max(cls["ms"][i], 1e-9) / 1e9,
                                                                "launches_per_step": cls["launches"][i] / args.steps}
                                      for i in range(3)},
                         "rank": 0},
            "e2e": {"value": head["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": head["h2d"],
                    "d2h_bytes_per_step": head["d2h"], "ms_per_step": head["e2e_ms_per_step"],
                    "api": ("PiperCUDARuntime.synthesize_packed (host ids -> pinned host audio)" if world == 1 else
                            "ShardedSynthesizer.synthesize_batch (host ids -> one /dev/shm pinned arena, all ranks; "
                            "frame-count all_gather and result views inside the timed region)")},
            "gpu_launches": launches, "clocks": sampler.result(),
        }
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            orc = CpuOracle(args.voice)
            try:
                orc.rate(utts, sids, 2 * orc.cores)                          # warm-up pass
                base, kept = orc.rate(utts, sids, 128, keep=True)
            finally:
                orc.close()
            line["cpu_baseline"] = base
            line["parity"] = parity_report(head["last"], kept, args.precision)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    head.pop("last", None)
    if world > 1:
        dist.barrier()
    for _, rt, syn in runtimes.values():
        if syn is not None:
            syn.close()
        rt.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
