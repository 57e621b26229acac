"""Multi-GPU synthesis: one batch of utterances sharded over the B200s of one box.

Utterances are independent units: every GPU runs the full graph on its shard with its own context and
stream, there is NO collective on the data path.  The final gather is the device->host copy itself:
every context writes its audio straight into its own region of ONE page-locked host arena
(piper_b200_set_output_buffer), and the caller receives per-utterance views into that arena in the
original utterance order - no second copy.  The reference has no counterpart (single MTLDevice,
MetalContext.swift:11-19; one utterance per call, GraphExecutor.swift:108,196).

Two drivers over the same C ABI:
  PiperCUDAPool       one process, one context per GPU, one host thread per context (ctypes releases the GIL)
  ShardedSynthesizer  one process per GPU (torchrun / torch.distributed); the arena is a /dev/shm mapping, every rank
                      pins its own region of it; the only exchange is one small all_gather of frame counts
"""
from __future__ import annotations

import functools
import heapq
import mmap
import os
from collections.abc import Sequence as _SequenceABC
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np


@functools.lru_cache(maxsize=64)
def _shard_cached(lengths: Tuple[int, ...], world_size: int) -> Tuple[Tuple[int, ...], ...]:
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    heap = [(0, r) for r in range(world_size)]          # (load, rank): the least-loaded rank, lowest rank on ties
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        load, r = heapq.heappop(heap)
        out[r].append(i)
        heapq.heappush(heap, (load + lengths[i], r))
    return tuple(tuple(sorted(s)) for s in out)


def shard_utterances(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """LPT-greedy: longest utterance first onto the least-loaded rank (work ~ frames ~ 3*T).
    Deterministic; every rank computes the same assignment.  Returns utterance indices per rank,
    each in ascending original order.  (A heap and a small cache keyed by the lengths: the plain O(n * world) form cost
    1 ms per call for 256 utterances on 8 ranks - a third of the 8-GPU step.)"""
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    return [list(s) for s in _shard_cached(tuple(int(x) for x in lengths), int(world_size))]


def unshard(per_rank_indices: Sequence[Sequence[int]], per_rank_values: Sequence[Sequence]) -> list:
    """Inverse of shard_utterances for per-utterance results gathered from all ranks."""
    n = sum(len(s) for s in per_rank_indices)
    out = [None] * n
    for idx, vals in zip(per_rank_indices, per_rank_values):
        if len(idx) != len(vals):
            raise ValueError("rank returned a different number of results than it was assigned")
        for i, v in zip(idx, vals):
            out[i] = v
    if any(v is None for v in out):
        raise ValueError("some utterances were not produced by any rank")
    return out


def gather_lengths(local_indices: Sequence[int], local_frame_lengths: Sequence[int], n_total: int,
                   group=None) -> List[int]:
    """Final host gather of per-utterance frame counts with torch.distributed (gloo or nccl
    process group; all_gather_object on the host side - audio stays with its rank)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    payload = (list(map(int, local_indices)), list(map(int, local_frame_lengths)))
    gathered: List[Tuple[List[int], List[int]]] = [None] * world
    dist.all_gather_object(gathered, payload, group=group)
    out = unshard([g[0] for g in gathered], [g[1] for g in gathered])
    if len(out) != n_total:
        raise ValueError("gathered result count mismatch")
    return out


# ---- arena layout ---------------------------------------------------------------------------------------------------
def region_layout(shards: Sequence[Sequence[int]], lengths: Sequence[int], bytes_per_id: int,
                  align: int = 4096) -> Tuple[List[int], List[int]]:
    """Byte offset and capacity of every rank's region in the shared arena: proportional to the phoneme ids of its
    shard (audio bytes ~ ids * frames-per-id * hop * sample size), page aligned."""
    offs, caps, at = [], [], 0
    for s in shards:
        n_ids = sum(int(lengths[i]) for i in s)
        cap = max(align, (n_ids * bytes_per_id + align - 1) // align * align)
        offs.append(at)
        caps.append(cap)
        at += cap
    return offs, caps


def arena_views_layout(shards: Sequence[Sequence[int]], table: np.ndarray, offs: Sequence[int], bytes_per_frame: int):
    """Where every utterance's audio lies in the shared arena: rank r wrote its utterances back to back from offs[r], in
    the order of shards[r]; table[r, 1 + j] = frames of its j-th utterance.  Returns (start byte, byte count, frames,
    rank) per utterance in the ORIGINAL order - rank-major arrays and one scatter, no per-utterance Python work."""
    world = len(shards)
    n = sum(len(s) for s in shards)
    counts = np.array([len(s) for s in shards], np.int64)
    starts = np.zeros(n, np.int64)
    nbytes = np.zeros(n, np.int64)
    frames_of = np.zeros(n, np.int64)
    rank_of = np.zeros(n, np.int32)
    if n == 0:
        return starts, nbytes, frames_of, rank_of
    order = np.concatenate([np.asarray(s, np.int64) for s in shards])
    frames = np.concatenate([np.asarray(table[r, 1:1 + len(s)], np.int64) for r, s in enumerate(shards)])
    nb = frames * int(bytes_per_frame)
    excl = np.cumsum(nb) - nb                                   # exclusive prefix over all ranks, rank-major
    first = np.minimum(np.concatenate([[0], np.cumsum(counts)[:-1]]), n - 1)   # first rank-major slot of every rank
    base = np.asarray(list(offs)[:world], np.int64) - excl[first]               # (unused for ranks without utterances)
    starts[order] = excl + np.repeat(base, counts)
    nbytes[order] = nb
    frames_of[order] = frames
    rank_of[order] = np.repeat(np.arange(world, dtype=np.int32), counts)
    return starts, nbytes, frames_of, rank_of


class ArenaViews(_SequenceABC):
    """audio[i] of a sharded result: a typed view of utterance i's bytes in the shared host arena, made on access
    (building 256 views per call cost 0.8 ms of host time on rank 0)."""
    __slots__ = ("whole", "starts", "nbytes", "dtype")

    def __init__(self, whole: np.ndarray, starts: np.ndarray, nbytes: np.ndarray, dtype):
        self.whole, self.dtype = whole, dtype
        self.starts, self.nbytes = starts.tolist(), nbytes.tolist()

    def __len__(self) -> int:
        return len(self.starts)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(len(self.starts)))]
        a = self.starts[i]
        return self.whole[a:a + self.nbytes[i]].view(self.dtype)


class ShardedResult:
    """Per-utterance results in the ORIGINAL order; `audio[i]` is a view into the shared host arena (valid until the
    next call on the same pool / synthesizer)."""

    def __init__(self, n: int):
        self.audio: List[Optional[np.ndarray]] = [None] * n
        self.pcm16: Optional[List[Optional[np.ndarray]]] = None
        self.frame_lengths = np.zeros(n, np.int64)
        self.durations: List[Optional[np.ndarray]] = [None] * n
        self.rank_of = np.zeros(n, np.int32)
        self.total_samples = 0
        self.total_frames = 0


def _is_short_buffer(e: Exception) -> bool:
    """PiperB200Error code 7 (capacity) raised because the caller-owned output buffer was too small."""
    return getattr(e, "code", None) == 7 and "output buffer" in str(e)


# ---- one process, N contexts ------------------------------------------------------------------------------------------
class PiperCUDAPool:
    """Several GPUs behind the PiperMetalRuntime-shaped API: `synthesize_batch` shards the utterances
    (shard_utterances), runs the shards concurrently - one context per device, one host thread each - and returns one
    result in the original utterance order whose audio lies in one pinned host arena."""

    def __init__(self, model_path: str, config_path: str, devices: Optional[Sequence[int]] = None, options=None,
                 runtime_factory: Optional[Callable] = None, bytes_per_id: int = 0, arena_factory: Optional[Callable] = None):
        from . import runtime as R
        if devices is None:
            devices = list(range(R.device_count()))
        if not devices:
            raise R.PiperB200Error(5, "no CUDA device visible (there is no CPU fallback)")
        base = options or R.Options()
        factory = runtime_factory or R.PiperCUDARuntime
        self._arena_factory = arena_factory or R.HostArena
        self.devices = list(devices)
        self.runtimes = [factory(model_path, config_path, R.Options(seed=base.seed, device=int(d), precision=base.precision))
                         for d in self.devices]
        self.info = self.runtimes[0].info
        self._threads = ThreadPoolExecutor(max_workers=len(self.devices), thread_name_prefix="piper_b200_gpu")
        self._arena = None
        # audio bytes reserved per phoneme id: float32 samples of 8 frames (stock voices speak ~3 frames per id);
        # doubled on demand when a call reports PIPER_B200_ERR_CAPACITY on its region
        self._bytes_per_id = bytes_per_id or 8 * int(self.info.hop) * 4
        self.last_shards: List[List[int]] = []

    @property
    def sample_rate(self) -> int:
        return int(self.info.sample_rate)

    def close(self):
        for rt in self.runtimes:
            rt.close()
        self.runtimes = []
        if self._arena is not None:
            self._arena.close()
            self._arena = None
        self._threads.shutdown(wait=True)

    def _ensure_arena(self, nbytes: int):
        if self._arena is None or self._arena.nbytes < nbytes:
            if self._arena is not None:
                for rt in self.runtimes:
                    rt.set_output_buffer(None)
                self._arena.close()
            self._arena = self._arena_factory(nbytes + nbytes // 4)
        return self._arena

    def synthesize(self, phoneme_ids: Sequence[int], noise_scale=None, length_scale=None, noise_w=None,
                   speaker_id: Optional[int] = None) -> np.ndarray:
        """PiperMetalRuntime.synthesize (PiperMetalRuntime.swift:62-80): one utterance, on the first device."""
        self.runtimes[0].set_output_buffer(None)      # a fresh array the caller owns, like the reference's [Float]
        return self.runtimes[0].synthesize(phoneme_ids, noise_scale, length_scale, noise_w, speaker_id)

    def synthesize_batch(self, utterances: Sequence[Sequence[int]], noise_scale=None, length_scale=None, noise_w=None,
                         speaker_ids: Optional[Sequence[int]] = None, seed: Optional[int] = None,
                         pcm16_only: bool = False) -> ShardedResult:
        n = len(utterances)
        lengths = [len(u) for u in utterances]
        shards = shard_utterances(lengths, len(self.runtimes))
        self.last_shards = shards
        while True:
            offs, caps = region_layout(shards, lengths, self._bytes_per_id // (2 if pcm16_only else 1))
            arena = self._ensure_arena(offs[-1] + caps[-1])

            def run(r: int):
                mine = shards[r]
                if not mine:
                    return None
                rt = self.runtimes[r]
                rt.set_output_buffer(arena.ptr + offs[r], caps[r])
                sids = None if speaker_ids is None else [int(speaker_ids[i]) for i in mine]
                return rt.synthesize_batch([utterances[i] for i in mine], noise_scale, length_scale, noise_w,
                                           speaker_ids=sids, seed=seed, copy=False, pcm16_only=pcm16_only)

            futures = [self._threads.submit(run, r) for r in range(len(self.runtimes))]
            results, short = [], False
            for f in futures:
                try:
                    results.append(f.result())
                except Exception as e:
                    if not _is_short_buffer(e):
                        raise
                    short = True
                    results.append(None)
            if not short:
                break
            self._bytes_per_id *= 2          # durations ran longer than the estimate: grow the regions, run again
        out = ShardedResult(n)
        if pcm16_only:
            out.pcm16 = [None] * n
        for r, (mine, res) in enumerate(zip(shards, results)):
            if res is None:
                continue
            for j, i in enumerate(mine):
                if pcm16_only:
                    out.pcm16[i] = res.pcm16[j]
                else:
                    out.audio[i] = res.audio[j]
                out.frame_lengths[i] = res.frame_lengths[j]
                out.durations[i] = res.durations[j]
                out.rank_of[i] = r
            out.total_samples += res.total_samples
            out.total_frames += res.total_frames
        return out


# ---- one process per GPU (torchrun) --------------------------------------------------------------------------------------
class _PlainRegion:
    """A region of the shared mapping WITHOUT CUDA pinning (host-logic tests on CPU)."""

    def __init__(self, view):
        import ctypes as C
        self._view = view
        self.nbytes = len(view)
        self.ptr = C.addressof(C.c_char.from_buffer(view))

    def close(self):
        self._view = None


class ShardedSynthesizer:
    """torch.distributed form: every rank holds one runtime (one GPU).  All ranks call `synthesize_batch` with the SAME
    global batch; each runs its shard and writes the audio into its region of a /dev/shm arena that all ranks map
    (rank r pins its own region, so its copies are asynchronous DMA).  One all_gather of the per-utterance frame counts
    (a few KB; NCCL or gloo, whatever the group is) tells rank 0 where each utterance lies; rank 0 returns the
    ShardedResult with views into the arena, the other ranks return their local result."""

    def __init__(self, runtime, hop: int, group=None, bytes_per_id: int = 0, arena_dir: str = "/dev/shm",
                 pin: bool = True, tag: Optional[str] = None):
        import torch.distributed as dist
        self._dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.runtime = runtime
        self.hop = int(hop)
        self._bytes_per_id = bytes_per_id or 8 * self.hop * 4
        self._pin = pin
        self._dir = arena_dir
        self._tag = tag or f"{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}"
        self._gen = 0
        self._fd = None
        self._map = None
        self._region = None
        self._arena_bytes = 0
        self._pinned: Optional[Tuple[int, int]] = None
        self.last_shards: List[List[int]] = []

    def _drop_region(self):
        if self._region is not None:
            self.runtime.set_output_buffer(None)
            self._region.close()
            self._region = None
        self._pinned = None

    def _drop_arena(self):
        self._drop_region()
        if self._map is not None:
            try:
                self._map.close()
            except BufferError:      # a caller still holds views into the old arena: leave it to the GC
                pass
            self._map = None
        if self._fd is not None:
            os.close(self._fd)
            self._fd = None
        self._arena_bytes = 0

    def _ensure_arena(self, nbytes: int, my_off: int, my_cap: int):
        """All ranks call this with the same nbytes.  Rank 0 creates and sizes the file, everyone maps it, every rank
        pins its own (page-aligned) region."""
        if self._arena_bytes < nbytes:
            self._drop_arena()
            self._gen += 1
            want = nbytes + nbytes // 4
            path = os.path.join(self._dir, f"piper_b200_arena_{self._tag}_{self._gen}")
            if self.rank == 0:
                fd = os.open(path, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
                os.ftruncate(fd, want)
            self._dist.barrier(self.group)
            if self.rank != 0:
                fd = os.open(path, os.O_RDWR)
            self._fd = fd
            self._map = mmap.mmap(fd, want)
            self._dist.barrier(self.group)
            if self.rank == 0:
                os.unlink(path)              # the mappings keep it alive; nothing is left behind if a rank dies
            self._arena_bytes = want
        if self._pinned != (my_off, my_cap):
            self._drop_region()
            view = memoryview(self._map)[my_off:my_off + my_cap]
            if self._pin:
                from . import runtime as R
                self._region = R.HostArena(my_cap, mapping=view)
            else:
                self._region = _PlainRegion(view)
            self._pinned = (my_off, my_cap)
        return self._region

    def close(self):
        self._drop_arena()

    def synthesize_batch(self, utterances: Sequence[Sequence[int]], noise_scale=None, length_scale=None, noise_w=None,
                         speaker_ids: Optional[Sequence[int]] = None, seed: Optional[int] = None,
                         pcm16_only: bool = False):
        import torch
        dist = self._dist
        n = len(utterances)
        lengths = [len(u) for u in utterances]
        shards = shard_utterances(lengths, self.world)
        self.last_shards = shards
        mine = shards[self.rank]
        itemsize = 2 if pcm16_only else 4
        dev = "cuda" if dist.get_backend(self.group) == "nccl" else "cpu"
        while True:
            offs, caps = region_layout(shards, lengths, self._bytes_per_id // (2 if pcm16_only else 1))
            region = self._ensure_arena(offs[-1] + caps[-1], offs[self.rank], caps[self.rank])
            self.runtime.set_output_buffer(region.ptr, caps[self.rank])
            res, short = None, 0
            if mine:
                sids = None if speaker_ids is None else [int(speaker_ids[i]) for i in mine]
                try:
                    res = self.runtime.synthesize_batch([utterances[i] for i in mine], noise_scale, length_scale, noise_w,
                                                        speaker_ids=sids, seed=seed, copy=False, pcm16_only=pcm16_only)
                except Exception as e:   # a short region must be learnt by every rank, or the gather below deadlocks
                    if not _is_short_buffer(e):
                        raise
                    short = 1
            # the one exchange: frame counts of every utterance (+ a "my region was too small" flag), padded to the
            # largest shard so that the tensor has the same shape on every rank
            width = max(len(s) for s in shards) + 1
            local = torch.zeros(width, dtype=torch.int64)
            local[0] = short
            if res is not None:
                local[1:1 + len(mine)] = torch.from_numpy(np.asarray(res.frame_lengths, np.int64))
            local = local.to(dev)
            gathered = [torch.empty_like(local) for _ in range(self.world)]
            dist.all_gather(gathered, local, group=self.group)
            table = torch.stack(gathered).cpu().numpy()
            if not table[:, 0].any():
                break
            self._bytes_per_id *= 2
        if self.rank != 0:
            return res
        out = ShardedResult(n)
        dt = np.int16 if pcm16_only else np.float32
        whole = np.frombuffer(self._map, dtype=np.uint8)
        starts, nbytes, frames_of, rank_of = arena_views_layout(shards, table, offs, self.hop * itemsize)
        out.frame_lengths[:] = frames_of
        out.rank_of[:] = rank_of
        out.total_frames = int(frames_of.sum())
        out.total_samples = out.total_frames * self.hop
        views = ArenaViews(whole, starts, nbytes, dt)
        if pcm16_only:
            out.pcm16 = views
        else:
            out.audio = views
        if res is not None:
            for j, i in enumerate(mine):
                out.durations[i] = res.durations[j]
        return out
