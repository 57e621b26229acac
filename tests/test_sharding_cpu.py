"""Host logic of the multi-GPU path: deterministic utterance sharding and the final host gather,
exercised with a world_size-2 gloo group on CPU."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from piper_b200.sharding import shard_utterances, unshard, gather_lengths
from piper_b200.synth_voice import synthetic_utterances


def test_shard_is_a_balanced_partition():
    utts = synthetic_utterances(256, 256)
    lens = [len(u) for u in utts]
    for world in (1, 2, 4, 8):
        sh = shard_utterances(lens, world)
        flat = sorted(i for s in sh for i in s)
        assert flat == list(range(256))
        loads = [sum(lens[i] for i in s) for s in sh]
        assert max(loads) - min(loads) <= max(lens)
        assert unshard(sh, [[lens[i] for i in s] for s in sh]) == lens


def test_ragged_and_empty_shards():
    assert shard_utterances([5], 4) == [[0], [], [], []]
    assert shard_utterances([], 2) == [[], []]
    sh = shard_utterances([3, 3, 3], 2)
    assert sorted(i for s in sh for i in s) == [0, 1, 2]


def _worker(rank, world, port, lens, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = shard_utterances(lens, world)
    mine = sh[rank]
    fake_frames = [3 * lens[i] + rank for i in mine]      # stands in for this rank's y_lengths
    got = gather_lengths(mine, fake_frames, len(lens))
    dist.barrier()
    if rank == 0:
        q.put(got)
    dist.destroy_process_group()


def test_world_size_2_gloo_gather():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    lens = [len(u) for u in synthetic_utterances(16, 64, 20, 40)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lens, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    sh = shard_utterances(lens, 2)
    expect = [None] * len(lens)
    for r, s_ in enumerate(sh):
        for i in s_:
            expect[i] = 3 * lens[i] + r
    assert got == expect


# ---- the product's multi-GPU drivers, host logic only (fake one-"GPU" runtimes; the real ones: tests/test_gpu_pool.py) ----
import ctypes as C
from types import SimpleNamespace

from piper_b200 import sharding

HOP = 4


class _Short(RuntimeError):
    code = 7


class FakeRuntime:
    """Stands in for PiperCUDARuntime behind ShardedSynthesizer / PiperCUDAPool: 3 frames per id (+1 for odd speakers),
    sample s of an utterance = sum(ids) + s, written to the caller-owned output buffer like the C ABI does."""

    def __init__(self, *a, **k):
        self.ptr, self.cap = None, 0
        self.info = SimpleNamespace(hop=HOP, sample_rate=22050)

    def set_output_buffer(self, ptr, nbytes=0):
        self.ptr, self.cap = ptr, nbytes

    def close(self):
        pass

    @staticmethod
    def expected(u, sid=0):
        frames = 3 * len(u) + (sid % 2)
        return (np.arange(frames * HOP, dtype=np.float32) + np.float32(sum(u))), frames

    def synthesize_batch(self, utts, ns=None, ls=None, nw=None, speaker_ids=None, seed=None, copy=False,
                         pcm16_only=False):
        sids = speaker_ids or [0] * len(utts)
        audio = [self.expected(u, s) for u, s in zip(utts, sids)]
        total = sum(a.size for a, _ in audio)
        if total * 4 > self.cap:
            raise _Short(f"[capacity] the output buffer holds {self.cap} bytes, this result needs {total * 4}")
        buf = np.frombuffer((C.c_char * (total * 4)).from_address(self.ptr), dtype=np.float32)
        out, at = [], 0
        for a, _ in audio:
            buf[at:at + a.size] = a
            out.append(buf[at:at + a.size])
            at += a.size
        return SimpleNamespace(audio=out, pcm16=None, frame_lengths=np.array([f for _, f in audio], np.int64),
                               durations=[np.full(len(u), 3, np.int32) for u in utts], total_samples=total,
                               total_frames=sum(f for _, f in audio))


def _sharded_worker(rank, world, port, utts, sids, bytes_per_id, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    syn = sharding.ShardedSynthesizer(FakeRuntime(), HOP, bytes_per_id=bytes_per_id, pin=False, tag=f"test{port}")
    ok = True
    for _ in range(2):                       # second call reuses the arena
        res = syn.synthesize_batch(utts, speaker_ids=sids)
        if rank == 0:
            for i, u in enumerate(utts):
                exp, frames = FakeRuntime.expected(u, sids[i])
                ok = ok and res.frame_lengths[i] == frames and np.array_equal(res.audio[i], exp)
            ok = ok and res.total_samples == sum(a.size for a in res.audio)
            ok = ok and sorted(i for s in syn.last_shards for i in s) == list(range(len(utts)))
    dist.barrier()
    if rank == 0:
        q.put(bool(ok))
    del res
    syn.close()
    dist.destroy_process_group()


def _run_sharded(bytes_per_id):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    utts = synthetic_utterances(13, 64, 5, 30)
    sids = list(range(13))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sharded_worker, args=(r, 2, port, utts, sids, bytes_per_id, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert ok


def test_sharded_synthesizer_world_size_2_gloo():
    """every utterance comes back in the original order with its own audio, out of one /dev/shm arena"""
    _run_sharded(bytes_per_id=8 * HOP * 4)


def test_sharded_synthesizer_grows_a_short_arena():
    """a region that is too small is reported by the runtime (ERR_CAPACITY), agreed on by all ranks, and retried"""
    _run_sharded(bytes_per_id=HOP * 4)          # 1 frame per id reserved, 3 needed


def test_region_layout_is_disjoint_and_aligned():
    lens = [7, 3, 9, 1, 4]
    sh = shard_utterances(lens, 3)
    offs, caps = sharding.region_layout(sh, lens, 100)
    for r in range(3):
        assert offs[r] % 4096 == 0 and caps[r] % 4096 == 0
        assert caps[r] >= 100 * sum(lens[i] for i in sh[r])
        if r:
            assert offs[r] == offs[r - 1] + caps[r - 1]


class _PlainArena:                              # pinned allocation replaced by plain memory (no CUDA here)
    def __init__(self, nbytes):
        self.buf = (C.c_char * nbytes)()
        self.ptr, self.nbytes = C.addressof(self.buf), nbytes

    def close(self):
        pass


def test_pool_gathers_in_original_order():
    """PiperCUDAPool over three fake devices: same contract, threads instead of ranks"""
    pool = sharding.PiperCUDAPool("m", "c", devices=[0, 1, 2], runtime_factory=FakeRuntime, bytes_per_id=HOP * 4,
                                  arena_factory=_PlainArena)
    utts = synthetic_utterances(20, 64, 5, 30)
    sids = list(range(20))
    res = pool.synthesize_batch(utts, speaker_ids=sids)
    for i, u in enumerate(utts):
        exp, frames = FakeRuntime.expected(u, sids[i])
        assert res.frame_lengths[i] == frames and np.array_equal(res.audio[i], exp)
        assert res.rank_of[i] == next(r for r, s in enumerate(pool.last_shards) if i in s)
    assert pool._bytes_per_id > HOP * 4          # the short first estimate was grown
    pool.close()


def test_arena_views_layout_equals_the_per_utterance_walk():
    """sharding.arena_views_layout (vectorised: rank-major arrays + one scatter) against the plain walk over every rank's
    utterances, random shards including empty ranks and empty batches."""
    import numpy as np
    from piper_b200.sharding import arena_views_layout, shard_utterances, ArenaViews
    rng = np.random.default_rng(7)
    for _ in range(100):
        world, n = int(rng.integers(1, 9)), int(rng.integers(0, 40))
        lengths = [int(x) for x in rng.integers(1, 200, n)]
        shards = shard_utterances(lengths, world)
        table = np.zeros((world, max(len(s) for s in shards) + 1), np.int64)
        for r, s in enumerate(shards):
            table[r, 1:1 + len(s)] = rng.integers(0, 500, len(s))
        offs = [int(x) * 4096 for x in np.cumsum(rng.integers(0, 10 ** 6, world))]
        starts, nbytes, frames, rank_of = arena_views_layout(shards, table, offs, 1024)
        for r, s in enumerate(shards):
            at = offs[r]
            for j, i in enumerate(s):
                f = int(table[r, 1 + j])
                assert (starts[i], nbytes[i], frames[i], rank_of[i]) == (at, f * 1024, f, r)
                at += f * 1024
    whole = np.arange(64, dtype=np.uint8)
    v = ArenaViews(whole, np.array([8, 0, 16]), np.array([8, 4, 0]), np.float32)
    assert len(v) == 3 and v[0].dtype == np.float32 and v[0].size == 2 and v[1].size == 1 and v[2].size == 0
    assert v[0].tobytes() == whole[8:16].tobytes() and [a.size for a in v] == [2, 1, 0]
