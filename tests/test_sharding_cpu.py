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
