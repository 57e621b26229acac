"""Utterance-level sharding across the GPUs of one box: independent units, no
collective on the data path, a final gather of per-utterance lengths / audio on the host.
The reference has no counterpart (single MTLDevice, MetalContext.swift:11-19)."""
from __future__ import annotations

from typing import List, Sequence, Tuple


def shard_utterances(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """LPT-greedy: longest utterance first onto the least-loaded rank (work ~ frames ~ 3*T).
    Deterministic; every rank computes the same assignment.  Returns utterance indices per rank,
    each in ascending original order."""
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    load = [0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += int(lengths[i])
    return [sorted(s) for s in out]


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
