"""TEST INFRASTRUCTURE (oracle).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.

numpy restatement of the reference's RandomNormalLike generator
(`random_normal_like_f32`, Sources/PiperMetal/Kernels/elementwise.metal:132-163):
state = seedLo ^ (gid*747796405 + 2891336453); two xorshift32 steps give u0,u1; both are
mapped to (0,1] in float32; z = sqrt(-2 ln u0) * cos(2 pi u1).  Only the low 32 bits of
the seed are used and both noise nodes of a run use the same seed
(GraphExecutor.swift:2657-2658), so gid is the flat index inside each noise tensor
([1,2,T] for the duration predictor, [1,192,F] for the prior).
"""
import numpy as np


def _xorshift32(x: np.ndarray) -> np.ndarray:
    x = x ^ ((x << np.uint32(13)) & np.uint32(0xFFFFFFFF))
    x = x ^ (x >> np.uint32(17))
    x = x ^ ((x << np.uint32(5)) & np.uint32(0xFFFFFFFF))
    return x


def random_normal_like(count: int, seed: int = 1234) -> np.ndarray:
    gid = np.arange(count, dtype=np.uint64)
    mixed = (gid * np.uint64(747796405) + np.uint64(2891336453)) & np.uint64(0xFFFFFFFF)
    state = (np.uint32(seed & 0xFFFFFFFF) ^ mixed.astype(np.uint32)).astype(np.uint32)
    state = _xorshift32(state)
    u0i = state
    state = _xorshift32(state)
    u1i = state
    one = np.float32(1.0)
    inv = np.float32(4294967296.0)
    u0 = (u0i.astype(np.float32) + one) / inv
    u1 = (u1i.astype(np.float32) + one) / inv
    r = np.sqrt(np.float32(-2.0) * np.log(u0))
    theta = np.float32(6.28318530718) * u1
    return (r * np.cos(theta)).astype(np.float32)


def noise_dp(T: int, seed: int = 1234) -> np.ndarray:
    """[2, T] noise of the duration predictor for one utterance."""
    return random_normal_like(2 * T, seed).reshape(2, T)


def noise_main(C: int, F: int, seed: int = 1234) -> np.ndarray:
    """[C, F] prior noise for one utterance."""
    return random_normal_like(C * F, seed).reshape(C, F)
