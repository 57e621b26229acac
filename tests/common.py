"""Shared helpers for the parity tests: synthetic voices, the oracle, error metrics."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "piper-swift_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from piper_b200 import spec as S          # noqa: E402
from piper_b200 import synth_voice       # noqa: E402

VOICE_DIR = os.environ.get("PIPER_B200_VOICE_DIR", "/tmp/piper_b200_voices")
_voices = {}
_oracles = {}


def voice_files(name: str):
    if name not in _voices:
        _voices[name] = synth_voice.write_voice(S.SPECS[name], VOICE_DIR)
    return _voices[name]


def oracle(name: str, dtype=None):
    import torch
    from oracle.vits_oracle import VitsOracle
    dtype = dtype or torch.float32
    key = (name, dtype)
    if key not in _oracles:
        _oracles[key] = VitsOracle.from_file(voice_files(name)[0], dtype)
    return _oracles[key]


def snr_db(ref: np.ndarray, x: np.ndarray) -> float:
    ref = ref.astype(np.float64)
    err = x.astype(np.float64) - ref
    return float(10 * np.log10((ref ** 2).sum() / max((err ** 2).sum(), 1e-300)))


def fixture_ids(factor: int = 1, n_vocab: int = 256):
    """np.tile scaling of the 14-id fixture (benchmark_onnxruntime.py:108-116)."""
    return [i % n_vocab for i in S.FIXTURE_IDS * factor]
