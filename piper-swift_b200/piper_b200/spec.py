"""Frozen hyper-parameter tables for the Piper-VITS voice architectures on the hot path.

The reference never names the architecture (it interprets the voice's ONNX graph,
GraphExecutor.swift:591-2663); the structure below is the rhasspy/piper VITS export the
reference's evidence points at (hidden 192 / 2 heads / 2T-1 relative
slots at GraphExecutor.swift:1872, initializer names at ONNXParsingTests.swift:31-32).
These tables are the single source of truth for the synthetic-voice generator, the
oracle and the CUDA runtime's shape checks.
"""
from __future__ import annotations

from dataclasses import dataclass, field, asdict
from typing import Tuple


@dataclass(frozen=True)
class VoiceSpec:
    name: str = "medium"
    n_vocab: int = 256
    sample_rate: int = 22050
    hidden: int = 192            # enc_p hidden == inter_channels
    filter_channels: int = 768   # FFN inner width
    n_heads: int = 2
    n_layers: int = 6
    ffn_kernel: int = 3
    window: int = 4              # relative-position window (emb_rel_* has 2*window+1 rows)
    # stochastic duration predictor
    dp_kernel: int = 3
    dp_layers: int = 3           # DDSConv layers
    dp_flows: int = 4            # ConvFlows built (one is dropped in reverse)
    dp_bins: int = 10
    dp_tail_bound: float = 5.0
    # residual coupling flow
    flow_n: int = 4
    flow_kernel: int = 5
    flow_wn_layers: int = 4
    # HiFi-GAN
    resblock: int = 2            # 1 -> ResBlock1, 2 -> ResBlock2
    dec_c0: int = 256
    ups_rates: Tuple[int, ...] = (8, 8, 4)
    ups_kernels: Tuple[int, ...] = (16, 16, 8)
    rb_kernels: Tuple[int, ...] = (3, 5, 7)
    rb_dilations: Tuple[Tuple[int, ...], ...] = ((1, 2), (2, 6), (3, 12))
    # speakers
    n_speakers: int = 1
    gin: int = 0

    @property
    def hop(self) -> int:
        h = 1
        for u in self.ups_rates:
            h *= u
        return h

    @property
    def k_channels(self) -> int:
        return self.hidden // self.n_heads

    def to_dict(self):
        return asdict(self)


MEDIUM = VoiceSpec()

HIGH = VoiceSpec(
    name="high",
    resblock=1,
    dec_c0=512,
    ups_rates=(8, 8, 2, 2),
    ups_kernels=(16, 16, 4, 4),
    rb_kernels=(3, 7, 11),
    rb_dilations=((1, 3, 5), (1, 3, 5), (1, 3, 5)),
)

MEDIUM_MULTISPEAKER = VoiceSpec(name="medium-multispeaker", n_speakers=904, gin=512)

# the x_low quality tier of rhasspy/piper (16 kHz, half-width text encoder / flow: hidden 96, FFN 384; SURVEY.md
# section 8f rank 3 "remaining voice families").  Head dim 48 and hidden 96 exercise the generic attention and
# fp32 conv kernels and the 96-wide tensor-core tiles.
X_LOW = VoiceSpec(name="x_low", sample_rate=16000, hidden=96, filter_channels=384)

# the `low` quality tier: the medium architecture trained on 16 kHz audio (piper_train --quality low keeps the default
# hidden 192 / ResBlock2 / (8, 8, 4) generator; only audio.sample_rate differs; VOICES.md:10-159 lists the tier)
LOW = VoiceSpec(name="low", sample_rate=16000)

# reduced-size variant used only to keep CPU unit tests fast (same structure, fewer channels)
TINY = VoiceSpec(
    name="tiny",
    n_vocab=64,
    hidden=32,
    filter_channels=64,
    n_layers=2,
    dec_c0=32,
    ups_rates=(4, 2),
    ups_kernels=(8, 4),
    rb_kernels=(3, 5),
    rb_dilations=((1, 2), (2, 3)),
)

SPECS = {s.name: s for s in (MEDIUM, HIGH, MEDIUM_MULTISPEAKER, LOW, X_LOW, TINY)}

# fixture of the reference: bench/fixtures/test_summary.json:8,16-18 and SmokeTests.swift:15-16
FIXTURE_IDS = [1, 20, 0, 120, 0, 61, 0, 24, 0, 59, 0, 100, 0, 2]
FIXTURE_SCALES = (0.667, 1.0, 0.8)   # noise_scale, length_scale, noise_w
DEFAULT_SEED = 1234                  # PiperMetalRuntime.swift:15, benchmark_onnxruntime.py:39

# The 50 ONNX op types of a Piper voice graph = the cases of the reference's executor switch
# (Sources/PiperMetal/Execution/GraphExecutor.swift:591-2663; anything else throws
# ExecutionError.unsupportedOp, :2661-2662).  tests/test_real_export_cpu.py checks that the upstream network restated
# in oracle/vits_modules.py, exported by torch's own ONNX exporter, uses exactly these (plus `Constant`, which the
# exporter of the voices' era wrote as initializers).
REFERENCE_GRAPH_OPS = (
    "Add", "And", "Cast", "Ceil", "Clip", "Concat", "ConstantOfShape", "Conv", "ConvTranspose", "CumSum", "Div",
    "Equal", "Erf", "Exp", "Expand", "Gather", "GatherElements", "GatherND", "GreaterOrEqual", "LeakyRelu", "Less",
    "LessOrEqual", "MatMul", "Mul", "Neg", "NonZero", "Not", "Pad", "Pow", "RandomNormalLike", "Range", "ReduceMax",
    "ReduceMean", "ReduceSum", "Relu", "Reshape", "ScatterND", "Shape", "Sigmoid", "Slice", "Softmax", "Softplus",
    "Split", "Sqrt", "Squeeze", "Sub", "Tanh", "Transpose", "Unsqueeze", "Where")
