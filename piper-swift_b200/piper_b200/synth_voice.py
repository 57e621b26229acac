"""Random-init ("synthetic") Piper voices of the named architectures.

Real voices are downloaded at run time by the reference (PiperVoices.swift:198-259) and
are unavailable offline, so every run uses random-init weights of the named voice
architecture (BASELINE.json north_star).  The generator is seed-fixed and calibrated so
that activations stay O(1) through the graph and durations come out at ~3 frames per
phoneme id.  It writes the same two files a Piper voice consists of:

  <voice>.onnx       initializers named like the upstream state-dict
                     (ONNXParsingTests.swift:31-32), graph signature
                     input/input_lengths/scales -> output, opset 15
                     (ONNXParsingTests.swift:22-27), first node Gather (:35), and
                     one Conv/ConvTranspose node per convolution carrying its
                     strides/dilations/pads/group attributes (the only place those
                     hyper-parameters live in a real export).
  <voice>.onnx.json  PiperConfig schema (PiperConfig.swift:3-46).
"""
from __future__ import annotations

import json
import math
import os
from typing import Dict, List, Tuple

import numpy as np

from . import onnx_proto as op
from .spec import VoiceSpec, SPECS, DEFAULT_SEED


class _Init:
    def __init__(self, seed: int):
        self.rng = np.random.Generator(np.random.PCG64(seed))

    def normal(self, shape, std):
        return (self.rng.standard_normal(shape, dtype=np.float32) * np.float32(std)).astype(np.float32)


def generate_state_dict(spec: VoiceSpec, seed: int = DEFAULT_SEED):
    """Returns (weights: name -> float32 ndarray, convs: list of conv descriptors)."""
    g = _Init(seed)
    w: Dict[str, np.ndarray] = {}
    convs: List[dict] = []
    H, Fc, kc = spec.hidden, spec.filter_channels, spec.k_channels

    def conv(name, cout, cin, k, std, bias=True, bias_std=0.02, dil=1, groups=1, pad=None):
        w[name + ".weight"] = g.normal((cout, cin // groups, k), std)
        if bias:
            w[name + ".bias"] = g.normal((cout,), bias_std)
        if pad is None:
            pad = (dil * (k - 1)) // 2
        convs.append(dict(name=name, op="Conv", k=k, stride=1, dil=dil, groups=groups,
                          pads=(pad, pad), bias=bias))

    def ln(name, c):
        w[name + ".gamma"] = (1.0 + g.normal((c,), 0.1)).astype(np.float32)
        w[name + ".beta"] = g.normal((c,), 0.1)

    # ---- text encoder -------------------------------------------------------------
    w["enc_p.emb.weight"] = g.normal((spec.n_vocab, H), H ** -0.5)
    for i in range(spec.n_layers):
        p = f"enc_p.encoder.attn_layers.{i}"
        w[p + ".emb_rel_k"] = g.normal((1, 2 * spec.window + 1, kc), kc ** -0.5)
        w[p + ".emb_rel_v"] = g.normal((1, 2 * spec.window + 1, kc), kc ** -0.5)
        for nm in ("conv_q", "conv_k", "conv_v", "conv_o"):
            conv(f"{p}.{nm}", H, H, 1, 1.0 / math.sqrt(H))
        ln(f"enc_p.encoder.norm_layers_1.{i}", H)
        f = f"enc_p.encoder.ffn_layers.{i}"
        k = spec.ffn_kernel
        # FFN uses asymmetric "same" padding done by an explicit Pad node; conv pads are 0
        conv(f"{f}.conv_1", Fc, H, k, 1.0 / math.sqrt(H * k), pad=0)
        conv(f"{f}.conv_2", H, Fc, k, 1.0 / math.sqrt(Fc * k), pad=0)
        ln(f"enc_p.encoder.norm_layers_2.{i}", H)
    conv("enc_p.proj", 2 * H, H, 1, 1.0 / math.sqrt(H))
    w["enc_p.proj.weight"][H:] *= 0.3
    w["enc_p.proj.bias"][H:] -= 0.7

    # ---- speaker embedding ----------------------------------------------------------
    gin = spec.gin if spec.n_speakers > 1 else 0
    if gin:
        w["emb_g.weight"] = g.normal((spec.n_speakers, gin), 1.0)

    # ---- stochastic duration predictor ----------------------------------------------
    def dds(prefix, c, k, n_layers):
        for i in range(n_layers):
            d = k ** i
            conv(f"{prefix}.convs_sep.{i}", c, c, k, 1.0 / math.sqrt(k), dil=d, groups=c)
            conv(f"{prefix}.convs_1x1.{i}", c, c, 1, 1.0 / math.sqrt(c))
            ln(f"{prefix}.norms_1.{i}", c)
            ln(f"{prefix}.norms_2.{i}", c)

    conv("dp.pre", H, H, 1, 1.0 / math.sqrt(H))
    conv("dp.proj", H, H, 1, 1.0 / math.sqrt(H))
    dds("dp.convs", H, spec.dp_kernel, spec.dp_layers)
    if gin:
        conv("dp.cond", H, gin, 1, 0.5 / math.sqrt(gin))
    w["dp.flows.0.m"] = np.array([[-2.5], [0.1]], dtype=np.float32)
    w["dp.flows.0.logs"] = np.array([[math.log(3.0)], [0.0]], dtype=np.float32)
    nparam = 3 * spec.dp_bins - 1
    for j in range(spec.dp_flows):
        p = f"dp.flows.{2 * j + 1}"
        conv(f"{p}.pre", H, 1, 1, 1.0)
        dds(f"{p}.convs", H, spec.dp_kernel, spec.dp_layers)
        conv(f"{p}.proj", nparam, H, 1, 2.0 / math.sqrt(H), bias_std=0.2)

    # ---- residual coupling flow -------------------------------------------------------
    half = H // 2
    for j in range(spec.flow_n):
        p = f"flow.flows.{2 * j}"
        conv(f"{p}.pre", H, half, 1, 1.0 / math.sqrt(half))
        for l in range(spec.flow_wn_layers):
            conv(f"{p}.enc.in_layers.{l}", 2 * H, H, spec.flow_kernel,
                 1.0 / math.sqrt(H * spec.flow_kernel))
            rs = 2 * H if l < spec.flow_wn_layers - 1 else H
            conv(f"{p}.enc.res_skip_layers.{l}", rs, H, 1, 0.7 / math.sqrt(H))
        if gin:
            conv(f"{p}.enc.cond_layer", 2 * H * spec.flow_wn_layers, gin, 1, 0.5 / math.sqrt(gin))
        conv(f"{p}.post", half, H, 1, 0.5 / math.sqrt(H))

    # ---- HiFi-GAN generator -------------------------------------------------------------
    conv("dec.conv_pre", spec.dec_c0, H, 7, 1.0 / math.sqrt(H * 7))
    if gin:
        conv("dec.cond", spec.dec_c0, gin, 1, 0.5 / math.sqrt(gin))
    ch = spec.dec_c0
    nk = len(spec.rb_kernels)
    for i, (u, k) in enumerate(zip(spec.ups_rates, spec.ups_kernels)):
        cin, ch = ch, ch // 2
        name = f"dec.ups.{i}"
        w[name + ".weight"] = g.normal((cin, ch, k), 1.4 / math.sqrt(cin * k / u))
        w[name + ".bias"] = g.normal((ch,), 0.02)
        convs.append(dict(name=name, op="ConvTranspose", k=k, stride=u, dil=1, groups=1,
                          pads=((k - u) // 2, (k - u) // 2), bias=True))
        for j, (rk, rd) in enumerate(zip(spec.rb_kernels, spec.rb_dilations)):
            p = f"dec.resblocks.{i * nk + j}"
            std = 0.6 / math.sqrt(ch * rk)
            if spec.resblock == 1:
                for m, d in enumerate(rd):
                    conv(f"{p}.convs1.{m}", ch, ch, rk, std, dil=d)
                    conv(f"{p}.convs2.{m}", ch, ch, rk, std, dil=1)
            else:
                for m, d in enumerate(rd):
                    conv(f"{p}.convs.{m}", ch, ch, rk, std, dil=d)
    conv("dec.conv_post", 1, ch, 7, 0.45 / math.sqrt(ch * 7), bias=False)
    return w, convs


def _config_json(spec: VoiceSpec) -> dict:
    id_map = {"_": [0], "^": [1], "$": [2]}
    alphabet = " !',-.:;?abcdefghijklmnopqrstuvwxyz"
    for i, chh in enumerate(alphabet):
        id_map[chh] = [3 + i]
    cfg = {
        "audio": {"sample_rate": spec.sample_rate, "quality": spec.name},
        "espeak": {"voice": "en-gb"},
        "inference": {"noise_scale": 0.667, "length_scale": 1.0, "noise_w": 0.8},
        "phoneme_type": "espeak",
        "phoneme_map": {},
        "phoneme_id_map": id_map,
        "num_symbols": spec.n_vocab,
        "num_speakers": spec.n_speakers,
        "speaker_id_map": ({f"spk{i}": i for i in range(spec.n_speakers)} if spec.n_speakers > 1 else {}),
        "piper_version": "synthetic-b200",
        "language": {"code": "en_GB"},
        "dataset": "synthetic",
    }
    return cfg


def build_onnx_bytes(spec: VoiceSpec, weights: Dict[str, np.ndarray], convs: List[dict]) -> bytes:
    nodes = [op.node("Gather", ["enc_p.emb.weight", "input"], ["/enc_p/emb/Gather_output_0"],
                     name="/enc_p/emb/Gather", attrs=[("axis", 0)])]
    for c in convs:
        nm = c["name"]
        path = "/" + nm.replace(".", "/")
        ins = [path + "/in", nm + ".weight"] + ([nm + ".bias"] if c["bias"] else [])
        attrs = [("dilations", [c["dil"]]), ("group", c["groups"]), ("kernel_shape", [c["k"]]),
                 ("pads", list(c["pads"])), ("strides", [c["stride"]])]
        nodes.append(op.node(c["op"], ins, [path + f"/{c['op']}_output_0"],
                             name=path + "/" + c["op"], attrs=attrs))
    inits = [op.tensor(k, v) for k, v in weights.items()]
    if spec.n_speakers <= 1:
        # single-speaker exports carry a dummy `sid` initializer (ONNXParsingTests.swift:31)
        inits.append(op.tensor("sid", np.zeros((1,), dtype=np.int64)))
        inputs = ["input", "input_lengths", "scales"]
    else:
        inputs = ["input", "input_lengths", "scales", "sid"]
    return op.model(nodes, inits, inputs, ["output"])


def write_voice(spec: VoiceSpec, out_dir: str, seed: int = DEFAULT_SEED) -> Tuple[str, str]:
    """Writes <out_dir>/<spec.name>-seed<seed>.onnx(.json); returns (model_path, config_path)."""
    os.makedirs(out_dir, exist_ok=True)
    base = os.path.join(out_dir, f"synthetic-{spec.name}-seed{seed}.onnx")
    cfgp = base + ".json"
    if not (os.path.exists(base) and os.path.exists(cfgp)):
        weights, convs = generate_state_dict(spec, seed)
        blob = build_onnx_bytes(spec, weights, convs)
        tmp = base + f".tmp{os.getpid()}"
        with open(tmp, "wb") as f:
            f.write(blob)
        os.replace(tmp, base)
        with open(cfgp + f".tmp{os.getpid()}", "w") as f:
            json.dump(_config_json(spec), f)
        os.replace(cfgp + f".tmp{os.getpid()}", cfgp)
    return base, cfgp


def synthetic_utterances(n: int, n_vocab: int, t_lo: int = 80, t_hi: int = 120, seed: int = 2024):
    """Config-3/5 style inputs: BOS 1, EOS 2, blank 0 interleaved
    (structure of ESpeakPhonemizer.swift:73-100), ids uniform over [3, n_vocab)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = []
    for _ in range(n):
        t = int(rng.integers(t_lo, t_hi + 1))
        n_sym = max(1, (t - 3) // 2)
        syms = rng.integers(3, n_vocab, size=n_sym)
        ids = [1]
        for s in syms:
            ids += [0, int(s)]
        ids += [0, 2]
        out.append(ids)
    return out


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--spec", default="medium", choices=sorted(SPECS))
    ap.add_argument("--out", default="/tmp/piper_b200_voices")
    ap.add_argument("--seed", type=int, default=DEFAULT_SEED)
    a = ap.parse_args()
    print(write_voice(SPECS[a.spec], a.out, a.seed))
