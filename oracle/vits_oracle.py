"""TEST INFRASTRUCTURE (oracle).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this; the product path never does and
fails loudly without its CUDA library.

CPU restatement (torch eager, fp32 or fp64) of what the reference computes for
`PiperMetalRuntime.synthesize` (PiperMetalRuntime.swift:62-80): the voice's ONNX graph
evaluated op by op (GraphExecutor.swift:591-2663).  The graph itself is a third-party
artefact that is NOT under /root/reference: a rhasspy/piper `piper_train` VITS export
(opset 15; `rhasspy/piper-voices` on Hugging Face, VOICES.md:10-159), normally evaluated by
onnxruntime==1.19.2 (bench/requirements.txt:2, call sites benchmark_onnxruntime.py:105,131).
Neither is installed/available offline, so this file restates the published VITS inference
algorithm with ONNX-op semantics as the reference implements them:
  - Conv / ConvTranspose indexing and weight layout: conv1d.metal:28-71, :97-142
  - Softmax over the last axis: softmax.metal:13-41; MatMul: matmul.metal:22-49
  - LayerNorm decomposed as ReduceMean/Sub/Pow/ReduceMean/Add/Sqrt/Div/Mul/Add
    (GraphExecutor.swift:2104-2125, 2071, 2089)
  - Ceil/Clip/Cast(f32->i64 trunc)/CumSum/Range/Less length regulator
    (elementwise.metal:224-232, cast.metal:8-29, CPUBackend.swift:1415-1498)
  - exact erf for GELU (the ORT-CPU target, not the A&S approximation of
    elementwise.metal:292-312); noise tensors are injected (the reference's
    `overrides` hook, GraphExecutor.swift:190-193, 2647-2651).

PARITY UNPINNED: the reference holds no golden output for this path (fixture audio is
empty, bench/fixtures/test_summary.json:20-32; SmokeTests.swift:18-19 checks finiteness
only).  What it pins - graph signature and the 14-id fixture - is checked in tests/.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1


@dataclass
class VoiceParams:
    """Hyper-parameters recovered from initializer shapes + Conv node attributes."""
    n_vocab: int
    hidden: int
    filter_channels: int
    n_heads: int
    n_layers: int
    ffn_kernel: int
    window: int
    dp_layers: int
    dp_kernel: int
    dp_flow_ids: List[int]       # indices i of dp.flows.i that are ConvFlows
    dp_bins: int
    flow_ids: List[int]          # indices of flow.flows.i that are coupling layers
    flow_kernel: int
    flow_wn_layers: int
    resblock: int
    dec_c0: int
    ups: List[Tuple[int, int]]   # (kernel, stride)
    rb_kernels: List[int]
    rb_dilations: List[List[int]]
    gin: int
    n_speakers: int

    @property
    def hop(self):
        h = 1
        for _, u in self.ups:
            h *= u
        return h


def canonicalize(inits: Dict[str, np.ndarray], nodes) -> Dict[str, np.ndarray]:
    """Real Piper exports do not name every weight after its module: piper_train removes weight-norm from
    `dec` only, and ONNX constant folding turns the flow's weight_g / weight_v pairs into initializers called
    `onnx::Conv_NNNN`; `exp(-logs)` of the duration predictor's ElementwiseAffine is folded the same way
    (`-logs` survives as `onnx::Exp_NNNN` or the whole factor as an anonymous Mul operand).  The reference never
    needs names - it walks the nodes (GraphExecutor.swift:227-264) - so this restates what the graph says:
    every Conv / ConvTranspose node's weight input is the `<module>.weight` of the module named by its bias
    input (`<module>.bias`) or by its node name (`/flow/flows.0/enc/in_layers.0/Conv`).  Returns a copy of
    `inits` with the canonical names added (and `dp.flows.0.inv_scale` when only the folded factor exists)."""
    out = dict(inits)
    for name in list(inits):          # unfolded weight-norm pairs: w = g * v / ||v|| over all axes but the first
        if name.endswith(".weight_v"):
            base = name[:-len(".weight_v")]
            if base + ".weight" in out or base + ".weight_g" not in inits:
                continue
            v = np.asarray(inits[name], np.float64)
            g = np.asarray(inits[base + ".weight_g"], np.float64).reshape(-1)
            norm = np.sqrt((v.reshape(v.shape[0], -1) ** 2).sum(axis=1))
            out[base + ".weight"] = (v * (g / norm).reshape((-1,) + (1,) * (v.ndim - 1))).astype(np.float32)
    for n in nodes:
        if n.op_type not in ("Conv", "ConvTranspose") or len(n.inputs) < 2:
            continue
        w_in = n.inputs[1]
        b_in = n.inputs[2] if len(n.inputs) >= 3 else ""
        if w_in.endswith(".weight") and w_in in inits:
            continue
        if b_in.endswith(".bias"):
            mod = b_in[:-len(".bias")]
        elif n.name.startswith("/") and n.name.rsplit("/", 1)[-1].startswith("Conv"):
            mod = n.name[1:].rsplit("/", 1)[0].replace("/", ".")
        else:
            continue
        if mod + ".weight" not in out and w_in in inits:
            out[mod + ".weight"] = inits[w_in]
    # The phoneme embedding: piper_train exports single-speaker voices with input_names = [input, input_lengths,
    # scales, sid] but only three real inputs, so torch hands the spare name "sid" to the next graph input - the first
    # parameter, enc_p.emb.weight.  A real single-speaker voice therefore has NO `enc_p.emb.weight`: its table is the
    # initializer called `sid` (which is what ONNXParsingTests.swift:31 finds), read by the first node,
    # Gather(sid, input) (ONNXParsingTests.swift:35).
    if "enc_p.emb.weight" not in out:
        for n in nodes:
            if n.op_type == "Gather" and len(n.inputs) == 2 and n.inputs[1] == "input" and n.inputs[0] in inits:
                out["enc_p.emb.weight"] = inits[n.inputs[0]]
                break
    if "dp.flows.0.logs" not in out:
        consts = {}
        for n in nodes:
            if n.op_type == "Constant" and n.outputs and isinstance(n.attrs.get("value"), np.ndarray):
                consts[n.outputs[0]] = n.attrs["value"]
        by_out = {o: n for n in nodes for o in n.outputs}
        diff = [n.outputs[0] for n in nodes if n.op_type == "Sub" and len(n.inputs) == 2 and n.inputs[1] == "dp.flows.0.m"]
        for n in nodes:
            if not diff or n.op_type != "Mul" or len(n.inputs) != 2 or diff[-1] not in n.inputs:
                continue
            other = n.inputs[1] if n.inputs[0] == diff[-1] else n.inputs[0]
            src = by_out.get(other)
            if src is not None and src.op_type == "Exp" and src.inputs[0] in inits:     # Exp(-logs): -logs survived
                out["dp.flows.0.logs"] = -np.asarray(inits[src.inputs[0]], np.float32)
                break
            val = inits.get(other, consts.get(other))
            if val is not None and np.asarray(val).size in (1, 2) and np.asarray(val).dtype == np.float32:
                out["dp.flows.0.inv_scale"] = np.asarray(val, np.float32)
                break
    return out


def derive_params(inits: Dict[str, np.ndarray], nodes) -> VoiceParams:
    attrs = {}
    for n in nodes:
        if n.op_type in ("Conv", "ConvTranspose") and len(n.inputs) >= 2:
            attrs[n.inputs[1]] = n.attrs
            b_in = n.inputs[2] if len(n.inputs) >= 3 else ""
            if b_in.endswith(".bias"):                         # canonical name of an anonymous weight (canonicalize)
                attrs.setdefault(b_in[:-5] + ".weight", n.attrs)
            elif n.name.startswith("/") and n.name.rsplit("/", 1)[-1].startswith("Conv"):
                attrs.setdefault(n.name[1:].rsplit("/", 1)[0].replace("/", ".") + ".weight", n.attrs)

    def count(fmt):
        i = 0
        while fmt.format(i) in inits:
            i += 1
        return i

    emb = inits["enc_p.emb.weight"]
    hidden = emb.shape[1]
    relk = inits["enc_p.encoder.attn_layers.0.emb_rel_k"]
    n_layers = count("enc_p.encoder.attn_layers.{}.conv_q.weight")
    c1 = inits["enc_p.encoder.ffn_layers.0.conv_1.weight"]
    dp_flow_ids = [i for i in range(32) if f"dp.flows.{i}.proj.weight" in inits]
    flow_ids = [i for i in range(32) if f"flow.flows.{i}.pre.weight" in inits]
    fl0 = f"flow.flows.{flow_ids[0]}"
    nparam = inits[f"dp.flows.{dp_flow_ids[0]}.proj.weight"].shape[0]
    n_ups = count("dec.ups.{}.weight")
    ups = []
    for i in range(n_ups):
        wname = f"dec.ups.{i}.weight"
        k = inits[wname].shape[2]
        a = attrs.get(wname, {})
        u = int(a["strides"][0]) if "strides" in a else k // 2
        ups.append((k, u))
    resblock = 1 if "dec.resblocks.0.convs1.0.weight" in inits else 2
    n_rb = count("dec.resblocks.{}.convs1.0.weight" if resblock == 1 else "dec.resblocks.{}.convs.0.weight")
    nk = n_rb // n_ups
    rb_kernels, rb_dil = [], []
    for j in range(nk):
        pre = f"dec.resblocks.{j}.convs1." if resblock == 1 else f"dec.resblocks.{j}.convs."
        rb_kernels.append(inits[pre + "0.weight"].shape[2])
        dl = []
        m = 0
        while pre + f"{m}.weight" in inits:
            a = attrs.get(pre + f"{m}.weight", {})
            dl.append(int(a["dilations"][0]) if "dilations" in a else 1)
            m += 1
        rb_dil.append(dl)
    gin = inits["emb_g.weight"].shape[1] if "emb_g.weight" in inits else 0
    return VoiceParams(
        n_vocab=emb.shape[0], hidden=hidden, filter_channels=c1.shape[0],
        n_heads=hidden // relk.shape[2], n_layers=n_layers, ffn_kernel=c1.shape[2],
        window=(relk.shape[1] - 1) // 2,
        dp_layers=count("dp.convs.convs_sep.{}.weight"),
        dp_kernel=inits["dp.convs.convs_sep.0.weight"].shape[2],
        dp_flow_ids=dp_flow_ids, dp_bins=(nparam + 1) // 3,
        flow_ids=flow_ids, flow_kernel=inits[fl0 + ".enc.in_layers.0.weight"].shape[2],
        flow_wn_layers=count(fl0 + ".enc.in_layers.{}.weight"),
        resblock=resblock, dec_c0=inits["dec.conv_pre.weight"].shape[0], ups=ups,
        rb_kernels=rb_kernels, rb_dilations=rb_dil, gin=gin,
        n_speakers=inits["emb_g.weight"].shape[0] if gin else 1)


def _layer_norm_c(x, gamma, beta, eps=1e-5):
    """Channel LayerNorm on [B,C,T], decomposed exactly like the opset-15 export
    (ReduceMean, Sub, Pow 2, ReduceMean, Add eps, Sqrt, Div, Mul, Add)."""
    xt = x.transpose(1, -1)
    mean = xt.mean(-1, keepdim=True)
    d = xt - mean
    var = (d * d).mean(-1, keepdim=True)
    y = d / torch.sqrt(var + eps)
    y = y * gamma + beta
    return y.transpose(1, -1)


def _gelu(x):
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def _searchsorted(bin_locations, inputs, eps=1e-6):
    bin_locations = bin_locations.clone()
    bin_locations[..., -1] += eps
    return torch.sum(inputs[..., None] >= bin_locations, dim=-1) - 1


def _rq_spline_inverse(inputs, uw, uh, ud, tail_bound, min_bin_width=1e-3, min_bin_height=1e-3,
                       min_derivative=1e-3):
    """Inverse unconstrained rational-quadratic spline with linear tails
    (piper_train/vits/transforms.py), evaluated for every element; elements outside
    [-tail_bound, tail_bound] pass through unchanged (the NonZero/GatherND/ScatterND
    masked-assignment subgraph, GraphExecutor.swift:2379-2602, computes the same thing)."""
    inside = (inputs >= -tail_bound) & (inputs <= tail_bound)
    num_bins = uw.shape[-1]
    ud = F.pad(ud, (1, 1))
    constant = math.log(math.exp(1 - min_derivative) - 1)
    ud[..., 0] = constant
    ud[..., -1] = constant
    left = bottom = -tail_bound
    right = top = tail_bound

    widths = F.softmax(uw, dim=-1)
    widths = min_bin_width + (1 - min_bin_width * num_bins) * widths
    cumwidths = torch.cumsum(widths, dim=-1)
    cumwidths = F.pad(cumwidths, (1, 0), value=0.0)
    cumwidths = (right - left) * cumwidths + left
    cumwidths[..., 0] = left
    cumwidths[..., -1] = right
    widths = cumwidths[..., 1:] - cumwidths[..., :-1]

    derivatives = min_derivative + F.softplus(ud)

    heights = F.softmax(uh, dim=-1)
    heights = min_bin_height + (1 - min_bin_height * num_bins) * heights
    cumheights = torch.cumsum(heights, dim=-1)
    cumheights = F.pad(cumheights, (1, 0), value=0.0)
    cumheights = (top - bottom) * cumheights + bottom
    cumheights[..., 0] = bottom
    cumheights[..., -1] = top
    heights = cumheights[..., 1:] - cumheights[..., :-1]

    x_in = torch.where(inside, inputs, torch.zeros_like(inputs))   # keep the gather in range
    bin_idx = _searchsorted(cumheights, x_in).clamp(0, num_bins - 1)[..., None]

    input_cumwidths = cumwidths.gather(-1, bin_idx)[..., 0]
    input_bin_widths = widths.gather(-1, bin_idx)[..., 0]
    input_cumheights = cumheights.gather(-1, bin_idx)[..., 0]
    delta = heights / widths
    input_delta = delta.gather(-1, bin_idx)[..., 0]
    input_derivatives = derivatives.gather(-1, bin_idx)[..., 0]
    input_derivatives_plus_one = derivatives[..., 1:].gather(-1, bin_idx)[..., 0]
    input_heights = heights.gather(-1, bin_idx)[..., 0]

    s = input_derivatives + input_derivatives_plus_one - 2 * input_delta
    a = (x_in - input_cumheights) * s + input_heights * (input_delta - input_derivatives)
    b = input_heights * input_derivatives - (x_in - input_cumheights) * s
    c = -input_delta * (x_in - input_cumheights)
    discriminant = b.pow(2) - 4 * a * c
    root = (2 * c) / (-b - torch.sqrt(discriminant))
    outputs = root * input_bin_widths + input_cumwidths
    return torch.where(inside, outputs, inputs), bin_idx[..., 0]


class VitsOracle:
    """B=1 evaluation per utterance, exactly what the reference can run
    (input hard-wired to [1, phonemes], GraphExecutor.swift:108,196)."""

    def __init__(self, inits: Dict[str, np.ndarray], params: VoiceParams, dtype=torch.float32):
        self.p = params
        self.dtype = dtype
        self.w = {k: torch.from_numpy(np.asarray(v)).to(dtype) for k, v in inits.items()
                  if np.asarray(v).dtype == np.float32}

    @classmethod
    def from_file(cls, path: str, dtype=torch.float32):
        from .onnx_reader import load_model
        m = load_model(path)
        inits = canonicalize(m.initializers, m.nodes)
        return cls(inits, derive_params(inits, m.nodes), dtype)

    # ---- helpers ------------------------------------------------------------------
    def _conv(self, x, name, dilation=1, padding=0, groups=1):
        return F.conv1d(x, self.w[name + ".weight"], self.w.get(name + ".bias"), 1, padding, dilation, groups)

    def _ln(self, x, name):
        return _layer_norm_c(x, self.w[name + ".gamma"], self.w[name + ".beta"])

    # ---- text encoder -----------------------------------------------
    def _rel_emb(self, emb, length):
        ws = self.p.window
        pad_length = max(length - (ws + 1), 0)
        slice_start = max((ws + 1) - length, 0)
        slice_end = slice_start + 2 * length - 1
        if pad_length > 0:
            emb = F.pad(emb, (0, 0, pad_length, pad_length))
        return emb[:, slice_start:slice_end]

    @staticmethod
    def _rel_to_abs(x):
        b, h, l, _ = x.shape
        x = F.pad(x, (0, 1))
        x_flat = x.view(b, h, l * 2 * l)
        x_flat = F.pad(x_flat, (0, l - 1))
        return x_flat.view(b, h, l + 1, 2 * l - 1)[:, :, :l, l - 1:]

    @staticmethod
    def _abs_to_rel(x):
        b, h, l, _ = x.shape
        x = F.pad(x, (0, l - 1))
        x_flat = x.view(b, h, l * l + l * (l - 1))
        x_flat = F.pad(x_flat, (l, 0))
        return x_flat.view(b, h, l, 2 * l)[:, :, :, 1:]

    def _attention(self, x, i):
        p = self.p
        pre = f"enc_p.encoder.attn_layers.{i}"
        q = self._conv(x, pre + ".conv_q")
        k = self._conv(x, pre + ".conv_k")
        v = self._conv(x, pre + ".conv_v")
        b, d, t = q.shape
        kc = d // p.n_heads
        q = q.view(b, p.n_heads, kc, t).transpose(2, 3)
        k = k.view(b, p.n_heads, kc, t).transpose(2, 3)
        v = v.view(b, p.n_heads, kc, t).transpose(2, 3)
        qs = q / math.sqrt(kc)
        scores = torch.matmul(qs, k.transpose(-2, -1))
        rel_k = self._rel_emb(self.w[pre + ".emb_rel_k"], t)
        rel_logits = torch.matmul(qs, rel_k.unsqueeze(0).transpose(-2, -1))
        scores = scores + self._rel_to_abs(rel_logits)
        # mask is all ones for B=1 (masked_fill(mask==0,-1e4) is a no-op)
        p_attn = F.softmax(scores, dim=-1)
        out = torch.matmul(p_attn, v)
        rel_w = self._abs_to_rel(p_attn)
        rel_v = self._rel_emb(self.w[pre + ".emb_rel_v"], t)
        out = out + torch.matmul(rel_w, rel_v.unsqueeze(0))
        out = out.transpose(2, 3).contiguous().view(b, d, t)
        return self._conv(out, pre + ".conv_o")

    def text_encoder(self, ids: torch.Tensor):
        p = self.p
        x = self.w["enc_p.emb.weight"][ids] * math.sqrt(p.hidden)   # [1,T,H]
        x = x.transpose(1, -1)
        k = p.ffn_kernel
        pl, pr = (k - 1) // 2, k // 2
        for i in range(p.n_layers):
            y = self._attention(x, i)
            x = self._ln(x + y, f"enc_p.encoder.norm_layers_1.{i}")
            f = f"enc_p.encoder.ffn_layers.{i}"
            y = self._conv(F.pad(x, (pl, pr)), f + ".conv_1")
            y = torch.relu(y)
            y = self._conv(F.pad(y, (pl, pr)), f + ".conv_2")
            x = self._ln(x + y, f"enc_p.encoder.norm_layers_2.{i}")
        stats = self._conv(x, "enc_p.proj")
        m, logs = torch.split(stats, p.hidden, dim=1)
        return x, m, logs

    # ---- stochastic duration predictor, reverse -------------------------
    def _dds(self, x, prefix, g=None):
        p = self.p
        if g is not None:
            x = x + g
        for i in range(p.dp_layers):
            d = p.dp_kernel ** i
            pad = (p.dp_kernel * d - d) // 2
            y = self._conv(x, f"{prefix}.convs_sep.{i}", dilation=d, padding=pad, groups=x.shape[1])
            y = _gelu(self._ln(y, f"{prefix}.norms_1.{i}"))
            y = self._conv(y, f"{prefix}.convs_1x1.{i}")
            y = _gelu(self._ln(y, f"{prefix}.norms_2.{i}"))
            x = x + y
        return x

    def _conv_flow_reverse(self, z, cond, idx):
        p = self.p
        pre = f"dp.flows.{idx}"
        x0, x1 = z[:, :1], z[:, 1:]
        h = self._conv(x0, pre + ".pre")
        h = self._dds(h, pre + ".convs", g=cond)
        h = self._conv(h, pre + ".proj")
        b, c, t = x0.shape
        h = h.reshape(b, c, -1, t).permute(0, 1, 3, 2)
        nb = p.dp_bins
        scale = math.sqrt(p.hidden)      # filter_channels of the SDP == in_channels
        uw = h[..., :nb] / scale
        uh = h[..., nb:2 * nb] / scale
        ud = h[..., 2 * nb:]
        x1n, bins = _rq_spline_inverse(x1, uw, uh, ud, 5.0)
        return torch.cat([x0, x1n], 1), bins

    def duration_predictor(self, x, noise_dp, noise_w, g=None):
        """x [1,H,T] (encoder output), noise_dp [1,2,T] standard normal -> logw [1,1,T]."""
        p = self.p
        h = self._conv(x, "dp.pre")
        if g is not None:
            h = h + self._conv(g, "dp.cond")
        h = self._dds(h, "dp.convs")
        h = self._conv(h, "dp.proj")
        # flows = reversed([EA, CF, Flip, CF, Flip, ...]) with the second-to-last dropped
        # (the dropped one is dp.flows.1; an exported graph no longer holds its weights, so it may be absent from the
        # initializers already - its Flip, dp.flows.2, stays either way)
        seq = [("ea", 0)]
        for i in sorted(set(p.dp_flow_ids) | {1}):
            seq += [("cf", i), ("flip", None)]
        seq = list(reversed(seq))
        seq = seq[:-2] + [seq[-1]]
        z = noise_dp * noise_w
        self.last_bins = []
        for kind, i in seq:
            if kind == "flip":
                z = torch.flip(z, [1])
            elif kind == "cf":
                z, bins = self._conv_flow_reverse(z, h, i)
                self.last_bins.append(bins)
            else:
                m = self.w["dp.flows.0.m"]
                if "dp.flows.0.logs" in self.w:
                    z = (z - m) * torch.exp(-self.w["dp.flows.0.logs"])
                else:                               # the exporter folded exp(-logs) into a constant
                    z = (z - m) * self.w["dp.flows.0.inv_scale"]
        return z[:, :1]

    # ---- length regulator: integer-exact part ----------------------------
    @staticmethod
    def length_regulate(logw: torch.Tensor, length_scale: float):
        """logw [1,1,T] -> (w_ceil [T] float, y_len int, frame->phoneme index [y_len] int64).
        path[t, f] = (f < cum[t]) - (f < cum[t-1]); the matmul path x (m_p, logs_p) is a
        gather by index (frames no phoneme covers get index -1 -> zero row)."""
        w = torch.exp(logw) * length_scale
        w_ceil = torch.ceil(w)
        y_len = int(torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long().item())
        cum = torch.cumsum(w_ceil, -1).view(-1)
        f = torch.arange(y_len, dtype=cum.dtype)
        idx = torch.sum(f[:, None] >= cum[None, :], dim=1)      # number of phonemes fully before f
        idx = torch.where(idx < cum.numel(), idx, torch.full_like(idx, -1))
        return w_ceil.view(-1), y_len, idx.to(torch.int64)

    # ---- residual coupling flow, reverse --------------------------------
    def _wn(self, x, prefix, g=None):
        p = self.p
        H = p.hidden
        out = torch.zeros_like(x)
        gc = self._conv(g, prefix + ".cond_layer") if g is not None else None
        pad = (p.flow_kernel - 1) // 2
        for i in range(p.flow_wn_layers):
            x_in = self._conv(x, f"{prefix}.in_layers.{i}", padding=pad)
            if gc is not None:
                x_in = x_in + gc[:, i * 2 * H:(i + 1) * 2 * H]
            acts = torch.tanh(x_in[:, :H]) * torch.sigmoid(x_in[:, H:])
            rs = self._conv(acts, f"{prefix}.res_skip_layers.{i}")
            if i < p.flow_wn_layers - 1:
                x = x + rs[:, :H]
                out = out + rs[:, H:]
            else:
                out = out + rs
        return out

    def flow_reverse(self, z, g=None):
        half = self.p.hidden // 2
        for i in reversed(self.p.flow_ids):
            z = torch.flip(z, [1])
            pre = f"flow.flows.{i}"
            x0, x1 = z[:, :half], z[:, half:]
            h = self._conv(x0, pre + ".pre")
            h = self._wn(h, pre + ".enc", g)
            m = self._conv(h, pre + ".post")
            z = torch.cat([x0, x1 - m], 1)
        return z

    # ---- HiFi-GAN generator -----------------------------------------------
    def decoder(self, z, g=None):
        p = self.p
        x = self._conv(z, "dec.conv_pre", padding=3)
        if g is not None:
            x = x + self._conv(g, "dec.cond")
        nk = len(p.rb_kernels)
        for i, (k, u) in enumerate(p.ups):
            x = F.leaky_relu(x, LRELU_SLOPE)
            x = F.conv_transpose1d(x, self.w[f"dec.ups.{i}.weight"], self.w[f"dec.ups.{i}.bias"],
                                   stride=u, padding=(k - u) // 2)
            xs = None
            for j in range(nk):
                pre = f"dec.resblocks.{i * nk + j}"
                rk = p.rb_kernels[j]
                y = x
                for m, d in enumerate(p.rb_dilations[j]):
                    xt = F.leaky_relu(y, LRELU_SLOPE)
                    if p.resblock == 1:
                        xt = self._conv(xt, f"{pre}.convs1.{m}", dilation=d, padding=d * (rk - 1) // 2)
                        xt = F.leaky_relu(xt, LRELU_SLOPE)
                        xt = self._conv(xt, f"{pre}.convs2.{m}", padding=(rk - 1) // 2)
                    else:
                        xt = self._conv(xt, f"{pre}.convs.{m}", dilation=d, padding=d * (rk - 1) // 2)
                    y = xt + y
                xs = y if xs is None else xs + y
            x = xs / nk
        x = F.leaky_relu(x)              # default slope 0.01
        x = self._conv(x, "dec.conv_post", padding=3)
        return torch.tanh(x)

    # ---- whole graph ------------------------------------------------------------------
    @torch.no_grad()
    def infer(self, ids, noise_scale, length_scale, noise_w, noise_dp, noise_main=None,
              sid: Optional[int] = None, noise_main_fn=None):
        """ids: list[int]; noise_dp [2,T]; noise_main [C, >=F] or noise_main_fn(C, F) -> [C,F].
        Returns dict of stage tensors (numpy) incl. 'audio' [S]."""
        p = self.p
        dt = self.dtype
        idt = torch.tensor(ids, dtype=torch.int64).view(1, -1)
        g = None
        if p.gin and sid is not None:
            g = self.w["emb_g.weight"][sid].view(1, -1, 1)
        x, m_p, logs_p = self.text_encoder(idt)
        nd = torch.as_tensor(np.asarray(noise_dp)).to(dt).view(1, 2, -1)
        logw = self.duration_predictor(x, nd, noise_w, g)
        w_ceil, y_len, idx = self.length_regulate(logw, length_scale)
        gather = idx.clamp_min(0)
        valid = (idx >= 0).to(dt).view(1, 1, -1)
        m_e = m_p[:, :, gather] * valid
        logs_e = logs_p[:, :, gather] * valid
        if noise_main is None:
            noise_main = noise_main_fn(p.hidden, y_len)
        nm = torch.as_tensor(np.asarray(noise_main)).to(dt)[:, :y_len].view(1, p.hidden, y_len)
        z_p = m_e + nm * torch.exp(logs_e) * noise_scale
        z = self.flow_reverse(z_p, g)
        audio = self.decoder(z, g)
        out = dict(x=x[0], m_p=m_p[0], logs_p=logs_p[0], logw=logw.view(-1), w_ceil=w_ceil,
                   y_len=y_len, frame_index=idx, z_p=z_p[0], z=z[0], audio=audio.view(-1))
        return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else v) for k, v in out.items()}
