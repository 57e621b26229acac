"""TEST INFRASTRUCTURE (oracle).  Only tests/ and tools/ may import this; the product path never does.

Module-structured restatement of the upstream VITS network a Piper voice is exported from
(rhasspy/piper `piper_train/vits`: models.py SynthesizerTrn.infer, attentions.py, modules.py,
transforms.py, commons.py - third-party, absent from /root/reference; the reference only ever sees
the exported graph, Tests/PiperONNXTests/ONNXParsingTests.swift:22-35, VOICES.md:10-159).

Why a second restatement next to oracle/vits_oracle.py:
  * it carries the upstream PARAMETER NAMES and module tree, weight_norm included (removed from `dec`
    only, exactly like piper_train's export script), so `torch.onnx.export` - the real exporter, not this
    repo's own protobuf writer - produces a file with the initializer / node naming of a real voice:
    anonymous `onnx::Conv_NNNN` weights for the weight-normed flow convolutions, a folded
    `exp(-logs)` constant, the decomposed LayerNorm, NonZero/GatherND/ScatterND spline masking, two
    RandomNormalLike nodes.  tests/test_real_export_cpu.py loads that file with the product's C++ reader
    and with oracle/onnx_reader.py and compares its counts with the ones the reference pins for the
    real medium voice (2755 nodes / 401 initializers, opset 15, first node Gather);
  * evaluated eagerly with the two random tensors injected it is an independent check of the functional
    oracle (same weights -> same audio).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils import remove_weight_norm, weight_norm

LRELU_SLOPE = 0.1


# ---- commons.py ------------------------------------------------------------------------------------
def convert_pad_shape(pad_shape):
    layer = pad_shape[::-1]
    return [item for sublist in layer for item in sublist]


def sequence_mask(length, max_length=None):
    if max_length is None:
        max_length = length.max()
    x = torch.arange(max_length, dtype=length.dtype, device=length.device)
    return x.unsqueeze(0) < length.unsqueeze(1)


def generate_path(duration, mask):
    b, _, t_y, t_x = mask.shape
    cum_duration = torch.cumsum(duration, -1)
    cum_duration_flat = cum_duration.view(b * t_x)
    path = sequence_mask(cum_duration_flat, t_y).type_as(mask)
    path = path.view(b, t_x, t_y)
    path = path - F.pad(path, convert_pad_shape([[0, 0], [1, 0], [0, 0]]))[:, :-1]
    path = path.unsqueeze(1).transpose(2, 3) * mask
    return path


def fused_add_tanh_sigmoid_multiply(input_a, input_b, n_channels_int: int):
    in_act = input_a + input_b
    t_act = torch.tanh(in_act[:, :n_channels_int, :])
    s_act = torch.sigmoid(in_act[:, n_channels_int:, :])
    return t_act * s_act


def get_padding(kernel_size, dilation=1):
    return int((kernel_size * dilation - dilation) / 2)


# ---- transforms.py -----------------------------------------------------------------------------------
def searchsorted(bin_locations, inputs, eps=1e-6):
    bin_locations[..., bin_locations.size(-1) - 1] += eps
    return torch.sum(inputs[..., None] >= bin_locations, dim=-1) - 1


def rational_quadratic_spline_inverse(inputs, uw, uh, ud, left, right, bottom, top, min_bin_width=1e-3,
                                      min_bin_height=1e-3, min_derivative=1e-3):
    num_bins = uw.shape[-1]
    widths = F.softmax(uw, dim=-1)
    widths = min_bin_width + (1 - min_bin_width * num_bins) * widths
    cumwidths = torch.cumsum(widths, dim=-1)
    cumwidths = F.pad(cumwidths, pad=(1, 0), mode="constant", value=0.0)
    cumwidths = (right - left) * cumwidths + left
    cumwidths[..., 0] = left
    cumwidths[..., -1] = right
    widths = cumwidths[..., 1:] - cumwidths[..., :-1]
    derivatives = min_derivative + F.softplus(ud)
    heights = F.softmax(uh, dim=-1)
    heights = min_bin_height + (1 - min_bin_height * num_bins) * heights
    cumheights = torch.cumsum(heights, dim=-1)
    cumheights = F.pad(cumheights, pad=(1, 0), mode="constant", value=0.0)
    cumheights = (top - bottom) * cumheights + bottom
    cumheights[..., 0] = bottom
    cumheights[..., -1] = top
    heights = cumheights[..., 1:] - cumheights[..., :-1]
    bin_idx = searchsorted(cumheights, inputs)[..., None]
    input_cumwidths = cumwidths.gather(-1, bin_idx)[..., 0]
    input_bin_widths = widths.gather(-1, bin_idx)[..., 0]
    input_cumheights = cumheights.gather(-1, bin_idx)[..., 0]
    delta = heights / widths
    input_delta = delta.gather(-1, bin_idx)[..., 0]
    input_derivatives = derivatives.gather(-1, bin_idx)[..., 0]
    input_derivatives_plus_one = derivatives[..., 1:].gather(-1, bin_idx)[..., 0]
    input_heights = heights.gather(-1, bin_idx)[..., 0]
    a = (inputs - input_cumheights) * (input_derivatives + input_derivatives_plus_one - 2 * input_delta) + \
        input_heights * (input_delta - input_derivatives)
    b = input_heights * input_derivatives - (inputs - input_cumheights) * (
        input_derivatives + input_derivatives_plus_one - 2 * input_delta)
    c = -input_delta * (inputs - input_cumheights)
    discriminant = b.pow(2) - 4 * a * c
    root = (2 * c) / (-b - torch.sqrt(discriminant))
    return root * input_bin_widths + input_cumwidths


def unconstrained_rq_spline_inverse(inputs, uw, uh, ud, tail_bound, min_derivative=1e-3):
    inside = (inputs >= -tail_bound) & (inputs <= tail_bound)
    outside = ~inside
    outputs = torch.zeros_like(inputs)
    ud = F.pad(ud, pad=(1, 1))
    constant = np.log(np.exp(1 - min_derivative) - 1)
    ud[..., 0] = constant
    ud[..., -1] = constant
    outputs[outside] = inputs[outside]
    outputs[inside] = rational_quadratic_spline_inverse(
        inputs[inside], uw[inside, :], uh[inside, :], ud[inside, :],
        left=-tail_bound, right=tail_bound, bottom=-tail_bound, top=tail_bound)
    return outputs


# ---- modules.py ----------------------------------------------------------------------------------------
class LayerNorm(nn.Module):
    def __init__(self, channels, eps=1e-5):
        super().__init__()
        self.channels, self.eps = channels, eps
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))

    def forward(self, x):
        x = x.transpose(1, -1)
        x = F.layer_norm(x, (self.channels,), self.gamma, self.beta, self.eps)
        return x.transpose(1, -1)


class DDSConv(nn.Module):
    def __init__(self, channels, kernel_size, n_layers):
        super().__init__()
        self.n_layers = n_layers
        self.convs_sep, self.convs_1x1 = nn.ModuleList(), nn.ModuleList()
        self.norms_1, self.norms_2 = nn.ModuleList(), nn.ModuleList()
        for i in range(n_layers):
            dilation = kernel_size ** i
            padding = (kernel_size * dilation - dilation) // 2
            self.convs_sep.append(nn.Conv1d(channels, channels, kernel_size, groups=channels, dilation=dilation,
                                            padding=padding))
            self.convs_1x1.append(nn.Conv1d(channels, channels, 1))
            self.norms_1.append(LayerNorm(channels))
            self.norms_2.append(LayerNorm(channels))

    def forward(self, x, x_mask, g=None):
        if g is not None:
            x = x + g
        for i in range(self.n_layers):
            y = self.convs_sep[i](x * x_mask)
            y = self.norms_1[i](y)
            y = F.gelu(y)
            y = self.convs_1x1[i](y)
            y = self.norms_2[i](y)
            y = F.gelu(y)
            x = x + y
        return x * x_mask


class WN(nn.Module):
    def __init__(self, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0):
        super().__init__()
        self.hidden_channels, self.n_layers = hidden_channels, n_layers
        self.in_layers, self.res_skip_layers = nn.ModuleList(), nn.ModuleList()
        if gin_channels != 0:
            self.cond_layer = weight_norm(nn.Conv1d(gin_channels, 2 * hidden_channels * n_layers, 1), name="weight")
        for i in range(n_layers):
            dilation = dilation_rate ** i
            padding = int((kernel_size * dilation - dilation) / 2)
            self.in_layers.append(weight_norm(nn.Conv1d(hidden_channels, 2 * hidden_channels, kernel_size,
                                                        dilation=dilation, padding=padding), name="weight"))
            rs = 2 * hidden_channels if i < n_layers - 1 else hidden_channels
            self.res_skip_layers.append(weight_norm(nn.Conv1d(hidden_channels, rs, 1), name="weight"))

    def forward(self, x, x_mask, g=None):
        output = torch.zeros_like(x)
        if g is not None:
            g = self.cond_layer(g)
        for i in range(self.n_layers):
            x_in = self.in_layers[i](x)
            if g is not None:
                off = i * 2 * self.hidden_channels
                g_l = g[:, off:off + 2 * self.hidden_channels, :]
            else:
                g_l = torch.zeros_like(x_in)
            acts = fused_add_tanh_sigmoid_multiply(x_in, g_l, self.hidden_channels)
            res_skip_acts = self.res_skip_layers[i](acts)
            if i < self.n_layers - 1:
                x = (x + res_skip_acts[:, :self.hidden_channels, :]) * x_mask
                output = output + res_skip_acts[:, self.hidden_channels:, :]
            else:
                output = output + res_skip_acts
        return output * x_mask


class ResBlock1(nn.Module):
    def __init__(self, channels, kernel_size=3, dilation=(1, 3, 5)):
        super().__init__()
        self.convs1 = nn.ModuleList([weight_norm(nn.Conv1d(channels, channels, kernel_size, 1, dilation=d,
                                                           padding=get_padding(kernel_size, d))) for d in dilation])
        self.convs2 = nn.ModuleList([weight_norm(nn.Conv1d(channels, channels, kernel_size, 1, dilation=1,
                                                           padding=get_padding(kernel_size, 1))) for _ in dilation])

    def forward(self, x):
        for c1, c2 in zip(self.convs1, self.convs2):
            xt = F.leaky_relu(x, LRELU_SLOPE)
            xt = c1(xt)
            xt = F.leaky_relu(xt, LRELU_SLOPE)
            xt = c2(xt)
            x = xt + x
        return x

    def remove_weight_norm(self):
        for l in list(self.convs1) + list(self.convs2):
            remove_weight_norm(l)


class ResBlock2(nn.Module):
    def __init__(self, channels, kernel_size=3, dilation=(1, 3)):
        super().__init__()
        self.convs = nn.ModuleList([weight_norm(nn.Conv1d(channels, channels, kernel_size, 1, dilation=d,
                                                          padding=get_padding(kernel_size, d))) for d in dilation])

    def forward(self, x):
        for c in self.convs:
            xt = F.leaky_relu(x, LRELU_SLOPE)
            xt = c(xt)
            x = xt + x
        return x

    def remove_weight_norm(self):
        for l in self.convs:
            remove_weight_norm(l)


class Flip(nn.Module):
    def forward(self, x, *args, reverse=False, **kwargs):
        return torch.flip(x, [1])


class ElementwiseAffine(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.m = nn.Parameter(torch.zeros(channels, 1))
        self.logs = nn.Parameter(torch.zeros(channels, 1))

    def forward(self, x, x_mask, reverse=False, **kwargs):
        return (x - self.m) * torch.exp(-self.logs) * x_mask


class ResidualCouplingLayer(nn.Module):
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0):
        super().__init__()
        self.half_channels = channels // 2
        self.pre = nn.Conv1d(self.half_channels, hidden_channels, 1)
        self.enc = WN(hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=gin_channels)
        self.post = nn.Conv1d(hidden_channels, self.half_channels, 1)   # mean_only=True

    def forward(self, x, x_mask, g=None, reverse=True):
        x0, x1 = torch.split(x, [self.half_channels] * 2, 1)
        h = self.pre(x0) * x_mask
        h = self.enc(h, x_mask, g=g)
        stats = self.post(h) * x_mask
        m = stats
        logs = torch.zeros_like(m)
        x1 = (x1 - m) * torch.exp(-logs) * x_mask
        return torch.cat([x0, x1], 1)


class ConvFlow(nn.Module):
    def __init__(self, in_channels, filter_channels, kernel_size, n_layers, num_bins=10, tail_bound=5.0):
        super().__init__()
        self.filter_channels, self.num_bins, self.tail_bound = filter_channels, num_bins, tail_bound
        self.half_channels = in_channels // 2
        self.pre = nn.Conv1d(self.half_channels, filter_channels, 1)
        self.convs = DDSConv(filter_channels, kernel_size, n_layers)
        self.proj = nn.Conv1d(filter_channels, self.half_channels * (num_bins * 3 - 1), 1)

    def forward(self, x, x_mask, g=None, reverse=True):
        x0, x1 = torch.split(x, [self.half_channels] * 2, 1)
        h = self.pre(x0)
        h = self.convs(h, x_mask, g=g)
        h = self.proj(h) * x_mask
        b, c, t = x0.shape
        h = h.reshape(b, c, -1, t).permute(0, 1, 3, 2)
        uw = h[..., :self.num_bins] / math.sqrt(self.filter_channels)
        uh = h[..., self.num_bins:2 * self.num_bins] / math.sqrt(self.filter_channels)
        ud = h[..., 2 * self.num_bins:]
        x1 = unconstrained_rq_spline_inverse(x1, uw, uh, ud, self.tail_bound)
        return torch.cat([x0, x1], 1) * x_mask


# ---- attentions.py ---------------------------------------------------------------------------------------
class MultiHeadAttention(nn.Module):
    def __init__(self, channels, out_channels, n_heads, window_size=4):
        super().__init__()
        self.channels, self.n_heads, self.window_size = channels, n_heads, window_size
        self.k_channels = channels // n_heads
        self.conv_q = nn.Conv1d(channels, channels, 1)
        self.conv_k = nn.Conv1d(channels, channels, 1)
        self.conv_v = nn.Conv1d(channels, channels, 1)
        self.conv_o = nn.Conv1d(channels, out_channels, 1)
        rel_stddev = self.k_channels ** -0.5
        self.emb_rel_k = nn.Parameter(torch.randn(1, window_size * 2 + 1, self.k_channels) * rel_stddev)
        self.emb_rel_v = nn.Parameter(torch.randn(1, window_size * 2 + 1, self.k_channels) * rel_stddev)

    def forward(self, x, c, attn_mask=None):
        q, k, v = self.conv_q(x), self.conv_k(c), self.conv_v(c)
        x = self.attention(q, k, v, mask=attn_mask)
        return self.conv_o(x)

    def attention(self, query, key, value, mask=None):
        b, d, t_s, t_t = (*key.size(), query.size(2))
        query = query.view(b, self.n_heads, self.k_channels, t_t).transpose(2, 3)
        key = key.view(b, self.n_heads, self.k_channels, t_s).transpose(2, 3)
        value = value.view(b, self.n_heads, self.k_channels, t_s).transpose(2, 3)
        scores = torch.matmul(query / math.sqrt(self.k_channels), key.transpose(-2, -1))
        key_relative_embeddings = self._get_relative_embeddings(self.emb_rel_k, t_s)
        rel_logits = torch.matmul(query / math.sqrt(self.k_channels),
                                  key_relative_embeddings.unsqueeze(0).transpose(-2, -1))
        scores = scores + self._relative_position_to_absolute_position(rel_logits)
        if mask is not None:
            scores = scores.masked_fill(mask == 0, -1e4)
        p_attn = F.softmax(scores, dim=-1)
        output = torch.matmul(p_attn, value)
        relative_weights = self._absolute_position_to_relative_position(p_attn)
        value_relative_embeddings = self._get_relative_embeddings(self.emb_rel_v, t_s)
        output = output + torch.matmul(relative_weights, value_relative_embeddings.unsqueeze(0))
        return output.transpose(2, 3).contiguous().view(b, d, t_t)

    def _get_relative_embeddings(self, relative_embeddings, length):
        pad_length = max(length - (self.window_size + 1), 0)
        slice_start_position = max((self.window_size + 1) - length, 0)
        slice_end_position = slice_start_position + 2 * length - 1
        if pad_length > 0:
            padded = F.pad(relative_embeddings, convert_pad_shape([[0, 0], [pad_length, pad_length], [0, 0]]))
        else:
            padded = relative_embeddings
        return padded[:, slice_start_position:slice_end_position]

    def _relative_position_to_absolute_position(self, x):
        batch, heads, length, _ = x.size()
        x = F.pad(x, convert_pad_shape([[0, 0], [0, 0], [0, 0], [0, 1]]))
        x_flat = x.view([batch, heads, length * 2 * length])
        x_flat = F.pad(x_flat, convert_pad_shape([[0, 0], [0, 0], [0, length - 1]]))
        return x_flat.view([batch, heads, length + 1, 2 * length - 1])[:, :, :length, length - 1:]

    def _absolute_position_to_relative_position(self, x):
        batch, heads, length, _ = x.size()
        x = F.pad(x, convert_pad_shape([[0, 0], [0, 0], [0, 0], [0, length - 1]]))
        x_flat = x.view([batch, heads, length ** 2 + length * (length - 1)])
        x_flat = F.pad(x_flat, convert_pad_shape([[0, 0], [0, 0], [length, 0]]))
        return x_flat.view([batch, heads, length, 2 * length])[:, :, :, 1:]


class FFN(nn.Module):
    def __init__(self, in_channels, out_channels, filter_channels, kernel_size):
        super().__init__()
        self.kernel_size = kernel_size
        self.conv_1 = nn.Conv1d(in_channels, filter_channels, kernel_size)
        self.conv_2 = nn.Conv1d(filter_channels, out_channels, kernel_size)

    def _same_padding(self, x):
        if self.kernel_size == 1:
            return x
        pad_l, pad_r = (self.kernel_size - 1) // 2, self.kernel_size // 2
        return F.pad(x, convert_pad_shape([[0, 0], [0, 0], [pad_l, pad_r]]))

    def forward(self, x, x_mask):
        x = self.conv_1(self._same_padding(x * x_mask))
        x = torch.relu(x)
        x = self.conv_2(self._same_padding(x * x_mask))
        return x * x_mask


class Encoder(nn.Module):
    def __init__(self, hidden_channels, filter_channels, n_heads, n_layers, kernel_size, window_size=4):
        super().__init__()
        self.n_layers = n_layers
        self.attn_layers, self.norm_layers_1 = nn.ModuleList(), nn.ModuleList()
        self.ffn_layers, self.norm_layers_2 = nn.ModuleList(), nn.ModuleList()
        for _ in range(n_layers):
            self.attn_layers.append(MultiHeadAttention(hidden_channels, hidden_channels, n_heads, window_size))
            self.norm_layers_1.append(LayerNorm(hidden_channels))
            self.ffn_layers.append(FFN(hidden_channels, hidden_channels, filter_channels, kernel_size))
            self.norm_layers_2.append(LayerNorm(hidden_channels))

    def forward(self, x, x_mask):
        attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
        x = x * x_mask
        for i in range(self.n_layers):
            y = self.attn_layers[i](x, x, attn_mask)
            x = self.norm_layers_1[i](x + y)
            y = self.ffn_layers[i](x, x_mask)
            x = self.norm_layers_2[i](x + y)
        return x * x_mask


# ---- models.py -----------------------------------------------------------------------------------------------
class StochasticDurationPredictor(nn.Module):
    def __init__(self, in_channels, kernel_size, n_flows=4, n_layers=3, num_bins=10, gin_channels=0):
        super().__init__()
        filter_channels = in_channels
        self.flows = nn.ModuleList([ElementwiseAffine(2)])
        for _ in range(n_flows):
            self.flows.append(ConvFlow(2, filter_channels, kernel_size, n_layers=n_layers, num_bins=num_bins))
            self.flows.append(Flip())
        self.pre = nn.Conv1d(in_channels, filter_channels, 1)
        self.proj = nn.Conv1d(filter_channels, filter_channels, 1)
        self.convs = DDSConv(filter_channels, kernel_size, n_layers=n_layers)
        if gin_channels != 0:
            self.cond = nn.Conv1d(gin_channels, filter_channels, 1)

    def forward(self, x, x_mask, g=None, noise_scale=1.0, noise=None):
        x = self.pre(x)
        if g is not None:
            x = x + self.cond(g)
        x = self.convs(x, x_mask)
        x = self.proj(x) * x_mask
        flows = list(reversed(self.flows))
        flows = flows[:-2] + [flows[-1]]   # remove a useless vflow
        if noise is None:
            noise = torch.randn(x.size(0), 2, x.size(2)).type_as(x)
        z = noise * noise_scale
        for flow in flows:
            z = flow(z, x_mask, g=x, reverse=True)
        z0, _ = torch.split(z, [1, 1], 1)
        return z0


class TextEncoder(nn.Module):
    def __init__(self, n_vocab, out_channels, hidden_channels, filter_channels, n_heads, n_layers, kernel_size,
                 window_size=4):
        super().__init__()
        self.out_channels, self.hidden_channels = out_channels, hidden_channels
        self.emb = nn.Embedding(n_vocab, hidden_channels)
        self.encoder = Encoder(hidden_channels, filter_channels, n_heads, n_layers, kernel_size, window_size)
        self.proj = nn.Conv1d(hidden_channels, out_channels * 2, 1)

    def forward(self, x, x_lengths):
        x = self.emb(x) * math.sqrt(self.hidden_channels)
        x = torch.transpose(x, 1, -1)
        x_mask = torch.unsqueeze(sequence_mask(x_lengths, x.size(2)), 1).type_as(x)
        x = self.encoder(x * x_mask, x_mask)
        stats = self.proj(x) * x_mask
        m, logs = torch.split(stats, self.out_channels, dim=1)
        return x, m, logs, x_mask


class ResidualCouplingBlock(nn.Module):
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, n_flows=4, gin_channels=0):
        super().__init__()
        self.flows = nn.ModuleList()
        for _ in range(n_flows):
            self.flows.append(ResidualCouplingLayer(channels, hidden_channels, kernel_size, dilation_rate, n_layers,
                                                    gin_channels=gin_channels))
            self.flows.append(Flip())

    def forward(self, x, x_mask, g=None):
        for flow in reversed(self.flows):
            x = flow(x, x_mask, g=g, reverse=True)
        return x


class Generator(nn.Module):
    def __init__(self, initial_channel, resblock, resblock_kernel_sizes, resblock_dilation_sizes, upsample_rates,
                 upsample_initial_channel, upsample_kernel_sizes, gin_channels=0):
        super().__init__()
        self.num_kernels, self.num_upsamples = len(resblock_kernel_sizes), len(upsample_rates)
        self.conv_pre = nn.Conv1d(initial_channel, upsample_initial_channel, 7, 1, padding=3)
        rb = ResBlock1 if resblock == 1 else ResBlock2
        self.ups = nn.ModuleList()
        for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
            self.ups.append(weight_norm(nn.ConvTranspose1d(upsample_initial_channel // (2 ** i),
                                                           upsample_initial_channel // (2 ** (i + 1)), k, u,
                                                           padding=(k - u) // 2)))
        self.resblocks = nn.ModuleList()
        ch = upsample_initial_channel
        for i in range(len(self.ups)):
            ch = upsample_initial_channel // (2 ** (i + 1))
            for k, d in zip(resblock_kernel_sizes, resblock_dilation_sizes):
                self.resblocks.append(rb(ch, k, d))
        self.conv_post = nn.Conv1d(ch, 1, 7, 1, padding=3, bias=False)
        if gin_channels != 0:
            self.cond = nn.Conv1d(gin_channels, upsample_initial_channel, 1)

    def forward(self, x, g=None):
        x = self.conv_pre(x)
        if g is not None:
            x = x + self.cond(g)
        for i in range(self.num_upsamples):
            x = F.leaky_relu(x, LRELU_SLOPE)
            x = self.ups[i](x)
            xs = None
            for j in range(self.num_kernels):
                if xs is None:
                    xs = self.resblocks[i * self.num_kernels + j](x)
                else:
                    xs += self.resblocks[i * self.num_kernels + j](x)
            x = xs / self.num_kernels
        x = F.leaky_relu(x)
        x = self.conv_post(x)
        return torch.tanh(x)

    def remove_weight_norm(self):
        for l in self.ups:
            remove_weight_norm(l)
        for l in self.resblocks:
            l.remove_weight_norm()


class SynthesizerTrn(nn.Module):
    """Inference half of upstream SynthesizerTrn (the posterior encoder and the training-only `post_*` modules of
    the duration predictor never reach an exported graph)."""

    def __init__(self, spec):
        super().__init__()
        self.spec = spec
        H = spec.hidden
        gin = spec.gin if spec.n_speakers > 1 else 0
        self.n_speakers = spec.n_speakers
        self.enc_p = TextEncoder(spec.n_vocab, H, H, spec.filter_channels, spec.n_heads, spec.n_layers,
                                 spec.ffn_kernel, spec.window)
        self.dec = Generator(H, spec.resblock, spec.rb_kernels, spec.rb_dilations, spec.ups_rates, spec.dec_c0,
                             spec.ups_kernels, gin_channels=gin)
        self.flow = ResidualCouplingBlock(H, H, spec.flow_kernel, 1, spec.flow_wn_layers, n_flows=spec.flow_n,
                                          gin_channels=gin)
        self.dp = StochasticDurationPredictor(H, spec.dp_kernel, n_flows=spec.dp_flows, n_layers=spec.dp_layers,
                                              num_bins=spec.dp_bins, gin_channels=gin)
        if spec.n_speakers > 1:
            self.emb_g = nn.Embedding(spec.n_speakers, gin)

    def infer(self, x, x_lengths, sid=None, noise_scale=0.667, length_scale=1.0, noise_scale_w=0.8,
              noise_dp=None, noise_main=None):
        x, m_p, logs_p, x_mask = self.enc_p(x, x_lengths)
        g = self.emb_g(sid).unsqueeze(-1) if self.n_speakers > 1 else None
        logw = self.dp(x, x_mask, g=g, noise_scale=noise_scale_w, noise=noise_dp)
        w = torch.exp(logw) * x_mask * length_scale
        w_ceil = torch.ceil(w)
        y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()
        y_mask = torch.unsqueeze(sequence_mask(y_lengths, y_lengths.max()), 1).type_as(x_mask)
        attn_mask = torch.unsqueeze(x_mask, 2) * torch.unsqueeze(y_mask, -1)
        attn = generate_path(w_ceil, attn_mask)
        m_p = torch.matmul(attn.squeeze(1), m_p.transpose(1, 2)).transpose(1, 2)
        logs_p = torch.matmul(attn.squeeze(1), logs_p.transpose(1, 2)).transpose(1, 2)
        eps = torch.randn_like(m_p) if noise_main is None else noise_main
        z_p = m_p + eps * torch.exp(logs_p) * noise_scale
        z = self.flow(z_p, y_mask, g=g)
        o = self.dec(z * y_mask, g=g)
        return o, w_ceil, y_lengths, z_p, z

    # ---- weights: the synthetic state dict (piper_b200/synth_voice.py) uses the post-folding names --------
    def load_plain_state(self, weights: Dict[str, np.ndarray]):
        """`weights` holds `<module>.weight` for every convolution.  Weight-normed modules receive it as
        weight_v with weight_g = ||v|| (so g * v / ||v|| == the given weight up to rounding)."""
        sd = self.state_dict()
        out = {}
        for name in sd:
            if name.endswith(".weight_g") or name.endswith(".weight_v"):
                base = name.rsplit(".", 1)[0]
                w = torch.from_numpy(np.asarray(weights[base + ".weight"]))
                if name.endswith(".weight_v"):
                    out[name] = w
                else:
                    out[name] = torch.sqrt((w.double() ** 2).sum(dim=tuple(range(1, w.dim())), keepdim=True)).float()
            else:
                out[name] = torch.from_numpy(np.asarray(weights[name])).reshape(sd[name].shape)
        self.load_state_dict(out)


def build_from_spec(spec, weights: Dict[str, np.ndarray]) -> SynthesizerTrn:
    m = SynthesizerTrn(spec)
    m.load_plain_state(weights)
    m.eval()
    return m


def export_like_piper(model: SynthesizerTrn, path: str, dummy_len: int = 50, seed: int = 1234):
    """piper_train/export_onnx.py: weight norm removed from `dec` only, forward = infer with
    scales = (noise_scale, length_scale, noise_w), opset 15, dynamic batch / phonemes / time axes, output
    [batch, 1, 1, time].  The exporter here is torch's own (TorchScript) ONNX exporter; the only deviation is that its
    final onnxscript hook is stubbed out, because the `onnx` Python package it imports is absent in this image (the hook
    is a no-op for graphs without onnxscript functions)."""
    from torch.onnx._internal.torchscript_exporter import onnx_proto_utils
    onnx_proto_utils._add_onnxscript_fn = lambda proto, custom_opsets: proto
    torch.manual_seed(seed)
    model.dec.remove_weight_norm()
    multi = model.n_speakers > 1

    def infer_forward(text, text_lengths, scales, sid=None):
        audio = model.infer(text, text_lengths, noise_scale=scales[0], length_scale=scales[1],
                            noise_scale_w=scales[2], sid=sid)[0].unsqueeze(1)
        return audio

    model.forward = infer_forward
    n_vocab = model.spec.n_vocab
    seq = torch.randint(low=0, high=n_vocab, size=(1, dummy_len), dtype=torch.long)
    lens = torch.LongTensor([seq.size(1)])
    sid = torch.LongTensor([0]) if multi else None
    scales = torch.FloatTensor([0.667, 1.0, 0.8])
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        torch.onnx.export(model=model, args=(seq, lens, scales, sid), f=path, verbose=False, opset_version=15,
                          input_names=["input", "input_lengths", "scales", "sid"], output_names=["output"],
                          dynamic_axes={"input": {0: "batch_size", 1: "phonemes"},
                                        "input_lengths": {0: "batch_size"},
                                        "output": {0: "batch_size", 1: "time"}}, dynamo=False)
    return path
