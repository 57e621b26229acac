"""TEST INFRASTRUCTURE (oracle).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this; the product path never does.

numpy restatement of the reference's node-by-node ONNX interpreter: `GraphExecutor.executeNode`
(Sources/PiperMetal/Execution/GraphExecutor.swift:591-2663) - the 50 op types a Piper-VITS export uses (opset 15), with
the op semantics of the reference's CPU / Metal kernels (broadcasting `elementwise.metal`, `conv1d.metal:28-142`,
`matmul.metal:22-49`, `softmax.metal:13-41`, `cumsum.metal:10-32`, `nonzero.metal`, `cast.metal:8-29`) and its
RandomNormalLike generator (oracle/refrng.py).  Two uses:
  * it evaluates the exported graph op by op, independently of oracle/vits_oracle.py (which restates the MODEL): the two
    must produce the same waveform from the same file - a cross-check of the oracle that needs no third-party runtime;
  * it is the checker of the CUDA generic op executor (csrc/graph_exec.cu), node by node.
Tensors are numpy arrays: float32, int64 or bool (TensorValue.swift:4-8)."""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from . import refrng
from .onnx_reader import Model

_ERF_VEC = None


def _erf(x: np.ndarray) -> np.ndarray:
    global _ERF_VEC
    if _ERF_VEC is None:
        import math
        _ERF_VEC = np.vectorize(math.erf, otypes=[np.float64])
    return _ERF_VEC(x.astype(np.float64)).astype(np.float32)


_DTYPES = {1: np.float32, 6: np.int64, 7: np.int64, 9: np.bool_, 11: np.float32}   # int32 widens, double narrows


def _ints(v) -> List[int]:
    if v is None:
        return []
    if isinstance(v, (list, tuple)):
        return [int(x) for x in v]
    return [int(x) for x in np.asarray(v).reshape(-1)]


def _conv1d(x, w, b, stride, dilation, pads, group):
    """x [N, Cin, L], w [Cout, Cin/g, K] (conv1d.metal:28-71: zero padding, one output per thread, fp32 accumulation)"""
    n, cin, length = x.shape
    cout, cg, k = w.shape
    xp = np.pad(x, ((0, 0), (0, 0), (pads[0], pads[1])))
    lout = (length + pads[0] + pads[1] - dilation * (k - 1) - 1) // stride + 1
    out = np.zeros((n, cout, lout), np.float32)
    og = cout // group
    for g in range(group):
        xs = xp[:, g * cg:(g + 1) * cg]
        ws = w[g * og:(g + 1) * og]
        for t in range(k):
            seg = xs[:, :, t * dilation: t * dilation + (lout - 1) * stride + 1: stride]     # [n, cg, lout]
            out[:, g * og:(g + 1) * og] += np.einsum("ncl,oc->nol", seg, ws[:, :, t], optimize=True).astype(np.float32)
    if b is not None:
        out += b.reshape(1, -1, 1)
    return out


def _conv_transpose1d(x, w, b, stride, dilation, pads, output_padding, group):
    """x [N, Cin, L], w [Cin, Cout/g, K] (conv1d.metal:97-142)"""
    n, cin, length = x.shape
    _, og, k = w.shape
    cout = og * group
    full = (length - 1) * stride + dilation * (k - 1) + 1 + output_padding
    out = np.zeros((n, cout, full), np.float32)
    cg = cin // group
    for g in range(group):
        xs = x[:, g * cg:(g + 1) * cg]
        ws = w[g * cg:(g + 1) * cg]
        for t in range(k):
            contrib = np.einsum("ncl,co->nol", xs, ws[:, :, t], optimize=True).astype(np.float32)
            out[:, g * og:(g + 1) * og, t * dilation: t * dilation + (length - 1) * stride + 1: stride] += contrib
    out = out[:, :, pads[0]: full - pads[1]]
    if b is not None:
        out = out + b.reshape(1, -1, 1)
    return out.astype(np.float32)


class NumpyGraph:
    """executeNode over the whole graph.  `run` returns the value table (every node output by name) so that the CUDA
    executor can be compared node by node."""

    def __init__(self, model: Model, seed: int = 1234):
        self.m = model
        self.seed = seed

    def run(self, feeds: Dict[str, np.ndarray], keep_all: bool = True, overrides: Optional[Dict[str, np.ndarray]] = None):
        v: Dict[str, np.ndarray] = dict(self.m.initializers)
        v.update(feeds)
        overrides = overrides or {}
        for node in self.m.nodes:
            if node.outputs and node.outputs[0] in overrides:       # GraphExecutor.swift:190-193 (injected noise)
                v[node.outputs[0]] = overrides[node.outputs[0]]
                continue
            ins = [v[i] if i else None for i in node.inputs]
            outs = self._op(node, ins)
            if not isinstance(outs, (list, tuple)):
                outs = [outs]
            for name, o in zip(node.outputs, outs):
                if name:
                    v[name] = o
        return v

    # ---- one node (GraphExecutor.swift:591-2663; line ranges per op in the comments) ----------------------------
    def _op(self, node, x):
        op, at = node.op_type, node.attrs
        f32 = np.float32
        if op == "Constant":                                   # :597-640
            return np.asarray(at["value"])
        if op in ("Add", "Sub", "Mul", "Div", "Pow"):          # :1237-1420 broadcasting binary ops
            a, b = x[0], x[1]
            if op == "Add":
                r = a + b
            elif op == "Sub":
                r = a - b
            elif op == "Mul":
                r = a * b
            elif op == "Pow":                                  # the result has the type of the base (ONNX Pow: T)
                r = np.power(a.astype(np.float64), b.astype(np.float64)).astype(a.dtype)
            elif a.dtype.kind == "i" and b.dtype.kind == "i":  # integer division truncates toward zero
                r = (np.trunc(a.astype(np.float64) / b.astype(np.float64))).astype(np.int64)
            else:
                r = a / b
            return r.astype(f32) if r.dtype.kind == "f" else r
        if op in ("Equal", "Less", "LessOrEqual", "GreaterOrEqual"):   # :1422-1530
            a, b = x[0], x[1]
            return {"Equal": np.equal, "Less": np.less, "LessOrEqual": np.less_equal, "GreaterOrEqual": np.greater_equal}[op](a, b)
        if op == "And":
            return np.logical_and(x[0], x[1])
        if op == "Not":
            return np.logical_not(x[0])
        if op == "Where":                                      # :1532-1600
            r = np.where(x[0], x[1], x[2])
            return r.astype(f32) if r.dtype.kind == "f" else r
        if op in ("Neg", "Exp", "Sqrt", "Tanh", "Sigmoid", "Softplus", "Relu", "Erf", "Ceil", "LeakyRelu"):   # :2017-2069, 2191-2260
            a = x[0]
            if op == "Neg":
                return -a
            a = a.astype(f32)
            if op == "Exp":
                return np.exp(a)
            if op == "Sqrt":
                return np.sqrt(a)
            if op == "Tanh":
                return np.tanh(a)
            if op == "Sigmoid":
                return (1.0 / (1.0 + np.exp(-a))).astype(f32)
            if op == "Softplus":
                return np.logaddexp(a, f32(0)).astype(f32)
            if op == "Relu":
                return np.maximum(a, 0)
            if op == "Erf":
                return _erf(a)
            if op == "Ceil":
                return np.ceil(a)
            alpha = f32(at.get("alpha", 0.01))
            return np.where(a >= 0, a, a * alpha).astype(f32)
        if op == "Clip":                                       # :2262-2300
            lo = x[1] if len(x) > 1 and x[1] is not None else None
            hi = x[2] if len(x) > 2 and x[2] is not None else None
            return np.clip(x[0], lo, hi).astype(x[0].dtype)
        if op == "Cast":                                       # cast.metal:8-29: float -> int truncates toward zero
            to = _DTYPES[int(at["to"])]
            a = x[0]
            if to is np.int64 and a.dtype.kind == "f":
                return np.trunc(a).astype(np.int64)
            if to is np.bool_:
                return a != 0
            return a.astype(to)
        if op == "Shape":
            return np.array(x[0].shape, np.int64)
        if op == "ConstantOfShape":
            val = np.asarray(at["value"]).reshape(-1)[0] if at.get("value") is not None else f32(0)
            return np.full(tuple(_ints(x[0])), val, dtype=np.asarray(val).dtype)
        if op == "Range":
            return np.arange(x[0].item(), x[1].item(), x[2].item()).astype(x[0].dtype)
        if op == "Reshape":                                    # :700-760 (0 = keep, -1 = infer)
            shp = _ints(x[1])
            shp = [x[0].shape[i] if s == 0 else s for i, s in enumerate(shp)]
            return x[0].reshape(shp)
        if op == "Transpose":
            perm = _ints(at.get("perm")) or list(range(x[0].ndim))[::-1]
            return np.ascontiguousarray(np.transpose(x[0], perm))
        if op == "Unsqueeze":
            axes = _ints(x[1]) if len(x) > 1 else _ints(at.get("axes"))
            out = x[0]
            rank = out.ndim + len(axes)
            for ax in sorted(a % rank for a in axes):
                out = np.expand_dims(out, ax)
            return out
        if op == "Squeeze":
            axes = _ints(x[1]) if len(x) > 1 and x[1] is not None else _ints(at.get("axes"))
            return np.squeeze(x[0], axis=tuple(axes) if axes else None)
        if op == "Concat":
            return np.concatenate(x, axis=int(at["axis"]))
        if op == "Split":                                      # :1100-1170
            axis = int(at.get("axis", 0))
            if len(x) > 1 and x[1] is not None:
                sizes = _ints(x[1])
            elif at.get("split") is not None:
                sizes = _ints(at["split"])
            else:
                n = len(node.outputs)
                sizes = [x[0].shape[axis] // n] * n
            return [np.ascontiguousarray(p) for p in np.split(x[0], np.cumsum(sizes)[:-1], axis=axis)]
        if op == "Slice":                                      # :940-1098 (clamped starts / ends, negative steps)
            data = x[0]
            starts, ends = _ints(x[1]), _ints(x[2])
            axes = _ints(x[3]) if len(x) > 3 and x[3] is not None else list(range(len(starts)))
            steps = _ints(x[4]) if len(x) > 4 and x[4] is not None else [1] * len(starts)
            sl = [slice(None)] * data.ndim
            for s, e, ax, st in zip(starts, ends, axes, steps):
                dim = data.shape[ax]
                if st > 0:
                    s = min(max(s + dim if s < 0 else s, 0), dim)
                    e = min(max(e + dim if e < 0 else e, 0), dim)
                    sl[ax] = slice(s, e, st)
                else:
                    s = min(max(s + dim if s < 0 else s, -1), dim - 1)
                    e = min(max(e + dim if e < 0 else e, -1), dim - 1)
                    sl[ax] = slice(s, e if e >= 0 else None, st)
            return np.ascontiguousarray(data[tuple(sl)])
        if op == "Expand":
            shp = _ints(x[1])
            return np.ascontiguousarray(np.broadcast_to(x[0], np.broadcast_shapes(x[0].shape, tuple(shp))))
        if op == "Pad":                                        # :1602-1700 (constant mode)
            pads = _ints(x[1])
            r = x[0].ndim
            cval = x[2].reshape(-1)[0] if len(x) > 2 and x[2] is not None and x[2].size else 0
            return np.pad(x[0], [(pads[i], pads[i + r]) for i in range(r)], constant_values=cval)
        if op == "Gather":                                     # :780-860
            return np.take(x[0], x[1], axis=int(at.get("axis", 0)))
        if op == "GatherElements":
            return np.take_along_axis(x[0], x[1], axis=int(at.get("axis", 0)))
        if op == "GatherND":                                   # :2120-2189 (batch_dims 0)
            idx = x[1]
            return x[0][tuple(np.moveaxis(idx, -1, 0))]
        if op == "ScatterND":
            out = x[0].copy()
            out[tuple(np.moveaxis(x[1], -1, 0))] = x[2]
            return out
        if op == "NonZero":                                    # nonzero.metal: row-major order, [rank, n]
            return np.array(np.nonzero(x[0]), np.int64).reshape(x[0].ndim, -1)
        if op == "CumSum":                                     # cumsum.metal:10-32
            axis = int(x[1].reshape(-1)[0])
            a = x[0]
            if int(at.get("reverse", 0)):
                a = np.flip(a, axis)
            r = np.cumsum(a, axis=axis, dtype=a.dtype)
            if int(at.get("exclusive", 0)):
                r = r - a
            if int(at.get("reverse", 0)):
                r = np.flip(r, axis)
            return r
        if op in ("ReduceMean", "ReduceSum", "ReduceMax"):     # :1930-2015
            axes = _ints(x[1]) if len(x) > 1 and x[1] is not None else _ints(at.get("axes"))
            keep = bool(int(at.get("keepdims", 1)))
            fn = {"ReduceMean": np.mean, "ReduceSum": np.sum, "ReduceMax": np.max}[op]
            r = fn(x[0], axis=tuple(axes) if axes else None, keepdims=keep)
            return np.asarray(r, dtype=x[0].dtype)
        if op == "Softmax":                                    # softmax.metal:13-41
            axis = int(at.get("axis", -1))
            a = x[0].astype(f32)
            e = np.exp(a - a.max(axis=axis, keepdims=True))
            return (e / e.sum(axis=axis, keepdims=True)).astype(f32)
        if op == "MatMul":                                     # matmul.metal:22-49
            return np.matmul(x[0].astype(f32), x[1].astype(f32)).astype(f32)
        if op in ("Conv", "ConvTranspose"):                    # :1739-1860
            w = x[1]
            b = x[2] if len(x) > 2 else None
            k = w.shape[2]
            stride = _ints(at.get("strides") or [1])[0]
            dil = _ints(at.get("dilations") or [1])[0]
            pads = _ints(at.get("pads") or [0, 0])
            group = int(at.get("group", 1))
            if op == "Conv":
                return _conv1d(x[0].astype(f32), w, b, stride, dil, pads, group)
            opad = _ints(at.get("output_padding") or [0])[0]
            return _conv_transpose1d(x[0].astype(f32), w, b, stride, dil, pads, opad, group)
        if op == "RandomNormalLike":                           # :2640-2663, elementwise.metal:132-163
            n = int(np.prod(x[0].shape))
            z = refrng.random_normal_like(n, self.seed).reshape(x[0].shape)
            return (z * f32(at.get("scale", 1.0)) + f32(at.get("mean", 0.0))).astype(f32)
        raise NotImplementedError(f"op {op}")


def synthesize(model: Model, phoneme_ids, noise_scale, length_scale, noise_w, seed: int = 1234, keep_all: bool = False):
    """The three feeds of a Piper export (ExecutionInputs, GraphExecutor.swift:5-15) -> waveform."""
    feeds = {"input": np.asarray(phoneme_ids, np.int64).reshape(1, -1),
             "input_lengths": np.array([len(phoneme_ids)], np.int64),
             "scales": np.array([noise_scale, length_scale, noise_w], np.float32)}
    v = NumpyGraph(model, seed).run(feeds)
    out = v[model.outputs[0]]
    return (out.reshape(-1), v) if keep_all else out.reshape(-1)
