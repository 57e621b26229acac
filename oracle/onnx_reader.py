"""TEST INFRASTRUCTURE (oracle).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this; the product path never does.

Independent pure-Python restatement of the reference's ONNX protobuf reader:
wire primitives follow Wire.swift:38-150 (varint, fixed32/64, length-delimited, packed
repeated scalars); message field numbers follow ONNXLoader.swift:34-54 (ModelProto),
:93-147 (GraphProto), :169-211 (NodeProto), :213-318 (AttributeProto), :320-385
(TensorProto); initializer decoding (raw little-endian data, int32 widened to int64)
follows TensorValue.swift:46-116.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np


class ONNXDecodingError(Exception):
    pass


def _read_varint(buf: memoryview, pos: int) -> Tuple[int, int]:
    result = 0
    shift = 0
    while True:
        if pos >= len(buf):
            raise ONNXDecodingError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not (b & 0x80):
            return result & 0xFFFFFFFFFFFFFFFF, pos
        shift += 7
        if shift > 63:
            raise ONNXDecodingError("varint too long")


def _fields(buf: memoryview):
    pos = 0
    n = len(buf)
    while pos < n:
        key, pos = _read_varint(buf, pos)
        fnum, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
            yield fnum, wt, v
        elif wt == 1:
            if pos + 8 > n:
                raise ONNXDecodingError("truncated fixed64")
            yield fnum, wt, bytes(buf[pos:pos + 8])
            pos += 8
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            if pos + ln > n:
                raise ONNXDecodingError("truncated length-delimited field")
            yield fnum, wt, buf[pos:pos + ln]
            pos += ln
        elif wt == 5:
            if pos + 4 > n:
                raise ONNXDecodingError("truncated fixed32")
            yield fnum, wt, bytes(buf[pos:pos + 4])
            pos += 4
        else:
            raise ONNXDecodingError(f"unsupported wire type {wt}")


def _signed(v: int) -> int:
    return v - (1 << 64) if v >= (1 << 63) else v


def _packed_varints(v) -> List[int]:
    out = []
    pos = 0
    while pos < len(v):
        x, pos = _read_varint(v, pos)
        out.append(_signed(x))
    return out


@dataclass
class Node:
    op_type: str = ""
    name: str = ""
    inputs: List[str] = field(default_factory=list)
    outputs: List[str] = field(default_factory=list)
    attrs: Dict[str, object] = field(default_factory=dict)


@dataclass
class Model:
    ir_version: int = 0
    opset: int = 0
    graph_name: str = ""
    nodes: List[Node] = field(default_factory=list)
    initializers: Dict[str, np.ndarray] = field(default_factory=dict)
    inputs: List[str] = field(default_factory=list)
    outputs: List[str] = field(default_factory=list)


def _tensor(buf) -> Tuple[str, np.ndarray]:
    dims: List[int] = []
    dt = 0
    name = ""
    raw = None
    floats: List[float] = []
    i64s: List[int] = []
    for f, wt, v in _fields(buf):
        if f == 1:
            dims += _packed_varints(v) if wt == 2 else [_signed(v)]
        elif f == 2:
            dt = v
        elif f == 4:
            if wt == 2:
                floats += list(struct.unpack(f"<{len(v) // 4}f", bytes(v)))
            else:
                floats.append(struct.unpack("<f", v)[0])
        elif f == 7:
            i64s += _packed_varints(v) if wt == 2 else [_signed(v)]
        elif f == 8:
            name = bytes(v).decode()
        elif f == 9:
            raw = bytes(v)
    if dt == 1:
        arr = np.frombuffer(raw, dtype="<f4").copy() if raw is not None else np.array(floats, np.float32)
    elif dt == 7:
        arr = np.frombuffer(raw, dtype="<i8").copy() if raw is not None else np.array(i64s, np.int64)
    elif dt == 6:  # int32 widens to int64 (TensorValue.swift:84-100)
        arr = np.frombuffer(raw, dtype="<i4").astype(np.int64) if raw is not None else np.array(i64s, np.int64)
    elif dt == 9:
        arr = np.frombuffer(raw, dtype=np.uint8).astype(bool) if raw is not None else np.array(i64s, bool)
    else:
        raise ONNXDecodingError(f"unsupported tensor data_type {dt} for '{name}'")
    return name, arr.reshape(dims) if dims else arr.reshape(())


def _attr(buf) -> Tuple[str, object]:
    name, val, ints, floats = "", None, [], []
    for f, wt, v in _fields(buf):
        if f == 1:
            name = bytes(v).decode()
        elif f == 2:
            val = struct.unpack("<f", v)[0]
        elif f == 3:
            val = _signed(v)
        elif f == 4:
            val = bytes(v)
        elif f == 5:
            val = _tensor(v)[1]
        elif f == 7:
            floats += list(struct.unpack(f"<{len(v) // 4}f", bytes(v))) if wt == 2 else [struct.unpack("<f", v)[0]]
        elif f == 8:
            ints += _packed_varints(v) if wt == 2 else [_signed(v)]
    if ints:
        val = ints
    elif floats:
        val = floats
    return name, val


def _node(buf) -> Node:
    n = Node()
    for f, wt, v in _fields(buf):
        if f == 1:
            n.inputs.append(bytes(v).decode())
        elif f == 2:
            n.outputs.append(bytes(v).decode())
        elif f == 3:
            n.name = bytes(v).decode()
        elif f == 4:
            n.op_type = bytes(v).decode()
        elif f == 5:
            k, a = _attr(v)
            n.attrs[k] = a
    return n


def _value_info_name(buf) -> str:
    for f, wt, v in _fields(buf):
        if f == 1:
            return bytes(v).decode()
    return ""


def load_model(path: str) -> Model:
    with open(path, "rb") as fh:
        data = memoryview(fh.read())
    m = Model()
    graph = None
    for f, wt, v in _fields(data):
        if f == 1:
            m.ir_version = v
        elif f == 7:
            graph = v
        elif f == 8:
            for f2, _, v2 in _fields(v):
                if f2 == 2:
                    m.opset = max(m.opset, v2)
    if graph is None:
        raise ONNXDecodingError("model has no graph")
    for f, wt, v in _fields(graph):
        if f == 1:
            m.nodes.append(_node(v))
        elif f == 2:
            m.graph_name = bytes(v).decode()
        elif f == 5:
            name, arr = _tensor(v)
            m.initializers[name] = arr
        elif f == 11:
            m.inputs.append(_value_info_name(v))
        elif f == 12:
            m.outputs.append(_value_info_name(v))
    # ONNX lists initializers among graph inputs in old IR versions; the reference's
    # signature test (ONNXParsingTests.swift:22-27) sees only the three real feeds.
    m.inputs = [i for i in m.inputs if i not in m.initializers]
    return m
