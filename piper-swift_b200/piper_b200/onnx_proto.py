"""Minimal ONNX protobuf *writer* (no `onnx` package in this image).

Emits exactly the subset of ModelProto the reference's loader decodes
(ONNXLoader.swift:34-54 ModelProto{1,7,8}; :93-147 GraphProto{1,2,5,11,12};
:169-211 NodeProto{1,2,3,4,5,7}; :213-318 AttributeProto{1,2,3,4,5,7,8,9,20};
:320-385 TensorProto{1,2,4,7,8,9}), so a file written here is readable by the
reference's PiperONNX, by csrc/onnx_reader.cpp and by oracle/onnx_reader.py.
"""
from __future__ import annotations

import struct
from typing import Iterable, List, Sequence, Tuple, Union

import numpy as np

FLOAT, INT64, INT32, BOOL = 1, 7, 6, 9
ATTR_FLOAT, ATTR_INT, ATTR_STRING, ATTR_TENSOR, ATTR_FLOATS, ATTR_INTS = 1, 2, 3, 4, 6, 7


def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def f_varint(field: int, v: int) -> bytes:
    return _key(field, 0) + _varint(v)


def f_bytes(field: int, b: bytes) -> bytes:
    return _key(field, 2) + _varint(len(b)) + b


def f_str(field: int, s: str) -> bytes:
    return f_bytes(field, s.encode("utf-8"))


def f_float(field: int, v: float) -> bytes:
    return _key(field, 5) + struct.pack("<f", v)


def tensor(name: str, arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr)
    if arr.dtype == np.float32:
        dt = FLOAT
    elif arr.dtype == np.int64:
        dt = INT64
    else:
        raise TypeError(f"unsupported initializer dtype {arr.dtype}")
    out = bytearray()
    for d in arr.shape:
        out += f_varint(1, int(d))
    out += f_varint(2, dt)
    out += f_str(8, name)
    out += f_bytes(9, arr.astype(arr.dtype.newbyteorder("<")).tobytes())
    return bytes(out)


AttrVal = Union[int, float, str, Sequence[int]]


def attribute(name: str, val: AttrVal) -> bytes:
    out = bytearray(f_str(1, name))
    if isinstance(val, bool):
        val = int(val)
    if isinstance(val, int):
        out += f_varint(3, val) + f_varint(20, ATTR_INT)
    elif isinstance(val, float):
        out += f_float(2, val) + f_varint(20, ATTR_FLOAT)
    elif isinstance(val, str):
        out += f_bytes(4, val.encode()) + f_varint(20, ATTR_STRING)
    else:
        for v in val:
            out += f_varint(8, int(v))
        out += f_varint(20, ATTR_INTS)
    return bytes(out)


def node(op_type: str, inputs: Iterable[str], outputs: Iterable[str], name: str = "",
         attrs: Sequence[Tuple[str, AttrVal]] = ()) -> bytes:
    out = bytearray()
    for i in inputs:
        out += f_str(1, i)
    for o in outputs:
        out += f_str(2, o)
    if name:
        out += f_str(3, name)
    out += f_str(4, op_type)
    for k, v in attrs:
        out += f_bytes(5, attribute(k, v))
    return bytes(out)


def value_info(name: str) -> bytes:
    return f_str(1, name)


def model(nodes: List[bytes], initializers: List[bytes], inputs: List[str], outputs: List[str],
          graph_name: str = "torch_jit", opset: int = 15, ir_version: int = 8,
          producer: str = "pytorch") -> bytes:
    g = bytearray()
    for n in nodes:
        g += f_bytes(1, n)
    g += f_str(2, graph_name)
    for t in initializers:
        g += f_bytes(5, t)
    for i in inputs:
        g += f_bytes(11, value_info(i))
    for o in outputs:
        g += f_bytes(12, value_info(o))
    m = bytearray()
    m += f_varint(1, ir_version)
    m += f_str(2, producer)
    m += f_bytes(7, bytes(g))
    m += f_bytes(8, f_str(1, "") + f_varint(2, opset))
    return bytes(m)
