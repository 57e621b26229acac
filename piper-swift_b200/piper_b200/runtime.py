"""Host-side mirror of the reference's `PiperMetalRuntime` (PiperMetalRuntime.swift:10-122)
over the C ABI of include/piper_b200.h (ctypes; the Swift twin is in swift/).

Same method names, argument meaning and error behaviour:
  PiperCUDARuntime(model_path, config_path, options)      ~ init(modelURL:configURL:options:) :37
  .synthesize(phoneme_ids, noise_scale=None, ...)         ~ synthesize(phonemeIDs:...) -> [Float] :62
  .synthesize_stream(..., chunk_size=2048)                ~ synthesizeStream(...) :84 (PiperAudioChunk)
  .last_run_timings                                       ~ lastRunTimings :33
plus the additive batch / speaker / injected-noise forms the B200 configs need.
There is NO CPU fallback: if the CUDA library is missing or no GPU is present the
constructor raises.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from collections.abc import Sequence as _SequenceABC
from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PIPER_B200_LIB: a diagnostic build of the same library (e.g. `make trace`); there is no non-CUDA alternative
LIB_PATH = os.environ.get("PIPER_B200_LIB") or os.path.join(os.path.dirname(_HERE), "libpiper_b200.so")

OK = 0
PRECISION = {"fp32": 0, "bf16": 1, "f16": 2}
FLAG_IDS_ON_DEVICE, FLAG_KEEP_ON_DEVICE, FLAG_KEEP_STAGES, FLAG_PCM16, FLAG_PROFILE, FLAG_STREAM = 1, 2, 4, 8, 16, 32
FLAG_GENERIC = 128
FLAG_PCM16_ONLY = 64
ABI_VERSION = 2
ERR_NAMES = {1: "invalid_argument", 2: "io", 3: "decode", 4: "unsupported", 5: "cuda", 6: "state", 7: "capacity"}


class PiperB200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[{ERR_NAMES.get(code, code)}] {message}")
        self.code = code


class VoiceInfo(C.Structure):
    _fields_ = [("sample_rate", C.c_int32), ("num_symbols", C.c_int32), ("num_speakers", C.c_int32),
                ("hidden", C.c_int32), ("hop", C.c_int32), ("dec_channels0", C.c_int32),
                ("resblock", C.c_int32), ("n_ups", C.c_int32),
                ("noise_scale", C.c_float), ("length_scale", C.c_float), ("noise_w", C.c_float),
                ("n_params", C.c_int64), ("dec_flops_per_frame", C.c_double),
                ("flow_flops_per_frame", C.c_double), ("enc_flops_per_id", C.c_double)]


class Request(C.Structure):
    _fields_ = [("phoneme_ids", C.c_void_p), ("lengths", C.c_void_p), ("speaker_ids", C.c_void_p),
                ("batch", C.c_int32), ("flags", C.c_uint32),
                ("noise_scale", C.c_float), ("length_scale", C.c_float), ("noise_w", C.c_float),
                ("noise_dp", C.c_void_p), ("noise_main", C.c_void_p), ("noise_main_stride", C.c_int64),
                ("seed", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [("audio", C.c_void_p), ("pcm16", C.c_void_p), ("sample_offsets", C.POINTER(C.c_int64)),
                ("frame_lengths", C.POINTER(C.c_int64)), ("durations", C.POINTER(C.c_int32)),
                ("total_samples", C.c_int64), ("total_frames", C.c_int64)]


class Timings(C.Structure):
    _fields_ = [("wall_ms", C.c_double), ("gpu_ms", C.c_double), ("front_ms", C.c_double),
                ("flow_ms", C.c_double), ("decoder_ms", C.c_double), ("h2d_ms", C.c_double),
                ("d2h_ms", C.c_double), ("kernel_launches", C.c_int64), ("h2d_bytes", C.c_int64),
                ("d2h_bytes", C.c_int64), ("dec_flops", C.c_double), ("flow_flops", C.c_double),
                ("class_ms", C.c_double * 3), ("class_flops", C.c_double * 3), ("class_launches", C.c_int64 * 3)]

    def as_dict(self):
        d = {}
        for k, _ in self._fields_:
            v = getattr(self, k)
            d[k] = list(v) if hasattr(v, "__len__") else v
        return d


_lib = None


def load_library(path: str = LIB_PATH):
    """dlopen the C-ABI library and declare every prototype of include/piper_b200.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise PiperB200Error(5, f"CUDA library not built: {path} (run __graft_entry__.build())")
    lib = C.CDLL(path)
    P = C.c_void_p
    sig = {
        "piper_b200_abi_version": (C.c_int, []),
        "piper_b200_last_error": (C.c_char_p, [P]),
        "piper_b200_device_count": (C.c_int, [C.POINTER(C.c_int)]),
        "piper_b200_create": (C.c_int, [C.c_int, C.POINTER(P)]),
        "piper_b200_destroy": (None, [P]),
        "piper_b200_load_voice": (C.c_int, [P, C.c_char_p, C.c_char_p]),
        "piper_b200_set_weight": (C.c_int, [P, C.c_char_p, C.c_int32, C.POINTER(C.c_int64), C.c_int32, P, C.c_size_t]),
        "piper_b200_set_conv_attr": (C.c_int, [P, C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
        "piper_b200_add_node": (C.c_int, [P, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), C.c_int32, C.POINTER(C.c_char_p),
                                          C.c_int32] + [C.c_int32] * 5),
        "piper_b200_set_config": (C.c_int, [P, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float]),
        "piper_b200_finalize": (C.c_int, [P]),
        "piper_b200_get_voice_info": (C.c_int, [P, C.POINTER(VoiceInfo)]),
        "piper_b200_set_precision": (C.c_int, [P, C.c_int32]),
        "piper_b200_set_stream": (C.c_int, [P, P]),
        "piper_b200_set_output_buffer": (C.c_int, [P, P, C.c_size_t]),
        "piper_b200_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(P)]),
        "piper_b200_host_free": (None, [P]),
        "piper_b200_host_register": (C.c_int, [P, C.c_size_t]),
        "piper_b200_host_unregister": (C.c_int, [P]),
        "piper_b200_synthesize": (C.c_int, [P, C.POINTER(Request), C.POINTER(Result)]),
        "piper_b200_get_timings": (C.c_int, [P, C.POINTER(Timings)]),
        "piper_b200_stream_next": (C.c_int, [P, C.c_int64, C.POINTER(Result), C.POINTER(C.c_int32)]),
        "piper_b200_get_stage": (C.c_int, [P, C.c_char_p, C.c_int32, P, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
        "piper_b200_length_regulate": (C.c_int, [P, P, P, C.c_int32, C.c_float, P, P, P, C.c_int64]),
        "piper_b200_length_regulate_noalloc": (C.c_int, [P, P, P, C.c_int32, C.c_float, P, P]),
        "piper_b200_random_normal": (C.c_int, [P, C.c_uint64, C.c_int64, P]),
        "piper_b200_selftest_conv": (C.c_int, [P] + [C.c_int32] * 8 + [C.POINTER(C.c_float)]),
        "piper_b200_run_graph": (C.c_int, [P, P, C.c_int64, C.c_int64, P, C.c_uint64, C.c_char_p, P, C.c_size_t,
                                           C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
        "piper_b200_conv1d": (C.c_int, [P, C.c_int32, P, C.c_int64, P, P] + [C.c_int32] * 6 + [P]),
        "piper_b200_ipa_to_ids": (C.c_int, [P, C.c_char_p, C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_int64)]),
        "piper_b200_config_ipa_to_ids": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_int64), C.c_int64,
                                                   C.POINTER(C.c_int64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


EXPORTED_SYMBOLS = [
    "piper_b200_abi_version", "piper_b200_last_error", "piper_b200_create", "piper_b200_destroy",
    "piper_b200_load_voice", "piper_b200_set_weight", "piper_b200_set_conv_attr", "piper_b200_set_config",
    "piper_b200_finalize", "piper_b200_get_voice_info", "piper_b200_set_precision", "piper_b200_set_stream",
    "piper_b200_synthesize", "piper_b200_get_timings", "piper_b200_get_stage", "piper_b200_length_regulate",
    "piper_b200_random_normal", "piper_b200_selftest_conv", "piper_b200_ipa_to_ids", "piper_b200_config_ipa_to_ids",
    "piper_b200_stream_next", "piper_b200_device_count", "piper_b200_set_output_buffer", "piper_b200_host_alloc",
    "piper_b200_host_free", "piper_b200_host_register", "piper_b200_host_unregister",
    "piper_b200_length_regulate_noalloc", "piper_b200_add_node", "piper_b200_conv1d", "piper_b200_run_graph",
]
NOALLOC_MAX_IDS, NOALLOC_MAX_BATCH = 65536, 1024


def device_count() -> int:
    """GPUs visible to the process (the reference has one MTLDevice, MetalContext.swift:11-19)."""
    lib = load_library()
    n = C.c_int(0)
    rc = lib.piper_b200_device_count(C.byref(n))
    if rc != OK:
        raise PiperB200Error(rc, lib.piper_b200_last_error(None).decode())
    return int(n.value)


class HostArena:
    """Page-locked host memory that one or several contexts write their audio into (piper_b200_set_output_buffer):
    allocated by the library (portable across the GPUs of the process), or an existing mapping - e.g. a /dev/shm
    file shared by the ranks of one box - pinned in place."""

    def __init__(self, nbytes: int, mapping=None):
        self._lib = load_library()
        self.nbytes = int(nbytes)
        self._owned = mapping is None
        self._mapping = mapping
        if mapping is None:
            p = C.c_void_p()
            rc = self._lib.piper_b200_host_alloc(self.nbytes, C.byref(p))
            if rc != OK:
                raise PiperB200Error(rc, self._lib.piper_b200_last_error(None).decode())
            self.ptr = int(p.value)
        else:
            self.ptr = C.addressof(C.c_char.from_buffer(mapping))
            rc = self._lib.piper_b200_host_register(C.c_void_p(self.ptr), self.nbytes)
            if rc != OK:
                raise PiperB200Error(rc, self._lib.piper_b200_last_error(None).decode())

    def close(self):
        if getattr(self, "ptr", 0):
            if self._owned:
                self._lib.piper_b200_host_free(C.c_void_p(self.ptr))
            else:
                self._lib.piper_b200_host_unregister(C.c_void_p(self.ptr))
            self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def config_ipa_to_ids(config_path: str, ipa: str) -> List[int]:
    """idsFromIPA (PiperCLI.swift:115-136) / ESpeakPhonemizer.phonemeIDs (ESpeakPhonemizer.swift:73-100) on a
    voice's `*.onnx.json`: host-only, no context and no GPU needed."""
    lib = load_library()
    raw = ipa.encode("utf-8")
    cap = 2 * len(raw) + 2
    buf = (C.c_int64 * cap)()
    n = C.c_int64(0)
    rc = lib.piper_b200_config_ipa_to_ids(config_path.encode(), raw, buf, cap, C.byref(n))
    if rc != 0:
        raise PiperB200Error(rc, lib.piper_b200_last_error(None).decode())
    return list(buf[: n.value])


@dataclass
class Options:
    """PiperMetalRuntime.Options (PiperMetalRuntime.swift:11-26).  preferCPU* have no meaning
    here (there is no CPU path) and are accepted only for source compatibility."""
    seed: int = 1234
    prefer_cpu_conv: bool = False
    prefer_cpu_elementwise: bool = False
    device: int = 0
    precision: str = "fp32"


@dataclass
class PiperAudioFormat:          # PiperAudio.swift:3-11
    sample_rate: int
    channels: int = 1


@dataclass
class PiperAudioChunk:           # PiperAudio.swift:13-25
    format: PiperAudioFormat
    start_sample_index: int
    samples: np.ndarray
    is_final: bool



class SliceViews(_SequenceABC):
    """Per-utterance views of one flat result array, made on access.  A list of 256 numpy slices per field cost 0.2-0.4 ms
    of host time inside every call - 3 % of the end-to-end step of the 256-utterance batch."""
    __slots__ = ("flat", "offs")

    def __init__(self, flat: np.ndarray, offs):
        self.flat = flat
        self.offs = offs.tolist() if isinstance(offs, np.ndarray) else list(offs)

    def __len__(self) -> int:
        return len(self.offs) - 1

    def __getitem__(self, i):
        n = len(self.offs) - 1
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(n))]
        i = int(i)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError("utterance index out of range")
        return self.flat[self.offs[i]:self.offs[i + 1]]

@dataclass
class BatchResult:
    audio: List[np.ndarray]
    frame_lengths: np.ndarray
    durations: List[np.ndarray]
    total_samples: int
    total_frames: int
    pcm16: Optional[List[np.ndarray]] = None
    device_ptr: Optional[int] = None


class PiperCUDARuntime:
    """B200 drop-in for PiperMetalRuntime (PiperMetalRuntime.swift:10)."""

    def __init__(self, model_path: str, config_path: str, options: Optional[Options] = None):
        self.options = options or Options()
        self._lib = load_library()
        self._ctx = C.c_void_p()
        rc = self._lib.piper_b200_create(self.options.device, C.byref(self._ctx))
        if rc != OK:
            raise PiperB200Error(rc, self._lib.piper_b200_last_error(None).decode())
        try:
            self._check(self._lib.piper_b200_load_voice(self._ctx, model_path.encode(), config_path.encode()))
            self.set_precision(self.options.precision)
            info = VoiceInfo()
            self._check(self._lib.piper_b200_get_voice_info(self._ctx, C.byref(info)))
            self.info = info
        except Exception:
            self.close()
            raise
        self.model_path, self.config_path = model_path, config_path

    @classmethod
    def from_parsed(cls, initializers, nodes, sample_rate: int, num_speakers: int = 1, noise_scale: float = 0.667,
                    length_scale: float = 1.0, noise_w: float = 0.8, options: Optional[Options] = None):
        """init(model:config:options:) (PiperMetalRuntime.swift:54-59), route B of the C ABI: the host has already
        parsed the model (Swift: PiperONNX's ONNXModel; here any reader) and pushes it - every initializer with
        piper_b200_set_weight, every graph node with piper_b200_add_node, the config scalars with piper_b200_set_config -
        then piper_b200_finalize binds and repacks.  `initializers`: name -> ndarray; `nodes`: objects with op_type, name,
        inputs, outputs and an `attrs` dict (strides / dilations / group / pads / value)."""
        self = cls.__new__(cls)
        self.options = options or Options()
        self._lib = load_library()
        self._ctx = C.c_void_p()
        rc = self._lib.piper_b200_create(self.options.device, C.byref(self._ctx))
        if rc != OK:
            raise PiperB200Error(rc, self._lib.piper_b200_last_error(None).decode())
        dtypes = {np.dtype(np.float32): 1, np.dtype(np.int32): 6, np.dtype(np.int64): 7, np.dtype(np.bool_): 9}

        def push(name, arr):
            arr = np.ascontiguousarray(arr)
            if arr.dtype not in dtypes:
                return
            dims = (C.c_int64 * max(arr.ndim, 1))(*arr.shape)
            self._check(self._lib.piper_b200_set_weight(self._ctx, name.encode(), dtypes[arr.dtype], dims, arr.ndim,
                                                        arr.ctypes.data, arr.nbytes))
        try:
            for name, arr in initializers.items():
                push(name, arr)
            for n in nodes:
                at = getattr(n, "attrs", {}) or {}
                if n.op_type == "Constant":       # its float value travels as a weight under the node's output name
                    v = at.get("value")
                    if isinstance(v, np.ndarray) and v.dtype == np.float32 and n.outputs:
                        push(n.outputs[0], v)
                    continue
                ins = (C.c_char_p * max(len(n.inputs), 1))(*[i.encode() for i in n.inputs])
                outs = (C.c_char_p * max(len(n.outputs), 1))(*[o.encode() for o in n.outputs])
                pads = at.get("pads") or [-1, -1]
                self._check(self._lib.piper_b200_add_node(
                    self._ctx, n.op_type.encode(), (n.name or "").encode(), ins, len(n.inputs), outs, len(n.outputs),
                    int((at.get("strides") or [1])[0]), int((at.get("dilations") or [1])[0]), int(at.get("group") or 1),
                    int(pads[0]), int(pads[1] if len(pads) > 1 else pads[0])))
            self._check(self._lib.piper_b200_set_config(self._ctx, sample_rate, num_speakers, noise_scale, length_scale,
                                                        noise_w))
            self._check(self._lib.piper_b200_finalize(self._ctx))
            self.set_precision(self.options.precision)
            info = VoiceInfo()
            self._check(self._lib.piper_b200_get_voice_info(self._ctx, C.byref(info)))
            self.info = info
        except Exception:
            self.close()
            raise
        self.model_path = self.config_path = None
        return self

    # ---- plumbing ---------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != OK:
            raise PiperB200Error(rc, self._lib.piper_b200_last_error(self._ctx).decode())

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx:
            self._lib.piper_b200_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_precision(self, precision: str):
        self._check(self._lib.piper_b200_set_precision(self._ctx, PRECISION[precision]))
        self.precision = precision

    def set_stream(self, cuda_stream_ptr: Optional[int]):
        self._check(self._lib.piper_b200_set_stream(self._ctx, C.c_void_p(cuda_stream_ptr or 0)))

    def set_output_buffer(self, ptr: Optional[int], nbytes: int = 0):
        """Caller-owned page-locked destination of the audio (HostArena); None restores the library's own buffer.
        Results then are views into that memory: they stay valid until the caller reuses it."""
        self._check(self._lib.piper_b200_set_output_buffer(self._ctx, C.c_void_p(ptr or 0), nbytes if ptr else 0))
        self._out_ptr = ptr or None

    @property
    def sample_rate(self) -> int:
        return int(self.info.sample_rate)

    @property
    def last_run_timings(self) -> dict:
        t = Timings()
        self._check(self._lib.piper_b200_get_timings(self._ctx, C.byref(t)))
        return t.as_dict()

    # ---- the reference's API --------------------------------------------------------------
    def synthesize(self, phoneme_ids: Sequence[int], noise_scale: Optional[float] = None,
                   length_scale: Optional[float] = None, noise_w: Optional[float] = None,
                   speaker_id: Optional[int] = None) -> np.ndarray:
        """Returns raw float32 samples (mono), like PiperMetalRuntime.synthesize (:62-80)."""
        r = self.synthesize_batch([phoneme_ids], noise_scale, length_scale, noise_w,
                                  speaker_ids=None if speaker_id is None else [speaker_id])
        return r.audio[0]

    def ipa_to_ids(self, ipa: str) -> List[int]:
        """IPA string -> phoneme ids with the loaded voice's phoneme_id_map (ESpeakPhonemizer.swift:73-100)."""
        raw = ipa.encode("utf-8")
        cap = 2 * len(raw) + 2
        buf = (C.c_int64 * cap)()
        n = C.c_int64(0)
        self._check(self._lib.piper_b200_ipa_to_ids(self._ctx, raw, buf, cap, C.byref(n)))
        return list(buf[: n.value])

    def synthesize_ipa(self, ipa: str, noise_scale: Optional[float] = None, length_scale: Optional[float] = None,
                       noise_w: Optional[float] = None) -> np.ndarray:
        """The CLI's text path minus espeak-ng (PiperCLI.swift:115-136 -> :173): IPA in, waveform out."""
        return self.synthesize(self.ipa_to_ids(ipa), noise_scale, length_scale, noise_w)

    def synthesize_stream(self, phoneme_ids: Sequence[int], noise_scale: Optional[float] = None,
                          length_scale: Optional[float] = None, noise_w: Optional[float] = None,
                          chunk_size: int = 2048) -> Iterator[PiperAudioChunk]:
        """synthesizeStream (:84-121).  The reference chunks the finished waveform; here the generator itself
        runs chunk by chunk (PIPER_B200_FLAG_STREAM + piper_b200_stream_next): the front end and the flow run
        once, then every chunk decodes `chunk_size` samples (rounded up to whole frames) plus the generator's
        receptive-field halo.  The concatenated chunks are bit-identical to synthesize()."""
        assert chunk_size > 0
        fmt = PiperAudioFormat(self.sample_rate, 1)
        hop = int(self.info.hop)
        frames = max(1, -(-chunk_size // hop))
        ids = np.asarray(phoneme_ids, dtype=np.int64)
        self.synthesize_packed(ids, np.array([len(ids)], np.int64), noise_scale, length_scale, noise_w,
                               stream=True)
        idx, final = 0, C.c_int32(0)
        pending = np.zeros(0, np.float32)
        while not final.value:
            rs = Result()
            self._check(self._lib.piper_b200_stream_next(self._ctx, frames, C.byref(rs), C.byref(final)))
            n = int(rs.total_samples)
            if n:
                got = np.ctypeslib.as_array(C.cast(rs.audio, C.POINTER(C.c_float)), shape=(n,))
                pending = np.concatenate([pending, got])
            while len(pending) >= chunk_size and not (final.value and len(pending) == chunk_size):
                yield PiperAudioChunk(fmt, idx, pending[:chunk_size].copy(), False)
                pending = pending[chunk_size:]
                idx += chunk_size
        if len(pending):
            yield PiperAudioChunk(fmt, idx, pending.copy(), True)

    # ---- additive: ragged batch, speakers, injected noise, device-resident inputs ---------------
    def synthesize_batch(self, utterances: Sequence[Sequence[int]], noise_scale: Optional[float] = None,
                         length_scale: Optional[float] = None, noise_w: Optional[float] = None,
                         speaker_ids: Optional[Sequence[int]] = None,
                         noise_dp: Optional[Sequence[np.ndarray]] = None,
                         noise_main: Optional[Sequence[np.ndarray]] = None,
                         seed: Optional[int] = None, keep_stages: bool = False, pcm16: bool = False,
                         copy: bool = True, pcm16_only: bool = False, generic: bool = False) -> BatchResult:
        B = len(utterances)
        lengths = np.array([len(u) for u in utterances], dtype=np.int64)
        ids = np.concatenate([np.asarray(u, dtype=np.int64) for u in utterances]) if B else np.zeros(0, np.int64)
        return self.synthesize_packed(ids, lengths, noise_scale, length_scale, noise_w, speaker_ids, noise_dp,
                                      noise_main, seed, keep_stages, pcm16, copy, pcm16_only=pcm16_only, generic=generic)

    def synthesize_packed(self, ids: np.ndarray, lengths: np.ndarray, noise_scale=None, length_scale=None,
                          noise_w=None, speaker_ids=None, noise_dp=None, noise_main=None, seed=None,
                          keep_stages=False, pcm16=False, copy=True, ids_device_ptr: Optional[int] = None,
                          keep_on_device: bool = False, profile: bool = False, stream: bool = False,
                          pcm16_only: bool = False, generic: bool = False) -> BatchResult:
        """generic=True: PIPER_B200_FLAG_GENERIC - the graph runs node by node through the generic op executor."""
        B = int(len(lengths))
        lengths = np.ascontiguousarray(lengths, dtype=np.int64)
        rq = Request()
        keep = []   # keep numpy buffers alive across the call
        if ids_device_ptr is not None:
            rq.phoneme_ids = ids_device_ptr
        else:
            ids = np.ascontiguousarray(ids, dtype=np.int64)
            keep.append(ids)
            rq.phoneme_ids = ids.ctypes.data
        rq.lengths = lengths.ctypes.data
        if speaker_ids is not None:
            sids = np.ascontiguousarray(speaker_ids, dtype=np.int64)
            keep.append(sids)
            rq.speaker_ids = sids.ctypes.data
        rq.batch = B
        nan = float("nan")
        rq.noise_scale = nan if noise_scale is None else noise_scale
        rq.length_scale = nan if length_scale is None else length_scale
        rq.noise_w = nan if noise_w is None else noise_w
        if noise_dp is not None:
            nd = np.ascontiguousarray(np.concatenate([np.asarray(n, np.float32).reshape(-1) for n in noise_dp]))
            if nd.size != 2 * int(lengths.sum()):
                raise PiperB200Error(1, "noise_dp must be [2, T_b] per utterance")
            keep.append(nd)
            rq.noise_dp = nd.ctypes.data
        if noise_main is not None:
            H = int(self.info.hidden)
            stride = max(int(n.shape[1]) for n in noise_main)
            nm = np.zeros((B, H, stride), np.float32)
            for b, n in enumerate(noise_main):
                nm[b, :, :n.shape[1]] = n
            keep.append(nm)
            rq.noise_main = nm.ctypes.data
            rq.noise_main_stride = stride
        rq.seed = self.options.seed if seed is None else seed
        rq.flags = ((FLAG_KEEP_STAGES if keep_stages else 0) | (FLAG_PCM16 if pcm16 else 0) |
                    (FLAG_PROFILE if profile else 0) | (FLAG_STREAM if stream else 0) |
                    (FLAG_PCM16_ONLY if pcm16_only else 0) | (FLAG_GENERIC if generic else 0) |
                    (FLAG_IDS_ON_DEVICE if ids_device_ptr is not None else 0) |
                    (FLAG_KEEP_ON_DEVICE if keep_on_device else 0))
        rs = Result()
        self._check(self._lib.piper_b200_synthesize(self._ctx, C.byref(rq), C.byref(rs)))
        offs = np.ctypeslib.as_array(rs.sample_offsets, shape=(B + 1,)).copy()
        fl = np.ctypeslib.as_array(rs.frame_lengths, shape=(B,)).copy()
        n_ids = int(lengths.sum())
        dur_all = np.ctypeslib.as_array(rs.durations, shape=(n_ids,)).copy()
        id_off = np.concatenate([[0], np.cumsum(lengths)])
        durs = SliceViews(dur_all, id_off)
        out = BatchResult([], fl, durs, int(rs.total_samples), int(rs.total_frames))
        if keep_on_device:
            out.device_ptr = int(rs.audio)
            return out
        if stream:      # audio follows through piper_b200_stream_next
            return out
        total = int(rs.total_samples)
        own = getattr(self, "_out_ptr", None) is None      # a caller-owned buffer is the caller's to keep: no copy
        if pcm16_only:      # int16 is the only host output (half the device->host bytes)
            p = np.ctypeslib.as_array(C.cast(rs.pcm16, C.POINTER(C.c_int16)), shape=(total,))
            if copy and own:
                p = p.copy()
            out.pcm16 = SliceViews(p, offs)
            return out
        flat = np.ctypeslib.as_array(C.cast(rs.audio, C.POINTER(C.c_float)), shape=(total,))
        if copy and own:
            flat = flat.copy()
        out.audio = SliceViews(flat, offs)
        if pcm16:
            p = np.ctypeslib.as_array(C.cast(rs.pcm16, C.POINTER(C.c_int16)), shape=(total,)).copy()
            out.pcm16 = SliceViews(p, offs)
        return out

    # ---- parity hooks -------------------------------------------------------------------------------
    def get_stage(self, name: str, utt: int = 0) -> np.ndarray:
        ch, ln = C.c_int64(), C.c_int64()
        self._check(self._lib.piper_b200_get_stage(self._ctx, name.encode(), utt, None, 0, C.byref(ch), C.byref(ln)))
        out = np.zeros((ch.value, ln.value), np.float32)
        self._check(self._lib.piper_b200_get_stage(self._ctx, name.encode(), utt, out.ctypes.data, out.size,
                                                   C.byref(ch), C.byref(ln)))
        return out

    def length_regulate(self, logw: Sequence[np.ndarray], length_scale: float):
        """(durations per utterance, frame_lengths, frame->phoneme index per utterance)"""
        B = len(logw)
        lengths = np.array([len(l) for l in logw], dtype=np.int64)
        flat = np.ascontiguousarray(np.concatenate([np.asarray(l, np.float32) for l in logw]))
        dur = np.zeros(int(lengths.sum()), np.int32)
        fl = np.zeros(B, np.int64)
        self._check(self._lib.piper_b200_length_regulate(self._ctx, flat.ctypes.data, lengths.ctypes.data, B,
                                                         length_scale, dur.ctypes.data, fl.ctypes.data, None, 0))
        cap = int(fl.max())
        idx = np.full((B, cap), -2, np.int32)
        self._check(self._lib.piper_b200_length_regulate(self._ctx, flat.ctypes.data, lengths.ctypes.data, B,
                                                         length_scale, dur.ctypes.data, fl.ctypes.data,
                                                         idx.ctypes.data, cap))
        off = np.concatenate([[0], np.cumsum(lengths)])
        return ([dur[off[b]:off[b + 1]] for b in range(B)], fl, [idx[b, :fl[b]] for b in range(B)])

    def length_regulate_noalloc(self, logw: Sequence[np.ndarray], length_scale: float):
        """piper_b200_length_regulate_noalloc: (durations per utterance, frame_lengths) out of the library's fixed
        device scratch array - no allocation on the device side; at most NOALLOC_MAX_IDS ids / NOALLOC_MAX_BATCH
        utterances.  Calls on one GPU must not overlap."""
        B = len(logw)
        lengths = np.array([len(l) for l in logw], dtype=np.int64)
        flat = np.ascontiguousarray(np.concatenate([np.asarray(l, np.float32) for l in logw]))
        dur = np.zeros(int(lengths.sum()), np.int32)
        fl = np.zeros(B, np.int64)
        self._check(self._lib.piper_b200_length_regulate_noalloc(self._ctx, flat.ctypes.data, lengths.ctypes.data, B,
                                                                 length_scale, dur.ctypes.data, fl.ctypes.data))
        off = np.concatenate([[0], np.cumsum(lengths)])
        return [dur[off[b]:off[b + 1]] for b in range(B)], fl

    def random_normal(self, count: int, seed: int = 1234) -> np.ndarray:
        out = np.zeros(count, np.float32)
        self._check(self._lib.piper_b200_random_normal(self._ctx, seed, count, out.ctypes.data))
        return out

    def run_graph(self, phoneme_ids: Sequence[int], noise_scale: float, length_scale: float, noise_w: float,
                  value_name: Optional[str] = None, speaker_id: int = 0, seed: Optional[int] = None) -> np.ndarray:
        """piper_b200_run_graph: the generic op executor on one utterance; returns the named value (default: the graph
        output) as float32 / int64 / bool - GraphExecutor.execute(inputs:maxNodeIndex:) (GraphExecutor.swift:75-152)."""
        ids = np.ascontiguousarray(phoneme_ids, np.int64)
        scales = np.array([noise_scale, length_scale, noise_w], np.float32)
        dt, rank = C.c_int32(0), C.c_int32(0)
        dims = (C.c_int64 * 8)()
        name = value_name.encode() if value_name else None
        sd = self.options.seed if seed is None else seed
        self._check(self._lib.piper_b200_run_graph(self._ctx, ids.ctypes.data, len(ids), speaker_id, scales.ctypes.data, sd, name,
                                                   None, 0, C.byref(dt), dims, C.byref(rank)))
        shape = tuple(int(dims[i]) for i in range(rank.value))
        out = np.empty(shape, {1: np.float32, 7: np.int64, 9: np.bool_}[dt.value])
        self._check(self._lib.piper_b200_run_graph(self._ctx, ids.ctypes.data, len(ids), speaker_id, scales.ctypes.data, sd, name,
                                                   out.ctypes.data, out.nbytes, C.byref(dt), dims, C.byref(rank)))
        return out

    def conv1d(self, x: np.ndarray, w: np.ndarray, bias: Optional[np.ndarray] = None, dilation: int = 1, stride: int = 1,
               transposed: bool = False, precision: str = "fp32") -> np.ndarray:
        """piper_b200_conv1d: one Conv / ConvTranspose node on caller-supplied operands.  x [rows][cin] time-major, w in
        the ONNX layout (Conv [cout][cin][k], ConvTranspose [cin][cout][k]); returns [rows*stride][cout]."""
        x = np.ascontiguousarray(x, np.float32)
        w = np.ascontiguousarray(w, np.float32)
        rows, cin = x.shape
        cout = w.shape[1] if transposed else w.shape[0]
        assert (w.shape[0] if transposed else w.shape[1]) == cin
        k = w.shape[2]
        b = None if bias is None else np.ascontiguousarray(bias, np.float32)
        out = np.empty((rows * (stride if transposed else 1), cout), np.float32)
        self._check(self._lib.piper_b200_conv1d(self._ctx, 1 if transposed else 0, x.ctypes.data, rows, w.ctypes.data,
                                                None if b is None else b.ctypes.data, cin, cout, k, dilation, stride,
                                                PRECISION[precision], out.ctypes.data))
        return out

    def selftest_conv(self, mode: int, cin: int, cout: int, k: int = 1, dilation: int = 1, stride: int = 1,
                      rows: int = 1000, precision: str = "bf16") -> dict:
        """piper_b200_selftest_conv: tcgen05 kernel vs fp32 CUDA-core kernel on identical operands."""
        out = (C.c_float * 4)()
        self._check(self._lib.piper_b200_selftest_conv(self._ctx, mode, cin, cout, k, dilation, stride, rows,
                                                       PRECISION[precision], out))
        return {"max_abs_diff": out[0], "max_abs_ref": out[1], "tc_ms": out[2], "f32_ms": out[3]}
