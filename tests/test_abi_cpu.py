"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import pytest
import torch

from common import ROOT, voice_files
from piper_b200 import runtime


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "piper_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(piper_b200_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = runtime.load_library()
    syms = _header_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/piper_b200.h but not exported"
    assert sorted(runtime.EXPORTED_SYMBOLS) == syms
    assert lib.piper_b200_abi_version() == runtime.ABI_VERSION == 2


def test_struct_sizes_match_header():
    # layout guard for the ctypes mirrors (x86-64 / aarch64 LP64)
    assert C.sizeof(runtime.Request) == 80
    assert C.sizeof(runtime.Result) == 56
    assert C.sizeof(runtime.Timings) == 168
    assert C.sizeof(runtime.VoiceInfo) == 80


def test_null_arguments_are_rejected_without_touching_the_gpu():
    lib = runtime.load_library()
    assert lib.piper_b200_create(0, None) == 1
    assert lib.piper_b200_synthesize(None, None, None) == 1
    assert lib.piper_b200_load_voice(None, b"x", b"y") == 1
    assert b"invalid argument" in lib.piper_b200_last_error(None)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly():
    lib = runtime.load_library()
    ctx = C.c_void_p()
    rc = lib.piper_b200_create(0, C.byref(ctx))
    assert rc == 5 and not ctx.value           # PIPER_B200_ERR_CUDA, no context, no fallback
    assert lib.piper_b200_last_error(None).startswith(b"cuda:")
    mp, cp = voice_files("tiny")
    with pytest.raises(runtime.PiperB200Error) as e:
        runtime.PiperCUDARuntime(mp, cp)
    assert e.value.code == 5


def test_product_path_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "piper-swift_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h", ".swift")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_ipa_to_ids_follows_the_reference_rule():
    """ESpeakPhonemizer.phonemeIDs (ESpeakPhonemizer.swift:73-100) / idsFromIPA (PiperCLI.swift:115-136):
    BOS, id + blank per scalar, EOS without trailing blank; CR/LF, Cf and ZW*/VS15/16 scalars skipped."""
    _, cfg = voice_files("tiny")
    id_of = {c: 3 + i for i, c in enumerate(" !',-.:;?abcdefghijklmnopqrstuvwxyz")}
    text = "hi, you"
    want = [1] + [v for ch in text for v in (id_of[ch], 0)] + [2]
    assert runtime.config_ipa_to_ids(cfg, text) == want
    assert len(want) == 2 * len(text) + 2 and want[-2:] == [0, 2]
    noisy = "h‍i️,\r\n ­you⁠"          # ZWJ, VS16, CR LF, soft hyphen (Cf), word joiner (Cf)
    assert runtime.config_ipa_to_ids(cfg, noisy) == want
    assert runtime.config_ipa_to_ids(cfg, "") == [1, 2]
    with pytest.raises(runtime.PiperB200Error) as e:
        runtime.config_ipa_to_ids(cfg, "hé")              # not in phoneme_id_map
    assert "é" in str(e.value)
    lib = runtime.load_library()
    n = C.c_int64(0)
    buf = (C.c_int64 * 4)()
    rc = lib.piper_b200_config_ipa_to_ids(cfg.encode(), b"hello", buf, 4, C.byref(n))
    assert rc == 7 and n.value == 12 and list(buf) == [1, id_of["h"], 0, id_of["e"]]   # capacity: count still reported
    assert lib.piper_b200_config_ipa_to_ids(b"/nonexistent.json", b"a", buf, 4, C.byref(n)) == 2


def test_slice_views_behave_like_the_list_they_replace():
    """BatchResult.audio / .pcm16 / .durations are views made on access (runtime.SliceViews): same indexing, iteration,
    slicing and emptiness rules as the list of per-utterance arrays, including empty utterances."""
    import numpy as np
    flat = np.arange(10, dtype=np.float32)
    v = runtime.SliceViews(flat, np.array([0, 3, 3, 10], dtype=np.int64))
    assert len(v) == 3 and bool(v)
    assert v[0].tolist() == [0, 1, 2] and v[1].size == 0 and v[-1].tolist() == list(range(3, 10))
    assert [a.size for a in v] == [3, 0, 7]
    assert [a.tolist() for a in v[0:2]] == [[0, 1, 2], []]
    assert np.array_equal(np.concatenate(v), flat)
    assert v[2].base is flat or v[2].base is flat.base          # views, not copies
    with pytest.raises(IndexError):
        v[3]
    assert len(runtime.SliceViews(flat[:0], [0])) == 0 and not runtime.SliceViews(flat[:0], [0])
