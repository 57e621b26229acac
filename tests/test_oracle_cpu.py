"""CPU tests: the oracle against its committed golden vectors, against what the reference
pins (graph signature, fixture), and against size-independent properties."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from common import S, ROOT, voice_files, oracle, fixture_ids
from oracle import onnx_reader, refrng
from oracle.vits_oracle import VitsOracle, derive_params

GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz")))


def test_fixture_matches_reference_fixture():
    # bench/fixtures/test_summary.json:8,16-18 and SmokeTests.swift:15-16
    assert S.FIXTURE_IDS == [1, 20, 0, 120, 0, 61, 0, 24, 0, 59, 0, 100, 0, 2]
    assert S.FIXTURE_SCALES == (0.667, 1.0, 0.8)
    assert S.DEFAULT_SEED == 1234


def test_graph_signature_facts():
    # what ONNXParsingTests.swift:22-35 pins for a Piper voice
    m = onnx_reader.load_model(voice_files("medium")[0])
    assert m.opset == 15
    assert m.inputs == ["input", "input_lengths", "scales"]
    assert m.outputs == ["output"]
    assert m.nodes[0].op_type == "Gather"
    assert "sid" in m.initializers
    assert "enc_p.encoder.attn_layers.0.conv_q.weight" in m.initializers
    p = derive_params(m.initializers, m.nodes)
    assert (p.hidden, p.n_heads, p.n_layers, p.window) == (192, 2, 6, 4)
    assert p.ups == [(16, 8), (16, 8), (8, 4)] and p.hop == 256
    assert p.rb_dilations == [[1, 2], [2, 6], [3, 12]]


def test_high_and_multispeaker_params():
    m = onnx_reader.load_model(voice_files("high")[0])
    p = derive_params(m.initializers, m.nodes)
    assert p.resblock == 1 and p.dec_c0 == 512 and p.hop == 256
    assert p.ups == [(16, 8), (16, 8), (4, 2), (4, 2)]
    m = onnx_reader.load_model(voice_files("medium-multispeaker")[0])
    assert m.inputs == ["input", "input_lengths", "scales", "sid"]
    assert derive_params(m.initializers, m.nodes).gin == 512


def test_config_json_schema():
    cfg = json.load(open(voice_files("medium")[1]))
    for k in ("audio", "inference", "phoneme_type", "phoneme_id_map", "num_symbols", "num_speakers"):
        assert k in cfg                       # required keys of PiperConfig.swift:30-39
    assert cfg["audio"]["sample_rate"] == 22050
    assert cfg["inference"] == {"noise_scale": 0.667, "length_scale": 1.0, "noise_w": 0.8}


def test_refrng_known_answers():
    # independent scalar restatement of elementwise.metal:132-163 with Python ints
    import math

    def xs(x):
        x ^= (x << 13) & 0xFFFFFFFF
        x ^= x >> 17
        x ^= (x << 5) & 0xFFFFFFFF
        return x

    got = refrng.random_normal_like(1000, 1234)
    for gid in (0, 1, 2, 17, 999):
        st = (1234 ^ ((gid * 747796405 + 2891336453) & 0xFFFFFFFF)) & 0xFFFFFFFF
        st = xs(st); u0 = st
        st = xs(st); u1 = st
        f0 = (np.float32(u0) + np.float32(1)) / np.float32(4294967296.0)
        f1 = (np.float32(u1) + np.float32(1)) / np.float32(4294967296.0)
        z = math.sqrt(-2.0 * math.log(float(f0))) * math.cos(6.28318530718 * float(f1))
        assert abs(got[gid] - z) < 2e-5 * max(1.0, abs(z))
    assert abs(got.mean()) < 0.15 and 0.85 < got.std() < 1.15


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_against_golden(path):
    g = np.load(path)
    name = os.path.basename(path).rsplit("_x", 1)[0]
    orc = oracle(name)
    ids = g["ids"].tolist()
    ns, ls, nw = (float(v) for v in g["scales"])
    sid = int(g["sid"])
    o = orc.infer(ids, ns, ls, nw, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main,
                  sid=None if sid < 0 else sid)
    assert np.array_equal(o["w_ceil"].astype(np.int32), g["w_ceil"])        # bit-exact integers
    assert o["y_len"] == int(g["y_len"])
    assert np.array_equal(o["frame_index"].astype(np.int32), g["frame_index"])
    assert np.abs(o["logw"] - g["logw"]).max() < 1e-4
    assert np.abs(o["audio"][:2048] - g["audio_head"]).max() < 1e-3
    assert np.abs(o["audio"] - g["audio"].astype(np.float32)).max() < 3e-3   # fp16-stored


def test_fp64_oracle_agrees_on_durations():
    # near-tie adjudication: the fixtures must not sit on a ceil() boundary
    for f in (1, 4):
        ids = fixture_ids(f)
        a = oracle("medium").infer(ids, 0.667, 1.0, 0.8, refrng.noise_dp(len(ids)), noise_main_fn=refrng.noise_main)
        b = oracle("medium", torch.float64).infer(ids, 0.667, 1.0, 0.8, refrng.noise_dp(len(ids)),
                                                  noise_main_fn=refrng.noise_main)
        assert np.array_equal(a["w_ceil"], b["w_ceil"])
        w = np.exp(b["logw"])
        assert np.abs(w - np.round(w)).min() > 1e-4
        assert np.abs(a["audio"] - b["audio"]).max() < 1e-3


def test_length_regulator_properties():
    rng = np.random.default_rng(0)
    for T in (1, 2, 17, 300):
        logw = torch.tensor(rng.normal(0.8, 0.7, T), dtype=torch.float32).view(1, 1, T)
        w_ceil, y_len, idx = VitsOracle.length_regulate(logw, 1.0)
        d = w_ceil.numpy().astype(int)
        assert y_len == max(1, d.sum())
        assert np.all(np.diff(idx.numpy()) >= 0)                      # monotonic path
        assert np.array_equal(np.bincount(idx.numpy(), minlength=T), d)   # each phoneme covers ceil(w) frames
    # degenerate: all durations zero -> one frame covered by nobody
    w_ceil, y_len, idx = VitsOracle.length_regulate(torch.full((1, 1, 3), -200.0), 1.0)
    assert y_len == 1 and idx.tolist() == [-1]


def test_batch_independence_of_oracle_inputs():
    # the synthetic batch generator is deterministic and structured BOS .. EOS with blanks
    from piper_b200.synth_voice import synthetic_utterances
    a = synthetic_utterances(8, 256)
    b = synthetic_utterances(8, 256)
    assert a == b
    for u in a:
        assert u[0] == 1 and u[-1] == 2 and 75 <= len(u) <= 120
        assert all(x == 0 for x in u[1::2])
