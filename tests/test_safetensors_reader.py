"""CPU checks of the checkpoint reader (theroundtaible_b200/csrc/safetensors.cpp, compiled here with g++ next to a
small dump program) and of the Hugging Face config mapping.  No GPU, no engine calls."""
from __future__ import annotations

import json
import os
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "theroundtaible_b200", "csrc")


@pytest.fixture(scope="module")
def st_dump(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("st") / "st_dump")
    subprocess.run(["g++", "-O1", "-std=c++17", "-I", CSRC, os.path.join(ROOT, "tests", "csrc", "st_dump.cpp"),
                    os.path.join(CSRC, "safetensors.cpp"), "-o", exe], check=True)
    return exe


def _fnv(b: bytes) -> str:
    h = 1469598103934665603
    for x in b:
        h = ((h ^ x) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


def _run(exe, *args):
    r = subprocess.run([exe, *args], capture_output=True, text=True)
    return r.returncode, r.stdout.strip().splitlines()


def test_reader_single_file_and_directory(st_dump, tmp_path):
    from safetensors.torch import save_file
    torch.manual_seed(0)
    a = torch.randn(6, 8).to(torch.bfloat16)
    b = torch.randn(5)
    c = torch.randn(3, 4).to(torch.float16)
    save_file({"model.a.weight": a, "b": b}, str(tmp_path / "m-1.safetensors"), metadata={"format": "pt", "note": "x\"y"})
    save_file({"weird \u00e9 name": c}, str(tmp_path / "m-2.safetensors"))
    rc, out = _run(st_dump, str(tmp_path / "m-1.safetensors"), "model.a.weight", "b", "nope")
    assert rc == 0
    assert out[0] == "COUNT 2"
    assert out[1] == f"model.a.weight BF16 6x8 {_fnv(a.view(torch.int16).numpy().tobytes())}"
    assert out[2] == f"b F32 5 {_fnv(b.numpy().tobytes())}"
    assert out[3] == "nope MISSING"
    rc, out = _run(st_dump, str(tmp_path), "model.a.weight", "weird \u00e9 name")
    assert rc == 0 and out[0] == "COUNT 3"
    assert out[2] == f"weird \u00e9 name F16 3x4 {_fnv(c.view(torch.int16).numpy().tobytes())}"


def test_reader_errors(st_dump, tmp_path):
    rc, out = _run(st_dump, str(tmp_path / "absent.safetensors"))
    assert rc == 1 and "not found" in out[0]
    rc, out = _run(st_dump, str(tmp_path))
    assert rc == 1 and "no *.safetensors file" in out[0]
    (tmp_path / "short.safetensors").write_bytes(b"abc")
    rc, out = _run(st_dump, str(tmp_path / "short.safetensors"))
    assert rc == 1 and "too short" in out[0]
    (tmp_path / "long.safetensors").write_bytes((1 << 40).to_bytes(8, "little") + b"{}")
    rc, out = _run(st_dump, str(tmp_path / "long.safetensors"))
    assert rc == 1 and "malformed" in out[0]
    hdr = json.dumps({"t": {"dtype": "BF16", "shape": [4, 4], "data_offsets": [0, 64]}}).encode()
    (tmp_path / "oob.safetensors").write_bytes(len(hdr).to_bytes(8, "little") + hdr + b"\0" * 32)
    rc, out = _run(st_dump, str(tmp_path / "oob.safetensors"))
    assert rc == 1 and "outside the file" in out[0]
    hdr = json.dumps({"t": {"dtype": "BF16", "shape": [4, 4], "data_offsets": [0, 16]}}).encode()
    (tmp_path / "dis.safetensors").write_bytes(len(hdr).to_bytes(8, "little") + hdr + b"\0" * 32)
    rc, out = _run(st_dump, str(tmp_path / "dis.safetensors"))
    assert rc == 1 and "disagree" in out[0]


def test_config_from_hf_refuses_what_has_no_kernel(tmp_path):
    from theroundtaible_b200.engine import EngineError, config_from_hf
    base = dict(model_type="llama", num_hidden_layers=2, hidden_size=256, num_attention_heads=4, num_key_value_heads=2,
                intermediate_size=512, vocab_size=512, rms_norm_eps=1e-5)
    cfg = config_from_hf(base)
    assert (cfg.head_dim, cfg.n_kv, cfg.tied, cfg.qkv_bias) == (64, 2, 0, 0)
    with pytest.raises(EngineError):
        config_from_hf({**base, "model_type": "mistral"})
    with pytest.raises(EngineError):
        config_from_hf({**base, "attention_bias": True})
    with pytest.raises(EngineError):
        config_from_hf({**base, "rope_scaling": {"rope_type": "yarn", "factor": 4.0}})
    with open(os.path.join(tmp_path, "config.json"), "w") as f:
        json.dump({**base, "model_type": "qwen2", "tie_word_embeddings": True, "rope_theta": 1000000.0}, f)
    q = config_from_hf(str(tmp_path))
    assert (q.family, q.qkv_bias, q.tied) == (1, 1, 1) and abs(q.rope_theta - 1e6) < 1


def test_config_from_hf_reads_a_saved_transformers_config(tmp_path):
    from transformers import LlamaConfig
    from theroundtaible_b200.engine import config_from_hf
    rope = dict(rope_type="llama3", rope_theta=500000.0, factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0,
                original_max_position_embeddings=8192)
    LlamaConfig(vocab_size=128256, hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=32,
                num_key_value_heads=8, head_dim=64, rms_norm_eps=1e-5, tie_word_embeddings=True,
                rope_parameters=rope).save_pretrained(tmp_path)
    got = config_from_hf(str(tmp_path))
    from theroundtaible_b200.engine import RtConfig
    import ctypes as C
    # the same numbers rt_arch_by_name("llama-3.2-1b") holds
    assert (got.layers, got.hidden, got.n_q, got.n_kv, got.head_dim, got.ffn, got.vocab, got.tied) == \
        (16, 2048, 32, 8, 64, 8192, 128256, 1)
    assert (got.rope_factor, got.rope_low_freq, got.rope_high_freq, got.rope_orig_ctx) == (32.0, 1.0, 4.0, 8192)
    assert abs(got.rope_theta - 500000.0) < 1 and got.struct_size == C.sizeof(RtConfig)
