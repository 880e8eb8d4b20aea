"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol include/rt_engine.h
declares, struct layouts agree with the ctypes mirror, the tokenizer equals the oracle's, and the product
fails LOUDLY without a GPU (no CPU fallback).  No compute calls."""
import ctypes as C
import os
import re

import pytest
import torch

from oracle import tokenizer as OT
from oracle.archs import ARCHS
from theroundtaible_b200 import engine as E

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NO_GPU = not torch.cuda.is_available()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "rt_engine.h")).read()
    declared = re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", hdr)
    assert len(declared) >= 19
    assert set(declared) == set(E.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layouts_match(lib):
    cfg = E.arch_config("llama-3.1-8b")
    assert cfg.struct_size == C.sizeof(E.RtConfig)
    assert (cfg.layers, cfg.hidden, cfg.n_q, cfg.n_kv, cfg.head_dim, cfg.ffn, cfg.vocab) == (32, 4096, 32, 8, 128, 14336, 128256)
    # the header must be valid C, and its struct sizes must equal the ctypes mirrors
    import subprocess, tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "s.c")
        open(src, "w").write('#include <stdio.h>\n#include "rt_engine.h"\nint main(void){printf("%zu %zu %zu %zu\\n",'
                             'sizeof(rt_config),sizeof(rt_sampling),sizeof(rt_result),sizeof(rt_model_info));return 0;}\n')
        exe = os.path.join(d, "s")
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(E.RtConfig), C.sizeof(E.RtSampling), C.sizeof(E.RtResult), C.sizeof(E.RtModelInfo)]


def test_arch_table_equals_oracle_table(lib):
    for name, a in ARCHS.items():
        c = E.arch_config(name)
        assert (c.family, c.layers, c.hidden, c.n_q, c.n_kv, c.head_dim, c.ffn, c.vocab, c.tied, c.qkv_bias) == (
            0 if a.family == "llama" else 1, a.layers, a.hidden, a.n_q, a.n_kv, a.head_dim, a.ffn, a.vocab, int(a.tied),
            int(a.qkv_bias)), name
        assert c.rope_theta == a.rope_theta and abs(c.rms_eps - a.rms_eps) < 1e-12 and c.rope_factor == a.rope_factor
        assert c.rope_orig_ctx == a.rope_orig_ctx, name
    assert E.arch_config("llama-3.1-70b@L4").layers == 4
    with pytest.raises(E.EngineError, match="not found"):
        E.arch_config("gpt-5")


def test_tokenizer_equals_oracle(lib):
    for s in ["", "Je naam is: Claude", "hé – ünï ✓ 🎯\n{\"a\": 1}"]:
        ids = E.tokenize_chat(s)
        assert ids == OT.encode_chat(s)
        assert E.detokenize(ids[2:-1]) == s
    weird = [5 + 0xC3, 300, 5 + 0x41, 2, 5 + 0xF0, 5 + 0x9F, 5 + 0x8E, 5 + 0xAF, 5 + 0xED, 5 + 0xA0, 5 + 0x80, 151000]
    assert E.detokenize(weird) == OT.decode(weird)


@pytest.mark.skipif(not NO_GPU, reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(lib):
    with pytest.raises(E.EngineError, match="no CUDA device"):
        E.Engine("tiny-llama")
    import numpy as np
    with pytest.raises(E.EngineError, match="no CUDA device"):
        E.k_gemv(np.zeros((16, 32), np.uint16), np.zeros((1, 32), np.uint16))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "theroundtaible_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, fn), encoding="utf-8").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn
