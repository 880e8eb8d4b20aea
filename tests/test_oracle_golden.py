"""Oracle vs the committed HF-transformers fixtures (tests/golden/make_golden.py) + oracle self-checks.

The reference holds no golden vector for generation, so these fixtures are what pins
the oracle.  CPU only."""
import os

import numpy as np
import pytest

from oracle import tokenizer as T
from oracle import weights as W
from oracle.archs import ARCHS, decode_weight_bytes, kv_bytes_per_token, param_count
from oracle.model import KVState, OracleModel, rope_inv_freq

GOLD = os.path.join(os.path.dirname(__file__), "golden")
TINY = ["tiny-llama", "tiny-llama-hd128", "tiny-qwen", "tiny-qwen-hd128"]


@pytest.mark.parametrize("name", TINY)
def test_oracle_matches_hf_logits(name):
    g = np.load(os.path.join(GOLD, f"hf_{name}.npz"))
    a = ARCHS[name]
    m = OracleModel(a, int(g["seed"]), max_pos=256)
    ref = g["logits"]
    scale = np.abs(ref).max()
    lg = m.forward(g["ids"], KVState(), round_kv=False, round_act=False).numpy()
    assert np.abs(lg - ref).max() <= 2e-5 * scale            # same algorithm, fp32 summation order only
    assert np.abs(rope_inv_freq(a) - g["inv_freq"]).max() <= 1.2e-7 * np.abs(g["inv_freq"]).max()
    # engine numerics (bf16 KV, bf16 matmul inputs) stay within the north-star tolerance of HF fp32
    lg2 = m.forward(g["ids"], KVState(), round_kv=True, round_act=True).numpy()
    assert np.abs(lg2 - ref).max() <= 1e-2 * scale


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-qwen-hd128"])
def test_oracle_incremental_equals_full(name):
    g = np.load(os.path.join(GOLD, f"hf_{name}.npz"))
    m = OracleModel(ARCHS[name], int(g["seed"]), max_pos=256)
    full = m.forward(g["ids"], KVState(), round_kv=False).numpy()
    kv, parts = KVState(), []
    for s in range(0, len(g["ids"]), 7):
        parts.append(m.forward(g["ids"][s:s + 7], kv, round_kv=False).numpy())
    assert np.abs(np.concatenate(parts) - full).max() <= 2e-5 * np.abs(full).max()


def test_oracle_greedy_is_deterministic_and_prefix_consistent():
    a = ARCHS["tiny-llama"]
    m = OracleModel(a, 7, max_pos=256)
    ids = np.random.default_rng(0).integers(0, a.vocab, 40).tolist()
    out1 = m.generate(ids, 12)
    out2, logits = m.generate(ids, 12, return_logits=True)
    assert out1 == out2 and len(logits) == 12
    # teacher forcing the generated ids reproduces the same arg-max chain
    kv = KVState()
    lg = m.forward(ids + out1[:-1], kv, round_act=True).numpy()
    assert [int(x) for x in lg[len(ids) - 1:].argmax(-1)] == out1


def test_weight_generator_known_answers():
    # pinned bit patterns: any change to the hash / scale / rounding breaks GPU<->CPU weight identity
    bits = W.gen_bits(0, W.tensor_id(W.Q, 0), 0, 8)
    assert bits.dtype == np.uint16 and len(bits) == 8
    again = W.gen_bits(0, W.tensor_id(W.Q, 0), 0, 8)
    assert (bits == again).all()
    assert (W.gen_bits(0, W.tensor_id(W.Q, 0), 3, 5) == bits[3:]).all()          # counter-based: random access
    assert (W.gen_bits(1, W.tensor_id(W.Q, 0), 0, 8) != bits).any()
    assert (W.gen_bits(0, W.tensor_id(W.K, 0), 0, 8) != bits).any()
    big = W.bf16_bits_to_f32(W.gen_bits(0, W.EMBED, 0, 1 << 18))
    assert abs(float(big.mean())) < 2e-4 and abs(float(big.std()) - 0.02) < 3e-4
    nrm = W.bf16_bits_to_f32(W.gen_bits(0, W.FINAL_NORM, 0, 4096, 0.02, 1.0))
    assert abs(float(nrm.mean()) - 1.0) < 2e-3
    # splitmix64 known answer (reference constants of the published generator)
    z = W._splitmix64(np.array([0], dtype=np.uint64))[0]
    assert int(z) == 0xE220A8397B1DCDAF


def test_bf16_rounding_is_rne():
    x = np.array([1.0, 1.00390625, 1.001953125, 1.005859375, -2.5, 3.0e-40], dtype=np.float32)
    got = W.round_bf16(x)
    assert got[0] == 1.0 and got[1] == np.float32(1.0)            # tie -> even (1.0)
    assert got[2] == 1.0 and got[3] == np.float32(1.0078125)     # tie -> even (1.0078125)
    assert got[4] == -2.5


def test_tokenizer_roundtrip_and_never_empty():
    for s in ["", "hello", "hé – ünï ✓ 🎯", "line1\nline2\t{\"consensus_score\": 9}"]:
        ids = T.encode_chat(s)
        assert ids[0] == T.BOS and ids[1] == T.USER and ids[-1] == T.ASSISTANT and len(ids) == len(s.encode()) + 3
        assert T.decode(ids[2:-1]) == s
    assert T.decode([5 + 0xC3]) == "<0xC3>"                     # truncated multi-byte sequence
    assert T.decode([5 + 0xED, 5 + 0xA0, 5 + 0x80]) == "<0xED><0xA0><0x80>"   # surrogate is ill-formed
    assert T.decode([70000]) == "<t70000>" and T.decode([2]) == "<eos>"


def test_arch_table_sizes_match_survey():
    # params, P_dec and KV/token of the named architectures
    exp = {"qwen2.5-0.5b": (0.494e9, 0.99e9, 12 << 10), "llama-3.2-1b": (1.236e9, 2.47e9, 32 << 10),
           "llama-3.1-8b": (8.03e9, 15.01e9, 128 << 10), "qwen2.5-32b": (32.76e9, 63.97e9, 256 << 10),
           "llama-3.1-70b": (70.55e9, 139.0e9, 320 << 10)}
    for name, (p, pdec, kv) in exp.items():
        a = ARCHS[name]
        assert abs(param_count(a) - p) / p < 2e-3, name
        assert abs(decode_weight_bytes(a) - pdec) / pdec < 5e-3, name
        assert kv_bytes_per_token(a) == kv, name
