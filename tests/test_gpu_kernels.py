"""Per-kernel parity on the B200, through the C-ABI hooks (rt_k_*), against numpy restatements.
Integer/bit work (weights) is bit-exact; floating point is within the stated tolerance."""
import numpy as np
import pytest

from oracle import weights as W
from oracle.archs import ARCHS
from oracle.model import make_state_dict
from theroundtaible_b200 import engine as E

pytestmark = pytest.mark.gpu


def _rand_bits(rng, shape, std=1.0):
    return W.f32_to_bf16_bits((rng.standard_normal(shape) * std).astype(np.float32))


def _rand_f16(rng, shape, std=1.0):
    """Projection inputs are fp16 in the engine (DESIGN.md 3 "numerics"); returns (bits, float64 values)."""
    h = (rng.standard_normal(shape) * std).astype(np.float16)
    return h.view(np.uint16), h.astype(np.float64)


@pytest.mark.parametrize("N,K,T", [(16, 32, 1), (128, 256, 3), (3072, 2048, 5), (2048, 8192, 8), (1152, 896, 7),
                                   (512, 4864, 12), (256, 2048, 16), (640, 1024, 32), (1024, 512, 4), (16384, 2048, 3), (9728, 896, 1)])
def test_gemv_matches_fp32_matmul(N, K, T):
    rng = np.random.default_rng(N * 7 + K + T)
    w, (x, xf) = _rand_bits(rng, (N, K), 0.05), _rand_f16(rng, (T, K))
    y = E.k_gemv(w, x)                               # fp16 inputs enter the MMAs as bf16 hi + lo: the products are exact
    ref = xf @ W.bf16_bits_to_f32(w).astype(np.float64).T
    tol = 2e-5 * np.abs(ref).max() + 1e-6            # fp32 accumulation, different summation order
    assert np.abs(y - ref).max() <= tol


@pytest.mark.parametrize("S,N,K", [(128, 256, 64), (100, 576, 448), (300, 512, 128), (1, 32, 64), (1216, 3072, 2048), (257, 4096, 8192)])
def test_tcgen05_prefill_gemm_matches_fp64_matmul(S, N, K):
    rng = np.random.default_rng(S + N + K)
    (a, af), w = _rand_f16(rng, (S, K)), _rand_bits(rng, (N, K), 0.05)
    c = E.k_gemm_tc5(a, w)                           # fp16 activations x the fp16 twin of the bf16 weights
    ref = af @ W.bf16_bits_to_f32(w).astype(np.float64).T
    assert np.abs(c - ref).max() <= 2e-5 * np.abs(ref).max() + 1e-6     # fp32 accumulation in TMEM


def test_gemv_is_deterministic_run_to_run():
    rng = np.random.default_rng(5)
    w, (x, _) = _rand_bits(rng, (2048, 2048), 0.05), _rand_f16(rng, (8, 2048))
    a, b = E.k_gemv(w, x), E.k_gemv(w, x)
    assert (a.view(np.uint32) == b.view(np.uint32)).all()        # fixed-order reduction: bit-identical


@pytest.mark.parametrize("T,H", [(1, 128), (3, 2048), (8, 896), (32, 8192)])
def test_rmsnorm(T, H):
    rng = np.random.default_rng(H + T)
    x = (rng.standard_normal((T, H)) * 3).astype(np.float32)
    w = _rand_bits(rng, (H,), 1.0)
    y = W.bf16_bits_to_f32(E.k_rmsnorm(x, w, 1e-5))
    xd = x.astype(np.float64)
    ref = xd / np.sqrt((xd * xd).mean(-1, keepdims=True) + 1e-5) * W.bf16_bits_to_f32(w)
    assert np.abs(y - ref).max() <= 2 ** -8 * np.abs(ref).max()       # one bf16 rounding of the result


def _attn_ref(q, k, v, lens):
    T, n_q, hd = q.shape
    n_kv = k.shape[2]
    G = n_q // n_kv
    out = np.zeros((T, n_q, hd))
    for t in range(T):
        L = lens[t]
        for h in range(n_q):
            s = (k[t, :L, h // G] @ q[t, h]) / np.sqrt(hd)
            p = np.exp(s - s.max())
            out[t, h] = (p / p.sum()) @ v[t, :L, h // G]
    return out


@pytest.mark.parametrize("hd,n_q,n_kv,lens,shared,chunk", [
    (64, 32, 8, [1], 0, 32), (64, 32, 8, [17, 33, 64], 0, 32), (128, 32, 8, [200, 129, 77, 256, 31], 0, 64),
    (128, 5, 1, [130, 131], 0, 16), (32, 14, 2, [95, 40, 1], 0, 32), (64, 32, 8, [100, 90, 120], 64, 32),
    (128, 32, 8, [300, 290, 310, 289, 333], 256, 64), (128, 64, 8, [80] * 8, 64, 32), (32, 14, 2, [50, 60], 48, 16),
    (128, 40, 8, [700, 650, 690, 701, 655, 699, 700, 651], 640, 128)])
def test_paged_attention_with_and_without_shared_prefix(hd, n_q, n_kv, lens, shared, chunk):
    rng = np.random.default_rng(hd + n_q + len(lens) + shared)
    T, ctx = len(lens), max(lens)
    qb = _rand_bits(rng, (T, n_q, hd))
    kb = _rand_bits(rng, (T, ctx, n_kv, hd))
    vb = _rand_bits(rng, (T, ctx, n_kv, hd))
    if shared:
        kb[:, :shared] = kb[0:1, :shared]
        vb[:, :shared] = vb[0:1, :shared]
    out = E.k_attention(qb, kb, vb, lens, shared=shared, chunk=chunk)
    ref = _attn_ref(W.bf16_bits_to_f32(qb).astype(np.float64), W.bf16_bits_to_f32(kb).astype(np.float64),
                    W.bf16_bits_to_f32(vb).astype(np.float64), lens)
    # P is rounded to bf16 before the PV product (as in every flash-attention kernel): 2^-9 relative per weight
    assert np.abs(out - ref).max() <= 4e-3 * np.abs(ref).max()
    if shared:   # the cascade phase must give what the plain per-knight read gives
        plain = E.k_attention(qb, kb, vb, lens, shared=0, chunk=chunk)
        assert np.abs(out - plain).max() <= 4e-3 * np.abs(ref).max()


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-qwen", "tiny-qwen-hd128"])
def test_gpu_weight_generator_is_bit_exact(name):
    a = ARCHS[name]
    sd = make_state_dict(a, 42)
    with E.Engine(name, seed=42, max_batch=1, max_ctx=64) as eng:
        hd = a.head_dim
        shapes = {0: ("embed", (a.vocab, a.hidden)), 2: ("final_norm", (a.hidden,))}
        if not a.tied:
            shapes[1] = ("lm_head", (a.vocab, a.hidden))
        for l in range(a.layers):
            b = 16 * (l + 1)
            shapes.update({b + 0: (f"l{l}.q", (a.n_q * hd, a.hidden)), b + 1: (f"l{l}.k", (a.n_kv * hd, a.hidden)),
                           b + 2: (f"l{l}.v", (a.n_kv * hd, a.hidden)), b + 3: (f"l{l}.o", (a.hidden, a.n_q * hd)),
                           b + 4: (f"l{l}.gate", (a.ffn, a.hidden)), b + 5: (f"l{l}.up", (a.ffn, a.hidden)),
                           b + 6: (f"l{l}.down", (a.hidden, a.ffn)), b + 7: (f"l{l}.in_norm", (a.hidden,)),
                           b + 8: (f"l{l}.post_norm", (a.hidden,))})
            if a.qkv_bias:
                shapes.update({b + 9: (f"l{l}.q_bias", (a.n_q * hd,)), b + 10: (f"l{l}.k_bias", (a.n_kv * hd,)),
                               b + 11: (f"l{l}.v_bias", (a.n_kv * hd,))})
        for tid, (key, shape) in shapes.items():
            got = eng.export_weight(tid, shape)
            assert (got == W.f32_to_bf16_bits(sd[key]).reshape(shape)).all(), key
