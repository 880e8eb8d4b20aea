import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import weights as W
from theroundtaible_b200 import engine as E
os.environ["RT_ATTN_BENCH_ITERS"] = "200"
rng = np.random.default_rng(0)
def run(T, n_q, n_kv, hd, L, shared, chunk):
    qb = W.f32_to_bf16_bits(rng.standard_normal((T, n_q, hd)).astype(np.float32))
    kb = W.f32_to_bf16_bits(rng.standard_normal((T, L, n_kv, hd)).astype(np.float32))
    vb = W.f32_to_bf16_bits(rng.standard_normal((T, L, n_kv, hd)).astype(np.float32))
    if shared:
        kb[:, :shared] = kb[0:1, :shared]; vb[:, :shared] = vb[0:1, :shared]
    E.k_attention(qb, kb, vb, [L] * T, shared=shared, chunk=chunk)
for chunk in (32, 64, 128):
    run(3, 32, 8, 64, 1344, 1024, chunk)
for chunk in (64, 96, 128, 192, 256):
    run(5, 32, 8, 128, 8448, 8192, chunk)
