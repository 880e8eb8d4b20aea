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
for fused in (0, 1):
    if fused: os.environ["RT_ATTN_BENCH_FUSED"] = "1"
    print("fused", fused, file=sys.stderr)
    for chunk in (64, 128, 256):
        run(3, 32, 8, 64, 1344, 1024, chunk)
    run(3, 32, 8, 64, 1344, 0, 128)
    run(1, 32, 8, 64, 128, 0, 128)
    run(5, 32, 8, 128, 8448, 8192, 128)
