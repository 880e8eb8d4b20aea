import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import weights as W
from theroundtaible_b200 import engine as E
rng = np.random.default_rng(0)
for (S, N, K) in [(128, 256, 64), (128, 256, 256), (100, 576, 448), (300, 512, 128), (1216, 3072, 2048)]:
    ah = rng.standard_normal((S, K)).astype(np.float16); a = ah.view(np.uint16)
    w = W.f32_to_bf16_bits((rng.standard_normal((N, K)) * 0.05).astype(np.float32))
    c = E.k_gemm_tc5(a, w)
    ref = ah.astype(np.float64) @ W.bf16_bits_to_f32(w).astype(np.float64).T
    err = np.abs(c - ref).max() / np.abs(ref).max()
    print(f"S={S} N={N} K={K}: rel err {err:.3e}", "OK" if err < 1e-4 else "MISMATCH", flush=True)
    if err >= 1e-4:
        bad = np.argwhere(np.abs(c - ref) > 1e-3 * np.abs(ref).max())
        print("  first bad", bad[:5].tolist(), "n bad", len(bad), "c", c[tuple(bad[0])], "ref", ref[tuple(bad[0])])
        # structure of the error: which rows / cols are wrong
        print("  bad rows uniq", np.unique(bad[:, 0])[:20].tolist(), "bad cols uniq", np.unique(bad[:, 1])[:40].tolist())
