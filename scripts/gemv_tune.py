"""GEMV microbench over the per-layer shapes of a model (re-streamed from HBM, CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from theroundtaible_b200 import engine as E
T = int(sys.argv[1]) if len(sys.argv) > 1 else 3
shapes = {"qkv": (3072, 2048), "o": (2048, 2048), "gateup": (16384, 2048), "down": (2048, 8192), "lm_head": (128256, 2048)}
out = []
for name, (N, K) in shapes.items():
    ms = E.k_gemv_bench(N, K, T, iters=200)
    out.append(f"{name}:{ms*1e3:6.1f}us {N*K*2/ms/1e6:5.0f}GB/s")
print(f"MT={os.environ.get('RT_GEMV_MT','-')} NW={os.environ.get('RT_GEMV_NW','-')} PDL={'off' if os.environ.get('RT_NO_PDL') else 'on'} | " + "  ".join(out))
