import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from theroundtaible_b200 import engine as E
T = 5
shapes = {"qkv": (6144, 4096), "o": (4096, 4096), "gateup": (28672, 4096), "down": (4096, 14336), "lm_head": (128256, 4096)}
out = []
for name, (N, K) in shapes.items():
    ms = E.k_gemv_bench(N, K, T, iters=100)
    out.append(f"{name}:{ms*1e3:6.1f}us {N*K*2/ms/1e6:5.0f}GB/s")
print(f"MT={os.environ.get('RT_GEMV_MT','-')} NW={os.environ.get('RT_GEMV_NW','-')} | " + "  ".join(out))
