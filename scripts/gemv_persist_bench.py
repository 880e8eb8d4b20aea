"""A/B of the persistent row-group loop of gemv_kernel on the multi-wave shapes (rt_k_gemv_bench, weights rotated
through > 3x L2).  usage: gemv_persist_bench.py ; env RT_GEMV_PERSIST=0/1, RT_GEMV_PMT=1/2/4 are read by the library."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from theroundtaible_b200 import engine as E
shapes = [("lm_head 1B", 128256, 2048, 3), ("lm_head 8B", 128256, 4096, 5), ("gate/up 8B", 28672, 4096, 5),
          ("gate/up 1B", 16384, 2048, 3), ("gate/up 32B/tp2", 27648, 5120, 8), ("qkv 8B", 6144, 4096, 5), ("down 8B", 4096, 14336, 5), ("down 1B", 2048, 8192, 3), ("down 32B/tp2", 5120, 13824, 8),
          ("down 70B/tp8", 8192, 3584, 8), ("o 70B/tp8", 8192, 1024, 8)]
for name, N, K, T in shapes:
    ms = E.k_gemv_bench(N, K, T, iters=100)
    print(f"{name:16s} N={N} K={K} T={T}: {ms*1e3:8.2f} us  {2*N*K/ms/1e6:8.1f} GB/s")
