import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from theroundtaible_b200 import engine as E
os.environ["RT_GEMM_BENCH_ITERS"] = "20"
for (S, N, K) in [(1216, 3072, 2048), (1216, 16384, 2048), (1216, 2048, 8192), (2048, 6144, 4096), (2048, 28672, 4096), (2048, 4096, 14336), (2048, 8192, 8192)]:
    a = np.zeros((S, K), np.uint16); w = np.zeros((N, K), np.uint16)
    E.k_gemm_tc5(a, w)
