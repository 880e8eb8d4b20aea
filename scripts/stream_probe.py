import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from theroundtaible_b200 import engine as E
N, K, T = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
print(E.k_gemv_bench(N, K, T, iters=int(sys.argv[4]) if len(sys.argv) > 4 else 5) * 1e3, "us")
