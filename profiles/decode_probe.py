"""Short decode run for ncu (B200_PROFILING.md recipe): config-2 shape, few new tokens.
RT_PROFILE_DECODE=1 brackets the decode loop with cudaProfilerStart/Stop (ncu --profile-from-start off)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from theroundtaible_b200 import engine as E

model = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
knights, shared, own, new = (3, 1024, 64, int(sys.argv[2]) if len(sys.argv) > 2 else 6)
flags = E.RT_FLAG_NO_GRAPH if os.environ.get("RT_NO_GRAPH") else 0
eng = E.Engine(model, seed=0, max_batch=knights, max_ctx=2048, flags=flags)
rng = np.random.default_rng(1236)
common = rng.integers(0, eng.vocab, shared).tolist()
prompts = [common + rng.integers(0, eng.vocab, own).tolist() for _ in range(knights)]
g = eng.generate_ids(prompts, max_new_tokens=new)
print("decode_ms", g[0].decode_ms, "steps", g[0].decode_steps, "launches", g[0].gpu_launches)
eng.close()
