"""Short decode run for ncu / timing probes (B200_PROFILING.md recipe).
usage: decode_probe.py [model] [new_tokens] [knights] [shared] [own]
RT_PROFILE_DECODE=1 brackets the decode loop with cudaProfilerStart/Stop (ncu --profile-from-start off)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from theroundtaible_b200 import engine as E

arg = lambda i, d: type(d)(sys.argv[i]) if len(sys.argv) > i else d
model, new, knights, shared, own = arg(1, "llama-3.2-1b"), arg(2, 6), arg(3, 3), arg(4, 1024), arg(5, 64)
flags = E.RT_FLAG_NO_GRAPH if os.environ.get("RT_NO_GRAPH") else 0
eng = E.Engine(model, seed=0, max_batch=knights, max_ctx=shared + own + new + 64, flags=flags)
rng = np.random.default_rng(1236)
common = rng.integers(0, eng.vocab, shared).tolist()
prompts = [common + rng.integers(0, eng.vocab, own).tolist() for _ in range(knights)]
g = eng.generate_ids(prompts, max_new_tokens=new)
print("prefill_ms", round(g[0].prefill_ms, 2), "decode_ms", g[0].decode_ms, "steps", g[0].decode_steps, "launches", g[0].gpu_launches,
      "us/step", round(1e3 * g[0].decode_ms / max(1, g[0].decode_steps), 1))
eng.close()
