"""Per-rank compute of a tensor-parallel configuration WITHOUT the exchanges: a single-GPU engine with the sharded
dimensions (heads, FFN rows and vocab divided by tp; hidden unchanged).  The difference to the real TP step time is what
the two exchanges per layer and the sampler exchange cost.   usage: tp_shard_probe.py <model> <tp> [knights shared own new]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from theroundtaible_b200 import engine as E
arg = lambda i, d: type(d)(sys.argv[i]) if len(sys.argv) > i else d
model, tp, knights, shared, own, new = arg(1, "qwen2.5-32b"), arg(2, 4), arg(3, 8), arg(4, 2048), arg(5, 128), arg(6, 48)
cfg = E.arch_config(model)
cfg.n_q //= tp; cfg.n_kv //= tp; cfg.ffn //= tp; cfg.vocab //= tp
eng = E.Engine(cfg, seed=0, max_batch=knights, max_ctx=shared + own + new + 64)
rng = np.random.default_rng(1)
common = rng.integers(0, eng.vocab, shared).tolist()
prompts = [common + rng.integers(0, eng.vocab, own).tolist() for _ in range(knights)]
g = eng.generate_ids(prompts, max_new_tokens=new)
print(model, "tp", tp, "shard alone: prefill_ms", round(g[0].prefill_ms, 1), "us/step", round(1e3 * g[0].decode_ms / max(1, g[0].decode_steps), 1),
      "weight bytes/step", eng.info.weight_bytes_per_step)
eng.close()
