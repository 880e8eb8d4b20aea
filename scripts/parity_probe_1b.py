"""Error budget of the full-depth Llama-3.2-1B logits: engine (both kernel paths) vs the oracle with and without the
bf16 activation rounding, and the two engine paths against each other (the rounding-noise floor)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle.archs import ARCHS
from oracle.model import KVState, OracleModel
from theroundtaible_b200 import engine as E
name = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
a = ARCHS[name]
oracle = OracleModel(a, 0, max_pos=256)
ids = np.random.default_rng(1234 + 2).integers(0, a.vocab, 72)
ref = oracle.forward(ids, KVState(), round_act=True).numpy()
ref32 = oracle.forward(ids, KVState(), round_act=False).numpy()
scale = np.abs(ref).max(-1)
with E.Engine(name, seed=0, max_batch=3, max_ctx=256) as eng:
    eng.seq_reset(); g1 = eng.step_logits(ids, mode=1)
    eng.seq_reset(); g0 = eng.step_logits(ids, mode=0)
def st(x, y, tag):
    e = np.abs(x - y).max(-1) / scale
    rms = np.sqrt(((x - y) ** 2).mean(-1)) / scale
    print(f"{tag:28s} max-over-vocab: median {np.median(e):.4f} max {e.max():.4f}   rms-over-vocab: median {np.median(rms):.5f}")
st(g1, ref, "prefill kernels vs oracle")
st(g0, ref, "decode kernels vs oracle")
st(g1, g0, "prefill vs decode kernels")
st(ref, ref32, "oracle bf16-act vs fp32-act")
st(g1, ref32, "prefill kernels vs fp32-act")
print("max|logit| median", np.median(scale))
