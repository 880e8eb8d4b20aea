"""Per-kernel timeline of a stretch of the decode step from the in-kernel trace (build.py --trace, RT_LIB=..., RT_TRACE_FILE=...):
for every launch -- first CTA start, last release from the PDL wait, last "inputs consumed" mark, last end -- in microseconds
from the first launch shown.  Kernel kinds (tag / 10): 0 store, 1 resid, 2 swiglu, 3 logits, 4 qkv, 5 attention (mma.sync),
6 merge, 7 tensor-parallel exchange (72 = every peer's flag / packet seen), 8 attention (tcgen05 decode), 9 push projection.
usage: trace_chain.py dump.bin [n_launches]"""
import sys
import numpy as np

rec = np.fromfile(sys.argv[1], dtype=np.dtype([("t", "<u8"), ("tag", "<i4"), ("cta", "<i4")]))
rec = rec[np.argsort(rec["t"], kind="stable")]
names = {0: "store", 1: "resid", 2: "swiglu", 3: "logits", 4: "qkv", 5: "attn", 6: "merge", 7: "tp-xchg", 8: "attn-tc5", 9: "push"}
kind, ev = rec["tag"] // 10, rec["tag"] % 10
launches = []
for k in np.unique(kind):
    m = kind == k
    t, e = rec["t"][m], ev[m]
    st = t[e == 0]
    if len(st) == 0:
        continue
    cuts = [0] + [i + 1 for i in range(len(st) - 1) if st[i + 1] - st[i] > 3000] + [len(st)]
    heads = [int(st[c]) for c in cuts[:-1]]
    for h in heads:
        launches.append({"kind": int(k), "t0": h, "ev": {}})
    idx = {h: i for i, h in enumerate(heads)}
    base = len(launches) - len(heads)
    pos = np.searchsorted(np.array(heads), t, side="right") - 1
    for ti, ei, pi in zip(t, e, pos):
        if pi >= 0:
            launches[base + pi]["ev"].setdefault(int(ei), []).append(int(ti))
launches.sort(key=lambda l: l["t0"])
n = int(sys.argv[2]) if len(sys.argv) > 2 else 16
mid = len(launches) // 2
while mid < len(launches) and launches[mid]["kind"] != 4:
    mid += 1
sel = launches[mid:mid + n]
t0 = sel[0]["t0"]
print(f"{'kernel':9s} {'ctas':>5s} {'start0':>8s} {'waitN':>8s} {'mark2N':>8s} {'endN':>8s} {'end-wait':>8s}   (us)")
for l in sel:
    e = l["ev"]
    f = lambda v, fn: (fn(e[v]) - t0) / 1e3 if v in e else float("nan")
    end = f(4, max) if 4 in e else f(6, max)
    print(f"{str(names.get(l['kind'], l['kind'])):9s} {len(e.get(0, [])):5d} {f(0, min):8.2f} {f(1, max):8.2f} {f(2, max):8.2f} {end:8.2f} {end - f(1, max):8.2f}")
