"""Reads the in-kernel timeline dump (build with RT_NVCC_EXTRA=-DRT_TRACE, run with RT_TRACE_FILE=...) and prints,
for one decode step in the middle of the run, per kernel launch: first/last CTA start, first/last release from the
PDL wait, last weights-consumed, last end -- all in microseconds relative to the step's first event.
tag = kernel*10 + event; kernels: 0 store, 1 resid, 2 swiglu, 3 logits, 4 qkv, 5 attention, 6 merge;
events: 0 start, 1 after griddepcontrol.wait, 2 k-loop done, 3 reduced, 4 end."""
import sys
import numpy as np

rec = np.fromfile(sys.argv[1], dtype=np.dtype([("t", "<u8"), ("tag", "<i4"), ("cta", "<i4")]))
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 16
rec = rec[np.argsort(rec["t"], kind="stable")]
names = {0: "store", 1: "resid", 2: "swiglu", 3: "logits", 4: "qkv", 5: "attn", 6: "merge"}
# split into steps at the lm_head (logits) end events
ends = np.nonzero(rec["tag"] == 34)[0]
t_end = rec["t"][ends]
# a step boundary = gap between consecutive logits-end events larger than 50 us
cut = [0] + [int(ends[i]) + 1 for i in range(len(ends) - 1) if t_end[i + 1] - t_end[i] > 50000] + [len(rec)]
steps = [(cut[i], cut[i + 1]) for i in range(len(cut) - 1)]
print("records", len(rec), "steps", len(steps))
lo, hi = steps[len(steps) // 2]
st = rec[lo:hi]
# the step really starts at the first qkv start after the previous logits: approximate by the first record
t0 = st["t"].min()
print("step duration us", (st["t"].max() - t0) / 1e3)
# group launches: consecutive kernels in the chain have distinct kinds, so split by kind change on event 0 ordering
kind = st["tag"] // 10
ev = st["tag"] % 10
# launches identified by sequence: sort start events, segment when kind changes
starts = st[ev == 0]
order = []
prev = None
for k, t in zip(starts["tag"] // 10, starts["t"]):
    if k != prev:
        order.append([k, t])
        prev = k
print("launches seen in step:", len(order))
# per launch statistics: assign each record to the launch whose start time window contains it (same kind, nearest
# preceding launch start of that kind)
import bisect
by_kind = {}
for k, t in order:
    by_kind.setdefault(k, []).append(t)
rows = {}
for r in st:
    k, e = r["tag"] // 10, r["tag"] % 10
    lst = by_kind.get(k)
    if not lst:
        continue
    i = bisect.bisect_right(lst, r["t"]) - 1
    if e == 0:
        i = bisect.bisect_right(lst, r["t"]) - 1
    i = max(i, 0)
    rows.setdefault((k, i), {}).setdefault(e, []).append(r["t"])
out = []
for (k, i), evs in rows.items():
    f = lambda e, fn: (fn(evs[e]) - t0) / 1e3 if e in evs else float("nan")
    out.append((by_kind[k][i], names.get(k, k), len(evs.get(0, [])), f(0, min), f(0, max), f(1, min), f(1, max), f(2, max), f(4, max)))
out.sort()
print(f"{'kernel':8s} {'ctas':>5s} {'start0':>8s} {'startN':>8s} {'wait0':>8s} {'waitN':>8s} {'kdoneN':>8s} {'endN':>8s} {'dur':>6s}")
n_show = int(sys.argv[3]) if len(sys.argv) > 3 else 14
mid = len(out) // 2
for _, nm, n, s0, sN, w0, wN, kd, en in out[mid:mid + n_show]:
    print(f"{nm:8s} {n:5d} {s0:8.2f} {sN:8.2f} {w0:8.2f} {wN:8.2f} {kd:8.2f} {en:8.2f} {en - w0:6.2f}")
