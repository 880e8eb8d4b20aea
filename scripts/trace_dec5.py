"""Timeline of ONE CTA of attn_decode_tc5_kernel from the in-kernel trace (build.py --trace): events of the TMA lane (82 =
K/V tile issued), the MMA lane (87 = K/V tile landed, 88 = P ready -> PV issued), one softmax warp (83 = scores ready,
84 = P written) and the CTA (80 start, 85 set-up done, 81 released from the PDL wait, 86 end), microseconds from the start."""
import sys
import numpy as np
rec = np.fromfile(sys.argv[1], dtype=np.dtype([("t", "<u8"), ("tag", "<i4"), ("cta", "<i4")]))
rec = rec[(rec["tag"] >= 80) & (rec["tag"] < 90)]
rec = rec[np.argsort(rec["t"], kind="stable")]
starts = rec[(rec["tag"] == 80)]
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 100
mine = starts[starts["cta"] == cta]
t0 = mine["t"][len(mine) // 2]                      # a launch in the middle of the run
sel = rec[(rec["cta"] == cta) & (rec["t"] >= t0)]
end = sel[sel["tag"] == 86]["t"][0]
sel = sel[sel["t"] <= end]
names = {80: "start", 85: "setup done", 81: "pdl released", 82: "tma issue", 87: "kv landed", 88: "P ready", 83: "S ready", 84: "P written", 86: "end"}
for r in sel:
    print(f"{(r['t'] - t0) / 1e3:8.2f} us  {names.get(int(r['tag']), r['tag'])}")
# all CTAs of that launch: duration spread
lau = rec[(rec["t"] >= t0 - 3000) & (rec["t"] <= end + 30000)]
s80 = {int(c): int(t) for t, c in zip(lau[lau["tag"] == 80]["t"], lau[lau["tag"] == 80]["cta"])}
e86 = {}
for t, c in zip(lau[lau["tag"] == 86]["t"], lau[lau["tag"] == 86]["cta"]):
    e86.setdefault(int(c), int(t))
d = [(e86[c] - s80[c]) / 1e3 for c in s80 if c in e86 and e86[c] > s80[c]]
print("CTAs", len(d), "duration us: min %.1f median %.1f max %.1f" % (min(d), float(np.median(d)), max(d)),
      "; first start -> last end %.1f us" % ((max(e86.values()) - min(s80.values())) / 1e3))
