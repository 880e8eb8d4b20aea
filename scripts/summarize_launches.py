import csv, collections, re, sys
path = sys.argv[1]; steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
hdr = rows[hi]; data = rows[hi + 1:]
kn, mn, mv, gs = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0]); total = 0
for r in data:
    if len(r) <= mv: continue
    name = re.sub(r'\(GemvArgs\)|\(AttnArgs\)|void ', '', r[kn]); key = name + ' ' + r[gs]
    v = float(r[mv].replace(',', ''))
    if r[mn] == 'gpu__time_duration.sum':
        unit = rows[hi+1][hdr.index('Metric Unit')] if False else r[hdr.index('Metric Unit')]
        t = v / 1000.0 if unit in ('ns', 'nsecond') else v
        agg[key][0] += 1; agg[key][1] += t; total += t
    elif r[mn] == 'dram__bytes_read.sum':
        unit = r[hdr.index('Metric Unit')]
        mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)
        agg[key][2] += v * mult
    elif r[mn] == 'launch__registers_per_thread':
        agg[key][3] = int(v)
print(f"total {total:.0f} us over {steps} steps = {total/steps:.0f} us/step")
for k, (n, t, b, regs) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{t/total*100:5.1f}%  n/step={n/steps:5.1f} avg={t/n:7.2f}us  dram_rd={b/n/1e6:7.2f}MB  regs={regs:3d}  {k}")
