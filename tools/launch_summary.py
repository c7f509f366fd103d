"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the LAST eval
(profiling aid).  usage: python tools/launch_summary.py file.csv [launches_per_eval]"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; out = []
for r in rows:
    if len(r) > 5 and r[0] == 'ID': hdr = r; continue
    if hdr and len(r) == len(hdr): out.append(dict(zip(hdr, r)))
ks = [(re.sub(r'\(.*', '', d['Kernel Name']).replace('void mb200::', '').replace('mb200::', ''), d['Grid Size'],
       float(d['Metric Value'].replace(',', '')) / 1000) for d in out if 'mb200' in d['Kernel Name']]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
ev = ks[-n:]
tot = sum(k[2] for k in ev)
agg = collections.OrderedDict()
for k in ev:
    a = agg.setdefault((k[0], k[1]), [0, 0.0]); a[0] += 1; a[1] += k[2]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[0]:40s} {k[1]:14s} n={v[0]:2d} sum={v[1]:7.1f}us avg={v[1]/v[0]:6.1f} {100*v[1]/tot:5.1f}%")
print(len(ks), "kernels in file; last", n, "sum", round(tot, 1), "us")
if len(sys.argv) > 3:
    for k in ev: print(k)
