"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel. usage: launch_summary.py <csv>"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr, agg = None, collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if r[0] == 'ID': hdr = r; continue
    if hdr is None: continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum': continue
    v = float(d['Metric Value'].replace(',', ''))
    v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3, 'second': 1e3}.get(d['Metric Unit'], 1e-6)
    k = d['Kernel Name'].split('(')[0]
    agg[k][0] += 1; agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':52s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k[:52]:52s} {v[0]:8d} {v[1]:10.3f} {100 * v[1] / tot:6.1f}%")
print(f"{'total':52s} {sum(v[0] for v in agg.values()):8d} {tot:10.3f}")
