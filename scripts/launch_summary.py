#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (shares, not absolutes).
usage: launch_summary.py launches.csv"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); ui = h.index('Metric Unit')
agg = collections.defaultdict(list)
for r in rows[1:]:
    try: v = float(r[vi].replace(',', ''))
    except ValueError: continue
    v = v / 1000 if r[ui] == 'ns' else v * 1000 if r[ui] == 'ms' else v
    agg[re.sub(r'\(.*', '', r[ki]).replace('void ', '')].append(v)
tot = sum(sum(v) for v in agg.values())
print(f"# total {tot:.1f} us over {sum(len(v) for v in agg.values())} launches (cold-cache, serialised: compare SHARES)")
print(f"{'kernel':44s} {'launches':>8s} {'total_us':>10s} {'share':>7s} {'avg_us':>9s} {'max_us':>9s}")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:44s} {len(v):8d} {sum(v):10.1f} {100*sum(v)/tot:6.2f}% {sum(v)/len(v):9.2f} {max(v):9.2f}")
