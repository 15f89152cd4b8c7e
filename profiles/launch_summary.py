#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel and per BD step."""
import csv, collections, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = None
seq = []
for r in rows:
    if hdr is None:
        if 'Kernel Name' in r:
            hdr = r
        continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    v = float(d['Metric Value'].replace(',', ''))
    u = d['Metric Unit']
    v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
    seq.append((d['Kernel Name'].split('(')[0].replace('void ', '').replace('bd::', '')[:28], v))
agg = collections.defaultdict(lambda: [0, 0.0, []])
for k, v in seq:
    agg[k][0] += 1
    agg[k][1] += v
    agg[k][2].append(v)
tot = sum(v[1] for v in agg.values())
steps = max([v[0] for k, v in agg.items() if 'k_step_fused' in k] or [1])
print(f"launches {len(seq)}, BD steps {steps}, device time {tot/steps:.1f} us per step")
for k, (n, t, l) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:30s} n={n:4d} avg={t/n:9.2f} min={min(l):9.2f} max={max(l):9.2f} us/step={t/steps:8.2f} share={t/tot:6.1%}")
