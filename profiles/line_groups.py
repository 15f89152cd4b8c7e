#!/usr/bin/env python
"""Group the per-CUDA-line instruction counts of an ncu source page (cuda,sass view) by function.

    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > X_cs.csv
    python profiles/line_groups.py X_cs.csv path/to/bd_step.cuh path/to/bd_device.cuh
"""
import csv, re, sys, collections

def functions(path):
    """[(first_line, name)] of top-level functions/kernels, by a simple signature scan."""
    out = []
    for n, line in enumerate(open(path), 1):
        m = re.match(r'^(?:__device__|__global__|static|inline|__host__).*?\b([A-Za-z_0-9]+)\s*\(', line)
        if m and not line.startswith(' '):
            out.append((n, m.group(1)))
    return out

def main():
    rows = list(csv.reader(open(sys.argv[1])))
    funcs = {p.split('/')[-1]: functions(p) for p in sys.argv[2:]}
    agg = collections.defaultdict(lambda: [0, 0, 0])
    f = None
    for r in rows:
        if len(r) == 2 and r[0] == 'File Path':
            f = r[1].split('/')[-1]
            continue
        if len(r) < 10 or r[0] in ('Line No', ''):
            continue
        try:
            line, inst, smp, thr = int(r[0]), int(r[7]), int(r[6]), int(r[8])
        except ValueError:
            continue
        name = '?'
        for first, fn in funcs.get(f, []):
            if first <= line:
                name = fn
        a = agg[(f, name)]
        a[0] += inst; a[1] += smp; a[2] += thr
    tot = sum(a[0] for a in agg.values())
    tots = sum(a[1] for a in agg.values())
    print(f"total warp instructions {tot}, stall samples {tots}")
    for (f, name), (inst, smp, thr) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{inst:11d} {inst/tot:6.1%}  samples {smp/max(tots,1):6.1%}  active {thr/max(inst,1):5.1f}  {f}:{name}")

main()
