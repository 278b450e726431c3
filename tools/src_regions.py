"""Aggregates an `ncu --page source --print-source cuda,sass --csv` dump by source-line ranges.
usage: src_regions.py dump.csv file.cu name:a-b [name:a-b ...] [--px N]"""
import csv, collections, sys
args = sys.argv[1:]
px = None
if '--px' in args:
    i = args.index('--px'); px = float(args[i + 1]); del args[i:i + 2]
path, fname, regs = args[0], args[1], [(r.split(':')[0], tuple(map(int, r.split(':')[1].split('-')))) for r in args[2:]]
rows = list(csv.reader(open(path)))
h = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
hdr = rows[h]
iLine, iEx, iS = hdr.index('Line No'), hdr.index('Instructions Executed'), hdr.index('# Samples')
cur = None
agg = collections.defaultdict(lambda: [0, 0])
lines = collections.defaultdict(lambda: [0, 0, ''])
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if len(r) < len(hdr): continue
    if r[iLine].strip().isdigit() and r[iEx].strip().isdigit():
        ln = int(r[iLine]); key = cur
        if cur == fname:
            key = fname + ':other'
            for nm, (a, b) in regs:
                if a <= ln <= b: key = nm
            lines[ln][0] += int(r[iEx]); lines[ln][1] += int(r[iS]) if r[iS].strip().isdigit() else 0; lines[ln][2] = r[hdr.index('Source')]
        agg[key][0] += int(r[iEx]); agg[key][1] += int(r[iS]) if r[iS].strip().isdigit() else 0
tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values()) or 1
print("total warp-instructions", tot, "samples", ts)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    extra = f"  thread-instr/px {v[0] * 32 / px:6.2f}" if px else ""
    print(f"{k:22s} inst {v[0] / tot * 100:5.1f}%  samples {v[1] / ts * 100:5.1f}%{extra}")
print("-- hottest lines")
for ln, v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{ln:4d} inst {v[0] / tot * 100:5.1f}% smp {v[1] / ts * 100:5.1f}%  {v[2][:90]}")
