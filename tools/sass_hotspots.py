"""Segments an `ncu --page source --csv` (SASS) dump by execution count to show hot regions."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
hdr = rows[h]
idx = {k: j for j, k in enumerate(hdr)}
data = []
for r in rows[h + 1:]:
    if len(r) < len(hdr) or not r[idx['Instructions Executed']].isdigit():
        continue
    data.append((r[idx['Source']].strip(), int(r[idx['Instructions Executed']]), int(r[idx['# Samples']]),
                 float(r[idx['Avg. Threads Executed']] or 0)))
tot = sum(d[1] for d in data)
tots = sum(d[2] for d in data) or 1
print("instructions", len(data), "executed", tot, "samples", tots)
seg, start = [], 0
for i in range(1, len(data) + 1):
    if i == len(data) or abs(data[i][1] - data[start][1]) > 0.25 * max(data[start][1], 1):
        seg.append((start, i))
        start = i
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
for a, b in seg:
    ie = sum(d[1] for d in data[a:b])
    sm = sum(d[2] for d in data[a:b])
    if ie / tot > thr or sm / tots > thr:
        ops = {}
        for d in data[a:b]:
            t = d[0].split()
            op = t[1] if t[0].startswith('@') else t[0]
            ops[op] = ops.get(op, 0) + 1
        top = sorted(ops.items(), key=lambda kv: -kv[1])[:7]
        print(f"[{a:5d},{b:5d}) n={b - a:4d} exec/instr={data[a][1]:9d} inst%={ie / tot * 100:5.1f} "
              f"smp%={sm / tots * 100:5.1f} thr={data[a][3]:.0f} {top}")
