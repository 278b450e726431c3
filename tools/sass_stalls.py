"""Per-kernel stall-reason totals and opcode histogram from an `ncu --page source --csv` dump."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
hdr = rows[h]
idx = {k: j for j, k in enumerate(hdr)}
stalls = [k for k in hdr if k.startswith('stall_') and 'Not Issued' not in k]
tot = {k: 0 for k in stalls}
ops = {}
texec = 0
first = None
for r in rows[h + 1:]:
    if len(r) < len(hdr) or not r[idx['Instructions Executed']].isdigit():
        continue
    if first is None:
        first = r[idx['Source']]
    elif r[idx['Source']] == first and len(sys.argv) < 3:
        pass
    n = int(r[idx['Instructions Executed']])
    texec += n
    t = r[idx['Source']].split()
    op = t[1] if t[0].startswith('@') else t[0]
    op = op.split('.')[0]
    ops[op] = ops.get(op, 0) + n
    for k in stalls:
        v = r[idx[k]]
        tot[k] += int(v) if v.isdigit() else 0
s = sum(tot.values()) or 1
print("warp instructions executed:", texec)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]:
    print(f"  {k:26s} {v / s * 100:5.1f}%")
ALU = {'LOP3', 'SHF', 'PRMT', 'IADD3', 'VIADD', 'ISETP', 'SEL', 'VABSDIFF4', 'VIMNMX3', 'VIMNMX', 'IDP', 'LEA', 'POPC',
       'FLO', 'BREV', 'IABS', 'IMNMX', 'FMNMX', 'FSETP', 'FSEL', 'PLOP3', 'SGXT', 'BMSK', 'I2I', 'F2I', 'I2F', 'MOV',
       'R2P', 'P2R', 'VOTE', 'FCHK', 'CS2R', 'S2R', 'IADD', 'UIADD3', 'ULOP3', 'USHF', 'UMOV', 'VIADDMNMX'}
FMA = {'IMAD', 'FFMA', 'FMUL', 'FADD', 'HFMA2', 'HADD2', 'HMUL2', 'DFMA', 'DADD', 'DMUL'}
LSU = {'LDS', 'STS', 'LDG', 'STG', 'LD', 'ST', 'ATOMS', 'ATOMG', 'RED', 'SHFL', 'LDSM', 'LDC', 'LDGSTS', 'ATOM', 'LDL', 'STL'}
a = sum(v for k, v in ops.items() if k in ALU)
f = sum(v for k, v in ops.items() if k in FMA)
l = sum(v for k, v in ops.items() if k in LSU)
print(f"  alu-pipe {a / texec * 100:.1f}%  fma-pipe {f / texec * 100:.1f}%  lsu/mio {l / texec * 100:.1f}%  other {(texec - a - f - l) / texec * 100:.1f}%")
for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:18]:
    print(f"  {k:12s} {v / texec * 100:5.1f}%")
