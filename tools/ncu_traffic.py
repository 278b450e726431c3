"""Per-launch DRAM traffic of each kernel from an `ncu --page raw --csv` dump -> JSON
(kernel -> {grid, dram_read_bytes, dram_write_bytes, time_us}); bench.py reads the committed file to
fill roofline.traffic.  Usage: ncu_traffic.py raw.csv batch out.json [steps captured, default 1]"""
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}


def val(r, name):
    v = float(r[idx[name]].replace(",", ""))
    u = units[idx[name]].lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u, 1)
    return v * mult


out = {"batch": int(sys.argv[2]), "steps_captured": int(sys.argv[4]) if len(sys.argv) > 4 else 1, "source": sys.argv[1],
       "kernels": {}}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0]
    out["kernels"].setdefault(name, []).append({
        "grid": r[idx["Grid Size"]], "dram_read_bytes": val(r, "dram__bytes_read.sum"),
        "dram_write_bytes": val(r, "dram__bytes_write.sum"), "time_us": val(r, "gpu__time_duration.sum")})
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps({k: len(v) for k, v in out["kernels"].items()}))
