pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']))"; }
for c in C1 C5; do for k in 0 1; do for sp in 2 4; do echo "== $c STAGGER=$k SPLIT=$sp"; ORBX_STAGGER=$k ORBX_DEVICE_SPLIT=$sp python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr; done; done; done
