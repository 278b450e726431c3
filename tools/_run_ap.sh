timeout 900 python -m pytest tests/test_gpu_match.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -2
pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
for c in C1 C3 C5; do echo "== $c"; python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr; done
