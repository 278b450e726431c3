mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -5
pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
for k in 1 4; do echo "== KPW=$k"; ORBX_DESC_KPW=$k python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr; done
for k in 1 4; do echo "== C4 KPW=$k"; ORBX_DESC_KPW=$k python bench.py --config C4 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr; done
