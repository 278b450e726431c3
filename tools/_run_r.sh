mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py tests/test_gpu_octree.py tests/test_gpu_match.py -m gpu -x -q 2>&1 | tail -3
pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
echo "== default (dp2a resize, quad search_area)"; python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr
echo "== ORBX_SA_LANES=1"; ORBX_SA_LANES=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr
for t in 128 512; do echo "== octree threads $t"; ORBX_OCTREE_THREADS=$t python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr; done
echo "== C3"; python bench.py --config C3 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr
