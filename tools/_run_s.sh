pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py tests/test_gpu_octree.py tests/test_gpu_match.py -m gpu -x -q 2>&1 | tail -2
for c in C1 C2 C3 C4; do for t in 0 128 256; do echo "== $c octree threads $t"; ORBX_OCTREE_THREADS=$t python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr; done; done
