mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_octree.py tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/r2o_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2o_gpu_tests.log
tail -5 gpurun_out/r2o_gpu_tests.log
for t in 256 1024; do
  echo "== C5 octree threads=$t" 
  ORBX_OCTREE_THREADS=$t timeout 300 python bench.py --config C5 --steps 5 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"
done
for t in 256 1024; do
  echo "== C1 octree threads=$t" 
  ORBX_OCTREE_THREADS=$t timeout 300 python bench.py --config C1 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"
done
