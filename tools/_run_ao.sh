mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2ao_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ao_gpu_tests.log
tail -4 gpurun_out/r2ao_gpu_tests.log
timeout 900 python -m pytest tests/test_ref_frame.py tests/test_ref_dbow.py tests/test_ref_matcher.py -q 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2ao_bench_C1.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2ao_bench_C1.json').read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['roofline'])"
