mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
SECONDS=0; timeout 900 python bench.py --impl reference > gpurun_out/r2final_bench_reference.json 2> gpurun_out/r2final_ref.err; echo "ref rc=$? wall=${SECONDS}s"
SECONDS=0; timeout 900 python bench.py > gpurun_out/r2final_bench.json 2> gpurun_out/r2final_bench.err; echo "bench rc=$? wall=${SECONDS}s"
python -c "
import json
d=json.loads(open('gpurun_out/r2final_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('metric','value','unit','n_gpus','steps','warmup','ms_per_step','dtype','gpu_launches')}); print(d['e2e']); print(d['roofline']); print(d['cpu_baseline']['value'], d['cpu_baseline']['kind']); print(d['clocks'])
r=json.loads(open('gpurun_out/r2final_bench_reference.json').read().strip().splitlines()[-1]); print('ref', r['value'], r['impl'], r['e2e'])"
