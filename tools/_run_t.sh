set -x
mkdir -p gpurun_out
pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
echo "== C1 octree 64"; ORBX_OCTREE_THREADS=64 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2t_bench_C1.json 2> gpurun_out/r2t_bench_C1.err; echo "C1 rc=$?"
for c in C2 C3 C4 C5; do
timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2t_bench_$c.json 2> gpurun_out/r2t_bench_$c.err; echo "$c rc=$?"
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2t_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2t_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2t_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_describe|k_resize|k_octree|k_blur|k_search_area' -c 36 -o gpurun_out/r2t_full python tools/profile_step.py 256 1 > gpurun_out/r2t_ncu.log 2>&1
ncu -i gpurun_out/r2t_full.ncu-rep --page raw --csv > gpurun_out/r2t_full_raw.csv 2>/dev/null
for k in k_resize_tile k_octree; do ncu -i gpurun_out/r2t_full.ncu-rep --page source --csv -k regex:$k > gpurun_out/r2t_src_$k.csv 2>/dev/null; done
rm -f gpurun_out/r2t_full.ncu-rep
ls -la gpurun_out | tail -12
