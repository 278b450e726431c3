set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2z_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2z_gpu_tests.log
tail -4 gpurun_out/r2z_gpu_tests.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2z_bench_C1.json 2> gpurun_out/r2z_bench_C1.err; echo "C1 rc=$?"; tail -3 gpurun_out/r2z_bench_C1.err
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2z_bench_C1_reference.json 2> gpurun_out/r2z_ref.err; echo "ref rc=$?"; cat gpurun_out/r2z_bench_C1_reference.json
for c in C2 C3 C4 C5; do
timeout 600 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2z_bench_$c.json 2> gpurun_out/r2z_bench_$c.err; echo "$c rc=$?"; tail -2 gpurun_out/r2z_bench_$c.err
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2z_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2z_launches_bench_C1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2z_ncu_list.log 2>&1
python tools/profile_step.py 512 1 > gpurun_out/r2z_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_describe|k_resize|k_octree|k_blur|k_search_area' -c 30 -o gpurun_out/r2z_full python tools/profile_step.py 512 1 > gpurun_out/r2z_ncu.log 2>&1
ncu -i gpurun_out/r2z_full.ncu-rep --page raw --csv > gpurun_out/r2z_full_raw.csv 2>/dev/null
python __graft_entry__.py smoke > gpurun_out/r2z_smoke.log 2>&1; tail -2 gpurun_out/r2z_smoke.log
ls -la gpurun_out | tail -16
