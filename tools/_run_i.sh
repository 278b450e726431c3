set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2i_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2i_gpu_tests.log
tail -4 gpurun_out/r2i_gpu_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2i_bench_C1.json 2> gpurun_out/r2i_bench_C1.err; echo "C1 rc=$?"; tail -3 gpurun_out/r2i_bench_C1.err; cat gpurun_out/r2i_bench_C1.json
python tools/stage_times.py level > gpurun_out/r2i_stage_level.txt 2>&1
python tools/stage_times.py fused > gpurun_out/r2i_stage_fused.txt 2>&1
cat gpurun_out/r2i_stage_*.txt
python tools/profile_step.py 512 2 > gpurun_out/r2i_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches.csv python tools/profile_step.py 512 2 > gpurun_out/r2i_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_describe|k_resize|k_octree|k_order' -c 30 -o gpurun_out/r2i_full python tools/profile_step.py 256 1 > gpurun_out/r2i_ncu.log 2>&1
ncu -i gpurun_out/r2i_full.ncu-rep --page raw --csv > gpurun_out/r2i_full_raw.csv 2>/dev/null
for k in k_fast_strip k_describe k_resize_tile k_octree; do ncu -i gpurun_out/r2i_full.ncu-rep --page source --csv -k regex:$k > gpurun_out/r2i_src_$k.csv 2>/dev/null; done
ls -la gpurun_out | tail -20
