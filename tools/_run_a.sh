set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_gpu_tests.log
tail -3 gpurun_out/r2a_gpu_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; cat gpurun_out/r2a_bench.json
python tools/stage_times.py level > gpurun_out/r2a_stage_level.txt 2>&1
python tools/stage_times.py fused > gpurun_out/r2a_stage_fused.txt 2>&1
cat gpurun_out/r2a_stage_*.txt
python tools/sweep_step.py auto
python tools/sweep_step.py level
python tools/sweep_step.py fused
python tools/h2d_rate.py > gpurun_out/r2a_h2d_rate_1rank.txt 2>&1
python tools/profile_step.py 256 2 > gpurun_out/r2a_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2a_launches.csv python tools/profile_step.py 256 2 > gpurun_out/r2a_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_describe|k_resize|k_blur|k_octree|k_order' -c 40 -o gpurun_out/r2a_full python tools/profile_step.py 256 1 > gpurun_out/r2a_ncu.log 2>&1
ncu -i gpurun_out/r2a_full.ncu-rep --page raw --csv > gpurun_out/r2a_full_raw.csv 2>/dev/null
for k in k_fast k_describe k_resize_tile k_blur_levels k_octree; do ncu -i gpurun_out/r2a_full.ncu-rep --page source --csv -k regex:$k > gpurun_out/r2a_src_$k.csv 2>/dev/null; done
ls -la gpurun_out | tail -20
