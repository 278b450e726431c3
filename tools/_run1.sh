set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_gpu_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
python tools/h2d_rate.py > gpurun_out/r2a_h2d_rate_1rank.txt 2>&1
python tools/profile_step.py 256 2 level > gpurun_out/r2a_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_blur_levels|k_describe|k_resize' -c 12 -o gpurun_out/r2a_level python tools/profile_step.py 256 1 level > gpurun_out/r2a_ncu.log 2>&1
ncu -i gpurun_out/r2a_level.ncu-rep --page raw --csv > gpurun_out/r2a_level_raw.csv 2>/dev/null
ncu -i gpurun_out/r2a_level.ncu-rep --page source --csv -k regex:k_blur_levels > gpurun_out/r2a_src_blur.csv 2>/dev/null
ncu -i gpurun_out/r2a_level.ncu-rep --page source --csv -k regex:k_resize_tile > gpurun_out/r2a_src_resize.csv 2>/dev/null
ncu -i gpurun_out/r2a_level.ncu-rep --page source --csv -k regex:k_describe > gpurun_out/r2a_src_describe_lite.csv 2>/dev/null
rm -f gpurun_out/r2a_level.ncu-rep
ls -la gpurun_out | tail -12
