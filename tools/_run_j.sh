set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2j_gpu_tests.log
tail -5 gpurun_out/r2j_gpu_tests.log
python tools/profile_step.py 256 1 > gpurun_out/r2j_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_blur_levels|k_describe_lite|k_search_area|k_octree' -c 8 -o gpurun_out/r2j_full python tools/profile_step.py 256 1 > gpurun_out/r2j_ncu.log 2>&1
ncu -i gpurun_out/r2j_full.ncu-rep --page raw --csv > gpurun_out/r2j_full_raw.csv 2>/dev/null
for k in k_blur_levels k_describe_lite k_search_area k_octree; do ncu -i gpurun_out/r2j_full.ncu-rep --page source --csv -k regex:$k > gpurun_out/r2j_src_$k.csv 2>/dev/null; done
rm -f gpurun_out/r2j_full.ncu-rep
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_step.py > gpurun_out/r2j_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r2j_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_step.py > gpurun_out/r2j_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/r2j_racecheck.log
tail -5 gpurun_out/r2j_memcheck.log gpurun_out/r2j_racecheck.log
