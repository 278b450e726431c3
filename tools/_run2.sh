set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_gpu_tests.log
python tools/stage_times.py level > gpurun_out/r2b_stage_level.txt 2>&1
python tools/stage_times.py fused > gpurun_out/r2b_stage_fused.txt 2>&1
