set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py tests/test_gpu_octree.py -m gpu -x -q > gpurun_out/r2f_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2f_tests.log
tail -15 gpurun_out/r2f_tests.log
timeout 300 python tools/stage_times.py level 2>&1 | tail -2
ORBX_FAST_TMA=0 timeout 300 python tools/stage_times.py level 2>&1 | tail -2
timeout 300 python tools/sweep_step.py level 2>&1 | tail -1
ORBX_FAST_TMA=0 timeout 300 python tools/sweep_step.py level 2>&1 | tail -1
ORBX_FAST_TMA=0 timeout 900 python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -3
