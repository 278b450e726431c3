python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
for fp in 4 2; do for mode in fused level; do ORBX_FAST_PAIRS=$fp python tools/sweep_step.py $mode; done; done
for sp in 1 2 4 8; do for ns in 2 4; do ORBX_DEVICE_SPLIT=$sp ORBX_DEVICE_STREAMS=$ns python tools/sweep_step.py level; done; done
ORBX_FAST_PAIRS=2 python tools/stage_times.py level
ORBX_FAST_PAIRS=4 python tools/stage_times.py level
