mkdir -p gpurun_out
for side in 0 1; do for cfgs in "2 2" "4 4" "4 2"; do set -- $cfgs
  echo "== C1 blur_side=$side split=$1 streams=$2"
  ORBX_BLUR_SIDE=$side ORBX_DEVICE_SPLIT=$1 ORBX_DEVICE_STREAMS=$2 python tools/sweep_step.py auto 2>&1 | tail -1
done; done
ORBX_BLUR_SIDE=1 timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -2
