mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_full_batch.py -m gpu -x -q > gpurun_out/r2ar_full_batch.txt 2>&1; echo "rc=$?" >> gpurun_out/r2ar_full_batch.txt
tail -25 gpurun_out/r2ar_full_batch.txt
