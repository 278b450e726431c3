mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2an_no_alloc_alone.txt 2>&1; echo "rc=$?" >> gpurun_out/r2an_no_alloc_alone.txt
tail -15 gpurun_out/r2an_no_alloc_alone.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2an_gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r2an_gpu_tests.log
tail -5 gpurun_out/r2an_gpu_tests.log
