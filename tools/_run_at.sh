mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2at_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2at_gpu_tests.log
tail -4 gpurun_out/r2at_gpu_tests.log
python __graft_entry__.py smoke 2>&1 | tail -1
