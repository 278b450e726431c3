mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_compat_cpp.py tests/test_tri_rig.py -m gpu -x -q > gpurun_out/r2m_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_gpu_tests.log
tail -40 gpurun_out/r2m_gpu_tests.log
