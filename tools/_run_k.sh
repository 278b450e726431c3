mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rig.py tests/test_frame_geometry.py tests/test_vocabulary_text.py -m gpu -x -q > gpurun_out/r2k_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2k_gpu_tests.log
tail -30 gpurun_out/r2k_gpu_tests.log
