timeout 900 python -m pytest tests/test_gpu_compat_cpp.py -m gpu -x -q -k orbmatcher_class 2>&1 | tail -25
