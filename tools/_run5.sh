python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
python tools/stage_times.py level
python tools/sweep_step.py level
