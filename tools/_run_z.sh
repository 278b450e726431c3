timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_compat_cpp.py -m gpu -x -q 2>&1 | tail -3
python -c "
import bench, json
print(json.dumps(bench.latency_batch1(0)))"
