mkdir -p gpurun_out
ORBX_STRESS_SEEDS=60 timeout 1500 python -m pytest tests/test_gpu_rig.py tests/test_gpu_bow_random.py -m gpu -q > gpurun_out/r2ab_stress_rig.txt 2>&1; echo "rc=$?" >> gpurun_out/r2ab_stress_rig.txt
tail -4 gpurun_out/r2ab_stress_rig.txt
