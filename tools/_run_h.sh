mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_gpu_tests.log
tail -4 gpurun_out/r2h_gpu_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2h_bench_C1.json 2> gpurun_out/r2h_bench_C1.err; echo "C1 rc=$?"; tail -3 gpurun_out/r2h_bench_C1.err; cat gpurun_out/r2h_bench_C1.json
for c in C2 C3 C4 C5; do
timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2h_bench_$c.json 2> gpurun_out/r2h_bench_$c.err; echo "$c rc=$?"; tail -3 gpurun_out/r2h_bench_$c.err; cat gpurun_out/r2h_bench_$c.json
done
