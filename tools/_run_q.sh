set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2q_gpu_tests.log
tail -4 gpurun_out/r2q_gpu_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2q_bench_C1.json 2> gpurun_out/r2q_bench_C1.err; echo "C1 rc=$?"; tail -3 gpurun_out/r2q_bench_C1.err
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2q_bench_C1_reference.json 2> gpurun_out/r2q_ref.err; echo "ref rc=$?"; cat gpurun_out/r2q_bench_C1_reference.json
for c in C2 C3 C4 C5; do
timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2q_bench_$c.json 2> gpurun_out/r2q_bench_$c.err; echo "$c rc=$?"; tail -2 gpurun_out/r2q_bench_$c.err
done
python tools/profile_step.py 512 2 > gpurun_out/r2q_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2q_launches.csv python tools/profile_step.py 512 2 > gpurun_out/r2q_ncu_list.log 2>&1
python __graft_entry__.py smoke > gpurun_out/r2q_smoke.log 2>&1; tail -2 gpurun_out/r2q_smoke.log
ls -la gpurun_out | tail -14
