mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2v_bench_C1_n2.json 2> gpurun_out/r2v_n2.err; echo "rc=$?"; tail -2 gpurun_out/r2v_n2.err; cat gpurun_out/r2v_bench_C1_n2.json | head -c 1500
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 3 --warmup 1 > gpurun_out/r2v_ref_n2.json 2>> gpurun_out/r2v_n2.err; echo "rc=$?"; head -c 600 gpurun_out/r2v_ref_n2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --config C4 --steps 5 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2v_bench_C4_n2.json 2>> gpurun_out/r2v_n2.err; echo "rc=$?"; head -c 700 gpurun_out/r2v_bench_C4_n2.json
timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q -k two_handles 2>&1 | tail -2
