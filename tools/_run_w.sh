pr() { grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1], round(d['value']), 'e2e', round(d['e2e']['value']))" "$1"; }
export ORBX_BENCH_VERBOSE=1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2> gpurun_out/w1.err | pr "N=2 T=2 setdev"; grep "rank" gpurun_out/w1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-latency --e2e-threads 1 2> gpurun_out/w2.err | pr "N=2 T=1"; grep "rank" gpurun_out/w2.err
(CUDA_VISIBLE_DEVICES=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr "solo gpu0 (concurrent)") &
(CUDA_VISIBLE_DEVICES=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | pr "solo gpu1 (concurrent)") &
wait
