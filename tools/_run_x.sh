mkdir -p gpurun_out
pr() { grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1], 'value', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']))" "$1"; }
N=$1
for c in C1 C4 C5; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --config $c --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2> gpurun_out/x_$c.err | tee gpurun_out/r2x_bench_${c}_n$N.json | pr "$c N=$N"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --impl reference --steps 3 --warmup 1 2>/dev/null | grep '^{' | head -c 300
