mkdir -p gpurun_out
: > gpurun_out/r2n_c5_split.txt
for sp in 1 2 4 8; do for st in 2 4; do
  [ $sp -eq 1 ] && [ $st -eq 4 ] && continue
  echo "== C5 split=$sp streams=$st" >> gpurun_out/r2n_c5_split.txt
  ORBX_DEVICE_SPLIT=$sp ORBX_DEVICE_STREAMS=$st timeout 300 python bench.py --config C5 --steps 5 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']))" >> gpurun_out/r2n_c5_split.txt
done; done
for cfg in C2 C4; do for sp in 2 4; do
  echo "== $cfg split=$sp streams=4" >> gpurun_out/r2n_c5_split.txt
  ORBX_DEVICE_SPLIT=$sp ORBX_DEVICE_STREAMS=4 timeout 300 python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']))" >> gpurun_out/r2n_c5_split.txt
done; done
cat gpurun_out/r2n_c5_split.txt
