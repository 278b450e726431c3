mkdir -p gpurun_out
for mode in 0 1 2; do
for cfg in C1 C5; do
  echo "== ORBX_OCT_PRIO=$mode $cfg" >> gpurun_out/r2ac_octprio.txt
  ORBX_OCT_PRIO=$mode timeout 300 python bench.py --no-cpu-baseline --config $cfg --steps 20 --warmup 5 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['e2e']['value'])
" >> gpurun_out/r2ac_octprio.txt
done; done
cat gpurun_out/r2ac_octprio.txt
