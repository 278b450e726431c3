cat tmp_variants/variants.txt
pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), [(s['kernel'], s['ms']) for s in d['stages']])"; }
for i in 0 1 2; do
  cp tmp_variants/liborbx_$i.so orb_slam3_vio_fixes_b200/liborbx.so
  echo "== variant $i"
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr
  python bench.py --config C5 --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr
done
cp tmp_variants/liborbx_1.so orb_slam3_vio_fixes_b200/liborbx.so
timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -2
