pr() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], [(s['kernel'], s['ms']) for s in d['stages'] if s['kernel']=='describe'])"; }
for r in 35 56 70 105 35 70; do cp tmp_variants/liborbx_$r.so orb_slam3_vio_fixes_b200/liborbx.so; echo "== BLUR_ROWS=$r"; for c in C1 C4; do python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline --no-latency 2>/dev/null | pr; done; done
cp tmp_variants/liborbx_70.so orb_slam3_vio_fixes_b200/liborbx.so
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -2
