mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__shared_mem_per_block_dynamic,sm__inst_executed.avg.per_cycle_active,l1tex__throughput.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:k_fast_strip -c 4 --csv --log-file gpurun_out/r2av_fast_c5.csv python bench.py --config C5 --steps 1 --warmup 1 --no-cpu-baseline --no-latency > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2av_fast_c5.csv')) if len(r)>10]
h=rows[0]; 
for r in rows[1:]:
    print(r[h.index('ID')], r[h.index('Grid Size')], r[h.index('Metric Name')], r[h.index('Metric Value')])
PY
