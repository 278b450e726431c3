set -x
mkdir -p gpurun_out
python tools/profile_step.py 256 2 level > gpurun_out/r2e_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_fast' -c 4 -o gpurun_out/r2e_fast python tools/profile_step.py 256 1 level > gpurun_out/r2e_ncu.log 2>&1
ncu -i gpurun_out/r2e_fast.ncu-rep --page raw --csv > gpurun_out/r2e_fast_raw.csv 2>/dev/null
ncu -i gpurun_out/r2e_fast.ncu-rep --page source --print-source cuda,sass --csv --launch-count 1 > gpurun_out/r2e_src_fast.csv 2>/dev/null
ls -la gpurun_out | tail -5
