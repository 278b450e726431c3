set -x
python tools/profile_step.py 256 2 level > gpurun_out/r2c_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_blur_levels|k_describe' -c 4 -o gpurun_out/r2c_level python tools/profile_step.py 256 1 level > gpurun_out/r2c_ncu.log 2>&1
ncu -i gpurun_out/r2c_level.ncu-rep --page raw --csv > gpurun_out/r2c_level_raw.csv 2>/dev/null
ncu -i gpurun_out/r2c_level.ncu-rep --page source --csv -k regex:k_blur_levels > gpurun_out/r2c_src_blur.csv 2>/dev/null
rm -f gpurun_out/r2c_level.ncu-rep
