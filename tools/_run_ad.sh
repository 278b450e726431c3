mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_describe|k_resize|k_blur|k_octree' -c 13 -o gpurun_out/r2ad_full python tools/profile_step.py 128 1 > gpurun_out/r2ad_ncu.log 2>&1
echo rc=$?
ls -la gpurun_out | tail -5
