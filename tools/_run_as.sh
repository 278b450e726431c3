for i in 0 1 0 1; do cp tmp_variants/liborbx_$i.so orb_slam3_vio_fixes_b200/liborbx.so; echo "== variant $i"; python tools/proj_timing.py 512 2>/dev/null | tail -1; done
cp tmp_variants/liborbx_1.so orb_slam3_vio_fixes_b200/liborbx.so
timeout 1200 python -m pytest tests/test_gpu_match.py tests/test_gpu_rig.py tests/test_gpu_compat_cpp.py tests/test_gpu_golden.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
