mkdir -p gpurun_out
python tools/stage_times.py level 2>&1 | tail -1
ORBX_NVCC_EXTRA="-DFS_WARPS=1" python -c "from orb_slam3_vio_fixes_b200 import build as b; b.build(force=True)"
python tools/stage_times.py level 2>&1 | tail -1
ORBX_NVCC_EXTRA="-DFS_WARPS=4" python -c "from orb_slam3_vio_fixes_b200 import build as b; b.build(force=True)"
python tools/stage_times.py level 2>&1 | tail -1
