for pad in 0 2048 4096 8192; do echo "== fast pad $pad"; ORBX_FAST_PAD=$pad python tools/stage_times.py level 2>&1 | tail -1; done
