python tools/stage_times.py level 1 2>&1 | tail -1
python tools/stage_times.py fused 1 2>&1 | tail -1
python - <<'PY'
import time, numpy as np, sys
sys.path.insert(0,'.')
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
img = S.make_frames("C1", 1)[0]
for mode in (None, "fused", "level"):
    ex = ORBextractor(1000, 1.2, 8, 20, 7, blur_mode=mode).plan(752, 480)
    for _ in range(10): ex(img, None, (0, 1000))
    t=[]
    for _ in range(100):
        a=time.perf_counter(); ex(img, None, (0,1000)); t.append(time.perf_counter()-a)
    print(mode, "orbx_extract median ms", round(np.median(t)*1e3,4), "min", round(min(t)*1e3,4))
PY
