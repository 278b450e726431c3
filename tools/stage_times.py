"""Per-stage device times (ORBX_FLAG_STAGE_TIMING) of one C1 batch: python tools/stage_times.py [fused|level] [B]"""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200 import _lib
_lib_stages = list(_lib.STAGES)

mode = sys.argv[1] if len(sys.argv) > 1 else None
B = int(sys.argv[2]) if len(sys.argv) > 2 else 512
cfg = sys.argv[3] if len(sys.argv) > 3 else "C1"
w, h, nf, _, _ = S.CONFIGS[cfg]
frames = torch.from_numpy(S.make_frames(cfg, min(B, 64))).cuda()
frames = frames.repeat((B + frames.shape[0] - 1) // frames.shape[0], 1, 1)[:B].contiguous()
ex = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B, blur_mode=mode, stage_timing=True).plan(w, h)
out = ex.alloc_outputs(B, frames.device)
acc = np.zeros(5)
N = 10
for i in range(3 + N):
    ex.extract_batch_device(frames, (0, 1000), out)
    t = ex.stage_times()
    if i >= 3:
        acc += np.array([t[k] for k in _lib_stages])
torch.cuda.synchronize()
names = _lib_stages
print(cfg, "B", B, "mode", mode, " ".join(f"{n}={v / N:.4f}" for n, v in zip(names, acc)), f"sum={acc.sum() / N:.4f} ms")
