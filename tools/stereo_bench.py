"""C2 (EuRoC stereo shape): both eyes extracted, then Frame::ComputeStereoMatches on the device
(row-band Hamming -> 11x11 SAD sub-pixel refinement -> median cut).  Prints pairs/s and stage times."""
import json
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
w, h, nf, _, _ = S.CONFIGS[cfg]
left = torch.from_numpy(S.make_frames(cfg, B)).cuda()
right = torch.from_numpy(S.make_frames(cfg, B, right=True)).cuda()
exl = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B).plan(w, h)
exr = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B).plan(w, h)
ol, orr = exl.alloc_outputs(B, left.device), exr.alloc_outputs(B, left.device)
sf = exl.GetScaleFactors()
fx, bf = (458.654, 47.90639384423901) if cfg == "C2" else (718.856, 386.1448)
mbf = np.float32(bf)
max_d = mbf / np.float32(bf / fx)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]


def step(timed=False):
    if timed: ev[0].record()
    exl.extract_batch_device(left, (0, 0), ol)
    exr.extract_batch_device(right, (0, 0), orr)
    if timed: ev[1].record()
    bi, bd = ORBmatcher.stereo_rowband(ol["kps"], ol["desc"], ol["n"], orr["kps"], orr["desc"], orr["n"], sf, max_d)
    if timed: ev[2].record()
    r = ORBmatcher.stereo_refine(exl, exr, ol["kps"], ol["n"], orr["kps"], bi, bd, mbf, max_d)
    if timed: ev[3].record()
    return r


for _ in range(3): r = step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 10
e0.record()
for _ in range(K): r = step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
r = step(True); torch.cuda.synchronize()
print(json.dumps({"config": cfg, "pairs_per_step": B, "ms_per_step": round(ms, 4), "pairs_per_s": round(B / ms * 1e3, 1),
                  "extract_both_eyes_ms": round(ev[0].elapsed_time(ev[1]), 4),
                  "rowband_ms": round(ev[1].elapsed_time(ev[2]), 4), "refine_mediancut_ms": round(ev[2].elapsed_time(ev[3]), 4),
                  "kp_per_eye": float(ol["n"].float().mean()), "stereo_matches_per_pair": float(r["kept"].float().mean())}))
