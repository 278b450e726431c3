"""One warm-up + N profiled steps of the bench pipeline (extract + match) for ncu."""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from bench import shifts_for, MATCH_TH
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
blur_mode = sys.argv[3] if len(sys.argv) > 3 else None
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B, blur_mode=blur_mode).plan(752, 480)
out = ex.alloc_outputs(B, frames.device)
sf = ex.GetScaleFactors()
shift = torch.from_numpy(shifts_for(0, B)).cuda()
src = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).cuda()
bounds = (0.0, 752.0, 0.0, 480.0)
qb = None
for _ in range(steps):
    ex.extract_batch_device(frames, (0, 1000), out)
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src, shift, MATCH_TH, sf, (-1, 1), qb)
    res = ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"], qb["qr"],
                                 qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"])
torch.cuda.synchronize()
print("ok", int(out["n"].sum()))
