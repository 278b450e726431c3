"""Small workload for compute-sanitizer (memcheck / racecheck): smoke + one 4-frame C1 batch in both descriptor modes
+ the frame-to-frame match, stereo row-band/refine on a C2 pair."""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
import __graft_entry__ as G
from bench import shifts_for, MATCH_TH
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

G.smoke()
B = 4
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
for mode in ("fused", "level"):
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B, blur_mode=mode).plan(752, 480)
    out = ex.alloc_outputs(B, frames.device)
    ex.extract_batch_device(frames, (0, 1000), out)
    sf = ex.GetScaleFactors()
    shift = torch.from_numpy(shifts_for(0, B)).cuda()
    src = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).cuda()
    bounds = (0.0, 752.0, 0.0, 480.0)
    grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
    qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src, shift, MATCH_TH, sf, (-1, 1), None)
    res = ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"], qb["qr"],
                                 qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"])
    torch.cuda.synchronize()
    print(mode, int(out["n"].sum()))
print("sanitize step ok")
