"""Times orbm_search_area vs the whole-function orbm_search_by_projection (last-frame mode) at C1, batch 512."""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from bench import shifts_for, MATCH_TH
from orb_slam3_vio_fixes_b200 import synthetic as S, _lib
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).plan(752, 480)
out = ex.alloc_outputs(B, frames.device)
ex.extract_batch_device(frames, (0, 1000), out)
sf = ex.GetScaleFactors()
shift = torch.from_numpy(shifts_for(0, B)).cuda()
src = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).cuda()
bounds = (0.0, 752.0, 0.0, 480.0)
grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src, shift, MATCH_TH, sf, (-1, 1))
qb["angle"] = out["kps"][src.long(), :, 3].contiguous()


def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): r = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, r

t_a, ra = timeit(lambda: ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"], qb["qr"],
                                                qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"]))
t_p, rp = timeit(lambda: ORBmatcher.search_by_projection(out["kps"], out["desc"], out["n"], grid, bounds, qb,
                                                         _lib.PROJ_LASTFRAME, 0.9, True))
print("search_area ms", round(t_a, 4), "search_by_projection ms", round(t_p, 4), "nmatches/frame",
      float(rp["nmatches"].float().mean()))
