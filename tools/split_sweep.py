"""Device-resident step time (extract + match, C1, batch 512) for several ORBX_DEVICE_SPLIT values."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, '.')
from bench import shifts_for, MATCH_TH
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

B = 512
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
shift = torch.from_numpy(shifts_for(0, B)).cuda()
src = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).cuda()
bounds = (0.0, 752.0, 0.0, 480.0)
ref = None
for split in [1, 2, 4, 8, 1, 4]:
    os.environ["ORBX_DEVICE_SPLIT"] = str(split)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).plan(752, 480)
    out = ex.alloc_outputs(B, frames.device)
    sf = ex.GetScaleFactors()
    qb = None
    def step():
        global qb
        ex.extract_batch_device(frames, (0, 1000), out)
        grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
        qb = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src, shift, MATCH_TH, sf, (-1, 1), qb)
        return ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qb["qx"], qb["qy"], qb["qr"],
                                      qb["qmin"], qb["qmax"], qb["qdesc"], qb["nq"])
    for _ in range(3): r = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): r = step()
    e1.record(); torch.cuda.synchronize()
    sig = (out["kps"].cpu().numpy().tobytes(), out["desc"].cpu().numpy().tobytes(), r.cpu().numpy().tobytes())
    if ref is None: ref = sig
    print("split", split, "ms/step", round(e0.elapsed_time(e1) / 20, 4), "same_as_split1", sig == ref, flush=True)
    del ex
