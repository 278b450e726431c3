import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from orb_slam3_vio_fixes_b200 import synthetic as S, _lib
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B, stage_timing=True).plan(752, 480)
out = ex.alloc_outputs(B, frames.device)
for _ in range(3): ex.extract_batch_device(frames, (0, 1000), out)
torch.cuda.synchronize()
for mode in ("nosync", "sync_each", "nosync"):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    acc = {k: 0.0 for k in _lib.STAGES}
    e0.record()
    for i in range(10):
        ex.extract_batch_device(frames, (0, 1000), out)
        if mode == "sync_each":
            for k, v in ex.stage_times().items(): acc[k] += v / 10
    e1.record(); torch.cuda.synchronize()
    last = ex.stage_times()
    print(mode, "ms/step", e0.elapsed_time(e1) / 10, "last-step stages", {k: round(v, 3) for k, v in last.items()}, "sum", round(sum(last.values()), 3),
          "avg" if mode == "sync_each" else "", {k: round(v, 3) for k, v in acc.items()} if mode == "sync_each" else "")
