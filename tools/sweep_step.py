"""Device-resident extract step time (C1, batch 512, CUDA events) for one setting, taken from the environment:
ORBX_FAST_PAIRS, ORBX_DEVICE_SPLIT, ORBX_DEVICE_STREAMS; argv[1] = blur mode (fused|level|auto).
Prints one line; also a checksum of the outputs so settings can be compared for identical bytes."""
import hashlib, os, sys
import numpy as np
import torch
sys.path.insert(0, '.')
from orb_slam3_vio_fixes_b200 import synthetic as S
from orb_slam3_vio_fixes_b200.extractor import ORBextractor

mode = sys.argv[1] if len(sys.argv) > 1 and sys.argv[1] != "auto" else None
B = 512
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B, blur_mode=mode).plan(752, 480)
out = ex.alloc_outputs(B, frames.device)
for _ in range(3):
    ex.extract_batch_device(frames, (0, 1000), out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ex.extract_batch_device(frames, (0, 1000), out)
e1.record(); torch.cuda.synchronize()
n = out["n"].cpu().numpy()
hsh = hashlib.sha1(out["kps"].cpu().numpy().tobytes() + out["desc"].cpu().numpy().tobytes()).hexdigest()[:10]
env = {k: os.environ.get(k, "-") for k in ("ORBX_FAST_PAIRS", "ORBX_DEVICE_SPLIT", "ORBX_DEVICE_STREAMS")}
print(f"blur={mode} {env} ms/step={e0.elapsed_time(e1) / 20:.4f} frames/s={B / (e0.elapsed_time(e1) / 20) * 1e3:.0f} n={int(n.sum())} sha={hsh}", flush=True)
