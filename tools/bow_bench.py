"""Device times of the vocabulary-node functions on C1 extractor output (512 frames / pairs per call):
DBoW2 transform, SearchByBoW (both variants), SearchForTriangulation, Fuse candidates, ComputeDistinctiveDescriptors.
Prints one line per function: ms per call, calls' units per second, algorithmic bytes and achieved GB/s."""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from orb_slam3_vio_fixes_b200 import synthetic as S, _lib
from orb_slam3_vio_fixes_b200.extractor import ORBextractor
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, ORBVocabulary

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
REP = 10
frames = torch.from_numpy(S.make_frames("C1", B)).cuda()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).plan(752, 480)
out = ex.alloc_outputs(B, frames.device)
ex.extract_batch_device(frames, (0, 0), out)
torch.cuda.synchronize()
kps, desc, n = out["kps"], out["desc"], out["n"]
cap = kps.shape[1]
nk = int(n.sum())
pool = desc[0, :int(n[0])].cpu().numpy()
voc = S.make_vocabulary(10, 6, seed=3, desc_pool=np.concatenate([pool] * 4))   # ORBvoc shape: k=10, L=6
V = ORBVocabulary(voc["child_off"], voc["child_idx"], voc["node_desc"], voc["node_word"], voc["node_weight"], voc["L"])
sf = ex.GetScaleFactors()


def timed(fn):
    for _ in range(3):
        r = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(REP):
        r = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / REP, r


def report(name, ms, units, unit_name, alg_bytes):
    print(f"{name:34s} {ms:8.3f} ms/call  {units / ms * 1e3:12.0f} {unit_name}/s  alg {alg_bytes / 1e6:8.1f} MB  "
          f"{alg_bytes / ms / 1e6:8.1f} GB/s")


ms, tr = timed(lambda: V.transform(desc, n, levelsup=4, node_stride=128))
# per feature: its descriptor + L levels x 10 children x 32 B node descriptors + 16 B of outputs
report("bow_transform (L=6,k=10)", ms, nk, "features", nk * (32 + 6 * 10 * 32 + 16))
fv = tr["fv"]
# pair p = (frame p, frame p+1): roll the second side by one frame
roll = lambda t: torch.roll(t, -1, 0).contiguous()
fv2 = {k: roll(v) for k, v in fv.items()}
k2, d2, n2 = roll(kps), roll(desc), roll(n)
nn = fv["nn"].cpu().numpy()
off = fv["off"].cpu().numpy()
off2 = fv2["off"].cpu().numpy()
node1, node2, nn2 = fv["node"].cpu().numpy(), fv2["node"].cpu().numpy(), fv2["nn"].cpu().numpy()
pairs_scored = 0
for p in range(B):
    a = dict(zip(node1[p, :nn[p]], np.diff(off[p, :nn[p] + 1])))
    b = dict(zip(node2[p, :nn2[p]], np.diff(off2[p, :nn2[p] + 1])))
    pairs_scored += sum(a[k_] * b[k_] for k_ in a.keys() & b.keys())
for mode, name in ((_lib.BOW_KF_FRAME, "search_by_bow KF-Frame"), (_lib.BOW_KF_KF, "search_by_bow KF-KF")):
    ms, r = timed(lambda: ORBmatcher.search_by_bow(mode, kps, desc, n, fv, k2, d2, n2, fv2, 0.7, True))
    report(name, ms, B, "pairs", 2 * nk * 32 + 2 * nk * 4 + 2 * nk * 4)
    print(f"    matches/pair {float(r['nmatches'].float().mean()):.1f}, descriptor pairs in shared nodes/pair "
          f"{pairs_scored / B:.0f}")
F = torch.zeros((B, 9), dtype=torch.float32, device="cuda")
F[:, 5], F[:, 7] = -0.894, 0.894
F[:, 2], F[:, 6] = 0.447, -0.447   # [t]x for t ~ (2,1,0)/|t|: the sequence's image motion
ep = torch.full((B, 2), -1e4, dtype=torch.float32, device="cuda")
ms, r = timed(lambda: ORBmatcher.search_for_triangulation(kps, desc, n, fv, k2, d2, n2, fv2, F, ep, sf, sf * sf))
report("search_for_triangulation", ms, B, "pairs", 2 * nk * (32 + 28) + 2 * nk * 4 + nk * 4)
print(f"    matches/pair {float(r['nmatches'].float().mean()):.1f}")
# Fuse: MapPoints = keypoints of the next frame projected back by the known (2,1) px motion
bounds = (0.0, 752.0, 0.0, 480.0)
grid = ORBmatcher.grid_build(kps, n, bounds)
k2f = k2.view(B, cap, 7)
qx = (k2f[:, :, 0] - 2.0).contiguous()
qy = (k2f[:, :, 1] - 1.0).contiguous()
ql = k2.view(torch.int32).view(B, cap, 7)[:, :, 5].to(torch.int8).contiguous()
qr = (3.0 * torch.from_numpy(sf).cuda()[ql.long().clamp(0, 7)]).contiguous()
inv = (1.0 / (sf * sf)).astype(np.float32)
ms, r = timed(lambda: ORBmatcher.fuse_candidates(kps, desc, n, grid, bounds, inv, qx, qy, qr, ql, d2, n2))
report("fuse_candidates", ms, nk, "map points", nk * (32 + 32 + 12 + 16))
# distinctive descriptors: 100k map points with 2..20 observations
rng = np.random.default_rng(0)
M = 100000
sizes = rng.integers(2, 21, M)
o = np.zeros(M + 1, np.int32)
o[1:] = np.cumsum(sizes)
idx = rng.integers(0, B * cap, o[-1]).astype(np.int32)
pool_t = desc.view(B * cap, 32)
ot, it = torch.from_numpy(o).cuda(), torch.from_numpy(idx).cuda()
ms, r = timed(lambda: ORBmatcher.distinctive_descriptors(pool_t, ot, it))
report("distinctive_descriptors", ms, M, "map points", int(o[-1]) * 36 + M * 8)
