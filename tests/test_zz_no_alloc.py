"""orbm_search_by_projection_static: the allocation-free variant of orbm_search_by_projection (its candidate lists
live in a static device array inside the library).  It has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc.  The test calls it through the C ABI, twice in
a row with different inputs (the array keeps the first call's contents), and compares every output with the existing
entry point on the same inputs: integer results, so the tolerance is zero, as in the other matcher tests.  The file is
named so that pytest collects it last."""
import numpy as np
import pytest

from conftest import has_cuda
from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200 import synthetic as S

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not has_cuda(), reason="needs CUDA")]

ORB = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)
BOUNDS = (0.0, 640.0, 0.0, 480.0)


def _queries(kps, desc, n, sf, rng, B, th, band, unlocked, drop):
    """MapPoints of frame f = keypoints of the other frames moved by the known shift (as tests/test_gpu_match.py)."""
    pos = np.array([S.frame_shift(t) for t in range(B)], np.float32)
    per = []
    for f in range(B):
        parts = []
        for s in range(B):
            if s == f:
                continue
            k, d = kps[s, :n[s]], desc[s, :n[s]]
            sh = pos[s] - pos[f]
            m = len(k)
            fl = np.full(m, 3, np.uint8)
            fl[rng.random(m) < unlocked] = 1
            fl[rng.random(m) < drop] &= 2
            parts.append(dict(qx=(k["x"] + sh[0]).astype(np.float32), qy=(k["y"] + sh[1]).astype(np.float32),
                              qr=(np.float32(th) * sf[k["octave"]]).astype(np.float32),
                              qmin=(k["octave"] + band[0]).astype(np.int8), qmax=(k["octave"] + band[1]).astype(np.int8),
                              qdesc=d, angle=k["angle"].astype(np.float32), flags=fl))
        per.append({k_: np.concatenate([p[k_] for p in parts]) for k_ in parts[0]})
    Q = max(len(q["qx"]) for q in per)
    dt = dict(qx=np.float32, qy=np.float32, qr=np.float32, qmin=np.int8, qmax=np.int8, angle=np.float32, flags=np.uint8)
    host = {k_: np.zeros((B, Q), v) for k_, v in dt.items()}
    host["qdesc"] = np.zeros((B, Q, 32), np.uint8)
    host["nq"] = np.zeros(B, np.int32)
    for f, q in enumerate(per):
        m = len(q["qx"])
        for k_ in list(dt) + ["qdesc"]:
            host[k_][f, :m] = q[k_]
        host["nq"][f] = m
    return host, Q


def test_static_scratch_entry_point_equals_the_caller_scratch_one():
    import torch
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, _stream
    L = _lib.lib()
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    B = 4
    ex = ORBextractor(1000, max_batch=B, **ORB)
    out = ex.extract_batch_device(t(S.make_frames("C3", B)), (0, 0))
    torch.cuda.synchronize()
    kps_h = ORBextractor.kps_from_tensor(out["kps"])
    desc_h, n_h = out["desc"].cpu().numpy(), out["n"].cpu().numpy()
    cap = out["kps"].shape[1]
    sf = ex.GetScaleFactors()
    grid = ORBmatcher.grid_build(out["kps"], out["n"], BOUNDS)
    rng = np.random.default_rng(77)
    locked = t((rng.random((B, cap)) < 0.15).astype(np.uint8))

    # two different calls: (MapPoints, ratio 0.8, locks) then (last frame, rotation check, other radius / band)
    calls = [dict(mode=_lib.PROJ_MAPPOINTS, nn=0.8, ori=0, th=12.0, band=(-1, 0), locked=locked, unlocked=0.2, drop=0.05),
             dict(mode=_lib.PROJ_LASTFRAME, nn=0.9, ori=1, th=15.0, band=(-1, 1), locked=None, unlocked=0.0, drop=0.1)]
    prepared = []
    for c in calls:
        host, Q = _queries(kps_h, desc_h, n_h, sf, rng, B, c["th"], c["band"], c["unlocked"], c["drop"])
        assert B * Q * 32 <= _lib.STATIC_SCRATCH_BYTES
        q = {k_: t(v) for k_, v in host.items()}
        # everything either entry point needs exists before the calls: outputs for both, scratch for the old one
        bufs = {name: dict(match=torch.full((B, Q), -7, dtype=torch.int32, device=dev),
                           feat_query=torch.full((B, cap), -7, dtype=torch.int32, device=dev),
                           nmatches=torch.full((B,), -7, dtype=torch.int32, device=dev)) for name in ("old", "new")}
        scratch = torch.empty((B, Q, 8), dtype=torch.int32, device=dev)
        prepared.append((c, q, Q, bufs, scratch))
    torch.cuda.synchronize()

    def common(c, q, Q):
        p = lambda x: x.data_ptr() if x is not None else None
        return [out["kps"].data_ptr(), out["desc"].data_ptr(), out["n"].data_ptr(), cap, B, grid[0].data_ptr(),
                grid[1].data_ptr(), BOUNDS[0], BOUNDS[1], BOUNDS[2], BOUNDS[3], None, p(c["locked"]),
                q["qx"].data_ptr(), q["qy"].data_ptr(), q["qr"].data_ptr(), q["qmin"].data_ptr(), q["qmax"].data_ptr(),
                q["qdesc"].data_ptr(), q["nq"].data_ptr(), Q, None, q["flags"].data_ptr(), q["angle"].data_ptr(),
                c["mode"], c["nn"], c["ori"]]

    st = _stream(dev)
    # set-up: the first use loads the module that holds the array (CUDA loads modules lazily) and looks the symbol's
    # address up; it also leaves another call's lists in the array
    c, q, Q, bufs, _ = prepared[1]
    o = bufs["new"]
    rc = L.orbm_search_by_projection_static(*common(c, q, Q), o["match"].data_ptr(), o["feat_query"].data_ptr(),
                                            o["nmatches"].data_ptr(), st)
    assert rc == _lib.ORBX_OK, (rc, L.orbx_last_error())
    torch.cuda.synchronize()
    for t_ in o.values():
        t_.fill_(-7)
    torch.cuda.synchronize()
    # nothing else may touch device memory inside the window: garbage of earlier tests (handles whose destructor calls
    # cudaFree) is collected now, and the collector stays off until the check is done
    import gc
    gc.collect()
    torch.cuda.synchronize()
    gc.disable()
    try:
        free0 = torch.cuda.mem_get_info()[0]
        # the two new-entry-point calls back to back on one stream: the second overwrites the first one's lists
        for c, q, Q, bufs, _ in prepared:
            o = bufs["new"]
            rc = L.orbm_search_by_projection_static(*common(c, q, Q), o["match"].data_ptr(), o["feat_query"].data_ptr(),
                                                    o["nmatches"].data_ptr(), st)
            assert rc == _lib.ORBX_OK, (rc, L.orbx_last_error())
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
    finally:
        gc.enable()
    assert free1 >= free0, "the allocation-free entry point took device memory"

    for c, q, Q, bufs, scratch in prepared:
        o = bufs["old"]
        rc = L.orbm_search_by_projection(*common(c, q, Q), scratch.data_ptr(), o["match"].data_ptr(),
                                         o["feat_query"].data_ptr(), o["nmatches"].data_ptr(), st)
        assert rc == _lib.ORBX_OK, (rc, L.orbx_last_error())
    torch.cuda.synchronize()

    for c, q, Q, bufs, _ in prepared:
        nq = q["nq"].cpu().numpy()
        new = {k_: v.cpu().numpy() for k_, v in bufs["new"].items()}
        old = {k_: v.cpu().numpy() for k_, v in bufs["old"].items()}
        assert np.array_equal(new["nmatches"], old["nmatches"])
        assert int(new["nmatches"].sum()) > B * 200          # the comparison is not between two empty results
        for f in range(B):
            assert np.array_equal(new["match"][f, :nq[f]], old["match"][f, :nq[f]])
            assert np.array_equal(new["feat_query"][f, :n_h[f]], old["feat_query"][f, :n_h[f]])

    # beyond the static array: refused with a clear message, nothing launched, outputs untouched
    c, q, Q, bufs, _ = prepared[0]
    too_many = _lib.STATIC_SCRATCH_BYTES // 32 // Q + 1      # frames: batch * q_stride * 32 > the array
    args = common(c, q, Q)
    args[4] = too_many
    o = bufs["new"]
    before = o["match"].clone()
    rc = L.orbm_search_by_projection_static(*args, o["match"].data_ptr(), o["feat_query"].data_ptr(),
                                            o["nmatches"].data_ptr(), st)
    assert rc == -3  # ORBX_E_CAPACITY
    msg = L.orbx_last_error().decode()
    assert "static" in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert torch.equal(o["match"], before)
