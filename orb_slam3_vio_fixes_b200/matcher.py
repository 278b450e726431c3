"""Host-side mirror of the ORBmatcher cores (reference include/ORBmatcher.h:35-105) on the C ABI.

`ORBmatcher.DescriptorDistance` keeps the reference's static signature (a single pair is a host
popcount, as in include/orbx_compat/ORBmatcher_core.h).  The batched functions are the
candidate-scoring loops behind SearchByProjection / SearchByBoW / SearchForTriangulation /
Frame::ComputeStereoMatches; they take torch CUDA tensors and run on the current stream.
"""
import numpy as np

from . import _lib
from ._lib import RES_DTYPE, TIE_FIRST_MIN, TIE_LAST_MIN  # noqa: F401

_POP8 = np.array([bin(i).count("1") for i in range(256)], np.int32)


def _stream(dev):
    import torch
    return torch.cuda.current_stream(dev).cuda_stream


def _res_tensor(n, dev):
    import torch
    return torch.zeros((n, RES_DTYPE.itemsize), dtype=torch.uint8, device=dev)


def results_to_numpy(t):
    a = np.ascontiguousarray(t.detach().cpu().numpy())
    return a.view(RES_DTYPE).reshape(a.shape[:-1])


class ORBmatcher:
    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30  # ORBmatcher.cc:35-37

    def __init__(self, nnratio=0.6, checkOri=True):
        self.mfNNratio, self.mbCheckOrientation = np.float32(nnratio), bool(checkOri)

    @staticmethod
    def DescriptorDistance(a, b):
        """Hamming distance of two 32-byte descriptors (ORBmatcher.cc:2058-2074)."""
        a = np.asarray(a, np.uint8).reshape(-1)
        b = np.asarray(b, np.uint8).reshape(-1)
        assert a.size == 32 and b.size == 32
        return int(_POP8[a ^ b].sum())

    # ---- batched cores ---------------------------------------------------------------------
    @staticmethod
    def descriptor_distance_pairs(a, b, idx_a=None, idx_b=None):
        """DescriptorDistance for many pairs: a [na,32], b [nb,32] u8 CUDA tensors, optional i32 index tensors.
        Returns an i32 tensor of Hamming distances."""
        import torch
        n = (idx_a if idx_a is not None else a).shape[0]
        out = torch.empty((n,), dtype=torch.int32, device=a.device)
        p = lambda t: t.data_ptr() if t is not None else None
        rc = _lib.lib().orbm_descriptor_distance_pairs(a.data_ptr(), b.data_ptr(), p(idx_a), p(idx_b), n,
                                                       out.data_ptr(), _stream(a.device))
        _lib.check(rc, "orbm_descriptor_distance_pairs")
        return out

    @staticmethod
    def argmin2_csr(q, r, cand_off, cand_idx, ref_level=None, tie_rule=TIE_FIRST_MIN):
        """q [nq,32] u8, r [nr,32] u8, cand_off [nq+1] i32, cand_idx [total] i32 (CUDA tensors).
        Returns a [nq,12] uint8 tensor of orbm_result records."""
        out = _res_tensor(q.shape[0], q.device)
        rc = _lib.lib().orbm_argmin2_csr(q.data_ptr(), q.shape[0], r.data_ptr(), r.shape[0], cand_off.data_ptr(),
                                         cand_idx.data_ptr(), 0 if ref_level is None else ref_level.data_ptr(),
                                         tie_rule, out.data_ptr(), _stream(q.device))
        _lib.check(rc, "orbm_argmin2_csr")
        return out

    @staticmethod
    def grid_build(kps, n, bounds, out=None):
        """kps [B,cap,7] f32 (cv::KeyPoint records), n [B] i32; bounds = (minX, maxX, minY, maxY).
        Returns (cell_off [B, 64*48+1] i32, cell_idx [B,cap] i32); `out` = a previous return value to refill
        (no allocation, no fill kernel on the per-frame path)."""
        import torch
        b, cap = kps.shape[0], kps.shape[1]
        if out is not None:
            cell_off, cell_idx = out
        else:
            cell_off = torch.zeros((b, _lib.GRID_COLS * _lib.GRID_ROWS + 1), dtype=torch.int32, device=kps.device)
            cell_idx = torch.zeros((b, cap), dtype=torch.int32, device=kps.device)
        rc = _lib.lib().orbm_grid_build(kps.data_ptr(), n.data_ptr(), cap, b, *[float(v) for v in bounds],
                                        cell_off.data_ptr(), cell_idx.data_ptr(), _stream(kps.device))
        _lib.check(rc, "orbm_grid_build")
        return cell_off, cell_idx

    @staticmethod
    def search_area(kps, desc, n, grid, bounds, qx, qy, qr, qmin, qmax, qdesc, nq, tie_rule=TIE_FIRST_MIN, out=None):
        """Fused GetFeaturesInArea + best/second scoring.  Queries [B,Q] (qx,qy,qr f32; qmin,qmax i8;
        qdesc [B,Q,32] u8; nq [B] i32).  Returns [B,Q,12] uint8 orbm_result records (`out` = a previous return
        value to refill; records of queries >= nq keep what they held)."""
        b, cap = kps.shape[0], kps.shape[1]
        qs = qx.shape[1]
        if out is None:
            out = _res_tensor(b * qs, kps.device).view(b, qs, RES_DTYPE.itemsize)
        cell_off, cell_idx = grid
        rc = _lib.lib().orbm_search_area(kps.data_ptr(), desc.data_ptr(), n.data_ptr(), cap, cell_off.data_ptr(),
                                         cell_idx.data_ptr(), *[float(v) for v in bounds], qx.data_ptr(),
                                         qy.data_ptr(), qr.data_ptr(), qmin.data_ptr(), qmax.data_ptr(),
                                         qdesc.data_ptr(), nq.data_ptr(), qs, b, tie_rule, out.data_ptr(),
                                         _stream(kps.device))
        _lib.check(rc, "orbm_search_area")
        return out

    @staticmethod
    def fuse_candidates(kps, desc, n, grid, bounds, inv_level_sigma2, qx, qy, qr, qlevel, qdesc, nq, u_right=None,
                        q_ur=None):
        """Candidate scoring of ORBmatcher::Fuse (ORBmatcher.cc:1148-1338).  Queries [B,Q]: qx,qy,qr (,q_ur) f32,
        qlevel i8 (predicted level), qdesc [B,Q,32] u8, nq [B] i32; u_right [B,cap] f32 = mvuRight or None.
        Returns [B,Q,12] uint8 orbm_result records (best_idx, best_dist)."""
        b, cap = kps.shape[0], kps.shape[1]
        qs = qx.shape[1]
        out = _res_tensor(b * qs, kps.device).view(b, qs, RES_DTYPE.itemsize)
        sg = np.ascontiguousarray(inv_level_sigma2, np.float32)
        p = lambda t: t.data_ptr() if t is not None else None
        rc = _lib.lib().orbm_fuse_candidates(kps.data_ptr(), desc.data_ptr(), n.data_ptr(), cap, grid[0].data_ptr(),
                                             grid[1].data_ptr(), *[float(v) for v in bounds], p(u_right),
                                             sg.ctypes.data, sg.size, qx.data_ptr(), qy.data_ptr(), qr.data_ptr(),
                                             p(q_ur), qlevel.data_ptr(), qdesc.data_ptr(), nq.data_ptr(), qs, b,
                                             out.data_ptr(), _stream(kps.device))
        _lib.check(rc, "orbm_fuse_candidates")
        return out

    @staticmethod
    def queries_shift(kps, desc, n, src_frame, shift_xy, th, scale_factors, band=(-1, 1), out=None):
        """Window queries of SearchByProjection(Frame, Frame) for a per-frame image shift.
        src_frame [B] i32, shift_xy [B,2] f32.  Returns dict(qx,qy,qr,qmin,qmax,qdesc,nq)."""
        import torch
        b, cap = kps.shape[0], kps.shape[1]
        dev = kps.device
        if out is None:
            out = {"qx": torch.zeros((b, cap), dtype=torch.float32, device=dev),
                   "qy": torch.zeros((b, cap), dtype=torch.float32, device=dev),
                   "qr": torch.zeros((b, cap), dtype=torch.float32, device=dev),
                   "qmin": torch.zeros((b, cap), dtype=torch.int8, device=dev),
                   "qmax": torch.zeros((b, cap), dtype=torch.int8, device=dev),
                   "qdesc": torch.zeros((b, cap, 32), dtype=torch.uint8, device=dev),
                   "nq": torch.zeros(b, dtype=torch.int32, device=dev)}
        sf = np.ascontiguousarray(scale_factors, np.float32)
        rc = _lib.lib().orbm_queries_shift(kps.data_ptr(), desc.data_ptr(), n.data_ptr(), cap, b,
                                           src_frame.data_ptr(), shift_xy.data_ptr(), float(th), sf.ctypes.data,
                                           sf.size, int(band[0]), int(band[1]), out["qx"].data_ptr(),
                                           out["qy"].data_ptr(), out["qr"].data_ptr(), out["qmin"].data_ptr(),
                                           out["qmax"].data_ptr(), out["qdesc"].data_ptr(), out["nq"].data_ptr(),
                                           _stream(dev))
        _lib.check(rc, "orbm_queries_shift")
        return out

    @staticmethod
    def search_by_projection(kps, desc, n, grid, bounds, q, mode, nn_ratio=0.6, check_orientation=True,
                             u_right=None, feat_locked=None, th_dist=100):
        """ORBmatcher::SearchByProjection as a whole function (ORBmatcher.cc:43-142 / :1676-1887, Nleft == -1).
        q: dict of CUDA tensors qx,qy,qr [B,Q] f32, qmin,qmax [B,Q] i8, qdesc [B,Q,32] u8, nq [B] i32 and
        th_dist = TH_HIGH, or ORBdist of the relocalisation variant (ORBmatcher.cc:1889-2010); optionally q_ur [B,Q] f32, flags [B,Q] u8 (bit0 takes part, bit1 MapPoint has observations),
        angle [B,Q] f32.  Returns dict(match [B,Q] i32, feat_query [B,cap] i32, nmatches [B] i32)."""
        import torch
        b, cap = kps.shape[0], kps.shape[1]
        qs = q["qx"].shape[1]
        dev = kps.device
        scratch = torch.empty((b, qs, 8), dtype=torch.int32, device=dev)
        out = {"match": torch.empty((b, qs), dtype=torch.int32, device=dev),
               "feat_query": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        p = lambda t: t.data_ptr() if t is not None else None
        rc = _lib.lib().orbm_search_by_projection_th(
            kps.data_ptr(), desc.data_ptr(), n.data_ptr(), cap, b, grid[0].data_ptr(), grid[1].data_ptr(),
            float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), p(u_right), p(feat_locked),
            q["qx"].data_ptr(), q["qy"].data_ptr(), q["qr"].data_ptr(), q["qmin"].data_ptr(), q["qmax"].data_ptr(),
            q["qdesc"].data_ptr(), q["nq"].data_ptr(), qs, p(q.get("q_ur")), p(q.get("flags")), p(q.get("angle")),
            int(mode), float(nn_ratio), int(bool(check_orientation)), int(th_dist), scratch.data_ptr(),
            out["match"].data_ptr(),
            out["feat_query"].data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_by_projection")
        return out

    @staticmethod
    def rig_camera(kps, desc, n, grid, feat_locked=None, partner=None):
        """One camera of a two-camera Frame (orbm_rig_camera): its own keypoints [B,cap,28] u8, descriptors, counts,
        orbm_grid_build CSR, optional entry locks [B,cap] u8 and pair table [B,cap] i32 (mvLeftToRightMatch /
        mvRightToLeftMatch).  Keeps the tensors alive."""
        c = _lib.OrbmRigCamera()
        p = lambda t: t.data_ptr() if t is not None else None
        c.d_kps, c.d_desc, c.d_n, c.kp_stride = kps.data_ptr(), desc.data_ptr(), n.data_ptr(), kps.shape[1]
        c.d_cell_off, c.d_cell_idx = grid[0].data_ptr(), grid[1].data_ptr()
        c.d_feat_locked, c.d_partner = p(feat_locked), p(partner)
        c._keep = (kps, desc, n, grid, feat_locked, partner)
        return c

    @staticmethod
    def search_by_projection_rig(mode, left, right, bounds, q, nn_ratio=0.6, check_orientation=True):
        """ORBmatcher::SearchByProjection on a two-camera (fisheye) Frame, whole function (ORBmatcher.cc:43-213 /
        :1676-1887 with their right-camera blocks).  left / right = rig_camera(...).  q: dict of CUDA tensors
        qx,qy,qr,qmin,qmax (left window), qx_r,qy_r,qr_r,qmin_r,qmax_r (right window), qdesc [B,Q,32], nq [B],
        optional flags [B,Q] u8 (bit0 left, bit1 observations, bit2 right) and angle [B,Q] f32.
        Returns dict(match_l, match_r [B,Q], feat_query_l, feat_query_r [B,cap], nmatches [B])."""
        import ctypes as C
        import torch
        b, qs = q["qx"].shape
        dev = q["qx"].device
        scratch = torch.empty((b, qs, 65), dtype=torch.uint8, device=dev)
        out = {"match_l": torch.empty((b, qs), dtype=torch.int32, device=dev),
               "match_r": torch.empty((b, qs), dtype=torch.int32, device=dev),
               "feat_query_l": torch.empty((b, left.kp_stride), dtype=torch.int32, device=dev),
               "feat_query_r": torch.empty((b, right.kp_stride), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        p = lambda t: t.data_ptr() if t is not None else None
        rc = _lib.lib().orbm_search_by_projection_rig(
            int(mode), C.byref(left), C.byref(right), b, float(bounds[0]), float(bounds[1]), float(bounds[2]),
            float(bounds[3]), q["qx"].data_ptr(), q["qy"].data_ptr(), q["qr"].data_ptr(), q["qmin"].data_ptr(),
            q["qmax"].data_ptr(), q["qx_r"].data_ptr(), q["qy_r"].data_ptr(), q["qr_r"].data_ptr(),
            q["qmin_r"].data_ptr(), q["qmax_r"].data_ptr(), q["qdesc"].data_ptr(), q["nq"].data_ptr(), qs,
            p(q.get("flags")), p(q.get("angle")), float(nn_ratio), int(bool(check_orientation)), scratch.data_ptr(),
            out["match_l"].data_ptr(), out["match_r"].data_ptr(), out["feat_query_l"].data_ptr(),
            out["feat_query_r"].data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_by_projection_rig")
        return out

    @staticmethod
    def stereo_rowband(kps_l, desc_l, n_l, kps_r, desc_r, n_r, scale_factors, max_d):
        """Hamming stage of Frame::ComputeStereoMatches.  Returns (best_idx_r [B,cap] i32, best_dist [B,cap] i32)."""
        import torch
        b, cap = kps_l.shape[0], kps_l.shape[1]
        bi = torch.full((b, cap), -1, dtype=torch.int32, device=kps_l.device)
        bd = torch.full((b, cap), 100, dtype=torch.int32, device=kps_l.device)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        rc = _lib.lib().orbm_stereo_rowband(kps_l.data_ptr(), desc_l.data_ptr(), n_l.data_ptr(), kps_r.data_ptr(),
                                            desc_r.data_ptr(), n_r.data_ptr(), cap, b, sf.ctypes.data, sf.size,
                                            float(max_d), bi.data_ptr(), bd.data_ptr(), _stream(kps_l.device))
        _lib.check(rc, "orbm_stereo_rowband")
        return bi, bd

    @staticmethod
    def stereo_refine(ex_l, ex_r, kps_l, n_l, kps_r, best_idx, best_dist, mbf, max_d):
        """Sub-pixel stage + median cut of Frame::ComputeStereoMatches (Frame.cc:896-981) on the pyramids the
        two extractors (ORBextractor objects) hold on the device.  Returns dict(u_right, depth [B,cap] f32,
        sad [B,cap] i32, kept [B] i32)."""
        import torch
        b, cap = kps_l.shape[0], kps_l.shape[1]
        dev = kps_l.device
        out = {"u_right": torch.empty((b, cap), dtype=torch.float32, device=dev),
               "depth": torch.empty((b, cap), dtype=torch.float32, device=dev),
               "sad": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "kept": torch.empty((b,), dtype=torch.int32, device=dev)}
        rc = _lib.lib().orbm_stereo_refine(ex_l._h, ex_r._h, kps_l.data_ptr(), n_l.data_ptr(),
                                           kps_r.data_ptr(), cap, b, best_idx.data_ptr(), best_dist.data_ptr(),
                                           float(mbf), float(max_d), out["u_right"].data_ptr(),
                                           out["depth"].data_ptr(), out["sad"].data_ptr(), out["kept"].data_ptr(),
                                           _stream(dev))
        _lib.check(rc, "orbm_stereo_refine")
        return out

    @staticmethod
    def stereo_fisheye_matches(kps_l, desc_l, n_l, mono_l, kps_r, desc_r, n_r, mono_r, cam_l, cam_r, Rlr, tlr,
                               level_sigma2):
        """Frame::ComputeStereoFishEyeMatches (Frame.cc:1126-1166).  mono_* [B] i32 = the extractors' monoIndex.
        Returns dict(left_to_right, right_to_left [B,cap] i32, depth [B,cap] f32, p3d [B,cap,3] f32)."""
        import ctypes as C
        import torch
        b, cap = kps_l.shape[0], kps_l.shape[1]
        dev = kps_l.device
        out = {"left_to_right": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "right_to_left": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "depth": torch.empty((b, cap), dtype=torch.float32, device=dev),
               "p3d": torch.empty((b, cap, 3), dtype=torch.float32, device=dev)}
        R = np.ascontiguousarray(Rlr, np.float32).reshape(9)
        t = np.ascontiguousarray(tlr, np.float32).reshape(3)
        sg = np.ascontiguousarray(level_sigma2, np.float32)
        rc = _lib.lib().orbm_stereo_fisheye_matches(
            kps_l.data_ptr(), desc_l.data_ptr(), n_l.data_ptr(), mono_l.data_ptr(), kps_r.data_ptr(), desc_r.data_ptr(),
            n_r.data_ptr(), mono_r.data_ptr(), cap, b, C.byref(cam_l), C.byref(cam_r), R.ctypes.data, t.ctypes.data,
            sg.ctypes.data, sg.size, out["left_to_right"].data_ptr(), out["right_to_left"].data_ptr(),
            out["depth"].data_ptr(), out["p3d"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_stereo_fisheye_matches")
        return out

    @staticmethod
    def rgbd_depth(kps, n, depth_img, mbf, kps_un=None):
        """Frame::ComputeStereoFromRGBD (Frame.cc:984-1005).  depth_img: float32 CUDA tensor [B,h,w].
        Returns (u_right, depth) [B,cap] f32."""
        import torch
        b, cap = kps.shape[0], kps.shape[1]
        assert depth_img.dtype == torch.float32 and depth_img.dim() == 3 and depth_img.stride(2) == 1
        ur = torch.empty((b, cap), dtype=torch.float32, device=kps.device)
        dp = torch.empty((b, cap), dtype=torch.float32, device=kps.device)
        rc = _lib.lib().orbm_rgbd_depth(kps.data_ptr(), kps_un.data_ptr() if kps_un is not None else None,
                                        n.data_ptr(), cap, b, depth_img.data_ptr(), depth_img.stride(1) * 4,
                                        depth_img.stride(0) * 4, depth_img.shape[2], depth_img.shape[1], float(mbf),
                                        ur.data_ptr(), dp.data_ptr(), _stream(kps.device))
        _lib.check(rc, "orbm_rgbd_depth")
        return ur, dp

    @staticmethod
    def knn2_dense(q, r):
        """cv::BFMatcher(NORM_HAMMING).knnMatch(k=2).  Returns (idx [nq,2] i32, dist [nq,2] i32)."""
        import torch
        idx = torch.zeros((q.shape[0], 2), dtype=torch.int32, device=q.device)
        dist = torch.zeros((q.shape[0], 2), dtype=torch.int32, device=q.device)
        rc = _lib.lib().orbm_knn2_dense(q.data_ptr(), q.shape[0], r.data_ptr(), r.shape[0], idx.data_ptr(),
                                        dist.data_ptr(), _stream(q.device))
        _lib.check(rc, "orbm_knn2_dense")
        return idx, dist

    # ---- vocabulary-node matching (SearchByBoW / SearchForTriangulation / DBoW2 transform) -----------
    @staticmethod
    def _fv_struct(fv):
        return _lib.OrbmFeatureVector(fv["node"].data_ptr(), fv["off"].data_ptr(), fv["idx"].data_ptr(),
                                      fv["nn"].data_ptr())

    @staticmethod
    def feature_vector_alloc(batch, cap, node_stride, dev):
        """Empty FeatureVector CSR tensors: node [B,S], off [B,S+1], idx [B,cap], nn [B] (all i32)."""
        import torch
        z = lambda *s: torch.zeros(s, dtype=torch.int32, device=dev)
        return {"node": z(batch, node_stride), "off": z(batch, node_stride + 1), "idx": z(batch, cap), "nn": z(batch)}

    @staticmethod
    def search_by_bow(mode, kps1, desc1, n1, fv1, kps2, desc2, n2, fv2, nn_ratio=0.6, check_orientation=True,
                      valid1=None, valid2=None):
        """ORBmatcher::SearchByBoW (mode _lib.BOW_KF_FRAME: ORBmatcher.cc:223-425; BOW_KF_KF: :765-905).
        kps* [B,cap,7] f32, desc* [B,cap,32] u8, n* [B] i32, fv* dicts from feature_vector_alloc / bow_transform,
        valid* [B,cap] u8 or None.  Returns dict(match12, match21 [B,cap] i32, nmatches [B] i32)."""
        import torch
        b, cap = kps1.shape[0], kps1.shape[1]
        dev = kps1.device
        ns = fv1["node"].shape[1]
        assert fv2["node"].shape[1] == ns and kps2.shape[1] == cap
        out = {"match12": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "match21": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        p = lambda t: t.data_ptr() if t is not None else None
        s1, s2 = ORBmatcher._fv_struct(fv1), ORBmatcher._fv_struct(fv2)
        rc = _lib.lib().orbm_search_by_bow(int(mode), kps1.data_ptr(), desc1.data_ptr(), n1.data_ptr(), p(valid1), s1,
                                           kps2.data_ptr(), desc2.data_ptr(), n2.data_ptr(), p(valid2), s2, cap, ns,
                                           b, float(nn_ratio), int(bool(check_orientation)),
                                           out["match12"].data_ptr(), out["match21"].data_ptr(),
                                           out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_by_bow")
        return out

    @staticmethod
    def search_by_bow_rig(kps1, desc1, n1, fv1, kps2, desc2, n2, nleft2, fv2, nn_ratio=0.6, check_orientation=True,
                          valid1=None):
        """ORBmatcher::SearchByBoW(KeyFrame*, Frame&) with F.Nleft != -1 (ORBmatcher.cc:295-323, 356-380).  Side 2 =
        the frame's left|right features in one numbering, nleft2 [B] i32 = F.Nleft.
        Returns dict(match12_left, match12_right, match21 [B,cap] i32, nmatches [B] i32)."""
        import torch
        b, cap = kps1.shape[0], kps1.shape[1]
        dev = kps1.device
        ns = fv1["node"].shape[1]
        assert fv2["node"].shape[1] == ns and kps2.shape[1] == cap
        out = {k: torch.empty((b, cap), dtype=torch.int32, device=dev) for k in ("match12_left", "match12_right", "match21")}
        out["nmatches"] = torch.empty((b,), dtype=torch.int32, device=dev)
        p = lambda t: t.data_ptr() if t is not None else None
        s1, s2 = ORBmatcher._fv_struct(fv1), ORBmatcher._fv_struct(fv2)
        rc = _lib.lib().orbm_search_by_bow_rig(kps1.data_ptr(), desc1.data_ptr(), n1.data_ptr(), p(valid1), s1,
                                               kps2.data_ptr(), desc2.data_ptr(), n2.data_ptr(), nleft2.data_ptr(), s2,
                                               cap, ns, b, float(nn_ratio), int(bool(check_orientation)),
                                               out["match12_left"].data_ptr(), out["match12_right"].data_ptr(),
                                               out["match21"].data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_by_bow_rig")
        return out

    @staticmethod
    def search_for_triangulation_rig(kps1, desc1, n1, nleft1, fv1, kps2, desc2, n2, nleft2, fv2, rig_pose, cams1, cams2,
                                     level_sigma2_1, level_sigma2_2, only_stereo=False, coarse=False,
                                     check_orientation=True, has_mp1=None, has_mp2=None):
        """ORBmatcher::SearchForTriangulation for keyframes of a two-camera rig (ORBmatcher.cc:907-1146, camera selection
        :1036-1062, KannalaBrandt8::epipolarConstrain).  rig_pose [B,4,12] f32 CUDA = (R12,t12) of Tll,Tlr,Trl,Trr;
        cams1 / cams2 = 2-tuples of frame.camera(KANNALA_BRANDT8, ...).  Returns dict(match12 [B,cap], nmatches [B])."""
        import torch
        b, cap = kps1.shape[0], kps1.shape[1]
        dev = kps1.device
        ns = fv1["node"].shape[1]
        out = {"match12": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        p = lambda t: t.data_ptr() if t is not None else None
        s1, s2 = ORBmatcher._fv_struct(fv1), ORBmatcher._fv_struct(fv2)
        c1 = (_lib.OrbmCamera * 2)(*cams1)
        c2 = (_lib.OrbmCamera * 2)(*cams2)
        g1 = np.ascontiguousarray(level_sigma2_1, np.float32)
        g2 = np.ascontiguousarray(level_sigma2_2, np.float32)
        rc = _lib.lib().orbm_search_for_triangulation_rig(
            kps1.data_ptr(), desc1.data_ptr(), n1.data_ptr(), nleft1.data_ptr(), p(has_mp1), s1, kps2.data_ptr(),
            desc2.data_ptr(), n2.data_ptr(), nleft2.data_ptr(), p(has_mp2), s2, cap, ns, b, rig_pose.data_ptr(), c1, c2,
            g1.ctypes.data, g2.ctypes.data, g1.size, int(bool(only_stereo)), int(bool(coarse)),
            int(bool(check_orientation)), out["match12"].data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_for_triangulation_rig")
        return out

    @staticmethod
    def search_for_triangulation(kps1, desc1, n1, fv1, kps2, desc2, n2, fv2, F12, ep, scale_factors2, level_sigma2_2,
                                 only_stereo=False, coarse=False, check_orientation=True, has_mp1=None,
                                 has_mp2=None, u_right1=None, u_right2=None):
        """ORBmatcher::SearchForTriangulation (ORBmatcher.cc:907-1146), pinhole single camera.  F12 [B,9] f32,
        ep [B,2] f32 (CUDA).  Returns dict(match12 [B,cap] i32, nmatches [B] i32)."""
        import torch
        b, cap = kps1.shape[0], kps1.shape[1]
        dev = kps1.device
        ns = fv1["node"].shape[1]
        assert fv2["node"].shape[1] == ns and kps2.shape[1] == cap
        out = {"match12": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        p = lambda t: t.data_ptr() if t is not None else None
        sf = np.ascontiguousarray(scale_factors2, np.float32)
        sg = np.ascontiguousarray(level_sigma2_2, np.float32)
        s1, s2 = ORBmatcher._fv_struct(fv1), ORBmatcher._fv_struct(fv2)
        rc = _lib.lib().orbm_search_for_triangulation(
            kps1.data_ptr(), desc1.data_ptr(), n1.data_ptr(), p(has_mp1), p(u_right1), s1, kps2.data_ptr(),
            desc2.data_ptr(), n2.data_ptr(), p(has_mp2), p(u_right2), s2, cap, ns, b, p(F12), ep.data_ptr(),
            sf.ctypes.data, sg.ctypes.data, sf.size, int(bool(only_stereo)), int(bool(coarse)),
            int(bool(check_orientation)), out["match12"].data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_for_triangulation")
        return out

    @staticmethod
    def search_for_initialization(kps1, desc1, n1, kps2, desc2, n2, grid2, bounds, prev_matched, window_size=100,
                                  nn_ratio=0.9, check_orientation=True):
        """ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763).  prev_matched [B,cap,2] f32 is updated in
        place (vbPrevMatched).  Returns dict(match12 [B,cap] i32, nmatches [B] i32)."""
        import torch
        b, cap = kps1.shape[0], kps1.shape[1]
        dev = kps1.device
        out = {"match12": torch.empty((b, cap), dtype=torch.int32, device=dev),
               "nmatches": torch.empty((b,), dtype=torch.int32, device=dev)}
        scratch = torch.empty((2, b, cap), dtype=torch.int32, device=dev)
        rc = _lib.lib().orbm_search_for_initialization(
            kps1.data_ptr(), desc1.data_ptr(), n1.data_ptr(), kps2.data_ptr(), desc2.data_ptr(), n2.data_ptr(), cap, b,
            grid2[0].data_ptr(), grid2[1].data_ptr(), *[float(v) for v in bounds], prev_matched.data_ptr(),
            int(window_size), float(nn_ratio), int(bool(check_orientation)), out["match12"].data_ptr(),
            scratch.data_ptr(), out["nmatches"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_search_for_initialization")
        return out

    @staticmethod
    def distinctive_descriptors(pool, obs_off, obs_idx):
        """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:367-401) batched.  pool [R,32] u8, obs_off [M+1] i32,
        obs_idx [total] i32.  Returns (best [M] i32, median [M] i32)."""
        import torch
        m = obs_off.shape[0] - 1
        best = torch.empty((m,), dtype=torch.int32, device=pool.device)
        med = torch.empty((m,), dtype=torch.int32, device=pool.device)
        rc = _lib.lib().orbm_distinctive_descriptors(pool.data_ptr(), obs_off.data_ptr(), obs_idx.data_ptr(), m,
                                                     best.data_ptr(), med.data_ptr(), _stream(pool.device))
        _lib.check(rc, "orbm_distinctive_descriptors")
        return best, med


class ORBVocabulary:
    """Device-resident DBoW2 vocabulary tree (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) as flat arrays and its
    transform() (:1127-1259).  Arrays: child_off [nnodes+1] i32, child_idx i32, node_desc [nnodes,32] u8,
    node_word [nnodes] i32, node_weight [nnodes] f64; node 0 is the root; L = tree depth."""

    def __init__(self, child_off, child_idx, node_desc, node_word, node_weight, L, device="cuda:0"):
        import torch
        t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dt)).to(device)
        self.child_off, self.child_idx = t(child_off, np.int32), t(child_idx, np.int32)
        self.node_desc, self.node_word = t(node_desc, np.uint8), t(node_word, np.int32)
        self.node_weight = t(node_weight, np.float64)
        self.L, self.device = int(L), device
        self._s = _lib.OrbmVocabulary(self.child_off.data_ptr(), self.child_idx.data_ptr(), self.node_desc.data_ptr(),
                                      self.node_word.data_ptr(), self.node_weight.data_ptr(),
                                      self.node_word.shape[0], self.L)

    @staticmethod
    def parse_text_file(path):
        """TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1423) on the host, through the C ABI
        (no GPU needed).  Returns the flat tree as a dict of numpy arrays + k, L, scoring, weighting, n_words."""
        import ctypes as C
        h = _lib.OrbmVocabularyHost()
        rc = _lib.lib().orbm_vocabulary_parse_text(str(path).encode(), C.byref(h))
        _lib.check(rc, "orbm_vocabulary_parse_text")
        try:
            n = h.n_nodes
            arr = lambda p, cnt, dt: np.ctypeslib.as_array(p, shape=(cnt,)).astype(dt).copy()
            return {"k": h.k, "L": h.L, "scoring": h.scoring, "weighting": h.weighting, "n_words": h.n_words,
                    "child_off": arr(h.child_off, n + 1, np.int32), "child_idx": arr(h.child_idx, n - 1, np.int32),
                    "node_desc": arr(h.node_desc, n * 32, np.uint8).reshape(n, 32),
                    "node_word": arr(h.node_word, n, np.int32), "node_weight": arr(h.node_weight, n, np.float64),
                    "node_parent": arr(h.node_parent, n, np.int32)}
        finally:
            _lib.lib().orbm_vocabulary_host_free(C.byref(h))

    @classmethod
    def loadFromTextFile(cls, path, device="cuda:0"):
        """ORBVocabulary::loadFromTextFile + upload (src/System.cc:118)."""
        v = cls.parse_text_file(path)
        return cls(v["child_off"], v["child_idx"], v["node_desc"], v["node_word"], v["node_weight"], v["L"], device)

    def transform(self, desc, n, levelsup=4, node_stride=None):
        """desc [B,cap,32] u8, n [B] i32 (CUDA).  Returns dict(feat_word, feat_node [B,cap] i32, feat_weight
        [B,cap] f64, fv (FeatureVector CSR dict), bow_word [B,cap] i32, bow_val [B,cap] f64, bow_n [B] i32)."""
        import torch
        b, cap = desc.shape[0], desc.shape[1]
        dev = desc.device
        ns = cap if node_stride is None else int(node_stride)
        fv = ORBmatcher.feature_vector_alloc(b, cap, ns, dev)
        out = {"feat_word": torch.zeros((b, cap), dtype=torch.int32, device=dev),
               "feat_node": torch.zeros((b, cap), dtype=torch.int32, device=dev),
               "feat_weight": torch.zeros((b, cap), dtype=torch.float64, device=dev),
               "fv": fv,
               "bow_word": torch.zeros((b, cap), dtype=torch.int32, device=dev),
               "bow_val": torch.zeros((b, cap), dtype=torch.float64, device=dev),
               "bow_n": torch.zeros((b,), dtype=torch.int32, device=dev)}
        rc = _lib.lib().orbm_bow_transform(self._s, desc.data_ptr(), n.data_ptr(), cap, b, int(levelsup),
                                           out["feat_word"].data_ptr(), out["feat_node"].data_ptr(),
                                           out["feat_weight"].data_ptr(), ORBmatcher._fv_struct(fv), ns,
                                           out["bow_word"].data_ptr(), out["bow_val"].data_ptr(),
                                           out["bow_n"].data_ptr(), _stream(dev))
        _lib.check(rc, "orbm_bow_transform")
        return out
