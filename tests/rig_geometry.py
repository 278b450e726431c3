"""fp64 numpy evaluation of KannalaBrandt8::epipolarConstrain (src/CameraModels/KannalaBrandt8.cpp:216-220, 306-375)
and a synthetic two-camera rig scene for the SearchForTriangulation tests.  Test helper, not a product path."""
import numpy as np

CAM_L = (1, 190.978, 190.973, 254.932, 256.897, 0.0034, 0.00071, -0.0020, 0.00020)   # TUM-VI 512x512 fisheye
CAM_R = (1, 190.442, 190.434, 252.597, 254.917, 0.0034, 0.00076, -0.0021, 0.00021)


def kb8_project(cam, X):
    X = np.asarray(X, np.float64)
    th = np.arctan2(np.hypot(X[..., 0], X[..., 1]), X[..., 2])
    psi = np.arctan2(X[..., 1], X[..., 0])
    t2 = th * th
    r = th * (1 + t2 * (cam[5] + t2 * (cam[6] + t2 * (cam[7] + t2 * cam[8]))))
    return np.stack([cam[1] * r * np.cos(psi) + cam[3], cam[2] * r * np.sin(psi) + cam[4]], -1)


def kb8_unproject(cam, uv):
    uv = np.asarray(uv, np.float64)
    pw = np.stack([(uv[..., 0] - cam[3]) / cam[1], (uv[..., 1] - cam[4]) / cam[2]], -1)
    thd = np.clip(np.hypot(pw[..., 0], pw[..., 1]), -np.pi / 2, np.pi / 2)
    th = thd.copy()
    for _ in range(10):
        t2, t4 = th * th, (th * th) ** 2
        t6, t8 = t4 * t2, t4 * t4
        k0, k1, k2, k3 = cam[5] * t2, cam[6] * t4, cam[7] * t6, cam[8] * t8
        th = th - (th * (1 + k0 + k1 + k2 + k3) - thd) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3)
    scale = np.where(thd > 1e-8, np.tan(th) / np.maximum(thd, 1e-300), 1.0)
    return np.stack([pw[..., 0] * scale, pw[..., 1] * scale, np.ones_like(scale)], -1)


def epipolar_gate64(cam1, cam2, uv1, uv2, R12, t12, sigma1, sigma2):
    """Returns (ok [n] bool, margin [n]): margin = the smallest relative distance of a gate quantity from its
    threshold along the path the decision took (small margin = a decision fp32 may legitimately flip)."""
    r1, r2 = kb8_unproject(cam1, uv1), kb8_unproject(cam2, uv2)
    R12, t12 = np.asarray(R12, np.float64), np.asarray(t12, np.float64)
    r21 = r2 @ R12.T
    cosp = (r1 * r21).sum(-1) / (np.linalg.norm(r1, axis=-1) * np.linalg.norm(r21, axis=-1))
    n = len(cosp)
    R21 = R12.T
    t21 = -R21 @ t12
    A = np.zeros((n, 4, 4))
    A[:, 0, 0] = -1; A[:, 0, 2] = r1[:, 0]
    A[:, 1, 1] = -1; A[:, 1, 2] = r1[:, 1]
    T2 = np.concatenate([R21, t21[:, None]], 1)
    A[:, 2, :] = r2[:, 0:1] * T2[2] - T2[0]
    A[:, 3, :] = r2[:, 1:2] * T2[2] - T2[1]
    _, _, Vt = np.linalg.svd(A)
    xh = Vt[:, 3, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        X = xh[:, :3] / xh[:, 3:4]
    z1 = X[:, 2]
    X2 = X @ R21.T + t21
    z2 = X2[:, 2]
    e1 = ((kb8_project(cam1, X) - uv1) ** 2).sum(-1)
    e2 = ((kb8_project(cam2, X2) - uv2) ** 2).sum(-1)
    th1, th2 = 5.991 * np.asarray(sigma1, np.float64), 5.991 * np.asarray(sigma2, np.float64)
    ok = np.ones(n, bool)
    margin = np.full(n, np.inf)
    rel = lambda v, t: np.abs(v - t) / np.maximum(np.abs(t), 1e-12)
    steps = [(cosp > 0.9998, np.abs(cosp - 0.9998) / 1e-3),   # fp32 resolves cos near 1 to ~6e-8: 1e-6 is 'on the gate'
             (~(z1 > 1e-4), np.abs(z1 - 1e-4) / np.maximum(np.abs(z1), 1e-4)),
             (~(z2 > 0), np.abs(z2) / np.maximum(np.abs(X2).max(-1), 1e-9)),
             (e1 > th1, rel(e1, th1)), (e2 > th2, rel(e2, th2))]
    alive = np.ones(n, bool)
    for fail, m in steps:
        m = np.where(np.isfinite(m), m, 0.0)
        margin = np.where(alive, np.minimum(margin, m), margin)
        ok &= ~(alive & fail)
        alive &= ~fail
    ok &= np.isfinite(X).all(-1)
    return ok, margin


def rot(axis, ang):
    axis = np.asarray(axis, np.float64) / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def rig_poses(rng):
    """World->camera (R, t) of KF1 left/right and KF2 left/right: a rig with a 10 cm baseline moved by ~30 cm."""
    Rrl, trl = rot([0.1, 1.0, 0.05], 0.03), np.array([-0.101, 0.002, 0.001])
    T1l = (rot([0.2, 1, 0.1], 0.05), np.array([0.02, -0.01, 0.03]))
    T2l = (rot([0.1, 1, -0.2], -0.08), np.array([-0.30, 0.04, 0.08]))
    right = lambda T: (Rrl @ T[0], Rrl @ T[1] + trl)
    return [T1l, right(T1l)], [T2l, right(T2l)]


def rel_pose(Ta, Tb):
    """T12 = Ta * Tb^-1 as (R12, t12): maps camera-b coordinates into camera-a coordinates (ORBmatcher.cc:929-937)."""
    R12 = Ta[0] @ Tb[0].T
    return R12, Ta[1] - R12 @ Tb[1]


def fisheye_ratio_candidates(desc_l, desc_r, mono_l, mono_r):
    """Index logic of Frame::ComputeStereoFishEyeMatches before the triangulation (Frame.cc:1126-1152): Hamming
    knnMatch(k=2) of the lapping-area descriptors (ties to the lower train index), Lowe ratio 0.7 evaluated in double
    (`float < float * 0.7`).  Returns (left indices, right indices) of the pairs handed to TriangulateMatches, in the
    order the reference visits them.  Pinned against the reference's own text by tests/test_ref_frame.py."""
    ham = (np.unpackbits(desc_l[mono_l:], axis=1)[:, None, :] != np.unpackbits(desc_r[mono_r:], axis=1)[None, :, :]).sum(-1)
    order = np.argsort(ham, axis=1, kind="stable")
    b0, b1 = order[:, 0], order[:, 1]
    d0, d1 = ham[np.arange(len(b0)), b0], ham[np.arange(len(b0)), b1]
    cand = d0.astype(np.float64) < d1.astype(np.float64) * 0.7
    return np.nonzero(cand)[0] + mono_l, b0[cand] + mono_r


def fisheye_tables(n_l, n_r, qi, ti, ok):
    """mvLeftToRightMatch / mvRightToLeftMatch from the accepted pairs (Frame.cc:1157-1163): a right feature keeps the
    last left index written to it."""
    l2r, r2l = np.full(n_l, -1, np.int64), np.full(n_r, -1, np.int64)
    for q, t, o in zip(qi, ti, ok):
        if o:
            l2r[q] = t
            r2l[t] = q
    return l2r, r2l
