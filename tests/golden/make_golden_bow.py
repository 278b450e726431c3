#!/usr/bin/env python3
"""Generates tests/golden/bow_c1.npz — expected outputs of the vocabulary-node functions on the two C1 frames of
c1_mono.npz (their keypoints / descriptors are already pinned against cv2 there).

Every expected array is written only after the C++ oracle and the independent pure-Python line-by-line restatement
below (written from the reference sources alone: ORBmatcher.cc:223-425, 765-905, 907-1146, 2012-2053;
TemplatedVocabulary.h:1127-1259; BowVector.cpp:33-88; MapPoint.cc:367-401) agreed on it.  The GPU box needs neither
to check the CUDA path against the file."""
import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import oracle  # noqa: E402
from orb_slam3_vio_fixes_b200 import synthetic as S  # noqa: E402

OUT = Path(__file__).resolve().parent
POP = np.array([bin(i).count("1") for i in range(256)], np.int32)
TH_LOW, HISTO_LENGTH = 50, 30


def dist(a, b):
    return int(POP[a ^ b].sum())


def py_transform(desc, voc, levelsup):
    co, ci, nd, nw, wt, L = (voc[k] for k in ("child_off", "child_idx", "node_desc", "node_word", "node_weight", "L"))
    v, fv = {}, {}
    words, nodes = [], []
    for f, d in enumerate(desc):
        final, level, nid = 0, 0, 0
        while True:
            level += 1
            ch = ci[co[final]:co[final + 1]]
            final = int(ch[0])
            best = float(dist(d, nd[final]))
            for c in ch[1:]:
                dd = float(dist(d, nd[c]))
                if dd < best:
                    best, final = dd, int(c)
            if level == L - levelsup:
                nid = final
            if co[final] == co[final + 1]:
                break
        words.append(int(nw[final])); nodes.append(nid)
        w = float(wt[final])
        if w > 0:
            v[int(nw[final])] = v[int(nw[final])] + w if int(nw[final]) in v else w
            fv.setdefault(nid, []).append(f)
    norm = 0.0
    for k in sorted(v):
        norm += abs(v[k])
    bw = sorted(v)
    bv = [v[k] / norm if norm > 0 else v[k] for k in bw]
    return np.array(words, np.int32), np.array(nodes, np.int32), np.array(bw, np.int32), np.array(bv, np.float64), fv


def rot_bin(a1, a2):
    rot = np.float32(a1) - np.float32(a2)
    if rot < 0.0:
        rot = np.float32(rot + np.float32(360.0))
    x = float(np.float32(rot * np.float32(1.0 / HISTO_LENGTH)))
    b = int(math.floor(x + 0.5)) if x >= 0 else -int(math.floor(-x + 0.5))   # C round(): half away from zero
    return 0 if b == HISTO_LENGTH else b


def three_maxima(hist):
    max1 = max2 = max3 = 0
    i1 = i2 = i3 = -1
    for i in range(HISTO_LENGTH):
        s = len(hist[i])
        if s > max1:
            max3, max2, max1, i3, i2, i1 = max2, max1, s, i2, i1, i
        elif s > max2:
            max3, max2, i3, i2 = max2, s, i2, i
        elif s > max3:
            max3, i3 = s, i
    if max2 < np.float32(0.1) * np.float32(max1):
        i2 = i3 = -1
    elif max3 < np.float32(0.1) * np.float32(max1):
        i3 = -1
    return i1, i2, i3


def py_search_by_bow(mode, k1, d1, fv1, v1, k2, d2, fv2, v2, ratio, ori):
    m12 = np.full(len(k1), -1, np.int32)
    m21 = np.full(len(k2), -1, np.int32)
    hist = [[] for _ in range(HISTO_LENGTH)]
    n = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if not v1[i1]:
                continue
            b1, b2, bi = 256, 256, -1
            for i2 in fv2[node]:
                if m21[i2] >= 0 or (mode == 1 and not v2[i2]):
                    continue
                dd = dist(d1[i1], d2[i2])
                if dd < b1:
                    b2, b1, bi = b1, dd, i2
                elif dd < b2:
                    b2 = dd
            ok = b1 <= TH_LOW if mode == 0 else b1 < TH_LOW
            if ok and np.float32(b1) < np.float32(ratio) * np.float32(b2):
                m12[i1], m21[bi] = bi, i1
                n += 1
                if ori:
                    hist[rot_bin(k1["angle"][i1], k2["angle"][bi])].append(i1)
    if ori:
        keep = three_maxima(hist)
        for b in range(HISTO_LENGTH):
            if b not in keep:
                for i1 in hist[b]:
                    m21[m12[i1]] = -1
                    m12[i1] = -1
                    n -= 1
    return m12, m21, n


def py_triangulation(k1, d1, fv1, h1, u1, k2, d2, fv2, h2, u2, F, ep, sc, sg, only_stereo, coarse, ori):
    f32 = np.float32
    m12 = np.full(len(k1), -1, np.int32)
    hist = [[] for _ in range(HISTO_LENGTH)]
    n = 0
    F = F.astype(np.float32)
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if h1[i1]:
                continue
            st1 = u1[i1] >= 0
            if only_stereo and not st1:
                continue
            best, bi = TH_LOW, -1
            x1, y1 = f32(k1["x"][i1]), f32(k1["y"][i1])
            for i2 in fv2[node]:
                if h2[i2]:
                    continue
                st2 = u2[i2] >= 0
                if only_stereo and not st2:
                    continue
                dd = dist(d1[i1], d2[i2])
                if dd > TH_LOW or dd > best:
                    continue
                x2, y2 = f32(k2["x"][i2]), f32(k2["y"][i2])
                if not st1 and not st2:
                    ex, ey = f32(ep[0]) - x2, f32(ep[1]) - y2
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(f32(100) * f32(sc[k2["octave"][i2]])):
                        continue
                if not coarse:
                    a = f32(f32(f32(x1 * F[0]) + f32(y1 * F[3])) + F[6])
                    b = f32(f32(f32(x1 * F[1]) + f32(y1 * F[4])) + F[7])
                    c = f32(f32(f32(x1 * F[2]) + f32(y1 * F[5])) + F[8])
                    num = f32(f32(f32(a * x2) + f32(b * y2)) + c)
                    den = f32(f32(a * a) + f32(b * b))
                    if den == 0:
                        continue
                    dsqr = f32(f32(num * num) / den)
                    if not float(dsqr) < 3.84 * float(f32(sg[k2["octave"][i2]])):
                        continue
                bi, best = i2, dd
            if bi >= 0:
                m12[i1] = bi
                n += 1
                if ori:
                    hist[rot_bin(k1["angle"][i1], k2["angle"][bi])].append(i1)
    if ori:
        keep = three_maxima(hist)
        for b in range(HISTO_LENGTH):
            if b not in keep:
                for i1 in hist[b]:
                    m12[i1] = -1
                    n -= 1
    return m12, n


def py_distinctive(desc):
    n = len(desc)
    best_med, best = 1 << 30, 0
    for i in range(n):
        row = sorted(0 if i == j else dist(desc[i], desc[j]) for j in range(n))
        med = row[int(0.5 * (n - 1))]
        if med < best_med:
            best_med, best = med, i
    return best


def fv_csr(fv):
    nodes = sorted(fv)
    off = np.cumsum([0] + [len(fv[k]) for k in nodes]).astype(np.int32)
    idx = np.concatenate([fv[k] for k in nodes]).astype(np.int32)
    return np.array(nodes, np.int32), off, idx


def main():
    g = np.load(OUT / "c1_mono.npz", allow_pickle=False)
    k1, d1, k2, d2 = g["kps0"], g["desc0"], g["kps1"], g["desc1"]
    voc = S.make_vocabulary(10, 4, seed=7, desc_pool=np.concatenate([d1, d2]))
    rng = np.random.default_rng(77)
    out = {"voc_k": np.array(10), "voc_L": np.array(4), "voc_seed": np.array(7), "pinned_by_python_restatement": np.array(True)}
    fvs = {}
    for levelsup in (2, 3):
        for tag, d in (("a", d1), ("b", d2)):
            w, nd, bw, bv, fv = py_transform(d, voc, levelsup)
            r = oracle.bow_transform(d, voc, levelsup)
            node, off, idx = fv_csr(fv)
            assert np.array_equal(w, r["feat_word"]) and np.array_equal(nd, r["feat_node"])
            assert np.array_equal(bw, r["bow_word"]) and np.array_equal(bv, r["bow_val"])        # fp64 bits
            assert all(np.array_equal(x, y) for x, y in zip((node, off, idx), r["fv"]))
            fvs[(levelsup, tag)] = (fv, (node, off, idx))
            out[f"word_{tag}_{levelsup}"], out[f"node_{tag}_{levelsup}"] = w, nd
            out[f"bow_word_{tag}_{levelsup}"], out[f"bow_val_{tag}_{levelsup}"] = bw, bv
    v1 = (rng.random(len(k1)) < 0.8).astype(np.uint8)
    v2 = (rng.random(len(k2)) < 0.8).astype(np.uint8)
    u1 = np.where(rng.random(len(k1)) < 0.5, k1["x"] - 20, -1).astype(np.float32)
    u2 = np.where(rng.random(len(k2)) < 0.5, k2["x"] - 20, -1).astype(np.float32)
    out.update(valid_a=v1, valid_b=v2, uright_a=u1, uright_b=u2)
    (fa, ca), (fb, cb) = fvs[(2, "a")], fvs[(2, "b")]
    for mode in (0, 1):
        for ratio, ori in ((0.6, True), (0.9, False)):
            m12, m21, n = py_search_by_bow(mode, k1, d1, fa, v1, k2, d2, fb, v2, ratio, ori)
            r12, r21, rn = oracle.search_by_bow(mode, k1, d1, ca, k2, d2, cb, ratio, ori, v1, v2)
            assert n == rn and np.array_equal(m12, r12) and np.array_equal(m21, r21), (mode, ratio)
            key = f"bow{mode}_{int(ratio * 100)}_{int(ori)}"
            out[key + "_m12"], out[key + "_m21"], out[key + "_n"] = m12, m21, np.array(n)
            print(key, n)
    t = np.array([2.3, 0.8, 0.02])
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    F = (tx / np.linalg.norm(t)).astype(np.float32).reshape(-1)
    ep = np.array([491.0, 280.0], np.float32)
    sc = g["scale_factors"].astype(np.float32)
    sg = (sc * sc).astype(np.float32)
    out.update(F12=F, ep=ep, scale=sc, sigma2=sg)
    h1, h2 = 1 - v1, 1 - v2
    for only_stereo, coarse in ((0, 0), (1, 0), (0, 1)):
        m12, n = py_triangulation(k1, d1, fa, h1, u1, k2, d2, fb, h2, u2, F, ep, sc, sg, only_stereo, coarse, True)
        r12, rn = oracle.search_for_triangulation(k1, d1, ca, k2, d2, cb, F, ep, sc, sg, only_stereo, coarse, True, h1,
                                                  h2, u1, u2)
        assert n == rn and np.array_equal(m12, r12), (only_stereo, coarse)
        out[f"tri_{only_stereo}{coarse}_m12"], out[f"tri_{only_stereo}{coarse}_n"] = m12, np.array(n)
        print("tri", only_stereo, coarse, n)
    sizes = [1, 2, 3, 4, 7, 16, 33]
    obs = [rng.integers(0, len(d1), s) for s in sizes]
    best = []
    for o in obs:
        b = py_distinctive(d1[o])
        assert b == oracle.distinctive_descriptor(d1[o])
        best.append(b)
    out["dd_off"] = np.cumsum([0] + sizes).astype(np.int32)
    out["dd_idx"] = np.concatenate(obs).astype(np.int32)
    out["dd_best"] = np.array(best, np.int32)
    np.savez_compressed(OUT / "bow_c1.npz", **out)
    print("wrote", OUT / "bow_c1.npz")


if __name__ == "__main__":
    main()
