"""Randomised GPU-vs-oracle sweep of the vocabulary-node functions: ragged sizes (including empty sides and empty
nodes lists), heavy distance ties, nodes larger than a warp, random object-state flags and geometry."""
import os

import numpy as np
import pytest

from orb_slam3_vio_fixes_b200 import _lib
from orb_slam3_vio_fixes_b200.matcher import ORBmatcher

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _side(rng, n, base, n_nodes, cap):
    d = base[rng.integers(0, len(base), n)].copy() if n else np.zeros((0, 32), np.uint8)
    for i in range(n):
        for _ in range(rng.integers(0, 4)):
            d[i, rng.integers(0, 32)] ^= np.uint8(1 << rng.integers(0, 8))
    k = np.zeros(n, _lib.KP_DTYPE)
    k["x"] = rng.uniform(0, 752, n).astype(np.float32)
    k["y"] = rng.uniform(0, 480, n).astype(np.float32)
    k["angle"] = rng.choice([10.0, 15.0, 200.0, 359.9, 0.0], n).astype(np.float32) + rng.uniform(0, 3, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n)
    member = rng.random(n) < 0.95            # stopped words: features missing from the FeatureVector
    nodes = rng.integers(0, n_nodes, n) * 7 + 3
    groups = {}
    for i in np.nonzero(member)[0]:
        groups.setdefault(int(nodes[i]), []).append(int(i))
    nl = sorted(groups)
    off = np.cumsum([0] + [len(groups[x]) for x in nl]).astype(np.int32)
    idx = np.concatenate([groups[x] for x in nl]).astype(np.int32) if nl else np.zeros(0, np.int32)
    flag = (rng.random(n) < 0.7).astype(np.uint8)
    ur = np.where(rng.random(n) < 0.4, rng.uniform(0, 700, n), -1.0).astype(np.float32)
    return dict(k=k, d=d, fv=(np.array(nl, np.int32), off, idx), flag=flag, ur=ur, n=n)


def _batch(sides, cap, ns):
    B = len(sides)
    kps = np.zeros((B, cap), _lib.KP_DTYPE)
    desc = np.zeros((B, cap, 32), np.uint8)
    flag = np.zeros((B, cap), np.uint8)
    ur = np.full((B, cap), -1, np.float32)
    n = np.zeros(B, np.int32)
    fv = {"node": np.zeros((B, ns), np.int32), "off": np.zeros((B, ns + 1), np.int32),
          "idx": np.zeros((B, cap), np.int32), "nn": np.zeros(B, np.int32)}
    for p, s in enumerate(sides):
        m = s["n"]
        kps[p, :m], desc[p, :m], flag[p, :m], ur[p, :m], n[p] = s["k"], s["d"], s["flag"], s["ur"], m
        node, off, idx = s["fv"]
        fv["node"][p, :len(node)] = node
        fv["off"][p, :len(off)] = off
        fv["off"][p, len(off):] = off[-1]
        fv["idx"][p, :len(idx)] = idx
        fv["nn"][p] = len(node)
    kt = _t(kps.view(np.float32).reshape(B, cap, 7))
    return kt, _t(desc), _t(n), {k_: _t(v) for k_, v in fv.items()}, _t(flag), _t(ur)


@pytest.mark.parametrize("seed", range(int(os.environ.get("ORBX_STRESS_SEEDS", "6"))))  # more seeds: stress run
def test_random_bow_and_triangulation(oracle_mod, seed):
    import torch
    rng = np.random.default_rng(1000 + seed)
    base = rng.integers(0, 256, (rng.integers(3, 60), 32), dtype=np.uint8)
    B = 12
    sizes = [(int(rng.integers(0, 900)), int(rng.integers(0, 900))) for _ in range(B)]
    sizes[0], sizes[1], sizes[2] = (0, 50), (50, 0), (1, 1)
    n_nodes = int(rng.choice([1, 4, 30, 200]))
    cap = 900
    s1 = [_side(rng, a, base, n_nodes, cap) for a, _ in sizes]
    s2 = [_side(rng, b, base, n_nodes, cap) for _, b in sizes]
    ns = max(1, max(len(s["fv"][0]) for s in s1 + s2))
    b1, b2 = _batch(s1, cap, ns), _batch(s2, cap, ns)
    sf = np.array([1.2 ** i for i in range(8)], np.float32)
    sg = (sf * sf).astype(np.float32)
    for mode in (0, 1):
        ratio, ori = float(rng.choice([0.6, 0.75, 0.9])), bool(rng.integers(0, 2))
        out = ORBmatcher.search_by_bow(mode, b1[0], b1[1], b1[2], b1[3], b2[0], b2[1], b2[2], b2[3], ratio, ori,
                                       b1[4], b2[4])
        torch.cuda.synchronize()
        for p in range(B):
            a, b = s1[p], s2[p]
            m12, m21, nm = oracle_mod.search_by_bow(mode, a["k"], a["d"], a["fv"], b["k"], b["d"], b["fv"], ratio, ori,
                                                    a["flag"], b["flag"])
            assert int(out["nmatches"][p]) == nm, (mode, p)
            assert np.array_equal(out["match12"][p, :a["n"]].cpu().numpy(), m12), (mode, p)
            assert np.array_equal(out["match21"][p, :b["n"]].cpu().numpy(), m21), (mode, p)
    F = rng.normal(0, 1, (B, 9)).astype(np.float32) * np.float32(1e-2)
    F[:, [5, 7]] = [-0.9, 0.9]
    ep = rng.uniform(0, 700, (B, 2)).astype(np.float32)
    for only_stereo, coarse in ((0, 0), (1, 0), (0, 1)):
        ori = bool(rng.integers(0, 2))
        out = ORBmatcher.search_for_triangulation(b1[0], b1[1], b1[2], b1[3], b2[0], b2[1], b2[2], b2[3], _t(F), _t(ep), sf,
                                                  sg, only_stereo, coarse, ori, b1[4], b2[4], b1[5], b2[5])
        torch.cuda.synchronize()
        for p in range(B):
            a, b = s1[p], s2[p]
            m12, nm = oracle_mod.search_for_triangulation(a["k"], a["d"], a["fv"], b["k"], b["d"], b["fv"], F[p], ep[p], sf,
                                                          sg, only_stereo, coarse, ori, a["flag"], b["flag"], a["ur"],
                                                          b["ur"])
            assert int(out["nmatches"][p]) == nm, (only_stereo, coarse, p)
            assert np.array_equal(out["match12"][p, :a["n"]].cpu().numpy(), m12), (only_stereo, coarse, p)
