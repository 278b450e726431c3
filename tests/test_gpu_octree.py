"""GPU parity of the octree keypoint distribution (DistributeOctTree, ORBextractor.cc:555-779) on
adversarial key sets through the orbx_debug_octree hook: response ties, clusters, regular lattices
(ties in compareNodes that only libstdc++'s std::sort order decides), tiny and huge N."""
import numpy as np
import pytest

from orb_slam3_vio_fixes_b200.extractor import ORBextractor

pytestmark = pytest.mark.gpu


def _check(oracle_mod, ex, orc, width, height, xys):
    N = int(orc.features_per_level[0])
    got = ex.debug_octree(0, xys)
    ref = oracle_mod.distribute(xys.astype(np.float32), 16, 16 + width, 16, 16 + height, N).astype(np.int32)
    assert len(got) == len(ref), (N, len(got), len(ref))
    assert np.array_equal(got, ref), N


@pytest.mark.parametrize("seed", range(8))
def test_octree_random_stress(oracle_mod, seed):
    rng = np.random.default_rng(100 + seed)
    W, H = [(752, 480), (640, 480), (1241, 376), (400, 400)][seed % 4]
    width, height = W - 32, H - 32
    n = int(rng.integers(1, 6000))
    if seed % 3 == 0:   # clustered
        cx, cy = rng.integers(0, width, 6), rng.integers(0, height, 6)
        pick = rng.integers(0, 6, n)
        x = np.clip(cx[pick] + rng.integers(-25, 25, n), 0, width - 5)
        y = np.clip(cy[pick] + rng.integers(-25, 25, n), 0, height - 5)
    else:
        x, y = rng.integers(0, width - 4, n), rng.integers(0, height - 4, n)
    flat = np.unique(y.astype(np.int64) * 8192 + x)   # distinct pixels like real FAST output
    rng.shuffle(flat)
    xys = np.ascontiguousarray(np.stack([flat % 8192, flat // 8192, rng.integers(7, 40, len(flat))], 1).astype(np.int32))
    for nfeat in [5, 60, 300, 1000, 3000]:
        ex = ORBextractor(nfeat, 1.2, 4, 20, 7).plan(W, H)
        orc = oracle_mod.Extractor(nfeat, 1.2, 4, 20, 7)
        _check(oracle_mod, ex, orc, width, height, xys)
        ex.close()


def test_octree_regular_grid_many_ties(oracle_mod):
    xs, ys = np.meshgrid(np.arange(4, 700, 9), np.arange(4, 440, 9))
    xys = np.ascontiguousarray(np.stack([xs.ravel(), ys.ravel(), np.full(xs.size, 20)], 1).astype(np.int32))
    for nfeat in [300, 1000, 2500, 5000]:
        ex = ORBextractor(nfeat, 1.2, 4, 20, 7).plan(752, 480)
        orc = oracle_mod.Extractor(nfeat, 1.2, 4, 20, 7)
        _check(oracle_mod, ex, orc, 720, 448, xys)
        ex.close()


def test_octree_degenerate_inputs(oracle_mod):
    ex = ORBextractor(500, 1.2, 4, 20, 7).plan(752, 480)
    orc = oracle_mod.Extractor(500, 1.2, 4, 20, 7)
    for xys in ([[10, 10, 30]], [[10, 10, 30], [11, 10, 30]], [[0, 0, 9], [719, 447, 9]],
                [[x, 200, 15] for x in range(0, 700)],          # one row of pixels
                [[300, y, 15] for y in range(0, 440)]):         # one column
        _check(oracle_mod, ex, orc, 720, 448, np.array(xys, np.int32))
    assert len(ex.debug_octree(0, np.zeros((0, 3), np.int32))) == 0
