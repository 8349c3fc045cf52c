"""CPU tests of the geometric-verification oracle: oracle/ransac_oracle.py against the committed
outputs of the reference's own enhanced_filter_matches_with_homography (tests/golden/ransac_golden.npz, written by
oracle/make_golden_ransac.py) and against live cv2.findHomography where cv2 is installed."""
import importlib
import os

import numpy as np
import pytest

from oracle import ransac_oracle as ro

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ransac_golden.npz")
synth = importlib.import_module("3dreconstruction_b200.synth")
# the pure-Python oracle spends ~5 ms per hypothesis: cases that run thousands of iterations are checked on the GPU side only
FAST = ["easy_500", "half_1000", "few_iters_200", "large_6000", "n5", "n4", "n3", "n12_outliers"]


def test_rng_stream_is_cvs():
    r = ro.RNG()
    first = [r.next() for _ in range(3)]
    # state_{k+1} = (uint32)state_k * 4164903690 + (state_k >> 32), from 2^64 - 1
    s = 0xFFFFFFFFFFFFFFFF
    want = []
    for _ in range(3):
        s = ((s & 0xFFFFFFFF) * 4164903690 + (s >> 32)) & 0xFFFFFFFFFFFFFFFF
        want.append(s & 0xFFFFFFFF)
    assert first == want
    assert ro.RNG(0).state == 0xFFFFFFFF


def test_update_num_iters():
    assert ro.ransac_update_num_iters(0.99, 0.5, 4, 2000) == 71
    assert ro.ransac_update_num_iters(0.99, 0.0, 4, 2000) == 0
    assert ro.ransac_update_num_iters(0.99, 1.0, 4, 2000) == 2000
    assert ro.ransac_update_num_iters(0.99, 0.9, 4, 2000) == 2000


def test_jacobi_against_numpy():
    rng = np.random.default_rng(1)
    for n in (8, 9):
        B = rng.normal(size=(n, n))
        A = B @ B.T
        w, V = ro.jacobi_eigen(A)
        assert np.all(np.diff(w) <= 0)
        assert np.allclose(np.sort(w), np.linalg.eigvalsh(A), rtol=1e-10, atol=1e-12)
        assert np.allclose(V @ A @ V.T, np.diag(w), atol=1e-9)


@pytest.mark.parametrize("name", FAST)
def test_oracle_reproduces_reference_golden(name):
    g = np.load(GOLD)
    c = g[name + "_corr"]
    thr, conf, iters = g[name + "_params"]
    kept, H = ro.reference_filter(c[:, :2], c[:, 2:], thr, conf, int(iters))
    assert np.array_equal(kept, g[name + "_kept"])
    Hg = g[name + "_H"]
    if np.isnan(Hg).any():
        assert H is None
    else:
        assert np.abs(H - Hg).max() <= 1e-8 * np.abs(Hg).max()


def test_oracle_against_live_cv2():
    cv2 = pytest.importorskip("cv2")
    for seed, (n, inl, thr) in enumerate([(300, 0.7, 5.0), (120, 0.55, 3.0), (60, 0.8, 4.0), (9, 0.9, 5.0)]):
        c = synth.homography_pair(seed=40 + seed, n=n, inlier_frac=inl)
        Hc, mc = cv2.findHomography(c[:, :2].reshape(-1, 1, 2), c[:, 2:].reshape(-1, 1, 2), cv2.RANSAC, thr,
                                    confidence=0.99, maxIters=2000)
        Ho, mo = ro.find_homography_ransac(c[:, :2], c[:, 2:], thr, 0.99, 2000)
        assert (Hc is None) == (Ho is None)
        if Hc is not None:
            assert np.array_equal(mc, mo)
            assert np.abs(Hc - Ho).max() <= 1e-8 * np.abs(Hc).max()


def test_c_oracle_reproduces_every_golden_case():
    """oracle/ransac_oracle.c on all 12 outputs of the reference's function, the iteration-capped ones included."""
    g = np.load(GOLD)
    for name in [str(n) for n in g["names"]]:
        c = g[name + "_corr"]
        thr, conf, iters = g[name + "_params"]
        H, mask, _ = ro.find_homography_ransac_c(c[:, :2], c[:, 2:], thr, conf, int(iters))
        Hg = g[name + "_H"]
        if np.isnan(Hg).any():
            assert H is None, name
            continue
        assert np.array_equal(np.flatnonzero(mask.ravel()), g[name + "_kept"]), name
        assert np.abs(H - Hg).max() <= 1e-8 * np.abs(Hg).max(), name


def test_c_oracle_equals_python_oracle():
    for seed, (n, inl, thr, iters) in enumerate([(150, 0.7, 4.0, 2000), (80, 0.5, 3.0, 300), (40, 0.9, 5.0, 2000), (6, 1.0, 5.0, 50)]):
        c = synth.homography_pair(seed=90 + seed, n=n, inlier_frac=inl)
        Hp, mp, itp = ro.ransac_homography(c[:, :2], c[:, 2:], thr, 0.99, iters)
        H2, m2 = ro.find_homography_ransac(c[:, :2], c[:, 2:], thr, 0.99, iters)
        Hc, mc, itc = ro.find_homography_ransac_c(c[:, :2], c[:, 2:], thr, 0.99, iters)
        assert itc == itp and np.array_equal(mc, m2) and np.abs(Hc - H2).max() <= 1e-6 * np.abs(H2).max()   # numpy sums in another order


def test_c_oracle_against_live_cv2_many():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for _ in range(150):
        c = synth.homography_pair(seed=int(rng.integers(1 << 30)), n=int(rng.integers(5, 900)), inlier_frac=float(rng.uniform(0.2, 0.95)))
        thr = float(rng.choice([1.5, 3.0, 5.0]))
        Hc, mc = cv2.findHomography(c[:, :2].reshape(-1, 1, 2), c[:, 2:].reshape(-1, 1, 2), cv2.RANSAC, thr, confidence=0.99, maxIters=2000)
        H, m, _ = ro.find_homography_ransac_c(c[:, :2], c[:, 2:], thr, 0.99, 2000)
        assert (Hc is None) == (H is None)
        if Hc is not None:
            assert np.array_equal(mc, m)
