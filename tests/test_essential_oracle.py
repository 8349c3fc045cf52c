"""Pins oracle/essential_oracle.{py,c} -- the CPU restatement of cv2.findEssentialMat(RANSAC) as the reference calls it
(/root/reference/CameraPoseEstimation/essential_estimation.py:118-126) -- against live cv2 piece by piece and as a whole,
and against tests/golden/essential_golden.npz (outputs of the reference's own EssentialMatrixEstimator.estimate,
oracle/make_golden_essential.py).  CPU only."""
import os

import numpy as np
import pytest

from conftest import ROOT, pkg
from oracle import essential_oracle as EO

cv2 = pytest.importorskip("cv2")
synth = pkg("synth")


def _same_E(a, b, tol=1e-6):
    a, b = a / np.linalg.norm(a), b / np.linalg.norm(b)
    return min(np.abs(a - b).max(), np.abs(a + b).max()) <= tol


def test_null_space_basis_is_opencvs_svd_completion():
    """The basis fixes the parametrisation and thereby the ORDER of the candidate models: rows 5..8 of cv's Vt."""
    rng = np.random.default_rng(0)
    for _ in range(20):
        Q = rng.normal(size=(5, 9))
        vt = cv2.SVDecomp(Q, flags=cv2.SVD_FULL_UV)[2]
        assert np.abs(vt[5:9] - EO.null_space_basis(Q)).max() < 1e-12


def test_solve_poly_is_opencvs_durand_kerner_bit_for_bit():
    rng = np.random.default_rng(1)
    for k in range(20):
        c = rng.normal(size=11) * 10.0 ** rng.integers(-3, 4)
        r = cv2.solvePoly(c.reshape(1, -1))[1].reshape(-1, 2)
        assert np.array_equal(r, np.array(EO.solve_poly(c))), k            # values AND order


def test_normalise_is_undistort_points_bit_for_bit():
    corr, K = synth.two_view_pair(seed=3, n=500)
    p = corr[:, :2].astype(np.float64)
    want = cv2.undistortPoints(p.reshape(-1, 1, 2), K, None).reshape(-1, 2)
    assert np.array_equal(EO.normalise(p, K), want)


def test_five_point_solver_c_equals_python_and_satisfies_the_constraints():
    rng = np.random.default_rng(2)
    for _ in range(12):
        m1, m2 = rng.uniform(-0.5, 0.5, size=(5, 2)), rng.uniform(-0.5, 0.5, size=(5, 2))
        a, c = EO.run_kernel(m1, m2), EO.five_point_c(m1, m2)
        assert len(a) == len(c)
        for x, y in zip(a, c):                                  # same models in the same order
            assert np.abs(x - y).max() < 1e-8
            x1, x2 = np.c_[m1, np.ones(5)], np.c_[m2, np.ones(5)]
            assert np.abs(np.einsum("ni,ij,nj->n", x2, y, x1)).max() < 1e-9                 # x2^T E x1 = 0 on the sample
            assert abs(np.linalg.det(y)) < 1e-9 and abs(np.linalg.norm(y) - 1) < 1e-12
            assert np.abs(2 * y @ y.T @ y - np.trace(y @ y.T) * y).max() < 1e-8


def _cases(count, base):
    for s in range(count):
        n = [6, 8, 20, 60, 150, 400, 1500][s % 7]
        inl = [0.9, 0.7, 0.5, 0.35, 0.2][s % 5]
        thr = [0.006, 0.002, 0.02, 1.5][s % 4] / 4.0
        iters = [300, 1000, 5000][s % 3]
        prob = [0.999, 0.99][s % 2]
        corr, K = synth.two_view_pair(seed=base + s, n=n, inlier_frac=inl, noise=0.8)
        yield s, EO.normalise(corr[:, :2], K), EO.normalise(corr[:, 2:], K), thr, prob, iters


def test_python_oracle_equals_cv2_on_small_cases():
    for s, m1, m2, thr, prob, iters in _cases(8, 5000):
        if len(m1) > 60:
            continue
        Ec, mc = cv2.findEssentialMat(m1, m2, focal=1.0, pp=(0.0, 0.0), method=cv2.RANSAC, prob=prob, threshold=thr, maxIters=min(iters, 60))
        Eo, mo, _ = EO.find_essential_mat(m1, m2, 1.0, (0.0, 0.0), prob, thr, min(iters, 60))
        assert (Ec is None) == (Eo is None)
        if Ec is not None:
            assert np.array_equal(mc.ravel(), mo) and _same_E(Ec[:3], Eo), s


def test_c_oracle_equals_cv2_masks_and_models():
    """The whole function against live cv2.findEssentialMat: inlier masks bit for bit, E up to sign -- including the
    all-inlier and n = 6 cases where candidate models TIE and the first one in cv's order must win."""
    bad = []
    for s, m1, m2, thr, prob, iters in _cases(140, 6000):
        Ec, mc = cv2.findEssentialMat(m1, m2, focal=1.0, pp=(0.0, 0.0), method=cv2.RANSAC, prob=prob, threshold=thr, maxIters=iters)
        Eo, mo, _ = EO.find_essential_mat_c(m1, m2, 1.0, (0.0, 0.0), prob, thr, iters)
        if (Ec is None) != (Eo is None) or (Ec is not None and not (np.array_equal(mc.ravel(), mo) and _same_E(Ec[:3], Eo))):
            bad.append(s)
    assert not bad, f"cases that differ from cv2: {bad}"


def test_golden_from_the_reference_estimator():
    z = np.load(os.path.join(ROOT, "tests", "golden", "essential_golden.npz"))
    for name in z["names"]:
        name = str(name)
        corr, size = z[name + "_corr"], tuple(int(v) for v in z[name + "_size"])
        thr, prob, iters = z[name + "_params"]
        c64 = corr.astype(np.float64)
        r = EO.reference_estimate(c64[:, :2], c64[:, 2:], size, size, float(thr), float(prob), int(iters))
        assert r["success"] == bool(z[name + "_success"]), name
        if r["success"]:
            assert np.array_equal(r["inlier_mask"], z[name + "_mask"]), name
            assert _same_E(r["essential_matrix"], z[name + "_E"]), name
        else:
            assert r["error"] == str(z[name + "_error"]), name            # the reference's message carries the inlier count / ratio
