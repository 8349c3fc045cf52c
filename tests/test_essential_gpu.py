"""GPU parity of the essential-matrix RANSAC (csrc/essential.cu through the C-ABI, b200m_ransac_essential) against
  * tests/golden/essential_golden.npz -- outputs of the reference's own EssentialMatrixEstimator.estimate
    (/root/reference/CameraPoseEstimation/essential_estimation.py:62-186; oracle/make_golden_essential.py),
  * oracle/essential_oracle.c, itself pinned against live cv2.findEssentialMat (tests/test_essential_oracle.py),
  * live cv2.findEssentialMat where cv2 is importable on the box.
Bar: inlier MASK bit for bit, inlier count and RANSAC iterations run equal, E equal up to sign to 1e-9 (the kernel keeps
the oracle's order of operations; the mask is an integer result of float32 comparisons)."""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, pkg
from oracle import essential_oracle as EO

pytestmark = pytest.mark.gpu
b = pkg()
pose = pkg("pose")
synth = b.synth


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return b.Engine()


def _same_E(a, c, tol=1e-9):
    a, c = a / np.linalg.norm(a), c / np.linalg.norm(c)
    return min(np.abs(a - c).max(), np.abs(a + c).max()) <= tol


def _cam8(K):
    return np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]] * 2, np.float64)


def _oracle(corr, K, thr, prob, iters):
    return EO.find_essential_mat_c(EO.normalise(corr[:, :2], K), EO.normalise(corr[:, 2:], K), 1.0, (0.0, 0.0), prob, thr, iters)


def test_golden_from_the_reference_estimator(eng):
    z = np.load(os.path.join(ROOT, "tests", "golden", "essential_golden.npz"))
    for name in z["names"]:
        name = str(name)
        corr, size = z[name + "_corr"], tuple(int(v) for v in z[name + "_size"])
        thr, prob, iters = z[name + "_params"]
        est = pose.EssentialMatrixEstimator(engine=eng)
        est.config.RANSAC_THRESHOLD, est.config.RANSAC_CONFIDENCE, est.config.RANSAC_MAX_ITERATIONS = float(thr), float(prob), int(iters)
        c64 = corr.astype(np.float64)
        r = est.estimate(c64[:, :2], c64[:, 2:], size, size, method="RANSAC")
        assert r["success"] == bool(z[name + "_success"]), name
        if r["success"]:
            assert np.array_equal(r["inlier_mask"].ravel(), z[name + "_mask"]), name
            assert _same_E(r["essential_matrix"], z[name + "_E"], 1e-6), name
            assert r["camera_estimated"] == [True, True] and r["num_inliers"] == int(z[name + "_mask"].sum())
        else:
            assert r["error"] == str(z[name + "_error"]), name


def _seeded(count, base):
    for s in range(count):
        n = [6, 8, 20, 60, 150, 400, 1500, 5][s % 8]
        inl = [0.9, 0.7, 0.5, 0.35, 0.2][s % 5]
        thr = [0.006, 0.002, 0.02, 1.5][s % 4] / 4.0
        iters = [300, 1000, 5000][s % 3]
        prob = [0.999, 0.99][s % 2]
        corr, K = synth.two_view_pair(seed=base + s, n=n, inlier_frac=inl, noise=0.8)
        yield s, corr, K, thr, prob, iters


def test_single_pairs_equal_the_oracle(eng):
    bad = []
    for s, corr, K, thr, prob, iters in _seeded(48, 8000):
        Eo, mo, ito = _oracle(corr, K, thr, prob, iters)
        r = pose.estimate_pairs(corr, np.array([0, len(corr)], np.int64), _cam8(K)[None], thr, prob, iters, engine=eng)
        if Eo is None:
            ok = r.inliers[0] < 0 and np.isnan(r.E[0]).all() and not r.mask.any()
        else:
            ok = (r.inliers[0] == int(mo.sum()) and np.array_equal(r.mask, mo) and _same_E(r.E[0], Eo) and int(r.iters[0]) == ito
                  and abs(np.linalg.norm(r.E[0]) - 1) < 1e-12)
        if not ok:
            bad.append((s, len(corr), int(r.inliers[0]), None if mo is None else int(mo.sum()), int(r.iters[0]), ito))
    assert not bad, f"(seed, n, gpu inliers, oracle inliers, gpu iters, oracle iters): {bad}"


def test_batch_in_one_launch_is_deterministic_and_independent_of_neighbours(eng):
    cases = list(_seeded(40, 9000))
    corr = np.concatenate([c[1] for c in cases])
    ptr = np.concatenate([[0], np.cumsum([len(c[1]) for c in cases])]).astype(np.int64)
    cams = np.stack([_cam8(c[2]) for c in cases])
    thr, prob, iters = 0.0012, 0.999, 800
    r1 = pose.estimate_pairs(corr, ptr, cams, thr, prob, iters, engine=eng)
    r2 = pose.estimate_pairs(corr, ptr, cams, thr, prob, iters, engine=eng)
    assert np.array_equal(r1.mask, r2.mask) and np.array_equal(r1.inliers, r2.inliers) and np.array_equal(r1.iters, r2.iters)
    assert np.array_equal(r1.E, r2.E, equal_nan=True)
    for p, (_, c, K, *_rest) in enumerate(cases):
        Eo, mo, ito = _oracle(c, K, thr, prob, iters)
        if Eo is None:
            assert r1.inliers[p] < 0, p
            continue
        assert np.array_equal(r1.inlier_mask(p), mo) and _same_E(r1.E[p], Eo) and int(r1.iters[p]) == ito, p
        solo = pose.estimate_pairs(c, np.array([0, len(c)], np.int64), cams[p][None], thr, prob, iters, engine=eng)
        assert np.array_equal(solo.mask, r1.inlier_mask(p)) and np.array_equal(solo.E[0], r1.E[p])


def test_degenerate_inputs(eng):
    corr, K = synth.two_view_pair(seed=5, n=30, inlier_frac=1.0, noise=0.1)
    ptr = np.array([0, 0, 3, 4, 9, 9, 30], np.int64)                  # empty, 3, 1, 5 (exactly the model size), empty, 21 points
    r = pose.estimate_pairs(corr, ptr, np.tile(_cam8(K), (6, 1)), 0.001, 0.999, 500, engine=eng)
    assert list(r.inliers[:3]) == [-1, -1, -1] and r.inliers[4] == -1 and np.isnan(r.E[0]).all()
    Eo, mo, _ = _oracle(corr[4:9], K, 0.001, 0.999, 500)              # 5 points: one solver run, its first model, all inliers
    assert r.inliers[3] == 5 and r.inlier_mask(3).all() and _same_E(r.E[3], Eo)
    Eo, mo, _ = _oracle(corr[9:30], K, 0.001, 0.999, 500)
    assert np.array_equal(r.inlier_mask(5), mo) and _same_E(r.E[5], Eo)
    same = np.tile(corr[:1], (12, 1))                                  # twelve copies of one correspondence: no sample has a model
    r = pose.estimate_pairs(same, np.array([0, 12], np.int64), _cam8(K)[None], 0.001, 0.999, 50, engine=eng)
    Eo, mo, _ = _oracle(same, K, 0.001, 0.999, 50)
    assert (r.inliers[0] < 0) == (Eo is None)
    already = np.concatenate([EO.normalise(corr[:, :2], K), EO.normalise(corr[:, 2:], K)], axis=1).astype(np.float32)
    r = pose.estimate_pairs(already, np.array([0, 30], np.int64), None, 0.001, 0.999, 500, engine=eng)     # cam = NULL: no normalisation
    Eo, mo, _ = EO.find_essential_mat_c(already[:, :2].astype(np.float64), already[:, 2:].astype(np.float64), 1.0, (0.0, 0.0), 0.999, 0.001, 500)
    assert np.array_equal(r.mask, mo) and _same_E(r.E[0], Eo)


def test_against_live_cv2(eng):
    cv2 = pytest.importorskip("cv2")
    bad = 0
    for s, corr, K, thr, prob, iters in _seeded(60, 10000):
        if len(corr) < 6:
            continue
        m1, m2 = EO.normalise(corr[:, :2], K), EO.normalise(corr[:, 2:], K)
        Ec, mc = cv2.findEssentialMat(m1, m2, focal=1.0, pp=(0.0, 0.0), method=cv2.RANSAC, prob=prob, threshold=thr, maxIters=iters)
        r = pose.estimate_pairs(corr, np.array([0, len(corr)], np.int64), _cam8(K)[None], thr, prob, iters, engine=eng)
        if Ec is None:
            bad += r.inliers[0] >= 0
        else:
            bad += not (np.array_equal(r.mask, mc.ravel()) and _same_E(r.E[0], Ec[:3], 1e-6))
    assert bad == 0


def test_pipeline_match_gather_essential_on_the_device(eng):
    """N1 -> N2 -> N3 without leaving the GPU: sfm.match_feature_set(..., essential=...) on a small image set; every pair's E
    and mask equal cv2's route on the host (the oracle on the gathered correspondences with the camera from the image size)."""
    sfm = pkg("sfm")
    rng = np.random.default_rng(3)
    n, size = 400, (1600, 1200)
    corr, K = synth.two_view_pair(seed=77, n=n, inlier_frac=0.8, noise=0.5, size=size)
    base = synth.binary_pair(seed=78, n=n, nbytes=32)[0]
    F = b.contracts.FeatureData
    feats = {}
    for name, xy in (("a", corr[:, :2]), ("b", corr[:, 2:]), ("c", corr[rng.permutation(n), :2])):
        feats[name] = F(keypoints=xy.copy(), descriptors=base.copy(), method="ORB")     # identical descriptors: match i <-> i
    fs = sfm.build_feature_set(eng, feats, b.NORM_HAMMING, {k: size for k in feats})
    res, est = sfm.match_feature_set(eng, fs, None, True, None, essential=(1.0 / 1920, 0.999, 1000))
    assert len(res) == 3 and len(est.inliers) == 3
    for p in range(3):
        c = res.correspondences[res.pair_slice(p)]
        Eo, mo, ito = _oracle(c, EO.camera_matrix_from_size(size), 1.0 / 1920, 0.999, 1000)
        if Eo is None:
            assert est.inliers[p] < 0
        else:
            assert np.array_equal(est.inlier_mask(p), mo) and _same_E(est.E[p], Eo) and int(est.iters[p]) == ito
    assert est.inliers[0] > 250                                   # images a and b are the two views of one scene
