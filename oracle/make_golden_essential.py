"""oracle/make_golden_essential.py — TEST INFRASTRUCTURE ONLY.

Generates tests/golden/essential_golden.npz: what the REFERENCE itself returns for the essential-matrix estimation that
opens its pose stage on seeded two-view correspondence sets, by running
  EssentialMatrixEstimator.estimate(pts1, pts2, size1, size2, method='RANSAC')
                                               /root/reference/CameraPoseEstimation/essential_estimation.py:62-186
(camera matrix estimated from the image size :36-60, cv2.undistortPoints :98-108, cv2.findEssentialMat RANSAC :118-126)
with the reference's own configuration (threshold 1.5, prob 0.999, 5000 iterations, :10-12) and with the threshold the
reference's refinement loop sets on the same config object (iterative_refinement_for_camera.py:227).  Points are passed as
float64 arrays holding float32 pixel values, which is what result_converters.py:763-766 builds from cv2.KeyPoint.pt.
Run in the build container only (needs /root/reference and cv2):

    python oracle/make_golden_essential.py
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dreconstruction_b200.synth")
REF = os.path.join(os.environ.get("B200M_REFERENCE_ROOT", "/root/reference"), "CameraPoseEstimation", "essential_estimation.py")

# (name, n, inlier fraction, noise px, threshold (None = the reference's 1.5), prob, max_iters, image size)
CASES = [
    ("ref_default_600", 600, 0.7, 0.7, None, None, None, (1600, 1200)),      # threshold 1.5 in NORMALISED units: every point is an
    ("ref_default_40", 40, 0.5, 1.0, None, None, None, (1024, 768)),         # inlier, the first model of the first sample wins
    ("px1_800", 800, 0.7, 0.6, 1.0 / 1920, None, None, (1600, 1200)),         # ~1 pixel
    ("px2_half_1000", 1000, 0.5, 1.0, 2.0 / 1920, None, None, (1600, 1200)),
    ("px1_hard_500", 500, 0.42, 0.8, 1.0 / 1920, None, None, (1600, 1200)),
    ("px1_capped_400", 400, 0.3, 0.8, 1.0 / 1920, None, 600, (1600, 1200)),  # runs into maxIters; fails the reference's quality gate
    ("px3_few_200", 200, 0.6, 1.5, 3.0 / 1228.8, 0.99, 150, (1024, 768)),
    ("px1_large_5000", 5000, 0.8, 0.5, 1.0 / 1920, None, None, (1600, 1200)),
    ("n8", 8, 1.0, 0.2, 1.0 / 1920, None, None, (1600, 1200)),
    ("n7_too_few", 7, 1.0, 0.2, 1.0 / 1920, None, None, (1600, 1200)),       # < 8 points: the reference refuses
    ("n60_low_ratio", 60, 0.25, 1.0, 1.0 / 1920, None, None, (1600, 1200)),
]


def main():
    spec = importlib.util.spec_from_file_location("ref_essential_estimation", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {"names": np.array([c[0] for c in CASES])}
    for seed, (name, n, inl, noise, thr, prob, iters, size) in enumerate(CASES):
        corr, K = synth.two_view_pair(seed=700 + seed, n=n, inlier_frac=inl, noise=noise, size=size)
        est = mod.EssentialMatrixEstimator()
        if thr is not None:
            est.config.RANSAC_THRESHOLD = thr
        if prob is not None:
            est.config.RANSAC_CONFIDENCE = prob
        if iters is not None:
            est.config.RANSAC_MAX_ITERATIONS = iters
        c64 = corr.astype(np.float64)
        r = est.estimate(c64[:, :2].copy(), c64[:, 2:].copy(), size, size, method="RANSAC")
        out[name + "_corr"] = corr
        out[name + "_size"] = np.array(size, np.int64)
        out[name + "_params"] = np.array([est.config.RANSAC_THRESHOLD, est.config.RANSAC_CONFIDENCE, est.config.RANSAC_MAX_ITERATIONS], np.float64)
        out[name + "_success"] = np.array(bool(r["success"]))
        out[name + "_error"] = np.array(r.get("error", ""))
        if r["success"]:
            out[name + "_E"] = np.asarray(r["essential_matrix"], np.float64)
            out[name + "_mask"] = np.asarray(r["inlier_mask"]).ravel().astype(np.uint8)
            assert np.allclose(r["camera_matrices"][0], K) and r["camera_estimated"] == [True, True]
        print(f"{name}: n={n} success={r['success']} inliers={r.get('num_inliers')} {r.get('error', '')}")
    path = os.path.join(ROOT, "tests", "golden", "essential_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
