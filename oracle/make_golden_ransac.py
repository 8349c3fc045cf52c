"""oracle/make_golden_ransac.py — TEST INFRASTRUCTURE ONLY.

Generates tests/golden/ransac_golden.npz: what the REFERENCE itself returns for the geometric verification after
matching on seeded correspondence sets, by running
  enhanced_filter_matches_with_homography       /root/reference/FeatureMatchingExtraction/utils.py:118-153
(cv2.KeyPoint lists + a MatchData of EnhancedDMatch, as FeatureProcessingPipeline._apply_filtering passes them,
pipeline.py:299-320).  Run in the build container only (needs /root/reference and cv2):

    python oracle/make_golden_ransac.py
"""
from __future__ import annotations

import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

synth = importlib.import_module("3dreconstruction_b200.synth")

# (name, n, inlier fraction, noise px, threshold, confidence, max_iters)
CASES = [
    ("easy_500", 500, 0.8, 0.7, 5.0, 0.99, 2000),
    ("half_1000", 1000, 0.5, 1.0, 5.0, 0.99, 2000),
    ("hard_300", 300, 0.25, 1.0, 4.0, 0.99, 2000),
    ("capped_400", 400, 0.12, 1.0, 5.0, 0.99, 2000),      # runs into max_iters
    ("few_iters_200", 200, 0.35, 1.5, 3.0, 0.995, 150),
    ("tight_thr_800", 800, 0.6, 2.0, 1.5, 0.99, 2000),
    ("large_6000", 6000, 0.7, 1.0, 4.0, 0.99, 2000),
    ("n5", 5, 1.0, 0.2, 5.0, 0.99, 2000),
    ("n4", 4, 1.0, 0.0, 5.0, 0.99, 2000),
    ("n3", 3, 1.0, 0.0, 5.0, 0.99, 2000),
    ("n12_outliers", 12, 0.0, 0.0, 5.0, 0.99, 2000),
    ("n40_low", 40, 0.2, 1.0, 5.0, 0.99, 2000),
]


def main():
    import cv2
    ref_loader._MODULES = tuple(dict.fromkeys(ref_loader._MODULES + ("utils",)))
    ref = ref_loader.load()
    cds, utils = ref.core_data_structures, ref.utils
    out = {"names": np.array([c[0] for c in CASES])}
    for seed, (name, n, inl, noise, thr, conf, iters) in enumerate(CASES):
        c = synth.homography_pair(seed=900 + seed, n=n, inlier_frac=inl, noise=noise)
        kp1 = [cv2.KeyPoint(float(x), float(y), 1.0) for x, y in c[:, :2]]
        kp2 = [cv2.KeyPoint(float(x), float(y), 1.0) for x, y in c[:, 2:]]
        ms = [cds.EnhancedDMatch(queryIdx=i, trainIdx=i, score=float(i), score_type=cds.ScoreType.DISTANCE) for i in range(n)]
        md = cds.MatchData(matches=ms, method="BF", score_type=cds.ScoreType.DISTANCE)
        kept, H = utils.enhanced_filter_matches_with_homography(kp1, kp2, ms, md, ransac_threshold=thr, confidence=conf,
                                                                max_iters=iters)
        out[name + "_corr"] = c
        out[name + "_params"] = np.array([thr, conf, iters], np.float64)
        out[name + "_kept"] = np.array([m.queryIdx for m in kept], np.int32)
        out[name + "_H"] = np.full((3, 3), np.nan) if H is None else np.asarray(H, np.float64)
        print(f"{name}: n={n} kept={len(kept)} H={'None' if H is None else 'ok'}")
    path = os.path.join(ROOT, "tests", "golden", "ransac_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
