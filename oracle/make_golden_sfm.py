"""oracle/make_golden_sfm.py — TEST INFRASTRUCTURE ONLY.

Generates tests/golden/sfm_golden.npz: what the REFERENCE itself produces for the hand-off from matching to
structure-from-motion on a small seeded image set, by running, for every pair (i < j),
  EnhancedBFMatcher.match                     /root/reference/FeatureMatchingExtraction/feature_matchers.py:201-252
  -> MethodResult(features1, features2, match_data)                                   result_types.py:58-80
  -> ResultConverter._convert_method  (query/train indices, scores, N x 4 correspondences)  result_converters.py:730-783
with cv2.KeyPoint lists as the reference's detectors produce them.  Run in the build container only:

    python oracle/make_golden_sfm.py
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
NORM_L2, NORM_HAMMING = 4, 6


def image_sets():
    """name -> (norm, cross_check, ratio, [descriptors per image], [keypoint xy float32 per image])"""
    rng = np.random.default_rng(77)
    out = {}
    sizes = [220, 131, 1, 300, 64]
    orb = [synth.binary_pair(seed=300 + i, n=max(s, 2), nbytes=32)[0][:s] for i, s in enumerate(sizes)]
    orb[3][:120] = orb[0][:120]
    orb[3][:120, 1] ^= 3
    orb[1][:60] = orb[0][60:120]
    sift = [synth.sift_pair(seed=320 + i, n=max(s, 2))[0][:s] for i, s in enumerate(sizes)]
    sift[3][:150] = np.clip(sift[0][:150] + rng.integers(-3, 4, size=(150, 128)), 0, 255)
    sift[4][:40] = np.clip(sift[1][:40] + rng.integers(-2, 3, size=(40, 128)), 0, 255)
    xy = [np.stack([rng.uniform(0, 1600, s), rng.uniform(0, 1200, s)], axis=1).astype(np.float32) for s in sizes]
    out["orb_cross"] = (NORM_HAMMING, True, None, orb, xy)
    out["orb_ratio07"] = (NORM_HAMMING, False, 0.7, orb, xy)
    out["sift_ratio075"] = (NORM_L2, False, 0.75, sift, xy)
    return out


def main():
    import cv2
    ref_loader._MODULES = tuple(dict.fromkeys(ref_loader._MODULES + ("utils", "result_types", "result_converters")))
    ref = ref_loader.load()
    cds, rt, rc = ref.core_data_structures, ref.result_types, ref.result_converters
    conv = rc.ResultConverter()
    blob = {"_sets": []}
    for name, (norm, cc, ratio, descs, xys) in image_sets().items():
        feats = [cds.FeatureData(keypoints=[cv2.KeyPoint(float(x), float(y), 1.0) for x, y in xy], descriptors=d, method=name)
                 for d, xy in zip(descs, xys)]
        matcher = ref.feature_matchers.EnhancedBFMatcher(norm_type=norm, cross_check=cc, ratio_threshold=ratio)
        blob["_sets"].append(name)
        blob[f"{name}/params"] = np.array([norm, int(cc), -1.0 if ratio is None else ratio], np.float64)
        blob[f"{name}/count"] = np.array([len(d) for d in descs], np.int64)
        blob[f"{name}/desc"] = np.concatenate(descs)
        blob[f"{name}/xy"] = np.concatenate(xys)
        I = len(descs)
        for i in range(I):
            for j in range(i + 1, I):
                md = matcher.match(feats[i], feats[j])
                mr = rt.MethodResult(method_name=name, features1=feats[i], features2=feats[j], match_data=md)
                r = conv._convert_method(mr)
                k = f"{name}/{i}_{j}"
                blob[k + "/q"] = np.asarray(r.query_indices, np.int32)
                blob[k + "/t"] = np.asarray(r.train_indices, np.int32)
                blob[k + "/s"] = np.asarray(r.match_scores, np.float64)
                c = r.correspondences
                blob[k + "/c"] = np.zeros((0, 4), np.float64) if c is None else np.asarray(c, np.float64)
    blob["_sets"] = np.array(blob["_sets"])
    path = os.path.join(ROOT, "tests", "golden", "sfm_golden.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
