"""oracle/make_golden.py — TEST INFRASTRUCTURE ONLY.

Generates tests/golden/bf_golden.npz by running the REFERENCE's own `EnhancedBFMatcher.match`
(/root/reference/FeatureMatchingExtraction/feature_matchers.py:184-252, with the installed cv2)
on small seeded inputs.  Run in the build container only:

    python oracle/make_golden.py

The reference ships no fixtures for this path; these vectors are "outputs of the
reference itself run here" and pin both the oracle (tests/test_oracle.py, CPU) and the CUDA path
(tests/test_parity_gpu.py) on a box where /root/reference does not exist.
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


def cases():
    rng = np.random.default_rng(1234)
    out = []
    # (name, desc1, desc2, norm, cross_check, ratio)
    a, b = synth.sift_pair(seed=10, n=300, m=257)
    out += [("sift_ratio075", a, b, NORM_L2, False, 0.75),
            ("sift_ratio07", a, b, NORM_L2, False, 0.7),
            ("sift_cross", a, b, NORM_L2, True, None),
            ("sift_nn", a, b, NORM_L2, False, None)]
    a, b = synth.binary_pair(seed=11, n=333, m=290, nbytes=32)
    out += [("orb_cross", a, b, NORM_HAMMING, True, None),
            ("orb_ratio07", a, b, NORM_HAMMING, False, 0.7),
            ("orb_nn", a, b, NORM_HAMMING, False, None)]
    a, b = synth.binary_pair(seed=12, n=200, m=260, nbytes=61)
    out += [("akaze61_cross", a, b, NORM_HAMMING, True, None),
            ("akaze61_ratio07", a, b, NORM_HAMMING, False, 0.7)]
    a, b = synth.binary_pair(seed=13, n=150, m=140, nbytes=64)
    out += [("brisk_cross", a, b, NORM_HAMMING, True, None)]
    # tie-heavy: 1-byte descriptors, duplicated rows, zero rows
    a = rng.integers(0, 4, size=(90, 1), dtype=np.uint8)
    b = rng.integers(0, 4, size=(70, 1), dtype=np.uint8)
    out += [("ties1byte_cross", a, b, NORM_HAMMING, True, None),
            ("ties1byte_ratio", a, b, NORM_HAMMING, False, 0.9),
            ("ties1byte_nn", a, b, NORM_HAMMING, False, None)]
    a = rng.integers(0, 3, size=(64, 8)).astype(np.float32)
    b = np.concatenate([a[:20], a[:20], np.zeros((5, 8), np.float32),
                        rng.integers(0, 3, size=(30, 8)).astype(np.float32)])
    out += [("ties_l2_cross", a, b, NORM_L2, True, None),
            ("ties_l2_ratio", a, b, NORM_L2, False, 0.95),
            ("ties_l2_nn", a, b, NORM_L2, False, None)]
    # degenerate shapes: M == 1 (k=2 yields one neighbour -> dropped by the len==2 test), N == 1
    a, b = synth.sift_pair(seed=14, n=17, m=17)
    out += [("m1_ratio", a, b[:1], NORM_L2, False, 0.75),
            ("m1_nn", a, b[:1], NORM_L2, False, None),
            ("m1_cross", a, b[:1], NORM_L2, True, None),
            ("n1_ratio", a[:1], b, NORM_L2, False, 0.75),
            ("m2_ratio", a, b[:2], NORM_L2, False, 0.99)]
    # non-integer float (SuperPoint-like unit rows): tolerance cases
    s = synth.superpoint_set(seed=15, images=2, n=200, d=256, pool=300)
    out += [("superpoint_cross", s[0], s[1], NORM_L2, True, None),
            ("superpoint_ratio08", s[0], s[1], NORM_L2, False, 0.8)]
    # error convention: dtype / width mismatch -> empty MatchData (feature_matchers.py:250-252)
    out += [("err_width", np.zeros((4, 16), np.uint8), np.zeros((4, 32), np.uint8), NORM_HAMMING, False, None)]
    return out


def main():
    if not ref_loader.available():
        raise SystemExit("reference not present; golden vectors can only be generated in the build container")
    import cv2
    blob = {"_cv2_version": np.array(cv2.__version__)}
    names = []
    for name, a, b, norm, cc, ratio in cases():
        q, t, d, _ = ref_loader.run_reference_matcher(a, b, norm, cc, ratio)
        names.append(name)
        blob[f"{name}/desc1"] = a
        blob[f"{name}/desc2"] = b
        blob[f"{name}/params"] = np.array([norm, int(cc), -1.0 if ratio is None else ratio], np.float64)
        blob[f"{name}/q"] = q
        blob[f"{name}/t"] = t
        blob[f"{name}/d"] = d
        print(f"{name:22s} N={len(a):4d} M={len(b):4d} -> {len(q)} matches")
    blob["_names"] = np.array(names)
    path = os.path.join(ROOT, "tests", "golden", "bf_golden.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
