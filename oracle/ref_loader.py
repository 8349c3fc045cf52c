"""oracle/ref_loader.py — TEST INFRASTRUCTURE ONLY.

Loads the reference's own hot-path modules from /root/reference WITHOUT modifying it, so the
oracle can be pinned against `EnhancedBFMatcher.match` itself and golden vectors can be generated.
Only usable in the build container (/root/reference does not exist on the GPU box): nothing in
`-m gpu` tests, smoke() or bench.py calls this at run time.

Two obstacles: `feature_matchers.py:13` does `from pyparsing import Path` (pyparsing
absent) and the package `__init__.py:161` imports matplotlib-based visualisation.  We register a
3-line `pyparsing` stub and load the needed modules through a synthetic package that skips
`__init__.py`.
"""
from __future__ import annotations

import importlib.util
import os
import pathlib
import sys
import types

REFERENCE_ROOT = os.environ.get("B200M_REFERENCE_ROOT", "/root/reference")
_PKG = "FeatureMatchingExtraction"
_MODULES = ("core_data_structures", "base_classes", "feature_matchers",
            "matcher_compatibility", "matcher_factory")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, _PKG, "feature_matchers.py"))


def load():
    """Return a namespace with the reference modules (core_data_structures, feature_matchers, ...)."""
    if not available():
        raise FileNotFoundError(f"reference not present under {REFERENCE_ROOT}")
    if "pyparsing" not in sys.modules:
        stub = types.ModuleType("pyparsing")
        stub.Path = pathlib.Path
        sys.modules["pyparsing"] = stub
    pkg_dir = os.path.join(REFERENCE_ROOT, _PKG)
    if _PKG not in sys.modules:
        pkg = types.ModuleType(_PKG)
        pkg.__path__ = [pkg_dir]
        pkg.__package__ = _PKG
        sys.modules[_PKG] = pkg
    ns = types.SimpleNamespace()
    for name in _MODULES:
        full = f"{_PKG}.{name}"
        if full not in sys.modules:
            spec = importlib.util.spec_from_file_location(full, os.path.join(pkg_dir, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[full] = mod
            spec.loader.exec_module(mod)
        setattr(ns, name, sys.modules[full])
    return ns


def run_reference_matcher(desc1, desc2, norm, cross_check=False, ratio_threshold=None):
    """Run the reference's EnhancedBFMatcher.match and flatten MatchData to arrays."""
    import cv2
    import numpy as np
    ref = load()
    cds = ref.core_data_structures

    def feats(d):
        n = 0 if d is None else len(d)
        return cds.FeatureData(keypoints=[cv2.KeyPoint(float(i), 0.0, 1.0) for i in range(n)],
                               descriptors=d, method="synthetic")
    md = ref.feature_matchers.EnhancedBFMatcher(norm_type=norm, cross_check=cross_check,
                                                ratio_threshold=ratio_threshold).match(feats(desc1), feats(desc2))
    q = np.array([m.queryIdx for m in md.matches], np.int32)
    t = np.array([m.trainIdx for m in md.matches], np.int32)
    d = np.array([m.score for m in md.matches], np.float32)
    return q, t, d, md
