"""`B200BFMatcher` — drop-in for the reference's `EnhancedBFMatcher`
(/root/reference/FeatureMatchingExtraction/feature_matchers.py:184-252): same constructor
arguments (:187-188), same `match(features1, features2) -> MatchData` (base_classes.py:80), same
failure convention (:204-205 invalid -> empty MatchData; :250-252 exceptions printed and swallowed),
same result order (:241).  The arithmetic runs in libb200match.so on the GPU.

`install()` registers it where the reference looks the class up at call time
(matcher_factory.py:234 `_create_bf_matcher`, feature_matchers.py:567-570 `create_traditional_matcher`,
:643/:666 `auto_select_matcher`) by rebinding the module globals — no edit to the reference tree.
"""
from __future__ import annotations

import sys
import time
from typing import Optional

import numpy as np

from . import contracts
from ._lib import ECUDA, ENODEVICE, EUNSUPPORTED, EWORKSPACE, B200MatchError
from .engine import NORM_HAMMING, NORM_L2, Engine, MatchArrays, UnsupportedNormError, default_engine


def _materialise(c: contracts.Contracts, arr: MatchArrays) -> list:
    EDM, dist_t = c.EnhancedDMatch, c.ScoreType.DISTANCE
    q = arr.query_idx.tolist()
    t = arr.train_idx.tolist()
    d = arr.distance.astype(np.float64).tolist()      # float32 widened to Python float, as cv2.DMatch.distance
    # Thousands of small objects per pair: the cyclic collector's generation-0 passes, triggered by the allocation count,
    # cost 3x the construction itself (measured 4.8 -> 1.4 ms for 4096 matches); nothing here can form a cycle.
    import gc
    paused = gc.isenabled()
    if paused:
        gc.disable()
    try:
        if _positional_ok(EDM):                       # positional construction is ~3x cheaper than keywords
            return [EDM(qi, ti, di, dist_t, 0, di) for qi, ti, di in zip(q, t, d)]
        return [EDM(queryIdx=qi, trainIdx=ti, score=di, score_type=dist_t, raw_distance=di)
                for qi, ti, di in zip(q, t, d)]
    finally:
        if paused:
            gc.enable()


_POSITIONAL: dict = {}


def _positional_ok(EDM) -> bool:
    """True when the class lays its fields out as core_data_structures.py:64-74 does (queryIdx, trainIdx, score, score_type,
    imgIdx, raw_distance, ...), so the first six can be passed by position."""
    ok = _POSITIONAL.get(EDM)
    if ok is None:
        import dataclasses
        try:
            names = [f.name for f in dataclasses.fields(EDM)][:6]
        except TypeError:
            names = []
        ok = _POSITIONAL[EDM] = names == ["queryIdx", "trainIdx", "score", "score_type", "imgIdx", "raw_distance"]
    return ok


class B200BFMatcher(contracts.BaseFeatureMatcher):
    """Brute-force matcher on B200: knnMatch k=2 + Lowe ratio / mutual cross-check / plain 1-NN."""

    def __init__(self, norm_type: int = NORM_L2, cross_check: bool = False,
                 ratio_threshold: Optional[float] = None, engine: Optional[Engine] = None):
        self.norm_type = int(norm_type)
        self.ratio_threshold = ratio_threshold
        self.cross_check = bool(cross_check)
        self._engine = engine
        self.last_arrays: Optional[MatchArrays] = None     # array view of the last result (no Python objects)

    @property
    def engine(self) -> Engine:
        if self._engine is None:
            self._engine = default_engine()     # raises without the CUDA library / a GPU: no CPU fallback
        return self._engine

    def match_arrays(self, desc1: np.ndarray, desc2: np.ndarray) -> MatchArrays:
        return self.engine.match_arrays(desc1, desc2, self.norm_type, self.cross_check, self.ratio_threshold)

    def match(self, features1, features2):
        c = contracts.resolve()
        start = time.time()
        if not self.validate_features(features1, features2):
            return c.MatchData([], method="BF", score_type=c.ScoreType.DISTANCE)
        engine = self.engine                     # environment problems (no GPU / no .so) are NOT swallowed
        try:
            arr = engine.match_arrays(np.asarray(features1.descriptors), np.asarray(features2.descriptors),
                                      self.norm_type, self.cross_check, self.ratio_threshold)
            self.last_arrays = arr
            return c.MatchData(matches=_materialise(c, arr), method="BF",
                               matching_time=time.time() - start, score_type=c.ScoreType.DISTANCE)
        except B200MatchError as e:
            # Shapes THIS library does not implement (L2 descriptors wider than 256, B200M_EUNSUPPORTED) and device /
            # environment failures are not input errors: the reference's cv2 path returns real matches for them, so an
            # empty MatchData here would silently lose them.  They propagate (limits: INTEGRATION.md 7).
            if e.code in (EUNSUPPORTED, ENODEVICE, ECUDA, EWORKSPACE):
                raise
            print(f"BF matching failed: {e}")
            return c.MatchData([], method="BF", score_type=c.ScoreType.DISTANCE)
        except UnsupportedNormError:
            raise                                 # NORM_L1 / NORM_L2SQR / NORM_HAMMING2: cv2 matches these, this library does not
        except Exception as e:                   # reference convention for genuine input errors (:250-252): print and return empty
            print(f"BF matching failed: {e}")
            return c.MatchData([], method="BF", score_type=c.ScoreType.DISTANCE)


def install(verbose: bool = False) -> list:
    """Rebind `EnhancedBFMatcher` in the already-imported reference modules to a subclass of the
    REFERENCE's BaseFeatureMatcher backed by this matcher.  Returns the patched module names."""
    c = contracts.resolve()
    if c.source != "reference":
        raise RuntimeError("reference package FeatureMatchingExtraction is not imported; nothing to patch")
    cls = type("EnhancedBFMatcher", (B200BFMatcher, c.BaseFeatureMatcher), {"__doc__": B200BFMatcher.__doc__})
    patched = []
    for name in ("FeatureMatchingExtraction.feature_matchers", "FeatureMatchingExtraction.matcher_factory"):
        mod = sys.modules.get(name)
        if mod is not None and hasattr(mod, "EnhancedBFMatcher"):
            if not hasattr(mod, "_CV2_EnhancedBFMatcher"):
                mod._CV2_EnhancedBFMatcher = mod.EnhancedBFMatcher
            mod.EnhancedBFMatcher = cls
            patched.append(name)
    if verbose:
        print("b200match installed into", patched)
    return patched


def uninstall() -> None:
    for name in ("FeatureMatchingExtraction.feature_matchers", "FeatureMatchingExtraction.matcher_factory"):
        mod = sys.modules.get(name)
        if mod is not None and hasattr(mod, "_CV2_EnhancedBFMatcher"):
            mod.EnhancedBFMatcher = mod._CV2_EnhancedBFMatcher
