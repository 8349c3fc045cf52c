"""Data contracts at the plugin boundary.

When the reference package is importable (a user who switched matchers inside the reference
pipeline), `resolve()` hands back the REFERENCE's own classes so `MatchData` instances pickle and
type-check exactly as before (/root/reference/FeatureMatchingExtraction/core_data_structures.py:
32-36 ScoreType, :39-61 FeatureData, :64-101 EnhancedDMatch, :104-173 MatchData; base class
base_classes.py:76-105).  Stand-alone (GPU box, tests) the mirrors below are used: same field
names, defaults and helper semantics, written from the interface description in SURVEY §8(a8).
"""
from __future__ import annotations

import sys
from dataclasses import dataclass
from enum import Enum
from typing import Any, List, Optional, Tuple

import numpy as np


class ScoreType(Enum):
    DISTANCE = "distance"        # lower is better
    CONFIDENCE = "confidence"    # higher is better
    SIMILARITY = "similarity"    # higher is better


@dataclass
class FeatureData:
    keypoints: List[Any]                       # cv2.KeyPoint in the reference; len() counts these
    descriptors: Optional[np.ndarray]
    method: str
    confidence_scores: Optional[List[float]] = None
    detection_time: float = 0.0
    raw_image: Optional[np.ndarray] = None

    def __len__(self) -> int:
        return len(self.keypoints)


@dataclass
class EnhancedDMatch:
    queryIdx: int
    trainIdx: int
    score: float
    score_type: ScoreType
    imgIdx: int = 0
    raw_distance: Optional[float] = None
    confidence: Optional[float] = None
    source_method: Optional[str] = None

    @property
    def distance(self) -> float:
        if self.score_type == ScoreType.CONFIDENCE:
            return 1.0 - self.score
        return self.score

    def to_cv2_dmatch(self):
        import cv2
        m = cv2.DMatch()
        m.queryIdx, m.trainIdx, m.distance, m.imgIdx = self.queryIdx, self.trainIdx, self.distance, self.imgIdx
        return m


@dataclass
class MatchData:
    matches: List[Any]
    filtered_matches: Optional[List[Any]] = None
    homography: Optional[np.ndarray] = None
    fundamental_matrix: Optional[np.ndarray] = None
    method: str = "unknown"
    matching_time: float = 0.0
    score_type: ScoreType = ScoreType.DISTANCE
    match_confidences: Optional[np.ndarray] = None
    keypoint_confidences: Optional[Tuple[np.ndarray, np.ndarray]] = None

    def get_best_matches(self):
        return self.filtered_matches if self.filtered_matches else self.matches

    def get_match_scores(self, use_filtered: bool = True) -> np.ndarray:
        ms = self.get_best_matches() if use_filtered else self.matches
        if not ms:
            return np.array([])
        return np.array([getattr(m, "score", getattr(m, "distance", None)) for m in ms])

    def filter_by_score(self, threshold: float, top_k: Optional[int] = None) -> "MatchData":
        score = lambda m: m.score if hasattr(m, "score") else m.distance   # noqa: E731
        if self.score_type == ScoreType.DISTANCE:
            kept = sorted((m for m in self.matches if score(m) <= threshold), key=score)
        else:
            kept = sorted((m for m in self.matches if score(m) >= threshold), key=score, reverse=True)
        if top_k:
            kept = kept[:top_k]
        return MatchData(matches=self.matches, filtered_matches=kept, method=self.method,
                         score_type=self.score_type, match_confidences=self.match_confidences,
                         keypoint_confidences=self.keypoint_confidences)

    def to_cv2_matches(self):
        return [m.to_cv2_dmatch() if hasattr(m, "to_cv2_dmatch") else m for m in self.get_best_matches()]


class BaseFeatureMatcher:
    """Mirror of the reference ABC (base_classes.py:76-105)."""

    def match(self, features1: FeatureData, features2: FeatureData) -> MatchData:   # pragma: no cover
        raise NotImplementedError

    def validate_features(self, features1, features2) -> bool:
        return (len(features1) > 0 and len(features2) > 0 and
                features1.descriptors is not None and features2.descriptors is not None)


class Contracts:
    """The set of classes the matcher instantiates (reference's or the mirrors)."""

    def __init__(self, ScoreType, FeatureData, EnhancedDMatch, MatchData, BaseFeatureMatcher, source: str):
        self.ScoreType, self.FeatureData, self.EnhancedDMatch = ScoreType, FeatureData, EnhancedDMatch
        self.MatchData, self.BaseFeatureMatcher, self.source = MatchData, BaseFeatureMatcher, source


_MIRROR = Contracts(ScoreType, FeatureData, EnhancedDMatch, MatchData, BaseFeatureMatcher, "mirror")


def resolve() -> Contracts:
    """Reference classes if `FeatureMatchingExtraction` is already imported, else the mirrors."""
    cds = sys.modules.get("FeatureMatchingExtraction.core_data_structures")
    bc = sys.modules.get("FeatureMatchingExtraction.base_classes")
    if cds is not None and bc is not None:
        return Contracts(cds.ScoreType, cds.FeatureData, cds.EnhancedDMatch, cds.MatchData,
                         bc.BaseFeatureMatcher, "reference")
    return _MIRROR
