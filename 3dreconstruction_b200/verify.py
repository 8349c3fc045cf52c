"""Geometric verification after matching: RANSAC homography filtering, batched on the GPU.

Mirrors /root/reference/FeatureMatchingExtraction/utils.py:
  * `enhanced_filter_matches_with_homography` (:118-153) — per pair: points of the best matches ->
    `cv2.findHomography(pts1, pts2, cv2.RANSAC, thr, confidence=..., maxIters=...)` -> matches with mask == 1, H;
    `([], None)` for fewer than 4 matches or when no model is found.  Same name, arguments and returns here.
  * `adaptive_match_filtering` (:155-194) — the wrapper `FeatureProcessingPipeline._apply_filtering` calls
    (pipeline.py:299-310); same triple back.
and adds the batched form the all-pairs driver uses: `verify_pairs` runs every pair of a CSR in ONE kernel launch
(csrc/ransac.cu, one CTA per pair, cv2's own random sequence) on the correspondences `sfm.match_feature_set` gathered.

Host side only: no GPU or no library -> the calls raise (no CPU path).
"""
from __future__ import annotations

from typing import Any, Dict, List, NamedTuple, Optional, Tuple

import numpy as np
import torch

from . import _lib as L
from .engine import Engine, default_engine


class DeviceVerification(NamedTuple):
    H: torch.Tensor          # float64 [P, 9]  (NaN rows where cv2 returns None)
    mask: torch.Tensor       # uint8 [K]
    inliers: torch.Tensor    # int32 [P]  (-1: the reference's ([], None))
    iters: torch.Tensor      # int32 [P]  RANSAC iterations run


class Verification(NamedTuple):
    """Host arrays for the pairs of one CSR."""
    ptr: np.ndarray          # int64 [P + 1]
    H: np.ndarray            # float64 [P, 3, 3]
    mask: np.ndarray         # uint8 [K]
    inliers: np.ndarray      # int32 [P]
    iters: np.ndarray        # int32 [P]

    def homography(self, p: int) -> Optional[np.ndarray]:
        return None if self.inliers[p] < 0 else self.H[p].copy()

    def kept(self, p: int) -> np.ndarray:
        """Positions (within pair p's match list) of the matches the reference keeps."""
        if self.inliers[p] < 0:
            return np.zeros(0, np.int64)
        return np.flatnonzero(self.mask[int(self.ptr[p]):int(self.ptr[p + 1])])


def verify_pairs_device(engine: Engine, corr: torch.Tensor, ptr: torch.Tensor, ransac_threshold: float = 5.0,
                        confidence: float = 0.99, max_iters: int = 2000) -> DeviceVerification:
    """corr float32 [K, 4] (x1, y1, x2, y2), ptr int64 [P + 1], both on the engine's device."""
    if corr.dtype != torch.float32 or ptr.dtype != torch.int64:
        raise ValueError("corr must be float32 [K, 4] and ptr int64 [P + 1]")
    P = int(ptr.numel()) - 1
    dev = engine.device
    K = int(corr.shape[0]) if corr.dim() == 2 else 0
    corr = corr.contiguous()
    with torch.cuda.device(dev):
        if K == 0:
            corr = torch.zeros((1, 4), dtype=torch.float32, device=dev)      # a valid pointer for pairs without matches
        H = torch.empty((max(P, 0), 9), dtype=torch.float64, device=dev)
        mask = torch.empty((max(K, 1),), dtype=torch.uint8, device=dev)
        inl = torch.empty((max(P, 1),), dtype=torch.int32, device=dev)
        its = torch.empty((max(P, 1),), dtype=torch.int32, device=dev)
        L.check(engine.lib.b200m_ransac_homography(corr.data_ptr(), ptr.data_ptr(), P, float(ransac_threshold),
                                                   float(confidence), int(max_iters), H.data_ptr(), mask.data_ptr(),
                                                   inl.data_ptr(), its.data_ptr(), engine._stream()),
                "b200m_ransac_homography")
    return DeviceVerification(H, mask[:K], inl[:P], its[:P])


def verify_pairs(correspondences: np.ndarray, ptr: np.ndarray, ransac_threshold: float = 5.0, confidence: float = 0.99,
                 max_iters: int = 2000, engine: Optional[Engine] = None) -> Verification:
    """Host arrays in, host arrays out: correspondences float32 [K, 4], ptr int64 [P + 1] (e.g. SfMMatches.correspondences / .ptr)."""
    eng = engine or default_engine()
    ptr = np.ascontiguousarray(ptr, np.int64)
    corr = np.ascontiguousarray(correspondences, np.float32).reshape(-1, 4)
    if ptr[-1] != len(corr):
        raise ValueError("ptr[-1] must equal the number of correspondences")
    d = verify_pairs_device(eng, eng.to_device(corr) if len(corr) else torch.empty((0, 4), dtype=torch.float32, device=eng.device),
                            eng.to_device(ptr), ransac_threshold, confidence, max_iters)
    return Verification(ptr, d.H.cpu().numpy().reshape(-1, 3, 3), d.mask.cpu().numpy(), d.inliers.cpu().numpy(),
                        d.iters.cpu().numpy())


def _points(kp, idx: List[int]) -> np.ndarray:
    out = np.empty((len(idx), 2), np.float32)
    for r, i in enumerate(idx):
        k = kp[i]
        out[r] = k.pt if hasattr(k, "pt") else k[:2]
    return out


def enhanced_filter_matches_with_homography(kp1, kp2, matches, match_data, ransac_threshold: float = 5.0,
                                            confidence: float = 0.99, max_iters: int = 2000,
                                            engine: Optional[Engine] = None) -> Tuple[list, Optional[np.ndarray]]:
    """Drop-in for utils.py:118-153.  As there, the matches filtered are `match_data`'s best ones (`get_filtered_matches()`
    for multi-method data), not the `matches` argument."""
    if type(match_data).__name__ == "MultiMethodMatchData":          # utils.py:132-135
        todo = match_data.get_filtered_matches()
    else:
        todo = match_data.get_best_matches()
    if len(todo) < 4:
        return [], None
    p1 = _points(kp1, [m.queryIdx for m in todo])
    p2 = _points(kp2, [m.trainIdx for m in todo])
    v = verify_pairs(np.concatenate([p1, p2], axis=1), np.array([0, len(todo)], np.int64), ransac_threshold, confidence,
                     max_iters, engine)
    if v.inliers[0] < 0:
        return [], None
    return [m for m, keep in zip(todo, v.mask) if keep], v.H[0].copy()


def adaptive_match_filtering(kp1, kp2, matches, match_data=None, ransac_threshold: float = 4.0,
                             engine: Optional[Engine] = None) -> Tuple[list, Optional[np.ndarray], Dict[str, Any]]:
    """Drop-in for utils.py:155-194."""
    if not matches or len(matches) < 4:
        return [], None, {"method": "none", "reason": "insufficient_matches"}
    kept, H = enhanced_filter_matches_with_homography(kp1, kp2, matches, match_data, ransac_threshold, engine=engine)
    return kept, H, {"method": "homography_ransac", "original_count": len(matches), "filtered_count": len(kept),
                     "ransac_threshold": ransac_threshold, "homography_found": H is not None}
