"""Essential-matrix estimation of matched pairs, batched on the GPU (second half).

Mirrors /root/reference/CameraPoseEstimation/essential_estimation.py:
  * `MatrixEstimationConfig` (:6-19) — the reference's defaults: RANSAC threshold 1.5, confidence 0.999, 5000 iterations,
    at least 30 inliers and an inlier ratio of 0.4, focal length 1.2 x image width when no camera is given.
  * `EssentialMatrixEstimator.estimate(pts1, pts2, image_size1, image_size2, method='RANSAC', camera_matrix1, camera_matrix2)`
    (:62-186) — fewer than 8 points refused, cameras estimated from the image sizes (:36-60), `cv2.undistortPoints`,
    `cv2.findEssentialMat(..., focal=1, pp=(0, 0), method=cv2.RANSAC, prob, threshold, maxIters)` (:118-126), the
    MIN_INLIERS / MIN_INLIER_RATIO gates; same result dictionary keys and error strings here.  Only method='RANSAC' is
    implemented (the reference's '5POINT' / '8POINT' branches pass flags cv2.findEssentialMat treats as LMedS).
and adds the batched form: `estimate_pairs` runs every pair of a CSR in ONE kernel launch (csrc/essential.cu, one CTA per
pair, cv2's own random sequence and model order) on the pixel correspondences `sfm.match_feature_set` gathered.

Host side only: no GPU or no library -> the calls raise (no CPU path).
"""
from __future__ import annotations

from typing import Dict, NamedTuple, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib as L
from .engine import Engine, default_engine


class MatrixEstimationConfig:
    """essential_estimation.py:6-19."""
    RANSAC_THRESHOLD = 1.5
    RANSAC_CONFIDENCE = 0.999
    RANSAC_MAX_ITERATIONS = 5000
    MIN_INLIERS = 30
    MIN_INLIER_RATIO = 0.4
    DEFAULT_FOCAL_RATIO = 1.2


class DevicePose(NamedTuple):
    E: torch.Tensor          # float64 [P, 9] row-major 3x3, unit Frobenius norm (NaN rows where cv2 returns None)
    mask: torch.Tensor       # uint8 [K]
    inliers: torch.Tensor    # int32 [P]  (-1: no model)
    iters: torch.Tensor      # int32 [P]  RANSAC iterations run


class PoseEstimates(NamedTuple):
    ptr: np.ndarray          # int64 [P + 1]
    E: np.ndarray            # float64 [P, 3, 3]
    mask: np.ndarray         # uint8 [K]
    inliers: np.ndarray      # int32 [P]
    iters: np.ndarray        # int32 [P]

    def essential(self, p: int) -> Optional[np.ndarray]:
        return None if self.inliers[p] < 0 else self.E[p].copy()

    def inlier_mask(self, p: int) -> np.ndarray:
        return self.mask[int(self.ptr[p]):int(self.ptr[p + 1])]


def camera_matrix_from_size(image_size, focal_ratio: float = MatrixEstimationConfig.DEFAULT_FOCAL_RATIO) -> np.ndarray:
    """EssentialMatrixEstimator.estimate_camera_matrix (:36-60)."""
    width, height = image_size[:2]
    f = width * focal_ratio
    return np.array([[f, 0, width / 2.0], [0, f, height / 2.0], [0, 0, 1]], dtype=np.float64)


def _cam8(K1: np.ndarray, K2: np.ndarray) -> np.ndarray:
    return np.array([K1[0, 0], K1[1, 1], K1[0, 2], K1[1, 2], K2[0, 0], K2[1, 1], K2[0, 2], K2[1, 2]], np.float64)


def estimate_pairs_device(engine: Engine, corr: torch.Tensor, ptr: torch.Tensor, cams: Optional[torch.Tensor],
                          threshold: float = 1.5, prob: float = 0.999, max_iters: int = 5000) -> DevicePose:
    """corr float32 [K, 4] pixel (x1, y1, x2, y2), ptr int64 [P + 1], cams float64 [P, 8] = (fx1, fy1, cx1, cy1, fx2, fy2,
    cx2, cy2) per pair (None: the points are already normalised), all on the engine's device.  `threshold` is in
    NORMALISED units, as the reference passes it (focal = 1)."""
    if corr.dtype != torch.float32 or ptr.dtype != torch.int64:
        raise ValueError("corr must be float32 [K, 4] and ptr int64 [P + 1]")
    P = int(ptr.numel()) - 1
    if cams is not None and (cams.dtype != torch.float64 or cams.numel() != 8 * P):
        raise ValueError("cams must be float64 [P, 8]")
    dev = engine.device
    K = int(corr.shape[0]) if corr.dim() == 2 else 0
    corr = corr.contiguous()
    with torch.cuda.device(dev):
        if K == 0:
            corr = torch.zeros((1, 4), dtype=torch.float32, device=dev)
        E = torch.empty((max(P, 0), 9), dtype=torch.float64, device=dev)
        mask = torch.empty((max(K, 1),), dtype=torch.uint8, device=dev)
        inl = torch.empty((max(P, 1),), dtype=torch.int32, device=dev)
        its = torch.empty((max(P, 1),), dtype=torch.int32, device=dev)
        L.check(engine.lib.b200m_ransac_essential(corr.data_ptr(), ptr.data_ptr(), P, 0 if cams is None else cams.contiguous().data_ptr(),
                                                  float(threshold), float(prob), int(max_iters), E.data_ptr(), mask.data_ptr(),
                                                  inl.data_ptr(), its.data_ptr(), engine._stream()),
                "b200m_ransac_essential")
    return DevicePose(E, mask[:K], inl[:P], its[:P])


def estimate_pairs(correspondences: np.ndarray, ptr: np.ndarray, cams: Optional[np.ndarray], threshold: float = 1.5,
                   prob: float = 0.999, max_iters: int = 5000, engine: Optional[Engine] = None) -> PoseEstimates:
    """Host arrays in, host arrays out: correspondences float32 [K, 4], ptr int64 [P + 1], cams float64 [P, 8] or None."""
    eng = engine or default_engine()
    ptr = np.ascontiguousarray(ptr, np.int64)
    corr = np.ascontiguousarray(correspondences, np.float32).reshape(-1, 4)
    if ptr[-1] != len(corr):
        raise ValueError("ptr[-1] must equal the number of correspondences")
    dc = eng.to_device(corr) if len(corr) else torch.empty((0, 4), dtype=torch.float32, device=eng.device)
    dk = None if cams is None else eng.to_device(np.ascontiguousarray(cams, np.float64).reshape(-1, 8))
    d = estimate_pairs_device(eng, dc, eng.to_device(ptr), dk, threshold, prob, max_iters)
    return PoseEstimates(ptr, d.E.cpu().numpy().reshape(-1, 3, 3), d.mask.cpu().numpy(), d.inliers.cpu().numpy(), d.iters.cpu().numpy())


class EssentialMatrixEstimator:
    """Drop-in for essential_estimation.py's class: same `config`, `estimate_camera_matrix` and `estimate` (RANSAC)."""

    def __init__(self, engine: Optional[Engine] = None):
        self.config = MatrixEstimationConfig()
        self._engine = engine

    def estimate_camera_matrix(self, image_size: Tuple[int, int], focal_ratio: Optional[float] = None) -> np.ndarray:
        return camera_matrix_from_size(image_size, self.config.DEFAULT_FOCAL_RATIO if focal_ratio is None else focal_ratio)

    def estimate(self, pts1: np.ndarray, pts2: np.ndarray, image_size1, image_size2, method: str = "RANSAC",
                 camera_matrix1: Optional[np.ndarray] = None, camera_matrix2: Optional[np.ndarray] = None) -> Dict:
        if len(pts1) < 8:
            return {"success": False,
                    "error": f"Insufficient correspondences: {len(pts1)} < 8 points needed for essential matrix"}
        if method != "RANSAC":
            return {"success": False, "error": f"Unknown method: {method}"} if method not in ("5POINT", "8POINT") else \
                {"success": False, "error": f"Essential matrix estimation failed: method {method} is not implemented on the GPU path"}
        est1, est2 = camera_matrix1 is None, camera_matrix2 is None
        K1 = self.estimate_camera_matrix(image_size1) if est1 else np.asarray(camera_matrix1, np.float64)
        K2 = self.estimate_camera_matrix(image_size2) if est2 else np.asarray(camera_matrix2, np.float64)
        corr = np.concatenate([np.asarray(pts1, np.float32).reshape(-1, 2), np.asarray(pts2, np.float32).reshape(-1, 2)], axis=1)
        r = estimate_pairs(corr, np.array([0, len(corr)], np.int64), _cam8(K1, K2)[None], self.config.RANSAC_THRESHOLD,
                           self.config.RANSAC_CONFIDENCE, self.config.RANSAC_MAX_ITERATIONS, self._engine)
        if r.inliers[0] < 0:
            return {"success": False, "error": "Essential matrix estimation returned None - possibly degenerate configuration"}
        E, mask = r.E[0].copy(), r.mask.reshape(-1, 1).copy()
        inlier_count = int(mask.sum())
        inlier_ratio = inlier_count / len(pts1)
        if inlier_count < self.config.MIN_INLIERS:
            return {"success": False,
                    "error": f"Too few inliers for reliable monument reconstruction: {inlier_count} < {self.config.MIN_INLIERS}"}
        if inlier_ratio < self.config.MIN_INLIER_RATIO:
            return {"success": False, "error": f"Low inlier ratio for monument reconstruction: {inlier_ratio:.1%} < "
                                               f"{self.config.MIN_INLIER_RATIO:.1%}"}
        return {"success": True, "essential_matrix": E, "inlier_mask": mask, "camera_matrices": [K1, K2],
                "camera_estimated": [est1, est2], "method": method, "num_points": len(pts1), "num_inliers": inlier_count,
                "inlier_ratio": inlier_ratio, "matrix_type": "essential"}
