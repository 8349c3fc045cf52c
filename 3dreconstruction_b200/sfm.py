"""Feature cache, all-pairs driver and SfM hand-off around the matcher.

N1  The reference's `match_folder` runs detection on both images again for every pair
    (/root/reference/FeatureMatchingExtraction/pipeline.py:538-606 -> :151-152).  `build_feature_set` takes each
    image's `FeatureData` ONCE, keeps descriptors (in the matcher's layout) and keypoint coordinates in HBM, and
    `match_feature_set` runs every requested pair as one batched job (allpairs.AllPairsMatcher).
N2  `SfMMatches` holds what CameraPoseEstimation consumes: per pair the N x 4 correspondences
    (x1, y1, x2, y2) the reference assembles one Python object at a time (result_converters.py:754-766), produced by
    one device gather over the matcher's CSR (csrc/adapter.cu), and `to_matches_pickle()` lays them out as
    {'matches_data': {(img_i, img_j): {...}}, 'image_info': {img: {...}}} (CameraPoseEstimation/pipeline.py:231-235,
    pair_selector.py:54-67, correspondence_manager.py:78-119).
N4  `SfMMatches.save` / `load`: the same content as flat arrays in one .npz (CSR), next to the reference's pickle of
    per-match objects (result_types.py:148-228).

Host side only: everything that computes is a kernel behind the C-ABI; without the library or a GPU the calls raise.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Mapping, NamedTuple, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib as L
from . import contracts
from .allpairs import AllPairsMatcher, Store, all_pairs, build_store
from .engine import Engine


def keypoints_xy(keypoints) -> np.ndarray:
    """float32 [n, 2] of cv2.KeyPoint.pt (or of anything with .pt, or an [n, 2] array)."""
    if isinstance(keypoints, np.ndarray):
        return np.ascontiguousarray(keypoints, dtype=np.float32).reshape(-1, 2)
    n = len(keypoints)
    out = np.empty((n, 2), np.float32)
    for i, kp in enumerate(keypoints):
        out[i] = kp.pt if hasattr(kp, "pt") else kp[:2]
    return out


class FeatureSet(NamedTuple):
    names: List[str]
    sizes: List[Optional[Tuple[int, int]]]     # (width, height) per image, as ImagePairInfo carries them (result_types.py:28-33)
    store: Store                               # descriptors of all images, resident in HBM
    kp_xy: torch.Tensor                        # float32 [rows, 2], same row order as the store
    method: str


def build_feature_set(engine: Engine, features: Mapping[str, object], norm_type: int,
                      image_sizes: Optional[Mapping[str, Tuple[int, int]]] = None) -> FeatureSet:
    """features: image id -> FeatureData (keypoints, descriptors, method).  Images without descriptors contribute 0 rows."""
    names = list(features.keys())
    descs, xys, method = [], [], "unknown"
    width, dtype = None, None
    for nm in names:
        f = features[nm]
        d = getattr(f, "descriptors", None)
        if d is not None and len(d) > 0:
            d = np.asarray(d)
            width, dtype = d.shape[1], d.dtype
            method = getattr(f, "method", method)
    if width is None:
        raise ValueError("no image has descriptors")
    for nm in names:
        f = features[nm]
        d = getattr(f, "descriptors", None)
        if d is None or len(d) == 0:
            descs.append(np.zeros((0, width), dtype))
            xys.append(np.zeros((0, 2), np.float32))
            continue
        d = np.asarray(d)
        if d.shape[1] != width or d.dtype != dtype:
            raise ValueError(f"{nm}: descriptor width/dtype differs from the other images")   # cv2.error in the reference
        xy = keypoints_xy(f.keypoints)
        if len(xy) != len(d):
            raise ValueError(f"{nm}: {len(xy)} keypoints but {len(d)} descriptors")
        descs.append(d)
        xys.append(xy)
    store = build_store(engine, descs, norm_type)
    kp = engine.to_device(np.concatenate(xys, axis=0))
    sizes = [None if image_sizes is None else image_sizes.get(nm) for nm in names]
    return FeatureSet(names, sizes, store, kp, method)


class SfMMatches:
    """All matched pairs of one run as flat host arrays (CSR over pairs)."""

    def __init__(self, names, sizes, pairs, ptr, query_idx, train_idx, distance, correspondences,
                 method: str = "BF", feature_type: str = "unknown"):
        self.names = list(names)
        self.sizes = list(sizes)
        self.pairs = np.asarray(pairs, np.int32).reshape(-1, 2)
        self.ptr = np.asarray(ptr, np.int64)
        self.query_idx = np.asarray(query_idx, np.int32)
        self.train_idx = np.asarray(train_idx, np.int32)
        self.distance = np.asarray(distance, np.float32)
        self.correspondences = np.asarray(correspondences, np.float32).reshape(-1, 4)
        self.method = method
        self.feature_type = feature_type
        assert len(self.ptr) == len(self.pairs) + 1 and self.ptr[-1] == len(self.query_idx) == len(self.correspondences)

    def __len__(self) -> int:
        return len(self.pairs)

    def pair_slice(self, p: int) -> slice:
        return slice(int(self.ptr[p]), int(self.ptr[p + 1]))

    def pair_dict(self, p: int) -> dict:
        """One entry of `matches_data`.  Arrays have the dtypes the reference's converter yields: float64 coordinates
        (np.array of Python-float .pt tuples) and scores, int32 indices (result_converters.py:754-766)."""
        s = self.pair_slice(p)
        return {"correspondences": self.correspondences[s].astype(np.float64),
                "num_matches": s.stop - s.start,
                "query_indices": self.query_idx[s].copy(), "train_indices": self.train_idx[s].copy(),
                "match_scores": self.distance[s].astype(np.float64),
                "score_type": "distance", "method": self.method}

    def to_matches_pickle(self, min_matches: int = 1) -> dict:
        """The dict `MainPosePipeline.process_monument_reconstruction` takes (CameraPoseEstimation/pipeline.py:231-235).
        Pairs with fewer than min_matches matches are left out (the consumers index correspondences[0])."""
        md = {}
        for p, (i, j) in enumerate(self.pairs):
            if self.ptr[p + 1] - self.ptr[p] >= min_matches:
                md[(self.names[i], self.names[j])] = self.pair_dict(p)
        info = {nm: {"name": nm, "size": sz} for nm, sz in zip(self.names, self.sizes)}
        total = int(sum(v["num_matches"] for v in md.values()))
        return {"matches_data": md, "image_info": info, "feature_type": self.feature_type, "total_images": len(self.names),
                "processing_stats": {"successful_pairs": len(md), "failed_pairs": 0, "total_matches": total}}

    def to_match_data(self, p: int):
        """Pair p as the reference's MatchData (what EnhancedBFMatcher.match returns for that pair)."""
        c = contracts.resolve()
        s = self.pair_slice(p)
        dist_t = c.ScoreType.DISTANCE
        ms = [c.EnhancedDMatch(queryIdx=q, trainIdx=t, score=d, score_type=dist_t, raw_distance=d)
              for q, t, d in zip(self.query_idx[s].tolist(), self.train_idx[s].tolist(),
                                 self.distance[s].astype(np.float64).tolist())]
        return c.MatchData(matches=ms, method=self.method, score_type=dist_t)

    # ---- N3: geometric verification (FeatureProcessingPipeline._apply_filtering, pipeline.py:289-320, for every pair at once)
    def verify_homography(self, ransac_threshold: float = 4.0, confidence: float = 0.99, max_iters: int = 2000, engine=None):
        """RANSAC homography per pair on the GPU (verify.verify_pairs): what `enhanced_filter_matches_with_homography`
        (utils.py:118-153) returns for each pair's matches — mask, H, inlier count (-1: the reference's ([], None))."""
        from . import verify
        return verify.verify_pairs(self.correspondences, self.ptr, ransac_threshold, confidence, max_iters, engine)

    def filtered(self, verification) -> "SfMMatches":
        """The matches the reference would store in `match_data.filtered_matches` (pipeline.py:309, :322): inliers of
        each pair's homography, order kept; pairs without a model keep nothing."""
        keep = verification.mask.astype(bool)
        no_model = np.repeat(verification.inliers < 0, np.diff(self.ptr))
        keep &= ~no_model
        ptr = np.concatenate([[0], np.cumsum(keep, dtype=np.int64)])[self.ptr]       # kept matches before each pair
        return SfMMatches(self.names, self.sizes, self.pairs, ptr, self.query_idx[keep], self.train_idx[keep],
                          self.distance[keep], self.correspondences[keep], self.method, self.feature_type)

    # ---- N4: array-native persistence
    def save(self, path) -> None:
        sizes = np.array([(-1, -1) if s is None else tuple(s) for s in self.sizes], np.int64).reshape(-1, 2)
        np.savez_compressed(path, format=np.array("b200match-sfm-1"), names=np.array(self.names), sizes=sizes,
                            pairs=self.pairs, ptr=self.ptr, query_idx=self.query_idx, train_idx=self.train_idx,
                            distance=self.distance, correspondences=self.correspondences,
                            method=np.array(self.method), feature_type=np.array(self.feature_type))

    @classmethod
    def load(cls, path) -> "SfMMatches":
        z = np.load(path, allow_pickle=False)
        if str(z["format"]) != "b200match-sfm-1":
            raise ValueError("not a b200match SfM match file")
        sizes = [None if w < 0 else (int(w), int(h)) for w, h in z["sizes"]]
        return cls([str(n) for n in z["names"]], sizes, z["pairs"], z["ptr"], z["query_idx"], z["train_idx"],
                   z["distance"], z["correspondences"], str(z["method"]), str(z["feature_type"]))


def match_feature_set(engine: Engine, fs: FeatureSet, pairs: Optional[np.ndarray] = None, cross_check: bool = False,
                      ratio_threshold: Optional[float] = None, chunk_pairs: int = 2048,
                      matcher: Optional[AllPairsMatcher] = None, ransac: Optional[Tuple[float, float, int]] = None,
                      essential: Optional[Tuple[float, float, int]] = None):
    """Match `pairs` (default: all i < j, the reference's pair_mode='all', image_manager.py:491-494) and gather the
    matched keypoint coordinates on the device; one D2H of the finished arrays.
    ransac = (threshold, confidence, max_iters): also verify every pair by RANSAC homography (N3) on the correspondences
    while they are still in HBM; returns (SfMMatches, verify.Verification) instead of SfMMatches alone.
    essential = (threshold, prob, max_iters): also estimate every pair's essential matrix (pose.estimate_pairs_device, what
    EssentialMatrixEstimator.estimate's RANSAC branch computes, essential_estimation.py:118-126) with each image's camera
    taken from its size as the reference does when none is given (:36-60; every image of `fs` needs a size); the result is
    appended to the return value as a pose.PoseEstimates."""
    if pairs is None:
        pairs = all_pairs(len(fs.names))
    pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
    m = matcher or AllPairsMatcher(engine, chunk_pairs)
    csr = m.match(fs.store, pairs, cross_check, ratio_threshold)
    K = int(csr.query_idx.numel())
    dev = engine.device
    corr = torch.empty((max(K, 1), 4), dtype=torch.float32, device=dev)
    if len(pairs):
        offs = fs.store.offsets
        q_off = torch.from_numpy(offs[pairs[:, 0]].astype(np.int32)).to(dev)
        t_off = torch.from_numpy(offs[pairs[:, 1]].astype(np.int32)).to(dev)
        with torch.cuda.device(dev):
            L.check(engine.lib.b200m_gather_correspondences(
                fs.kp_xy.data_ptr(), int(fs.kp_xy.shape[0]), q_off.data_ptr(), t_off.data_ptr(), csr.ptr.data_ptr(),
                len(pairs), csr.query_idx.data_ptr(), csr.train_idx.data_ptr(), corr.data_ptr(), engine._stream()),
                "b200m_gather_correspondences")
    ver = None
    if ransac is not None:
        from . import verify
        ver = verify.verify_pairs_device(engine, corr[:K], csr.ptr, *ransac)
    est = None
    if essential is not None:
        from . import pose
        if any(sz is None for sz in fs.sizes):
            raise ValueError("essential=...: every image needs its (width, height) to estimate a camera matrix")
        cam4 = np.array([[1.2 * sz[0], 1.2 * sz[0], sz[0] / 2.0, sz[1] / 2.0] for sz in fs.sizes], np.float64)   # fx, fy, cx, cy
        cams = np.concatenate([cam4[pairs[:, 0]], cam4[pairs[:, 1]]], axis=1) if len(pairs) else np.zeros((0, 8))
        est = pose.estimate_pairs_device(engine, corr[:K], csr.ptr, torch.from_numpy(np.ascontiguousarray(cams)).to(dev), *essential)
    h = csr.to_host()
    res = SfMMatches(fs.names, fs.sizes, pairs, h.ptr, h.query_idx, h.train_idx, h.distance, corr[:K].cpu().numpy(),
                     method="BF", feature_type=fs.method)
    out = [res]
    if ver is not None:
        from . import verify
        out.append(verify.Verification(h.ptr, ver.H.cpu().numpy().reshape(-1, 3, 3), ver.mask.cpu().numpy(),
                                       ver.inliers.cpu().numpy(), ver.iters.cpu().numpy()))
    if est is not None:
        from . import pose
        out.append(pose.PoseEstimates(h.ptr, est.E.cpu().numpy().reshape(-1, 3, 3), est.mask.cpu().numpy(),
                                      est.inliers.cpu().numpy(), est.iters.cpu().numpy()))
    return out[0] if len(out) == 1 else tuple(out)


def match_all_pairs(features: Mapping[str, object], norm_type: int, cross_check: bool = False,
                    ratio_threshold: Optional[float] = None, image_sizes: Optional[Mapping[str, Tuple[int, int]]] = None,
                    engine: Optional[Engine] = None) -> SfMMatches:
    """features of every image in, SfM-ready matches of every pair out."""
    from .engine import default_engine
    eng = engine or default_engine()
    return match_feature_set(eng, build_feature_set(eng, features, norm_type, image_sizes), None, cross_check, ratio_threshold)
