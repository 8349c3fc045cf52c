"""oracle/bf_oracle.py — TEST INFRASTRUCTURE ONLY (the checker; never imported by the product path).

CPU restatement of the reference's brute-force matching hot path
(/root/reference/FeatureMatchingExtraction/feature_matchers.py:184-252, `EnhancedBFMatcher`), whose
arithmetic lives in OpenCV's `cv::BFMatcher` / `cv::batchDistance` (third-party; reference pin
`opencv-python>=4.5.0`, requirements.txt:2; Dockerfile:20 pins 4.12.0.88; pinned here against 4.13.0).

Parity status: PINNED — against (a) live cv2.BFMatcher in tests/test_oracle.py and (b) outputs of
the reference's own `EnhancedBFMatcher.match` run in the build container, committed under
tests/golden/ by oracle/make_golden.py.  The reference itself ships no tests/golden vectors for
this path.

Three layers, all returning flat arrays:
  * `knn2_numpy`   — pure numpy (small cases, tie semantics spelled out)
  * `knn2_c`       — oracle/bf_oracle.c through ctypes (OpenMP; seconds at 8k x 8k)
  * `reference_match` — the wrapper logic of `EnhancedBFMatcher.match` (modes A/B/C, fp64 ratio
                        test, stable sort by distance) on top of either knn2.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference may import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import NamedTuple, Optional

import numpy as np

NORM_L2 = 4        # cv2.NORM_L2
NORM_HAMMING = 6   # cv2.NORM_HAMMING

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libbf_oracle.so")
_lib = None


class Knn2(NamedTuple):
    idx1: np.ndarray   # int32 [N]  best train index (-1 if M == 0)
    idx2: np.ndarray   # int32 [N]  second best (-1 if M < 2)
    d1: np.ndarray     # float32 [N] distance as cv2.DMatch.distance would hold it
    d2: np.ndarray     # float32 [N]


class Matches(NamedTuple):
    query_idx: np.ndarray   # int32 [K]
    train_idx: np.ndarray   # int32 [K]
    distance: np.ndarray    # float32 [K]


def build(force: bool = False) -> str:
    """Compile oracle/bf_oracle.c -> oracle/libbf_oracle.so (gcc -O3 -fopenmp)."""
    src = os.path.join(_HERE, "bf_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        # No -march=native: the .so is built here and travels to the GPU box's (different) CPU.
        cmd = ["gcc", "-O3", "-mpopcnt", "-fopenmp", "-fPIC", "-shared", "-fvisibility=hidden",
               "-o", _SO, src, "-lm"]
        subprocess.run(cmd, check=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        lib = ctypes.CDLL(_SO)
        i32p = ctypes.POINTER(ctypes.c_int32)
        lib.oracle_knn2_hamming.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                            ctypes.c_int, i32p, i32p, i32p, i32p]
        lib.oracle_knn2_l2.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                       ctypes.c_int, i32p, i32p, ctypes.POINTER(ctypes.c_float),
                                       ctypes.POINTER(ctypes.c_float)]
        lib.oracle_dist_hamming.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                            ctypes.c_int, i32p]
        lib.oracle_set_threads.argtypes = [ctypes.c_int]
        _lib = lib
    return _lib


def num_threads() -> int:
    return int(_load().oracle_num_threads())


def set_threads(n: int) -> None:
    _load().oracle_set_threads(int(n))


def _check(desc1: np.ndarray, desc2: np.ndarray, norm: int):
    if desc1.ndim != 2 or desc2.ndim != 2 or desc1.shape[1] != desc2.shape[1]:
        raise ValueError("descriptor width mismatch")          # cv2.error in the reference (a7)
    if desc1.dtype != desc2.dtype:
        raise ValueError("descriptor dtype mismatch")
    if norm == NORM_HAMMING and desc1.dtype != np.uint8:
        raise ValueError("NORM_HAMMING needs uint8 descriptors")
    if norm == NORM_L2 and desc1.dtype != np.float32:
        raise ValueError("NORM_L2 path needs float32 descriptors")
    if norm not in (NORM_L2, NORM_HAMMING):
        raise ValueError("unsupported norm")


# ----------------------------------------------------------------------------- numpy restatement
_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)


def distance_matrix_numpy(desc1: np.ndarray, desc2: np.ndarray, norm: int) -> np.ndarray:
    """Full N x M matrix of *pre-sqrt* distances: int32 Hamming, or float32 squared L2."""
    _check(desc1, desc2, norm)
    if norm == NORM_HAMMING:
        x = desc1[:, None, :] ^ desc2[None, :, :]
        return _POP8[x].sum(axis=2, dtype=np.int32)
    diff = desc1[:, None, :].astype(np.float64) - desc2[None, :, :].astype(np.float64)
    return (diff * diff).sum(axis=2).astype(np.float32)


def _top2_lowest_index(dm: np.ndarray):
    """Per-row two smallest entries, ties -> lowest column (cv2 batchDistance order)."""
    n, m = dm.shape
    idx1 = np.full(n, -1, np.int32)
    idx2 = np.full(n, -1, np.int32)
    if m == 0:
        return idx1, idx2
    idx1 = dm.argmin(axis=1).astype(np.int32)            # argmin returns the first (lowest) index
    if m >= 2:
        big = dm.astype(np.float64)
        big[np.arange(n), idx1] = np.inf
        idx2 = big.argmin(axis=1).astype(np.int32)
    return idx1, idx2


def knn2_numpy(desc1: np.ndarray, desc2: np.ndarray, norm: int) -> Knn2:
    dm = distance_matrix_numpy(desc1, desc2, norm)
    n, m = dm.shape
    idx1, idx2 = _top2_lowest_index(dm)
    rows = np.arange(n)
    d1 = np.full(n, -1, np.float32)
    d2 = np.full(n, -1, np.float32)
    if m >= 1:
        d1 = dm[rows, idx1].astype(np.float32)
    if m >= 2:
        d2 = dm[rows, idx2].astype(np.float32)
    if norm == NORM_L2:
        if m >= 1:
            d1 = np.sqrt(d1, dtype=np.float32)
        if m >= 2:
            d2 = np.sqrt(d2, dtype=np.float32)
    return Knn2(idx1, idx2, d1, d2)


# ----------------------------------------------------------------------------- C restatement
def knn2_c(desc1: np.ndarray, desc2: np.ndarray, norm: int) -> Knn2:
    _check(desc1, desc2, norm)
    lib = _load()
    a = np.ascontiguousarray(desc1)
    b = np.ascontiguousarray(desc2)
    n, m, d = a.shape[0], b.shape[0], a.shape[1]
    idx1 = np.empty(n, np.int32)
    idx2 = np.empty(n, np.int32)
    i32p = ctypes.POINTER(ctypes.c_int32)
    if norm == NORM_HAMMING:
        d1 = np.empty(n, np.int32)
        d2 = np.empty(n, np.int32)
        rc = lib.oracle_knn2_hamming(a.ctypes.data, n, b.ctypes.data, m, d,
                                     idx1.ctypes.data_as(i32p), idx2.ctypes.data_as(i32p),
                                     d1.ctypes.data_as(i32p), d2.ctypes.data_as(i32p))
        d1 = d1.astype(np.float32)
        d2 = d2.astype(np.float32)
    else:
        d1 = np.empty(n, np.float32)
        d2 = np.empty(n, np.float32)
        f32p = ctypes.POINTER(ctypes.c_float)
        rc = lib.oracle_knn2_l2(a.ctypes.data, n, b.ctypes.data, m, d,
                                idx1.ctypes.data_as(i32p), idx2.ctypes.data_as(i32p),
                                d1.ctypes.data_as(f32p), d2.ctypes.data_as(f32p))
    if rc != 0:
        raise RuntimeError(f"oracle knn2 failed rc={rc}")
    return Knn2(idx1, idx2, d1, d2)


# ----------------------------------------------------------------------------- wrapper restatement
def _empty() -> Matches:
    return Matches(np.empty(0, np.int32), np.empty(0, np.int32), np.empty(0, np.float32))


def reference_match(desc1: Optional[np.ndarray], desc2: Optional[np.ndarray], norm: int = NORM_L2,
                    cross_check: bool = False, ratio_threshold: Optional[float] = None,
                    knn2=None) -> Matches:
    """Array form of `EnhancedBFMatcher.match` (feature_matchers.py:201-252).

    Mode A (:208-224)  ratio_threshold is not None and not cross_check: knn k=2, keep m iff the
                       pair has two entries and float64(m.distance) < ratio * float64(n.distance).
    Mode B (:225-238, crossCheck=True matcher): strict mutual first-argmin.
    Mode C (:225-238, crossCheck=False, no ratio): plain 1-NN for every query.
    Then the stable sort by score (:241): order = (distance asc, queryIdx asc).
    Failure convention (:204-205, :250-252): invalid inputs -> empty result, never raises.
    """
    knn2 = knn2 or knn2_c
    if desc1 is None or desc2 is None or len(desc1) == 0 or len(desc2) == 0:
        return _empty()
    try:
        _check(desc1, desc2, norm)
    except ValueError:
        return _empty()
    n = desc1.shape[0]
    fwd = knn2(desc1, desc2, norm)
    q = np.arange(n, dtype=np.int32)
    if ratio_threshold is not None and not cross_check:
        has2 = fwd.idx2 >= 0
        keep = has2 & (fwd.d1.astype(np.float64) < float(ratio_threshold) * fwd.d2.astype(np.float64))
    elif cross_check:
        bwd = knn2(desc2, desc1, norm)
        keep = bwd.idx1[fwd.idx1] == q
    else:
        keep = np.ones(n, bool)
    q, t, d = q[keep], fwd.idx1[keep], fwd.d1[keep]
    order = np.argsort(d, kind="stable")       # Python's list.sort(key=score) is stable
    return Matches(q[order].astype(np.int32), t[order].astype(np.int32), d[order].astype(np.float32))


# ----------------------------------------------------------------------------- live cv2 (when present)
def cv2_match(desc1, desc2, norm=NORM_L2, cross_check=False, ratio_threshold=None, threads=None) -> Matches:
    """The same three modes driven by the real cv2.BFMatcher, exactly as the reference drives it.
    Used (a) to pin the restatement in tests and (b) as the timed CPU arm of bench.py: cv2 is the
    reference's own arithmetic engine for this path."""
    import cv2
    if threads:
        cv2.setNumThreads(int(threads))
    if desc1 is None or desc2 is None or len(desc1) == 0 or len(desc2) == 0:
        return _empty()
    try:
        matcher = cv2.BFMatcher(norm, crossCheck=cross_check)
        out = []
        if ratio_threshold is not None and not cross_check:
            for pair in matcher.knnMatch(desc1, desc2, k=2):
                if len(pair) == 2:
                    m, n2 = pair
                    if m.distance < ratio_threshold * n2.distance:
                        out.append((m.queryIdx, m.trainIdx, m.distance))
        else:
            for m in matcher.match(desc1, desc2):
                out.append((m.queryIdx, m.trainIdx, m.distance))
        out.sort(key=lambda x: x[2])
    except Exception:
        return _empty()
    if not out:
        return _empty()
    arr = np.array(out, dtype=np.float64)
    return Matches(arr[:, 0].astype(np.int32), arr[:, 1].astype(np.int32), arr[:, 2].astype(np.float32))


def cv2_knn2(desc1, desc2, norm) -> Knn2:
    import cv2
    n = desc1.shape[0]
    idx1 = np.full(n, -1, np.int32); idx2 = np.full(n, -1, np.int32)
    d1 = np.full(n, -1, np.float32); d2 = np.full(n, -1, np.float32)
    for i, pair in enumerate(cv2.BFMatcher(norm).knnMatch(desc1, desc2, k=2)):
        if len(pair) >= 1:
            idx1[i], d1[i] = pair[0].trainIdx, pair[0].distance
        if len(pair) >= 2:
            idx2[i], d2[i] = pair[1].trainIdx, pair[1].distance
    return Knn2(idx1, idx2, d1, d2)
