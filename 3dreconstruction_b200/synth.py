"""Seeded synthetic descriptor sets for the BASELINE.json configs.

Uniform-random descriptors give zero Lowe-ratio survivors, so every generator plants
noisy true correspondences.  Shapes/dtypes follow the reference's descriptor producers
(/root/reference/FeatureMatchingExtraction/traditional_detectors.py:42 SIFT float32[N,128]
integer-valued, :80 ORB uint8[N,32], :107-117 AKAZE uint8[N,61], :144-154 BRISK uint8[N,64];
deep_learning_detectors.py:124 SuperPoint float32 256-d).
"""
from __future__ import annotations

import numpy as np


def sift_like(rng: np.random.Generator, n: int, d: int = 128) -> np.ndarray:
    """SIFT-style rows: gamma(0.6) -> L2-normalise -> clip 0.2 -> renormalise -> rint(512 x) in 0..255."""
    x = rng.gamma(0.6, 1.0, size=(n, d))
    x /= np.linalg.norm(x, axis=1, keepdims=True) + 1e-12
    x = np.minimum(x, 0.2)
    x /= np.linalg.norm(x, axis=1, keepdims=True) + 1e-12
    return np.clip(np.rint(512.0 * x), 0, 255).astype(np.float32)


def sift_pair(seed: int = 0, n: int = 8192, m: int | None = None, d: int = 128, sigma: float = 6.0):
    """Config 1: image 2 = permutation of image 1; first half + N(0, sigma) re-rounded (true
    matches), second half replaced by fresh descriptors."""
    rng = np.random.default_rng(seed)
    m = n if m is None else m
    a = sift_like(rng, n, d)
    src = rng.permutation(n)[:m] if m <= n else rng.integers(0, n, size=m)
    b = a[src].copy()
    half = m // 2
    b[:half] = np.clip(np.rint(b[:half] + rng.normal(0.0, sigma, size=(half, d))), 0, 255)
    b[half:] = sift_like(rng, m - half, d)
    return a, np.ascontiguousarray(b.astype(np.float32))


def binary_pair(seed: int = 1, n: int = 10000, m: int | None = None, nbytes: int = 32, flip: float = 0.08):
    """Config 2 (ORB 256-bit) / config 5 (AKAZE 61 B, BRISK 64 B): image 2 = permutation; first half
    with each bit flipped w.p. `flip`, second half fresh."""
    rng = np.random.default_rng(seed)
    m = n if m is None else m
    a = rng.integers(0, 256, size=(n, nbytes), dtype=np.uint8)
    src = rng.permutation(n)[:m] if m <= n else rng.integers(0, n, size=m)
    b = a[src].copy()
    half = m // 2
    flips = rng.random(size=(half, nbytes * 8)) < flip
    b[:half] ^= np.packbits(flips, axis=1)
    b[half:] = rng.integers(0, 256, size=(m - half, nbytes), dtype=np.uint8)
    return a, np.ascontiguousarray(b)


def _unit(x: np.ndarray) -> np.ndarray:
    return (x / (np.linalg.norm(x, axis=1, keepdims=True) + 1e-12)).astype(np.float32)


def superpoint_set(seed: int = 2, images: int = 100, n: int = 4096, d: int = 256, pool: int = 8192,
                   shared: float = 0.5, sigma: float = 0.05) -> np.ndarray:
    """Config 3: `images` x float32[n, d] unit-norm rows; each image shares `shared` of a common
    pool of scene descriptors perturbed by N(0, sigma) then renormalised."""
    rng = np.random.default_rng(seed)
    scene = _unit(rng.normal(size=(pool, d)))
    out = np.empty((images, n, d), np.float32)
    k = int(n * shared)
    for i in range(images):
        pick = rng.choice(pool, size=k, replace=False)
        out[i, :k] = _unit(scene[pick] + rng.normal(0.0, sigma, size=(k, d)))
        out[i, k:] = _unit(rng.normal(size=(n - k, d)))
        out[i] = out[i][rng.permutation(n)]
    return out


def sift_set(seed: int = 3, images: int = 200, n: int = 8192, d: int = 128, pool: int | None = None,
             shared: float = 0.3, sigma: float = 6.0) -> np.ndarray:
    """Config 4: `images` x float32[n,128] SIFT-like, `shared` of each image drawn from a shared
    pool (+noise sigma), the rest fresh."""
    rng = np.random.default_rng(seed)
    pool = pool or 2 * n
    scene = sift_like(rng, pool, d)
    out = np.empty((images, n, d), np.float32)
    k = int(n * shared)
    for i in range(images):
        pick = rng.choice(pool, size=k, replace=False)
        out[i, :k] = np.clip(np.rint(scene[pick] + rng.normal(0.0, sigma, size=(k, d))), 0, 255)
        out[i, k:] = sift_like(rng, n - k, d)
        out[i] = out[i][rng.permutation(n)]
    return out


def binary_set(seed: int = 4, images: int = 1000, n: int = 16384, nbytes: int = 61, pool: int | None = None,
               shared: float = 0.5, flip: float = 0.08) -> np.ndarray:
    """Config 5: `images` x uint8[n, nbytes] with planted correspondences through a shared pool."""
    rng = np.random.default_rng(seed)
    pool = pool or 2 * n
    scene = rng.integers(0, 256, size=(pool, nbytes), dtype=np.uint8)
    out = np.empty((images, n, nbytes), np.uint8)
    k = int(n * shared)
    for i in range(images):
        pick = rng.choice(pool, size=k, replace=False)
        flips = rng.random(size=(k, nbytes * 8)) < flip
        out[i, :k] = scene[pick] ^ np.packbits(flips, axis=1)
        out[i, k:] = rng.integers(0, 256, size=(n - k, nbytes), dtype=np.uint8)
        out[i] = out[i][rng.permutation(n)]
    return out


def all_pairs(images: int) -> np.ndarray:
    """The (i<j) list in itertools.combinations order
    (/root/reference/FeatureMatchingExtraction/image_manager.py:491-494)."""
    i, j = np.triu_indices(images, k=1)
    return np.stack([i, j], axis=1).astype(np.int32)


def homography_pair(seed: int = 5, n: int = 1000, inlier_frac: float = 0.5, noise: float = 1.0,
                    width: float = 1920.0, height: float = 1080.0):
    """Matched keypoint coordinates of one image pair for the geometric-verification stage: float32 [n, 4] rows
    (x1, y1, x2, y2).  A fraction of the matches follows one random perspective map (+ Gaussian pixel noise), the
    rest pairs uniformly random points (wrong matches); rows are shuffled."""
    rng = np.random.default_rng(seed)
    p1 = rng.random((n, 2)) * [width, height]
    H = np.array([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]])
    H += rng.normal(0.0, 1.0, (3, 3)) * [[0.08, 0.08, 40.0], [0.08, 0.08, 40.0], [2e-5, 2e-5, 0.0]]
    q = np.c_[p1, np.ones(n)] @ H.T
    p2 = q[:, :2] / q[:, 2:]
    k = int(round(n * inlier_frac))
    p2[:k] += rng.normal(0.0, noise, (k, 2))
    p2[k:] = rng.random((n - k, 2)) * [width, height]
    out = np.concatenate([p1, p2], axis=1).astype(np.float32)
    return out[rng.permutation(n)]


def homography_set(seed: int = 6, pairs: int = 64, n_lo: int = 50, n_hi: int = 2000, inl_lo: float = 0.3, inl_hi: float = 0.9):
    """`pairs` correspondence sets of varying size and inlier ratio as one CSR: (float32 [K, 4], int64 ptr [pairs + 1])."""
    rng = np.random.default_rng(seed)
    sets = [homography_pair(seed * 1000 + p, int(rng.integers(n_lo, n_hi + 1)), float(rng.uniform(inl_lo, inl_hi)))
            for p in range(pairs)]
    ptr = np.zeros(pairs + 1, np.int64)
    ptr[1:] = np.cumsum([len(s) for s in sets])
    return np.concatenate(sets, axis=0), ptr


def two_view_pair(seed: int = 7, n: int = 1000, inlier_frac: float = 0.6, noise: float = 0.7, size=(1600, 1200),
                  focal_ratio: float = 1.2):
    """Correspondences of a rigid scene seen by two cameras, for the essential-matrix stage: float32 [n, 4] pixel
    coordinates (x1, y1, x2, y2); `inlier_frac` of them follow the epipolar geometry (+ Gaussian pixel noise), the rest
    are uniformly random in the second image.  Camera as the reference estimates it when none is given
    (essential_estimation.py:36-60): f = focal_ratio * width, principal point at the image centre.  Returns (corr, K)."""
    rng = np.random.default_rng(seed)
    w, h = size
    f = w * focal_ratio
    K = np.array([[f, 0, w / 2.0], [0, f, h / 2.0], [0, 0, 1]], np.float64)
    X = np.concatenate([rng.uniform(-2.5, 2.5, size=(n, 2)), rng.uniform(5, 12, size=(n, 1))], axis=1)
    a = rng.uniform(-0.25, 0.25, size=3)
    th = np.linalg.norm(a) + 1e-12
    k = a / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx
    t = rng.uniform(-1.0, 1.0, size=3)
    Y = X @ R.T + t
    p1 = X[:, :2] / X[:, 2:] * f + K[:2, 2]
    p2 = Y[:, :2] / Y[:, 2:] * f + K[:2, 2]
    p1 += rng.normal(size=p1.shape) * noise
    p2 += rng.normal(size=p2.shape) * noise
    bad = rng.permutation(n)[:n - int(round(n * inlier_frac))]
    p2[bad] = np.stack([rng.uniform(0, w, len(bad)), rng.uniform(0, h, len(bad))], axis=1)
    return np.concatenate([p1, p2], axis=1).astype(np.float32), K
