"""CPU tests: the oracle (C + numpy restatements) against the golden vectors produced by the
reference's own EnhancedBFMatcher, against live cv2.BFMatcher, and against each other."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from conftest import is_integer_valued, pkg
from oracle import bf_oracle as O

synth = pkg("synth")


def _same(m, q, t, d, exact):
    assert np.array_equal(m.query_idx, q)
    assert np.array_equal(m.train_idx, t)
    if exact:
        assert np.array_equal(m.distance.view(np.uint32), d.view(np.uint32))
    else:
        np.testing.assert_allclose(m.distance, d, rtol=1e-5)


@pytest.mark.parametrize("impl", ["c", "numpy"])
def test_oracle_matches_reference_golden(golden, impl):
    knn2 = O.knn2_c if impl == "c" else O.knn2_numpy
    for name, g in golden.items():
        m = O.reference_match(g["desc1"], g["desc2"], g["norm"], g["cross_check"], g["ratio"], knn2=knn2)
        exact = is_integer_valued(g["desc1"]) and is_integer_valued(g["desc2"])
        try:
            _same(m, g["q"], g["t"], g["d"], exact)
        except AssertionError as e:
            raise AssertionError(f"golden case {name} ({impl}): {e}")


def test_oracle_vs_live_cv2_configs():
    pytest.importorskip("cv2")
    a, b = synth.sift_pair(seed=0, n=1024)
    _same(O.reference_match(a, b, O.NORM_L2, False, 0.75), *O.cv2_match(a, b, O.NORM_L2, False, 0.75), True)
    _same(O.reference_match(a, b, O.NORM_L2, True, None), *O.cv2_match(a, b, O.NORM_L2, True, None), True)
    a, b = synth.binary_pair(seed=1, n=1500)
    _same(O.reference_match(a, b, O.NORM_HAMMING, True, None), *O.cv2_match(a, b, O.NORM_HAMMING, True, None), True)
    _same(O.reference_match(a, b, O.NORM_HAMMING, False, 0.7), *O.cv2_match(a, b, O.NORM_HAMMING, False, 0.7), True)
    k = O.knn2_c(a, b, O.NORM_HAMMING)
    c = O.cv2_knn2(a, b, O.NORM_HAMMING)
    for x, y in zip(k, c):
        assert np.array_equal(x, y)


def test_planted_survivor_counts():
    a, b = synth.sift_pair(seed=0, n=2048)
    m = O.reference_match(a, b, O.NORM_L2, False, 0.75)
    assert 0.4 * 2048 < len(m.query_idx) <= 0.55 * 2048      # ~half planted
    a, b = synth.binary_pair(seed=1, n=2000)
    m = O.reference_match(a, b, O.NORM_HAMMING, True, None)
    assert len(m.query_idx) > 1000


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 40), m=st.integers(1, 40), nbytes=st.sampled_from([1, 2, 5, 32, 61, 64]),
       hi=st.sampled_from([2, 256]), seed=st.integers(0, 10 ** 6),
       mode=st.sampled_from([(False, None), (False, 0.7), (False, 1.0), (True, None)]))
def test_hamming_property_c_numpy_cv2(n, m, nbytes, hi, seed, mode):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, hi, size=(n, nbytes), dtype=np.uint8)
    b = rng.integers(0, hi, size=(m, nbytes), dtype=np.uint8)
    rc = O.reference_match(a, b, O.NORM_HAMMING, *mode, knn2=O.knn2_c)
    rn = O.reference_match(a, b, O.NORM_HAMMING, *mode, knn2=O.knn2_numpy)
    _same(rc, *rn, True)
    try:
        import cv2  # noqa: F401
    except ImportError:
        return
    _same(rc, *O.cv2_match(a, b, O.NORM_HAMMING, *mode), True)


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 40), m=st.integers(1, 40), d=st.sampled_from([1, 3, 8, 128]),
       hi=st.sampled_from([2, 4, 256]), seed=st.integers(0, 10 ** 6),
       mode=st.sampled_from([(False, None), (False, 0.75), (True, None)]))
def test_l2_integer_property_c_numpy_cv2(n, m, d, hi, seed, mode):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, hi, size=(n, d)).astype(np.float32)
    b = rng.integers(0, hi, size=(m, d)).astype(np.float32)
    rc = O.reference_match(a, b, O.NORM_L2, *mode, knn2=O.knn2_c)
    rn = O.reference_match(a, b, O.NORM_L2, *mode, knn2=O.knn2_numpy)
    _same(rc, *rn, True)
    try:
        import cv2  # noqa: F401
    except ImportError:
        return
    _same(rc, *O.cv2_match(a, b, O.NORM_L2, *mode), True)


def test_error_convention_returns_empty():
    e = O.reference_match(np.zeros((3, 4), np.uint8), np.zeros((3, 5), np.uint8), O.NORM_HAMMING)
    assert len(e.query_idx) == 0
    assert len(O.reference_match(None, np.zeros((3, 5), np.uint8), O.NORM_HAMMING).query_idx) == 0
    assert len(O.reference_match(np.zeros((0, 5), np.uint8), np.zeros((3, 5), np.uint8), O.NORM_HAMMING).query_idx) == 0
