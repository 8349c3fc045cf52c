"""CPU tests of the boundary: the C-ABI library loads and exports every symbol include/b200match.h
declares (no compute calls without a GPU), host-side logic, and the loud failure without a device."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, pkg

b = pkg()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200match.h")).read()
    return sorted(set(re.findall(r"B200M_API[^;]*?\b(b200m_\w+)\s*\(", text)))


def test_header_symbols_all_exported():
    lib = b._lib.load()
    syms = _declared_symbols()
    assert len(syms) >= 11
    assert set(syms) == set(b._lib.EXPORTS)
    for s in syms:
        assert getattr(lib, s) is not None, s


def test_version_and_error_strings():
    lib = b._lib.load()
    assert lib.b200m_version() == 200
    assert lib.b200m_error_string(0) == b"ok"
    assert b"workspace" in lib.b200m_error_string(-2)
    assert b"CPU" in lib.b200m_error_string(-4)


def test_workspace_query_needs_no_device():
    lib = b._lib.load()
    n = ctypes.c_size_t(0)
    assert lib.b200m_workspace_bytes(0, 1, 3, 30000, 30000, 10000, 10000, 32, ctypes.byref(n)) == 0
    assert n.value > 3 * 10000 * 16
    assert lib.b200m_workspace_bytes(0, 1, -1, 0, 0, 0, 0, 32, ctypes.byref(n)) == b._lib.EINVAL


def test_pair_struct_layout_matches_header():
    assert ctypes.sizeof(b._lib.Pair) == 24
    assert b._lib.Pair.out_off.offset == 16


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_gpu():
    with pytest.raises(b._lib.B200MatchError):
        b.Engine()
    m = b.B200BFMatcher(norm_type=b.NORM_HAMMING)
    f = b.contracts.FeatureData(keypoints=[0, 1], descriptors=np.zeros((2, 32), np.uint8), method="ORB")
    with pytest.raises(b._lib.B200MatchError):       # environment failure is loud, not an empty MatchData
        m.match(f, f)


def test_validate_features_convention():
    m = b.B200BFMatcher()
    F = b.contracts.FeatureData
    empty = F(keypoints=[], descriptors=None, method="SIFT")
    md = m.match(empty, empty)        # invalid input -> empty MatchData before any device work
    assert md.matches == [] and md.method == "BF" and md.score_type == b.contracts.ScoreType.DISTANCE


def test_mode_selection_mirrors_reference_branches():
    L = b._lib
    assert b.mode_of(False, 0.7) == L.MODE_RATIO
    assert b.mode_of(True, 0.7) == L.MODE_CROSS      # ratio ignored when cross_check (feature_matchers.py:208)
    assert b.mode_of(True, None) == L.MODE_CROSS
    assert b.mode_of(False, None) == L.MODE_NN


def test_synth_shapes_and_dtypes():
    s = b.synth
    a, c = s.sift_pair(seed=0, n=64)
    assert a.dtype == np.float32 and a.shape == (64, 128) and np.all(a == np.rint(a)) and a.max() <= 255
    a, c = s.binary_pair(seed=1, n=50, nbytes=61)
    assert a.dtype == np.uint8 and c.shape == (50, 61)
    assert s.all_pairs(5).shape == (10, 2) and tuple(s.all_pairs(5)[0]) == (0, 1)


def test_install_into_reference_modules():
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present on this box")
    ref = ref_loader.load()
    try:
        patched = b.install()
        assert len(patched) == 2
        cls = ref.matcher_factory.EnhancedBFMatcher
        assert issubclass(cls, ref.base_classes.BaseFeatureMatcher) and issubclass(cls, b.B200BFMatcher)
        # the reference factory path (matcher_factory.py:210-234) now builds our matcher, same ctor args
        m = ref.matcher_factory.MatcherFactory()._create_bf_matcher("ORB", {})
        assert isinstance(m, b.B200BFMatcher) and m.norm_type == 6 and m.ratio_threshold == 0.7 and not m.cross_check
        m2 = ref.feature_matchers.create_traditional_matcher("BF", norm_type=4, cross_check=True)
        assert isinstance(m2, b.B200BFMatcher) and m2.cross_check
    finally:
        b.uninstall()
    assert ref.matcher_factory.EnhancedBFMatcher is ref.matcher_factory._CV2_EnhancedBFMatcher
