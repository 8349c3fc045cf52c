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


def test_reference_class_path_end_to_end_with_the_reference_imported():
    """VERDICT r1 parity hole 2: everything that only runs when the REFERENCE's package is imported -- install(), the factory
    building our class, validate_features inherited from the reference's ABC, contracts.resolve() handing back the
    reference's own MatchData / EnhancedDMatch, the positional fast path of matcher._materialise keyed on the reference's
    field order, ResultConverter._convert_method and a pickle round trip -- against the reference's own EnhancedBFMatcher on
    the same inputs, field by field.  The reference tree cannot travel to the GPU box (its setup.py installs no modules:
    DESIGN.md 2), so the kernels are replaced here by the oracle's arrays behind Engine.match_arrays; the kernels themselves
    are compared with the same oracle in tests/test_parity_gpu.py."""
    import pickle
    import cv2
    from oracle import bf_oracle as O
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present on this box")
    ref_loader._MODULES = tuple(dict.fromkeys(ref_loader._MODULES + ("utils", "result_types", "result_converters")))
    ref = ref_loader.load()
    cds, rt = ref.core_data_structures, ref.result_types
    conv = ref.result_converters.ResultConverter()

    class OracleEngine:                                  # stands in for the GPU: same signature as Engine.match_arrays
        def match_arrays(self, d1, d2, norm, cc, ratio):
            m = O.reference_match(d1, d2, norm, cc, ratio)
            return b.MatchArrays(m.query_idx, m.train_idx, m.distance)

    rng = np.random.default_rng(5)
    xy = lambda n: [cv2.KeyPoint(float(x), float(y), 1.0) for x, y in rng.uniform(0, 900, size=(n, 2))]   # noqa: E731
    a8, c8 = b.synth.binary_pair(seed=21, n=400, m=350, nbytes=32)
    af, cf = b.synth.sift_pair(seed=22, n=300)
    cases = [("ORB", a8, c8, {}), ("ORB", a8, c8, {"cross_check": True}), ("SIFT", af, cf, {"ratio_threshold": 0.8}),
             ("SIFT", af, cf[:1], {"ratio_threshold": 0.8}), ("ORB", a8, c8, {"ratio_threshold": None})]
    try:
        assert len(b.install()) == 2
        for method, d1, d2, params in cases:
            f1 = cds.FeatureData(keypoints=xy(len(d1)), descriptors=d1, method=method)
            f2 = cds.FeatureData(keypoints=xy(len(d2)), descriptors=d2, method=method)
            ours = ref.matcher_factory.MatcherFactory()._create_bf_matcher(method, dict(params))
            theirs_params = {"norm_type": ours.norm_type, "cross_check": ours.cross_check, "ratio_threshold": ours.ratio_threshold}
            theirs = ref.matcher_factory._CV2_EnhancedBFMatcher(**theirs_params)
            assert isinstance(ours, b.B200BFMatcher) and isinstance(ours, ref.base_classes.BaseFeatureMatcher)
            ours._engine = OracleEngine()
            got, want = ours.match(f1, f2), theirs.match(f1, f2)
            assert type(got) is cds.MatchData and type(want) is cds.MatchData
            assert got.method == want.method == "BF" and got.score_type == want.score_type
            assert len(got.matches) == len(want.matches) and isinstance(got.matches, list)
            for g, w in zip(got.matches, want.matches):
                assert type(g) is cds.EnhancedDMatch
                for field in ("queryIdx", "trainIdx", "score", "score_type", "imgIdx", "raw_distance", "confidence", "source_method"):
                    assert getattr(g, field) == getattr(w, field), (method, params, field)
                assert g.distance == w.distance and type(g.score) is float and type(g.queryIdx) is int
                dm = g.to_cv2_dmatch()
                assert (dm.queryIdx, dm.trainIdx) == (w.queryIdx, w.trainIdx) and dm.distance == np.float32(w.distance)
            assert [m.queryIdx for m in got.get_best_matches()] == [m.queryIdx for m in want.get_best_matches()]
            # the consumers of MatchData: the converter of the SfM hand-off and the per-pair pickle (result_types.py:173)
            rg = conv._convert_method(rt.MethodResult(method_name=method, features1=f1, features2=f2, match_data=got))
            rw = conv._convert_method(rt.MethodResult(method_name=method, features1=f1, features2=f2, match_data=want))
            assert np.array_equal(rg.query_indices, rw.query_indices) and np.array_equal(rg.train_indices, rw.train_indices)
            assert np.array_equal(rg.match_scores, rw.match_scores)
            assert (rg.correspondences is None) == (rw.correspondences is None)
            if rw.correspondences is not None:
                assert np.array_equal(rg.correspondences, rw.correspondences)
            back = pickle.loads(pickle.dumps(got))
            assert type(back) is cds.MatchData and back.matches == got.matches
            got.filtered_matches = got.matches[:3]            # callers mutate the result (pipeline.py:309-310)
            assert len(got.get_best_matches()) == min(3, len(want.matches))
    finally:
        b.uninstall()
