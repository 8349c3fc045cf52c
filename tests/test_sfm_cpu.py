"""CPU tests of the SfM hand-off: the golden vectors made by the reference's own converter
(oracle/make_golden_sfm.py) against the oracle + a numpy gather, and the host-side container (dict layout the
CameraPoseEstimation stages read, npz round trip).  No GPU, no compute through the C-ABI."""
import numpy as np
import pytest

from conftest import pkg, sfm_golden_sets
from oracle import bf_oracle as O

sfm = pkg("sfm")


def oracle_sfm(g):
    """SfMMatches from the oracle: per pair reference_match + keypoint gather (restates result_converters.py:754-766)."""
    offs = np.concatenate([[0], np.cumsum(g["count"])])
    I = len(g["count"])
    pairs, ptr, q, t, d, c = [], [0], [], [], [], []
    for i in range(I):
        for j in range(i + 1, I):
            a, b = g["desc"][offs[i]:offs[i + 1]], g["desc"][offs[j]:offs[j + 1]]
            m = O.reference_match(a, b, g["norm"], g["cross_check"], g["ratio"])
            pairs.append((i, j)); ptr.append(ptr[-1] + len(m.query_idx))
            q.append(m.query_idx); t.append(m.train_idx); d.append(m.distance)
            c.append(np.hstack([g["xy"][offs[i]:offs[i + 1]][m.query_idx], g["xy"][offs[j]:offs[j + 1]][m.train_idx]]))
    names = [f"img_{i:02d}.jpg" for i in range(I)]
    return sfm.SfMMatches(names, [(1600, 1200)] * I, pairs, ptr, np.concatenate(q), np.concatenate(t),
                          np.concatenate(d), np.concatenate(c).reshape(-1, 4), feature_type="synthetic")


@pytest.mark.parametrize("name", ["orb_cross", "orb_ratio07", "sift_ratio075"])
def test_oracle_handoff_equals_reference_converter(name):
    g = sfm_golden_sets()[name]
    r = oracle_sfm(g)
    assert len(r) == len(g["pairs"])
    for p, (i, j) in enumerate(r.pairs):
        want = g["pairs"][(int(i), int(j))]
        got = r.pair_dict(p)
        assert np.array_equal(got["query_indices"], want["q"]) and np.array_equal(got["train_indices"], want["t"])
        assert got["match_scores"].dtype == np.float64 and np.array_equal(got["match_scores"], want["s"])
        assert got["correspondences"].dtype == np.float64 and np.array_equal(got["correspondences"], want["c"])


def test_matches_pickle_layout_is_what_the_pose_stages_read():
    r = oracle_sfm(sfm_golden_sets()["sift_ratio075"])
    blob = r.to_matches_pickle()
    assert set(blob) >= {"matches_data", "image_info"}
    assert set(blob["image_info"]) == set(r.names) and blob["image_info"][r.names[0]]["size"] == (1600, 1200)
    for key, pd in blob["matches_data"].items():
        # pair_selector._validate_matches_data (CameraPoseEstimation/pair_selector.py:54-67)
        assert isinstance(key, tuple) and len(key) == 2 and "correspondences" in pd
        c = pd["correspondences"]
        assert c.ndim == 2 and c.shape[1] == 4 and len(c) == pd["num_matches"] >= 1
        # MatchExtractor._extract_point_arrays (correspondence_manager.py:78-119)
        pts1, pts2 = np.array(c)[:, :2], np.array(c)[:, 2:4]
        assert pts1.shape == pts2.shape
        assert pd["score_type"] == "distance" and pd["method"] == "BF"
    # pairs without matches are left out; everything else is there
    kept = sum(1 for p in range(len(r)) if r.ptr[p + 1] > r.ptr[p])
    assert len(blob["matches_data"]) == kept and 0 < kept < len(r)
    assert blob["processing_stats"]["total_matches"] == sum(v["num_matches"] for v in blob["matches_data"].values())


def test_npz_round_trip(tmp_path):
    r = oracle_sfm(sfm_golden_sets()["sift_ratio075"])
    path = tmp_path / "matches.npz"
    r.save(path)
    s = sfm.SfMMatches.load(path)
    assert s.names == r.names and s.sizes == r.sizes and s.method == r.method and s.feature_type == r.feature_type
    for f in ("pairs", "ptr", "query_idx", "train_idx", "distance", "correspondences"):
        a, b = getattr(r, f), getattr(s, f)
        assert a.dtype == b.dtype and np.array_equal(a, b), f
    md = s.to_match_data(1)
    sl = s.pair_slice(1)
    assert [m.queryIdx for m in md.matches] == s.query_idx[sl].tolist()
    assert md.method == "BF" and md.score_type.value == "distance"


def test_keypoints_xy_accepts_cv2_keypoints_and_arrays():
    class KP:                      # duck-typed cv2.KeyPoint
        def __init__(self, x, y): self.pt = (x, y)
    xy = sfm.keypoints_xy([KP(1.5, 2.0), KP(3.0, 4.25)])
    assert xy.dtype == np.float32 and xy.tolist() == [[1.5, 2.0], [3.0, 4.25]]
    assert np.array_equal(sfm.keypoints_xy(np.array([[1, 2], [3, 4]], np.float64)), np.array([[1, 2], [3, 4]], np.float32))


def test_filtered_by_verification_host_logic():
    """SfMMatches.filtered (N3 hand-off): inliers kept in order, pairs without a model emptied, empty pairs stay empty."""
    verify = pkg("verify")
    ptr = np.array([0, 3, 3, 8, 10, 10], np.int64)
    K = 10
    m = sfm.SfMMatches(list("abcdef"), [None] * 6, [(0, 1), (0, 2), (0, 3), (0, 4), (0, 5)], ptr, np.arange(K), np.arange(K) + 100,
                       np.arange(K, dtype=np.float32), np.arange(4 * K).reshape(K, 4))
    mask = np.array([1, 0, 1, 1, 1, 0, 0, 1, 1, 1], np.uint8)
    v = verify.Verification(ptr, np.zeros((5, 3, 3)), mask, np.array([2, -1, 3, -1, -1], np.int32), np.zeros(5, np.int32))
    f = m.filtered(v)
    assert f.ptr.tolist() == [0, 2, 2, 5, 5, 5]
    assert f.query_idx.tolist() == [0, 2, 3, 4, 7] and f.train_idx.tolist() == [100, 102, 103, 104, 107]
    assert np.array_equal(f.correspondences, m.correspondences[[0, 2, 3, 4, 7]])
    assert v.kept(2).tolist() == [0, 1, 4] and v.kept(1).tolist() == [] and v.homography(3) is None
