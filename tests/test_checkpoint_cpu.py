"""progress.json compatibility: 3dreconstruction_b200/checkpoint.py against the reference's own
BatchProcessor (/root/reference/FeatureMatchingExtraction/batch_processor.py:101-130), both directions, plus the shard
store's round trip.  The reference module is loaded from /root/reference where that exists (this container); the format
test against a literal copy of the documented layout runs everywhere."""
import importlib.util
import itertools
import json
import os

import numpy as np
import pytest

from conftest import pkg

ck = pkg("checkpoint")
REF = os.path.join(os.environ.get("B200M_REFERENCE_ROOT", "/root/reference"), "FeatureMatchingExtraction", "batch_processor.py")


def _ref_batch_processor():
    if not os.path.isfile(REF):
        pytest.skip("reference tree not present")
    spec = importlib.util.spec_from_file_location("ref_batch_processor", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.BatchProcessor


def test_pair_index_is_combinations_order():
    for I in (2, 3, 7, 40):
        combos = list(itertools.combinations(range(I), 2))
        arr = np.array(combos)
        assert np.array_equal(ck.pair_index(arr[:, 0], arr[:, 1], I), np.arange(len(combos)))
    assert ck.pair_id(7) == "pair_007" and ck.pair_id(12345) == "pair_12345"        # f"pair_{i:03d}", pipeline.py:572


def test_progress_json_layout(tmp_path):
    p = ck.ProgressFile(tmp_path, batch_size=10)
    p.save_many(["pair_002", "pair_000"])
    p.save_progress("pair_001")
    data = json.loads((tmp_path / "progress.json").read_text())
    assert list(data) == ["completed_pairs", "total_completed", "last_updated", "metadata"]      # batch_processor.py:114-119
    assert data["completed_pairs"] == ["pair_000", "pair_001", "pair_002"] and data["total_completed"] == 3
    assert set(data["metadata"]) == {"batch_size", "started"} and data["metadata"]["batch_size"] == 10
    again = ck.ProgressFile(tmp_path)
    assert again.is_completed("pair_001") and not again.is_completed("pair_003")
    assert again.metadata == data["metadata"]                       # a resumed job keeps the original start time
    (tmp_path / "progress.json").write_text("{not json")
    assert ck.ProgressFile(tmp_path).completed_pairs == set()       # corrupt file: start fresh, as the reference does
    again.reset()
    assert not (tmp_path / "progress.json").exists()


def test_reference_batch_processor_reads_our_file_and_we_read_its(tmp_path):
    BP = _ref_batch_processor()
    ours = ck.ProgressFile(tmp_path / "a", batch_size=7)
    ours.save_many(ck.pair_id(p) for p in (0, 5, 41, 1234))
    ref = BP(tmp_path / "a")
    assert ref.is_completed("pair_041") and ref.is_completed("pair_1234") and not ref.is_completed("pair_002")
    assert ref.completed_pairs == ours.completed_pairs and ref.metadata == ours.metadata
    assert ref.get_progress_info()["completed_pairs"] == 4
    ref.save_progress("pair_002")                                   # the reference appends to a job we started ...
    assert ck.ProgressFile(tmp_path / "a").completed_pairs == ours.completed_pairs | {"pair_002"}
    (tmp_path / "b").mkdir()
    ref2 = BP(tmp_path / "b", batch_size=3)                         # ... and we resume a job the reference started
    for p in (3, 1, 2):
        ref2.save_progress(f"pair_{p:03d}")
    mine = ck.AllPairsCheckpoint(tmp_path / "b", num_images=5)
    assert np.array_equal(np.nonzero(mine.done_mask())[0], [1, 2, 3])
    assert mine.progress.metadata["batch_size"] == 3


def test_shards_round_trip_and_resume_bookkeeping(tmp_path):
    I = 6
    combos = np.array(list(itertools.combinations(range(I), 2)), np.int32)
    rng = np.random.default_rng(0)
    dt = np.dtype([("q", "<u2"), ("t", "<u2"), ("d", "<u2")])
    counts = rng.integers(0, 5, size=len(combos)).astype(np.int32)
    recs = [np.array([(p, k, 7 * p + k) for k in range(c)], dt) for p, c in enumerate(counts)]
    c = ck.AllPairsCheckpoint(tmp_path, I)
    order = rng.permutation(len(combos))                              # blocks arrive in processing order, not pair order
    for blk in (order[:4], order[4:9]):
        c.save_block(combos[blk], counts[blk], np.concatenate([recs[p] for p in blk]))
    c2 = ck.AllPairsCheckpoint(tmp_path, I)                           # a new process
    assert np.array_equal(np.nonzero(c2.done_mask())[0], np.sort(order[:9]))
    c2.save_block(combos[order[7:]], counts[order[7:]], np.concatenate([recs[p] for p in order[7:]]))   # overlaps pairs 7, 8: taken once
    pairs, ptr, rec = c2.load()
    assert np.array_equal(pairs, combos) and np.array_equal(np.diff(ptr), counts)
    assert np.array_equal(rec, np.concatenate(recs))
    assert c2.done_mask().all() and len(c2.progress.completed_pairs) == len(combos)
