"""Resumable all-pairs jobs: `progress.json` in the reference's own layout plus array shards of the finished pairs
(SURVEY 8f row N4).

The reference checkpoints its per-pair loop with `BatchProcessor`
(/root/reference/FeatureMatchingExtraction/batch_processor.py:101-128 `save_progress`, :130 `is_completed`): after every
pair it rewrites `<output_dir>/progress.json` = {"completed_pairs": sorted ids, "total_completed", "last_updated",
"metadata": {"batch_size", "started"}}, pair ids being `pair_{i:03d}` with i the index in the pair list
(pipeline.py:572; the list is itertools.combinations, image_manager.py:491-494).  `ProgressFile` reads and writes exactly
that file, so a job started by either side can be inspected -- and its finished pairs skipped -- by the other:
the reference's `BatchProcessor(output_dir).is_completed('pair_042')` answers from a file written here and vice versa.

What the reference stores per pair is a pickle of Python match objects (result_types.py:148-228); here a finished BLOCK of
pairs is one `.npz` shard of flat arrays (pair indices, counts, packed match records), written BEFORE its ids are added
to progress.json: an id in the file means the pair's matches are on disk.  A 45-600 s `configs[4]` job interrupted at
any point resumes with the unfinished pairs only."""
from __future__ import annotations

import json
import os
import time
from pathlib import Path
from typing import Iterable, Optional, Sequence

import numpy as np


def pair_id(p: int) -> str:
    """The reference's id of the p-th pair of the list (pipeline.py:572)."""
    return f"pair_{int(p):03d}"


def pair_index(i, j, num_images: int):
    """Index of pair (i, j), i < j, in itertools.combinations(range(num_images), 2) order (scalar or arrays)."""
    i = np.asarray(i, np.int64)
    j = np.asarray(j, np.int64)
    return i * (2 * num_images - i - 1) // 2 + (j - i - 1)


class ProgressFile:
    """`progress.json` of batch_processor.py, field for field.  Differences in behaviour, none in format: the file is
    replaced atomically (write + rename), and `save_many` adds a whole block of ids with one rewrite (the reference
    rewrites the file once per pair, which is O(P^2) bytes for a 499 500-pair job)."""

    def __init__(self, output_dir, batch_size: int = 10):
        self.output_dir = Path(output_dir)
        self.progress_file = self.output_dir / "progress.json"
        self.batch_size = batch_size
        self.completed_pairs = set()
        self.metadata = {}
        if self.progress_file.exists():
            try:
                data = json.loads(self.progress_file.read_text())
                self.completed_pairs = set(data.get("completed_pairs", []))
                self.metadata = dict(data.get("metadata", {}))
            except (json.JSONDecodeError, OSError) as e:      # batch_processor.py:81-86: a corrupt file means "start fresh"
                print(f"Warning: could not read {self.progress_file}, starting fresh: {e}")
        if not self.metadata:                                  # batch_processor.py:61-65
            self.metadata = {"batch_size": batch_size, "started": time.strftime("%Y-%m-%d %H:%M:%S")}

    def is_completed(self, pid: str) -> bool:
        return pid in self.completed_pairs

    def save_progress(self, pid: str) -> None:
        self.save_many([pid])

    def save_many(self, ids: Iterable[str]) -> None:
        self.completed_pairs.update(ids)
        data = {"completed_pairs": sorted(self.completed_pairs), "total_completed": len(self.completed_pairs),
                "last_updated": time.strftime("%Y-%m-%d %H:%M:%S"), "metadata": self.metadata}
        self.output_dir.mkdir(parents=True, exist_ok=True)
        tmp = self.progress_file.with_suffix(".json.tmp")
        tmp.write_text(json.dumps(data, indent=2))
        os.replace(tmp, self.progress_file)

    def reset(self) -> None:
        self.completed_pairs.clear()
        if self.progress_file.exists():
            self.progress_file.unlink()
        self.metadata = {"batch_size": self.batch_size, "started": time.strftime("%Y-%m-%d %H:%M:%S")}

    def get_progress_info(self) -> dict:                       # batch_processor.py:166-180
        return {"completed_pairs": len(self.completed_pairs), "pairs_list": sorted(self.completed_pairs),
                "last_updated": self.metadata.get("last_updated", "Never"),
                "batch_size": self.metadata.get("batch_size", self.batch_size),
                "started": self.metadata.get("started", "Unknown")}


class AllPairsCheckpoint:
    """Shards + progress.json of ONE all-pairs job over `num_images` images.

        ck = AllPairsCheckpoint(out_dir, num_images=I)
        dap.match_host(host, counts, ..., consumer=ck.consumer, keep=False, done_mask=ck.done_mask())   # resumes
        pairs, ptr, rec = ck.load()                                                                     # everything so far

    `consumer` has the signature distributed.DistributedAllPairs hands blocks over with (pairs [K, 2], counts [K], rec);
    `save_block` takes the same from any other driver (run_checkpointed below uses it with AllPairsMatcher)."""

    def __init__(self, output_dir, num_images: int, batch_size: int = 10):
        self.dir = Path(output_dir)
        self.num_images = int(num_images)
        self.total_pairs = self.num_images * (self.num_images - 1) // 2
        self.progress = ProgressFile(self.dir, batch_size)
        self.shard_dir = self.dir / "matching_results"            # the directory name the reference uses (pipeline.py:630)
        self._next = len(list(self.shard_dir.glob("block_*.npz"))) if self.shard_dir.exists() else 0

    def done_mask(self) -> np.ndarray:
        """bool [P] over itertools.combinations order: True = finished (listed in progress.json)."""
        m = np.zeros(self.total_pairs, bool)
        for pid in self.progress.completed_pairs:
            try:
                p = int(pid.split("_", 1)[1])
            except (IndexError, ValueError):
                continue
            if 0 <= p < self.total_pairs:
                m[p] = True
        return m

    def save_block(self, pairs: np.ndarray, counts: np.ndarray, rec: np.ndarray) -> None:
        if len(pairs) == 0:
            return
        pairs = np.asarray(pairs)
        idx = pair_index(pairs[:, 0], pairs[:, 1], self.num_images)
        self.shard_dir.mkdir(parents=True, exist_ok=True)
        path = self.shard_dir / f"block_{self._next:06d}.npz"
        tmp = self.shard_dir / f".part_{self._next:06d}.npz"       # not matched by the block_*.npz glob
        np.savez(tmp, pair_index=idx, pairs=pairs.astype(np.int32), counts=np.asarray(counts, np.int32), rec=np.asarray(rec))
        os.replace(tmp, path)                                     # the shard is complete on disk ...
        self._next += 1
        self.progress.save_many(pair_id(p) for p in idx)         # ... before its pairs count as done

    consumer = save_block

    def load(self):
        """(pairs [K, 2], ptr int64 [K + 1], rec) of every finished pair, in itertools.combinations order.  A pair that
        appears in more than one shard (a block was written, the job died before progress.json was) is taken once."""
        idxs, prs, cnts, recs = [], [], [], []
        for path in sorted(self.shard_dir.glob("block_*.npz")) if self.shard_dir.exists() else []:
            z = np.load(path)
            idxs.append(z["pair_index"]); prs.append(z["pairs"]); cnts.append(z["counts"]); recs.append(z["rec"])
        if not idxs:
            return np.zeros((0, 2), np.int32), np.zeros(1, np.int64), np.zeros(0, np.uint8)
        idx, pr, cnt, rec = np.concatenate(idxs), np.concatenate(prs), np.concatenate(cnts), np.concatenate(recs)
        start = np.concatenate([[0], np.cumsum(cnt, dtype=np.int64)])[:-1]
        uniq, first = np.unique(idx, return_index=True)            # sorted by pair index; first occurrence wins
        ptr = np.zeros(len(uniq) + 1, np.int64)
        np.cumsum(cnt[first], out=ptr[1:])
        take = np.concatenate([np.arange(start[k], start[k] + cnt[k]) for k in first]) if len(first) else np.zeros(0, np.int64)
        return pr[first], ptr, rec[take]


def run_checkpointed(matcher, store, num_images: int, ck: AllPairsCheckpoint, cross_check: bool = False,
                     ratio_threshold: Optional[float] = None, block_pairs: int = 4096, stop_after_blocks: Optional[int] = None) -> int:
    """Single-GPU driver: match the unfinished pairs of the job in blocks of `block_pairs` with allpairs.AllPairsMatcher,
    writing a shard and updating progress.json after each block's results have reached the host.  Returns the number of
    pairs matched by THIS call (`stop_after_blocks` ends the call early: tests and time-boxed runs)."""
    from .allpairs import all_pairs
    pairs = all_pairs(num_images)
    todo = pairs[~ck.done_mask()]
    done = 0
    for k, c0 in enumerate(range(0, len(todo), block_pairs)):
        if stop_after_blocks is not None and k >= stop_after_blocks:
            break
        blk = todo[c0:c0 + block_pairs]
        r = matcher.match(store, blk, cross_check, ratio_threshold).to_host()
        rec = np.zeros(len(r.query_idx), dtype=[("q", "<i4"), ("t", "<i4"), ("d", "<f4")])
        rec["q"], rec["t"], rec["d"] = r.query_idx, r.train_idx, r.distance
        ck.save_block(blk, np.diff(r.ptr).astype(np.int32), rec)
        done += len(blk)
    return done
