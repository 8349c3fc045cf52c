"""world_size-2/3 `gloo` tests (CPU) of the multi-GPU host logic in 3dreconstruction_b200/allpairs.py:
balanced partition of the (i<j) pair list, broadcast of the packed descriptor store, variable-length
gather of the per-rank CSRs to rank 0.  The per-rank matching itself needs a B200, so here each rank
fills its shard's CSR with the oracle (test infrastructure) — what is under test is that the sharded
job, reassembled on rank 0, equals the unsharded one pair for pair and in the reference's pair order
(itertools.combinations, /root/reference/FeatureMatchingExtraction/image_manager.py:491-494)."""
import itertools
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import pkg

b = pkg()
ap = pkg("allpairs")
SIZES = [40, 17, 1, 64, 23, 50, 9]


def _images():
    return [b.synth.binary_pair(seed=300 + i, n=max(s, 2), nbytes=32)[0][:s] for i, s in enumerate(SIZES)]


def _oracle_csr(imgs, pairs):
    from oracle import bf_oracle as O
    ptr, q, t, d = [0], [], [], []
    for i, j in pairs:
        m = O.reference_match(imgs[i], imgs[j], O.NORM_HAMMING, True, None)
        q.append(m.query_idx); t.append(m.train_idx); d.append(m.distance)
        ptr.append(ptr[-1] + len(m.query_idx))
    cat = lambda xs, dt: torch.from_numpy(np.concatenate(xs).astype(dt)) if xs else torch.empty(0, dtype=torch.from_numpy(np.empty(0, dt)).dtype)  # noqa: E731
    return ap.PairCSR(np.asarray(pairs, np.int32).reshape(-1, 2), torch.tensor(ptr, dtype=torch.int64),
                      cat(q, np.int32), cat(t, np.int32), cat(d, np.float32))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dev = torch.device("cpu")
        store = None
        if rank == 0:                                   # only rank 0 holds the descriptors before the broadcast
            imgs = _images()
            cat = torch.from_numpy(np.concatenate(imgs))
            offsets = np.zeros(len(imgs) + 1, np.int64); np.cumsum([len(x) for x in imgs], out=offsets[1:])
            store = ap.Store(b.DeviceDescriptors(b._lib.KIND_HAMMING, cat, None, int(cat.shape[0]), 32), offsets, b.NORM_HAMMING)
        store = ap.broadcast_store(store, b.NORM_HAMMING, dev)
        assert store.dd.rows == sum(SIZES) and store.dd.norm is None
        imgs = [store.dd.desc[store.offsets[i]:store.offsets[i + 1]].numpy() for i in range(len(SIZES))]
        pairs = ap.all_pairs(len(SIZES))
        shards = ap.partition_pairs(pairs, store.offsets, world)
        local = _oracle_csr(imgs, [tuple(p) for p in shards[rank]])
        full = ap.gather_csr(local, dev)
        if rank == 0:
            h = full.to_host()
            np.savez(out_path, pairs=h.pairs, ptr=h.ptr, q=h.query_idx, t=h.train_idx, d=h.distance)
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_allpairs_equals_unsharded(world, tmp_path):
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    z = np.load(out)
    imgs = _images()
    want_pairs = list(itertools.combinations(range(len(SIZES)), 2))
    assert [tuple(p) for p in z["pairs"]] == want_pairs            # reference pair order survives the round trip
    want = _oracle_csr(imgs, want_pairs).to_host()
    assert np.array_equal(z["ptr"], want.ptr)
    assert np.array_equal(z["q"], want.query_idx) and np.array_equal(z["t"], want.train_idx)
    assert np.array_equal(z["d"], want.distance)


def test_partition_is_balanced_contiguous_and_complete():
    rng = np.random.default_rng(0)
    counts = rng.integers(1, 9000, size=40)
    offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    pairs = ap.all_pairs(40)
    work = counts[pairs[:, 0]].astype(np.float64) * counts[pairs[:, 1]]
    for world in (1, 2, 4, 8):
        shards = ap.partition_pairs(pairs, offsets, world)
        assert len(shards) == world
        assert np.array_equal(np.concatenate(shards), pairs)        # contiguous, ordered, nothing lost or duplicated
        loads = [float((counts[s[:, 0]].astype(np.float64) * counts[s[:, 1]]).sum()) for s in shards]
        assert max(loads) <= work.sum() / world + work.max()         # within one pair of the ideal share
    assert all(len(s) == 0 for s in ap.partition_pairs(pairs[:0], offsets, 4))
    assert [len(s) for s in ap.partition_pairs(pairs[:2], offsets, 4)].count(0) >= 2   # more ranks than pairs


def _board_worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        D = pkg("distributed")
        bd = D._HostBoard(world, rank, dist.new_group(backend="gloo"))
        for k in range(3 * D._HostBoard.SLOTS):                 # more exchanges than slots: the ring wraps
            got = bd.exchange(rank * 100000 + k)
            if rank == 0:
                assert got == [r * 100000 + k for r in range(world)]
            else:
                assert got is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_host_board_rendezvous(world):
    """distributed._HostBoard: the shared-memory count exchange of the streamed return path (rank 0 learns every rank's
    record count per image group without a device collective)."""
    mp.spawn(_board_worker, args=(world, _free_port()), nprocs=world, join=True)


def test_group_plan_covers_every_pair_once_and_blocks_tile_groups():
    D = pkg("distributed")
    for I, G in [(1, 4), (2, 1), (7, 3), (10, 4), (25, 24), (50, 8)]:
        plan = D.group_plan(I, G)
        assert [(a, c) for a, c, _ in plan] == [(j, min(I, j + G)) for j in range(0, I, G)]
        got = sorted(tuple(p) for _, _, gp in plan for p in gp)
        assert got == list(itertools.combinations(range(I), 2))
        for j0, j1, gp in plan:
            assert all(j0 <= j < j1 and i < j for i, j in gp)       # a group's pairs need only images < j1: resident by then
    for world in (1, 2, 8):
        cuts = D.block_cuts(24, 48, world)
        assert cuts[0] == 24 and cuts[-1] == 48 and len(cuts) == world + 1 and cuts == sorted(cuts)
