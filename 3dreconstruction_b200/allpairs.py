"""All-pairs (i<j) exhaustive matching over a descriptor store resident in HBM — the batched form of
the per-pair loop in the reference's `match_folder`
(/root/reference/FeatureMatchingExtraction/pipeline.py:538-606; pair list = itertools.combinations,
image_manager.py:491-494), with detection hoisted out: descriptors are uploaded once, every pair is a
work item of one C-ABI call, and results come back as a CSR (pair_ptr, query_idx, train_idx, distance).

Multi-GPU: pairs are independent, so the (i<j) list is cut into balanced contiguous
blocks, one per rank; the packed store is broadcast once from rank 0 (NCCL over NVLink) and the
per-pair match lists return to rank 0 by a variable-length gather.  No collective sits inside the
matching itself.
"""
from __future__ import annotations

import ctypes
from typing import List, NamedTuple, Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from .engine import DeviceDescriptors, Engine, NORM_HAMMING, NORM_L2, mode_of


_PAIR_DTYPE = np.dtype([("q_off", "<i4"), ("n", "<i4"), ("t_off", "<i4"), ("m", "<i4"), ("out_off", "<i8")])
assert _PAIR_DTYPE.itemsize == ctypes.sizeof(L.Pair)


class Store(NamedTuple):
    dd: DeviceDescriptors            # all images' rows, concatenated
    offsets: np.ndarray              # int64 [I+1] first row of each image
    norm_type: int


class PairCSR(NamedTuple):
    pairs: np.ndarray                # int32 [P, 2] image indices (i, j)
    ptr: "torch.Tensor | np.ndarray"     # int64 [P+1]
    query_idx: "torch.Tensor | np.ndarray"
    train_idx: "torch.Tensor | np.ndarray"
    distance: "torch.Tensor | np.ndarray"

    def pair(self, p: int):
        s, e = int(self.ptr[p]), int(self.ptr[p + 1])
        return self.query_idx[s:e], self.train_idx[s:e], self.distance[s:e]

    def to_host(self) -> "PairCSR":
        f = lambda x: x.cpu().numpy() if isinstance(x, torch.Tensor) else x   # noqa: E731
        return PairCSR(self.pairs, f(self.ptr), f(self.query_idx), f(self.train_idx), f(self.distance))


def build_store(engine: Engine, images: Sequence[np.ndarray] | np.ndarray | torch.Tensor, norm_type: int,
                counts: Optional[Sequence[int]] = None) -> Store:
    """Upload and re-encode all images' descriptors.  `images` is a list of [n_i, D] arrays, an
    [I, n, D] array, or an already concatenated [rows, D] tensor plus `counts`."""
    if counts is not None:
        cat = images
        counts = list(counts)
    elif isinstance(images, np.ndarray) and images.ndim == 3:
        counts = [images.shape[1]] * images.shape[0]
        cat = images.reshape(-1, images.shape[2])
    else:
        counts = [len(x) for x in images]
        cat = np.concatenate(list(images), axis=0)
    if norm_type == NORM_HAMMING and not isinstance(cat, torch.Tensor) and cat.shape[1] % 16 != 0:
        # AKAZE's 61-byte rows: zero-pad to a 16-byte multiple once (adds 0 to every Hamming distance)
        pad = (-cat.shape[1]) % 16
        cat = np.pad(cat, ((0, 0), (0, pad)))
    offsets = np.zeros(len(counts) + 1, np.int64)
    np.cumsum(counts, out=offsets[1:])
    return Store(engine.prepare(cat, norm_type), offsets, norm_type)


def all_pairs(num_images: int) -> np.ndarray:
    i, j = np.triu_indices(num_images, k=1)
    return np.stack([i, j], axis=1).astype(np.int32)


def partition_pairs(pairs: np.ndarray, offsets: np.ndarray, world: int) -> List[np.ndarray]:
    """Balanced contiguous blocks of the pair list: block boundaries are placed on the prefix sum of
    per-pair work n_i * n_j, so every rank gets ~1/world of the distance evaluations while keeping
    the (i<j) order (consecutive pairs share image i: L2 locality)."""
    if len(pairs) == 0:
        return [pairs[:0] for _ in range(world)]
    n = np.diff(offsets)
    w = n[pairs[:, 0]].astype(np.float64) * n[pairs[:, 1]].astype(np.float64)
    cum = np.concatenate([[0.0], np.cumsum(w)])
    cuts = np.searchsorted(cum, np.linspace(0.0, cum[-1], world + 1)[1:-1], side="left")
    bounds = np.concatenate([[0], cuts, [len(pairs)]]).astype(np.int64)
    return [pairs[bounds[r]:bounds[r + 1]] for r in range(world)]


class AllPairsMatcher:
    """Chunked CSR matching of an arbitrary pair list over one Store (single GPU)."""

    def __init__(self, engine: Engine, chunk_pairs: int = 2048, pipeline: bool = True):
        self.engine = engine
        self.chunk_pairs = int(chunk_pairs)
        self.pipeline = bool(pipeline)     # False: every chunk on the caller's stream, one after the other (kernel timing)
        self._bufs = {}
        self._streams = None

    def _buf(self, name: str, numel: int, dtype) -> torch.Tensor:
        t = self._bufs.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype:
            t = torch.empty(max(numel, 1), dtype=dtype, device=self.engine.device)
            self._bufs[name] = t
        return t

    def _enqueue_chunk(self, store: Store, chunk: np.ndarray, mode: int, ratio: float, slot: int, stream_handle: int):
        """Plan + enqueue one chunk of pairs on `stream_handle` into the buffers of `slot`; returns (C, optr, oq, ot, od)."""
        eng, lib, dd = self.engine, self.engine.lib, store.dd
        n_of = np.diff(store.offsets)
        C = len(chunk)
        rec = np.zeros(C, dtype=_PAIR_DTYPE)
        rec["q_off"] = store.offsets[chunk[:, 0]]; rec["n"] = n_of[chunk[:, 0]]
        rec["t_off"] = store.offsets[chunk[:, 1]]; rec["m"] = n_of[chunk[:, 1]]
        arr = rec.ctypes.data_as(ctypes.POINTER(L.Pair))
        tn, tm = int(rec["n"].sum(dtype=np.int64)), int(rec["m"].sum(dtype=np.int64))
        mn, mm = int(rec["n"].max()), int(rec["m"].max())
        nbytes = ctypes.c_size_t(0)
        L.check(lib.b200m_workspace_bytes(dd.kind, mode, C, tn, tm, mn, mm, dd.width, ctypes.byref(nbytes)),
                "b200m_workspace_bytes")
        ws = self._buf(f"ws{slot}", max(nbytes.value, 1 << 20), torch.uint8)
        oq = self._buf(f"q{slot}", tn, torch.int32); ot = self._buf(f"t{slot}", tn, torch.int32)
        od = self._buf(f"d{slot}", tn, torch.float32); oc = self._buf(f"c{slot}", C, torch.int32)
        optr = self._buf(f"p{slot}", C + 1, torch.int64)
        with torch.cuda.device(eng.device):
            rc = lib.b200m_match_pairs_csr(dd.kind, dd.desc.data_ptr(), 0 if dd.norm is None else dd.norm.data_ptr(),
                                           dd.rows, dd.desc.data_ptr(), 0 if dd.norm is None else dd.norm.data_ptr(),
                                           dd.rows, dd.width, arr, C, mode, ratio, oq.data_ptr(), ot.data_ptr(),
                                           od.data_ptr(), oc.data_ptr(), optr.data_ptr(), ws.data_ptr(), ws.numel(),
                                           stream_handle)
        L.check(rc, "b200m_match_pairs_csr")      # the pair table was staged inside the call: `rec` may go away
        return C, optr, oq, ot, od

    def match(self, store: Store, pairs: np.ndarray, cross_check: bool = False,
              ratio_threshold: Optional[float] = None) -> PairCSR:
        """Returns a device-resident CSR over `pairs` (order preserved).

        More than one chunk: chunks alternate between two side streams with their own workspaces, so that the
        tail of chunk k (finalize, scan, sort, result copy, the host's size read-back) runs while the knn kernel
        of chunk k+1 already occupies the SMs; the caller's stream waits for both at the end."""
        eng = self.engine
        mode = mode_of(cross_check, ratio_threshold)
        ratio = float(ratio_threshold or 0.0)
        P = len(pairs)
        dev = eng.device
        if P == 0:
            z = torch.zeros(1, dtype=torch.int64, device=dev)
            e = torch.empty(0, dtype=torch.int32, device=dev)
            return PairCSR(pairs, z, e, e.clone(), torch.empty(0, dtype=torch.float32, device=dev))
        if P <= self.chunk_pairs:              # single chunk on the caller's stream
            C, optr, oq, ot, od = self._enqueue_chunk(store, pairs, mode, ratio, 0, eng._stream())
            total = int(optr[C].item())
            return PairCSR(pairs, optr[:C + 1].clone(), oq[:total].clone(), ot[:total].clone(), od[:total].clone())
        cur = torch.cuda.current_stream(dev)
        if not self.pipeline:
            streams = [cur, cur]
        else:
            if self._streams is None:
                self._streams = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]
            streams = self._streams
            for st in streams:
                st.wait_stream(cur)
        outs_q, outs_t, outs_d, ptrs = [], [], [], []
        base = 0

        def finish(slot, C, optr, oq, ot, od):
            nonlocal base
            with torch.cuda.stream(streams[slot]):
                total = int(optr[C].item())    # waits for this chunk only; the next one is already enqueued
                for lst, src in ((outs_q, oq), (outs_t, ot), (outs_d, od)):
                    t = src[:total].clone()
                    t.record_stream(cur)
                    lst.append(t)
                p = optr[:C] + base
                p.record_stream(cur)
                ptrs.append(p)
            base += total

        pending = None
        for ci, c0 in enumerate(range(0, P, self.chunk_pairs)):
            slot = ci & 1
            with torch.cuda.stream(streams[slot]):
                job = self._enqueue_chunk(store, pairs[c0:c0 + self.chunk_pairs], mode, ratio, slot,
                                          streams[slot].cuda_stream)
            if pending is not None:
                finish(*pending)
            pending = (slot, *job)
        finish(*pending)
        if self.pipeline:
            for st in streams:
                cur.wait_stream(st)
        ptrs.append(torch.tensor([base], dtype=torch.int64, device=dev))
        return PairCSR(pairs, torch.cat(ptrs), torch.cat(outs_q), torch.cat(outs_t), torch.cat(outs_d))


class StreamingAllPairs:
    """All pairs of ONE host-resident set of float32 descriptors with the upload hidden behind the matching.
    Kept as the round-1 name of the end-to-end path: it now delegates to distributed.DistributedAllPairs (any
    descriptor kind, 1..N GPUs, packed match records); see that module for the data path.

    The result's `pairs` are in PROCESSING order (by group of j) -- the same set as itertools.combinations, permuted
    (look a pair up through `pairs`).  Its arrays are views of a pinned arena owned by this object: valid until the
    next match_host() call."""

    def __init__(self, engine: Engine, chunk_pairs: int = 2048, group_images: int = 25, **kw):
        from .distributed import DistributedAllPairs
        self.engine = engine
        self.impl = DistributedAllPairs(engine, chunk_pairs=chunk_pairs, group_images=group_images, **kw)

    def match_host(self, host: Optional[torch.Tensor], counts: Sequence[int], ratio_threshold: Optional[float] = None,
                   cross_check: bool = False, D: Optional[int] = None, norm_type: int = NORM_L2, **kw):
        dtype = torch.uint8 if norm_type == NORM_HAMMING else torch.float32
        return self.impl.match_host(host, counts, norm_type, cross_check, ratio_threshold, D=D, dtype=dtype, **kw)


# ---------------------------------------------------------------------------------- multi-GPU
def broadcast_store(store: Optional[Store], norm_type: int, device: torch.device, src: int = 0) -> Store:
    """Broadcast the packed store (descriptors, norms, offsets) from `src` to every rank."""
    import torch.distributed as dist
    rank = dist.get_rank()
    meta = [None]
    if rank == src:
        dd = store.dd
        meta = [(dd.kind, tuple(dd.desc.shape), str(dd.desc.dtype), None if dd.norm is None else (str(dd.norm.dtype), int(dd.norm.numel())),
                 dd.rows, dd.width, store.offsets.tolist())]
    dist.broadcast_object_list(meta, src=src)
    kind, shape, dt, ndt, rows, width, offsets = meta[0]
    tdt = getattr(torch, dt.split(".")[-1])
    desc = store.dd.desc if rank == src else torch.empty(shape, dtype=tdt, device=device)
    dist.broadcast(desc, src=src)
    norm = None
    if kind in (L.KIND_HAMMING_TC, L.KIND_HAMMING_FP4):
        # the +-1 expansion is 8x / 4x the packed bits: ship the bits only, every rank expands its own copy
        if rank == src:
            norm = store.dd.norm
        else:
            lib = L.load()
            fp4 = kind == L.KIND_HAMMING_FP4
            wpad = ctypes.c_int(0)
            L.check((lib.b200m_hamming_fp4_width if fp4 else lib.b200m_hamming_tc_width)(width, ctypes.byref(wpad)), "b200m_hamming_*_width")
            norm = torch.empty((rows, wpad.value), dtype=torch.uint8 if fp4 else torch.int8, device=device)
            with torch.cuda.device(device):
                L.check((lib.b200m_prep_hamming_fp4 if fp4 else lib.b200m_prep_hamming_tc)(
                    desc.data_ptr(), rows, width, norm.data_ptr(), wpad.value, torch.cuda.current_stream(device).cuda_stream),
                    "b200m_prep_hamming_*")
    elif ndt is not None:
        norm = store.dd.norm if rank == src else torch.empty((ndt[1],), dtype=getattr(torch, ndt[0].split(".")[-1]), device=device)
        dist.broadcast(norm, src=src)
    return Store(DeviceDescriptors(kind, desc, norm, rows, width), np.asarray(offsets, np.int64), norm_type)


def gather_csr(local: PairCSR, device: torch.device, dst: int = 0) -> Optional[PairCSR]:
    """Variable-length gather of every rank's CSR to `dst` (all_gather of sizes, then one grouped send/recv exchange)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    P = len(local.pairs)
    sizes = torch.tensor([P, int(local.query_idx.numel())], dtype=torch.int64, device=device)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    all_sizes = [tuple(int(v) for v in s.tolist()) for s in all_sizes]
    pairs_dev = torch.from_numpy(np.ascontiguousarray(local.pairs)).to(device)
    # one grouped exchange (ncclGroup): posted one by one, the 5 x (world - 1) transfers ran serialised behind each other
    if rank != dst:
        ops = [dist.P2POp(dist.isend, t.contiguous(), dst)
               for t in (pairs_dev, local.ptr, local.query_idx, local.train_idx, local.distance) if t.numel()]
        for req in (dist.batch_isend_irecv(ops) if ops else []):
            req.wait()
        return None
    bufs, ops = {}, []
    for r in range(world):
        Pr, Kr = all_sizes[r]
        if r == dst:
            continue
        bufs[r] = (torch.empty((Pr, 2), dtype=torch.int32, device=device), torch.empty(Pr + 1, dtype=torch.int64, device=device),
                   torch.empty(Kr, dtype=torch.int32, device=device), torch.empty(Kr, dtype=torch.int32, device=device),
                   torch.empty(Kr, dtype=torch.float32, device=device))
        ops += [dist.P2POp(dist.irecv, buf, r) for buf in bufs[r] if buf.numel()]
    for req in (dist.batch_isend_irecv(ops) if ops else []):
        req.wait()
    parts_pairs, parts_ptr, parts_q, parts_t, parts_d = [], [], [], [], []
    base = 0
    for r in range(world):
        Pr, Kr = all_sizes[r]
        pr, ptr, q, t, d = (pairs_dev, local.ptr, local.query_idx, local.train_idx, local.distance) if r == dst else bufs[r]
        parts_pairs.append(pr); parts_ptr.append(ptr[:Pr] + base); parts_q.append(q); parts_t.append(t); parts_d.append(d)
        base += Kr
    parts_ptr.append(torch.tensor([base], dtype=torch.int64, device=device))
    return PairCSR(torch.cat(parts_pairs).cpu().numpy(), torch.cat(parts_ptr), torch.cat(parts_q), torch.cat(parts_t),
                   torch.cat(parts_d))
