"""All pairs (i<j) of ONE host-resident image set on 1..N GPUs of one box, end to end: host descriptors in,
host match records out — the batched, sharded form of the reference's per-pair loop
(/root/reference/FeatureMatchingExtraction/pipeline.py:538-606; pair list = itertools.combinations,
image_manager.py:491-494).  One process per GPU (`torch.distributed`, NCCL over NVLink).

Data path (every step is a stream operation, nothing waits for the whole store):

  ingest    images go up in GROUPS.  Inside a group every rank uploads its 1/world block of the images over its own PCIe
            link (pinned host -> HBM), re-encodes it (int8 for integer-valued float32, packed bits for uint8) and ONE
            NCCL all-gather per group puts every block into every rank's store, in place.  What each kernel needs
            besides the rows (norms, fold digits, the +-1 expansion of binary rows) is derived locally from the
            gathered rows, so the rows are the only thing that crosses NVLink.  (`upload="rank0"`: rank 0 uploads
            whole groups and broadcasts them — for callers whose descriptors live in one process.)
  match     as soon as group g is resident, the pairs (i<j, j in g) are cut into `world` balanced slices; every rank
            matches its slice chunk by chunk on two compute streams (b200m_match_pairs_rec: one packed record per
            match).  No collective inside.
  return    per group, lagging one group behind the matching: ranks send their records to rank 0 (one NCCL message
            per rank and group) and rank 0 copies the group to pinned host memory, or hands it to a consumer, while
            the next group is being matched.  `gather="local"` keeps every rank's records in its own host memory.

Non-integer float descriptors cannot be encoded before the whole store is known (the fp16 scale is a property of the
store): they take the same path with a single group — upload, all-gather of the raw rows, local re-encoding."""
from __future__ import annotations

import ctypes
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from .allpairs import _PAIR_DTYPE, Store, partition_pairs
from .engine import DeviceDescriptors, Engine, NORM_HAMMING, NORM_L2, mode_of

REC_DTYPES = {
    L.REC_Q32T32D32: np.dtype([("q", "<i4"), ("t", "<i4"), ("d", "<f4")]),
    L.REC_Q16T16D32: np.dtype([("q", "<u2"), ("t", "<u2"), ("d", "<f4")]),
    L.REC_Q16T16D16: np.dtype([("q", "<u2"), ("t", "<u2"), ("d", "<u2")]),
}
assert [REC_DTYPES[f].itemsize for f in (1, 2, 3)] == [12, 8, 6]


class RecCSR:
    """Host result: `pairs` [P, 2] image indices in PROCESSING order (by group of j, then rank slice — the same set as
    itertools.combinations, permuted), `ptr` int64 [P+1], `rec` one packed record per match (REC_DTYPES)."""

    def __init__(self, pairs: np.ndarray, ptr: np.ndarray, rec: np.ndarray, rec_format: int):
        self.pairs, self.ptr, self.rec, self.rec_format = pairs, ptr, rec, rec_format

    query_idx = property(lambda self: self.rec["q"])
    train_idx = property(lambda self: self.rec["t"])
    distance = property(lambda self: self.rec["d"])

    def __len__(self) -> int:
        return len(self.pairs)

    @property
    def nbytes(self) -> int:
        return int(self.ptr.nbytes + self.rec.nbytes)

    def pair(self, p: int):
        """(queryIdx int32, trainIdx int32, distance float32) of pair p, contiguous copies."""
        r = self.rec[int(self.ptr[p]):int(self.ptr[p + 1])]
        return r["q"].astype(np.int32), r["t"].astype(np.int32), r["d"].astype(np.float32)

    def sorted_by_pair(self) -> "RecCSR":
        """The same result with `pairs` in itertools.combinations order (what the reference's loop produces)."""
        order = np.lexsort((self.pairs[:, 1], self.pairs[:, 0]))
        cnt = np.diff(self.ptr)[order]
        ptr = np.zeros(len(order) + 1, np.int64)
        np.cumsum(cnt, out=ptr[1:])
        idx = np.concatenate([np.arange(self.ptr[p], self.ptr[p + 1]) for p in order]) if len(order) else np.zeros(0, np.int64)
        return RecCSR(self.pairs[order], ptr, self.rec[idx], self.rec_format)


def group_sizes(num_images: int, group_images: int, first: Optional[int] = None) -> List[int]:
    """Images per upload group.  With `first` the groups ramp up from `first` (first, 2 first, 4 first, ... until
    group_images) AND taper down to `first` again at the end:
      * front: matching starts after a SMALL first upload (the step's exposed front is the first group's H2D + re-encoding
        + all-gather);
      * back: the records of the LAST group come home after the last kernel (its pairs are (i, j) with j in the group and
        every i below: a 48-image last group of a 200-image set holds 42 % of all pairs, 25 MB of records) -- a small last
        group keeps that exposed tail small;
      * the groups in between are large, so launches, collectives and returns stay few."""
    if not first or first >= group_images:
        sizes, left = [], num_images
        while left > 0:
            sizes.append(min(group_images, left))
            left -= sizes[-1]
        return sizes
    front, back, left, g = [], [], num_images, max(1, first)
    while left > 0 and g < group_images:           # ramp up from the front and down towards the back, symmetrically
        front.append(min(g, left))
        left -= front[-1]
        if left > 0:
            back.append(min(g, left))
            left -= back[-1]
        g *= 2
    mid = []
    while left > 0:
        mid.append(min(group_images, left))
        left -= mid[-1]
    return front + mid + back[::-1]


def group_plan(num_images: int, group_images: int, first: Optional[int] = None):
    """[(j0, j1, pairs[K, 2])]: for every group of images [j0, j1), the pairs (i, j), i < j, j in the group."""
    out, j0 = [], 0
    for size in group_sizes(num_images, group_images, first):
        j1 = j0 + size
        js = np.arange(max(j0, 1), j1)
        if len(js):
            i = np.concatenate([np.arange(0, j) for j in js])
            gp = np.stack([i, np.repeat(js, js)], axis=1).astype(np.int32)
        else:
            gp = np.zeros((0, 2), np.int32)
        out.append((j0, j1, gp))
        j0 = j1
    return out


def block_cuts(j0: int, j1: int, world: int) -> List[int]:
    """Image indices cutting group [j0, j1) into `world` contiguous upload blocks."""
    return [j0 + (j1 - j0) * k // world for k in range(world + 1)]


class _Phases:
    """Per-phase CUDA-event timing of one match_host call (bench.py's `e2e.phases`): every phase is a list of
    (start, end) event pairs recorded on the stream that runs it."""

    def __init__(self, on: bool):
        self.on = on
        self.spans = {}
        self.t0 = None

    def mark(self, stream) -> Optional[torch.cuda.Event]:
        if not self.on:
            return None
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(stream)
        return ev

    def add(self, name: str, e0, e1) -> None:
        if self.on:
            self.spans.setdefault(name, []).append((e0, e1))

    def summary(self, t_end) -> dict:
        """ms per phase (sum of its spans) and where the step's ends are exposed: `front_ms` = start of the call to the
        first matching kernel, `tail_ms` = end of the last matching kernel to the last byte on the host."""
        if not self.on:
            return {}
        out = {k + "_ms": float(sum(a.elapsed_time(b) for a, b in v)) for k, v in self.spans.items()}
        m = self.spans.get("match", [])
        if m:
            first = min(self.t0.elapsed_time(a) for a, _ in m)
            last = max(self.t0.elapsed_time(b) for _, b in m)
            out["front_ms"] = float(first)
            out["match_span_ms"] = float(last - first)
            out["tail_ms"] = float(self.t0.elapsed_time(t_end) - last)
        out["total_ms"] = float(self.t0.elapsed_time(t_end))
        return out


class _HostBoard:
    """Host-side rendezvous of the ranks of ONE box through a shared-memory page (/dev/shm): every rank posts a number
    under a sequence tag, rank 0 collects them and posts "go".  Microseconds (measured 19 us for 4 ranks) instead of the
    ~0.6 ms of a gloo collective over loopback TCP, and no kernel on any GPU -- the persistent matching kernels own every
    SM, so a device-side size exchange would wait behind the running chunk and stall the thread that launches the next one."""
    SLOTS = 64

    def __init__(self, world: int, rank: int, group):
        import os
        import tempfile
        import torch.distributed as dist
        self.world, self.rank = world, rank
        path = [None]
        if rank == 0:
            d = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
            fd, name = tempfile.mkstemp(prefix="b200match_board_", dir=d)
            os.write(fd, bytes(8 * 2 * self.SLOTS * (world + 1)))
            os.close(fd)
            path[0] = name
        dist.broadcast_object_list(path, src=0, group=group)
        self.path = path[0]
        self.a = np.memmap(self.path, dtype=np.int64, mode="r+", shape=(self.SLOTS, world + 1, 2))
        dist.barrier(group=group)                                # everyone has mapped the file ...
        if rank == 0:
            os.unlink(self.path)                                 # ... so the name can go; the pages live as long as the maps
        self.seq = 0

    def exchange(self, value: int) -> Optional[List[int]]:
        """Every rank posts `value`; rank 0 gets the list of all values (the others None) once ALL ranks have posted and
        rank 0 has released them."""
        self.seq += 1
        seq, slot = self.seq, self.seq % self.SLOTS
        a = self.a
        a[slot, self.rank, 0] = value
        a[slot, self.rank, 1] = seq                              # tag last (x86 keeps the store order)
        if self.rank == 0:
            while not bool((a[slot, :self.world, 1] == seq).all()):
                pass
            out = [int(v) for v in a[slot, :self.world, 0]]
            a[slot, self.world, 1] = seq                         # go
            return out
        while int(a[slot, self.world, 1]) != seq:
            pass
        return None


class DistributedAllPairs:
    """See the module docstring.  One instance per process; buffers (device stores, pinned arenas) are kept between calls."""

    def __init__(self, engine: Engine, chunk_pairs: int = 2048, group_images: int = 24, rec_format: Optional[int] = None,
                 gather: str = "rank0", upload: str = "auto", ramp: bool = True):
        if gather not in ("rank0", "local"):
            raise ValueError("gather must be rank0 or local")
        if upload not in ("auto", "sharded", "rank0"):
            raise ValueError("upload must be auto, sharded or rank0")
        self.engine = engine
        self.chunk_pairs = int(chunk_pairs)
        self.group_images = int(group_images)
        self.rec_format = rec_format
        self.gather, self.upload, self.ramp = gather, upload, bool(ramp)
        self._dev = {}
        self._pinned = {}
        self._streams = None
        self.last_phases: dict = {}

    # ------------------------------------------------------------------ buffers
    def _dbuf(self, name: str, numel: int, dtype) -> torch.Tensor:
        t = self._dev.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype:
            self._dev[name] = t = torch.empty(max(int(numel), 1), dtype=dtype, device=self.engine.device)
        return t

    def _hbuf(self, name: str, nbytes: int) -> torch.Tensor:
        """Pinned byte arena that keeps its content when it grows."""
        t = self._pinned.get(name)
        if t is None or t.numel() < nbytes:
            new = torch.empty(max(int(nbytes), 1 << 22, 0 if t is None else 2 * t.numel()), dtype=torch.uint8, pin_memory=True)
            if t is not None:
                torch.cuda.synchronize(self.engine.device)     # copies into the old arena must have landed
                new[:t.numel()].copy_(t)
            self._pinned[name] = t = new
        return t

    def _get_streams(self):
        if self._streams is None:
            dev = self.engine.device
            self._streams = {"copy": torch.cuda.Stream(device=dev), "gather": torch.cuda.Stream(device=dev),
                             "compute": [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]}
        return self._streams

    # ------------------------------------------------------------------ ingest
    def _ingest(self, host, offsets, D, dtype, norm_type, groups, world, rank, sharded, float_path, ph):
        """Set up the store and return (Store, ingest_group, events, not-integer flag tensor or None): ingest_group(gi)
        enqueues, on the copy stream, upload + re-encoding + all-gather (+ local derivations) of group gi and returns its
        event.  Float descriptors are ingested here at once (one group + the store-wide re-encoding): then ingest_group is
        None and `events` holds the event of every group."""
        import torch.distributed as dist
        eng, dev = self.engine, self.engine.device
        lib = eng.lib
        cs = self._get_streams()["copy"]
        rows = int(offsets[-1])
        hamming = norm_type == NORM_HAMMING
        events = []
        flag = None
        with torch.cuda.stream(cs):
            if hamming:
                Dp = (D + 15) // 16 * 16                        # AKAZE's 61-byte rows: zero-padded (adds 0 to every distance)
                bits = self._dbuf("bits", rows * Dp, torch.uint8)[:rows * Dp].view(rows, Dp)
                if Dp != D:
                    bits.zero_()
                want = {"popc": L.KIND_HAMMING, "int8": L.KIND_HAMMING_TC}.get(eng.hamming_path, L.KIND_HAMMING_FP4)
                wpad = ctypes.c_int(0)
                if want != L.KIND_HAMMING:
                    fp4 = want == L.KIND_HAMMING_FP4
                    rc = (lib.b200m_hamming_fp4_width if fp4 else lib.b200m_hamming_tc_width)(Dp, ctypes.byref(wpad))
                    if rc == L.EUNSUPPORTED and eng.hamming_path == "auto":
                        want = L.KIND_HAMMING                    # wider than 512 bits: INT-pipe kernel
                    else:
                        L.check(rc, "b200m_hamming_*_width")
                ex = None
                if want != L.KIND_HAMMING:
                    ex = self._dbuf("expand", rows * wpad.value, torch.uint8)[:rows * wpad.value].view(rows, wpad.value)
                    if want == L.KIND_HAMMING_TC:
                        ex = ex.view(torch.int8)
                dd = DeviceDescriptors(want, bits, ex, rows, Dp)
                wire, stage_dt = bits, torch.uint8
            elif float_path:
                Dp = (D + 31) // 32 * 32
                f32 = self._dbuf("f32rows", rows * Dp, torch.float32)[:rows * Dp].view(rows, Dp)
                if Dp != D:
                    f32.zero_()
                dd = None                                        # built after the whole store is resident
                wire, stage_dt = f32, torch.float32
            else:
                Dp = (D + 127) // 128 * 128
                d8 = self._dbuf("i8rows", rows * Dp, torch.int8)[:rows * Dp].view(rows, Dp)
                nb = ctypes.c_size_t(0)
                L.check(lib.b200m_l2_i8_aux_bytes(rows, ctypes.byref(nb)), "b200m_l2_i8_aux_bytes")
                aux = self._dbuf("i8aux", nb.value // 4, torch.int32)[:nb.value // 4]
                flag = self._dbuf("flag", 1, torch.int32)
                flag.zero_()
                self._flag_pin = self._hbuf("flagpin", 4)[:4].view(torch.int32)
                self._flag_pin.zero_()
                with torch.cuda.device(dev):
                    L.check(lib.b200m_l2_i8_aux_init(aux.data_ptr(), rows, D, cs.cuda_stream), "b200m_l2_i8_aux_init")
                dd = DeviceDescriptors(L.KIND_L2_I8, d8, aux, rows, Dp)
                wire, stage_dt = d8, torch.float32
            stage = None
            if host is not None and not (hamming and Dp == D) and not (float_path and Dp == D):
                stage = self._dbuf("stage", rows * D, stage_dt)[:rows * D].view(rows, D)   # raw rows of the blocks this rank uploads

            def upload_rows(a0: int, a1: int) -> None:
                """host rows [a0, a1) -> wire rows [a0, a1) (H2D + re-encoding) on the copy stream."""
                if a1 <= a0:
                    return
                e0 = ph.mark(cs)
                if stage is None:
                    wire[a0:a1].copy_(host[a0:a1], non_blocking=True)
                    ph.add("h2d", e0, ph.mark(cs))
                    return
                stage[a0:a1].copy_(host[a0:a1], non_blocking=True)
                e1 = ph.mark(cs)
                ph.add("h2d", e0, e1)
                if hamming or float_path:
                    wire[a0:a1, :D].copy_(stage[a0:a1])
                else:
                    eng.prepare_l2_i8_rows(dd, flag, stage[a0:a1], a0)
                ph.add("prep", e1, ph.mark(cs))

        def ingest_group(gi: int):
            """Enqueue group gi on the copy stream; returns the event that fires when its rows and aux data are resident."""
            j0, j1, _ = groups[gi]
            with torch.cuda.stream(cs):
                r0, r1 = int(offsets[j0]), int(offsets[j1])
                if world == 1:
                    upload_rows(r0, r1)
                elif not sharded:
                    if rank == 0:
                        upload_rows(r0, r1)
                    e0 = ph.mark(cs)
                    if r1 > r0:
                        dist.broadcast(wire[r0:r1], src=0)
                    ph.add("bcast", e0, ph.mark(cs))
                else:
                    cuts = [int(offsets[c]) for c in block_cuts(j0, j1, world)]
                    upload_rows(cuts[rank], cuts[rank + 1])
                    e0 = ph.mark(cs)
                    sizes = [cuts[k + 1] - cuts[k] for k in range(world)]
                    if r1 > r0:
                        if len(set(sizes)) == 1:               # equal blocks: one in-place NCCL all-gather
                            dist.all_gather_into_tensor(wire[r0:r1], wire[cuts[rank]:cuts[rank + 1]])
                        else:                                   # ragged blocks: one coalesced group of broadcasts
                            dist.all_gather([wire[cuts[k]:cuts[k + 1]] for k in range(world)], wire[cuts[rank]:cuts[rank + 1]])
                    ph.add("bcast", e0, ph.mark(cs))
                # local derivations from the resident rows of this group
                e0 = ph.mark(cs)
                if r1 > r0:
                    with torch.cuda.device(dev):
                        if hamming and dd.kind != L.KIND_HAMMING:
                            fn = lib.b200m_prep_hamming_fp4 if dd.kind == L.KIND_HAMMING_FP4 else lib.b200m_prep_hamming_tc
                            L.check(fn(bits[r0:r1].data_ptr(), r1 - r0, Dp, dd.norm[r0:r1].data_ptr(), dd.norm.shape[1],
                                       cs.cuda_stream), "b200m_prep_hamming_*")
                        elif not hamming and not float_path and world > 1:
                            L.check(lib.b200m_l2_i8_aux_rows(d8.data_ptr(), r0, r1 - r0, rows, Dp, aux.data_ptr(),
                                                             cs.cuda_stream), "b200m_l2_i8_aux_rows")
                ph.add("derive", e0, ph.mark(cs))
                if flag is not None:                             # host mirror of "a non-integer element was seen" (read without a sync)
                    self._flag_pin.copy_(flag, non_blocking=True)
                if flag is not None and world > 1 and gi == 0:
                    dist.all_reduce(flag, op=dist.ReduceOp.MAX)   # every rank must take the same branch on "integer-valued?"
                ev = torch.cuda.Event()
                ev.record(cs)
            return ev
        if float_path:                                   # one group: ingest it now, then the store-wide re-encoding
            ev0 = ingest_group(0)
            with torch.cuda.stream(cs):
                e0 = ph.mark(cs)
                dd = eng.prepare(f32, NORM_L2, force_kind=L.KIND_L2_F16C if eng.float_path == "f16" else L.KIND_L2_F32)
                ph.add("prep", e0, ph.mark(cs))
                ev = torch.cuda.Event()
                ev.record(cs)
            return Store(dd, offsets, norm_type), None, [ev] * len(groups), flag
        return Store(dd, offsets, norm_type), ingest_group, None, flag

    # ------------------------------------------------------------------ one chunk of pairs -> packed records (device)
    def _enqueue_chunk(self, store: Store, chunk: np.ndarray, mode: int, ratio: float, fmt: int, slot: int, stream):
        """Plan + enqueue b200m_match_pairs_rec for `chunk` on `stream` into the buffers of `slot`; returns (C, ptr, rec)."""
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
        rb = REC_DTYPES[fmt].itemsize
        ws = self._dbuf(f"ws{slot}", max(nbytes.value, 1 << 20), torch.uint8)
        out = self._dbuf(f"rec{slot}", (tn * rb + 7) // 8, torch.int64)          # 8-byte aligned record buffer
        oc = self._dbuf(f"cnt{slot}", C, torch.int32)
        optr = self._dbuf(f"ptr{slot}", C + 1, torch.int64)
        norm = 0 if dd.norm is None else dd.norm.data_ptr()
        with torch.cuda.device(eng.device):
            rc = lib.b200m_match_pairs_rec(dd.kind, dd.desc.data_ptr(), norm, dd.rows, dd.desc.data_ptr(), norm, dd.rows,
                                           dd.width, arr, C, mode, ratio, fmt, out.data_ptr(), oc.data_ptr(), optr.data_ptr(),
                                           ws.data_ptr(), ws.numel(), stream.cuda_stream)
        L.check(rc, "b200m_match_pairs_rec")
        return C, optr, out.view(torch.uint8)

    # ------------------------------------------------------------------ the call
    def match_host(self, host: Optional[torch.Tensor], counts: Sequence[int], norm_type: int = NORM_L2,
                   cross_check: bool = False, ratio_threshold: Optional[float] = None, D: Optional[int] = None,
                   dtype: Optional[torch.dtype] = None, consumer: Optional[Callable] = None, keep: bool = True,
                   timing: bool = False, assume_float: bool = False, done_mask: Optional[np.ndarray] = None) -> Optional[RecCSR]:
        """host: pinned [sum(counts), D] float32 (NORM_L2) or uint8 (NORM_HAMMING) rows of all images, image after image.
        Under torch.distributed every rank may pass them (sharded upload: rank k only reads the rows of the image
        blocks it owns) or only rank 0 (the others pass None plus D and dtype; rank 0 then uploads and broadcasts).
        consumer(pairs[K, 2], counts[K], rec) is called on the returning rank for every block of pairs as soon as its
        records are in host memory (`rec` is only valid during the call unless keep=True).
        Returns the RecCSR on rank 0 (gather="rank0"; other ranks: None) or this rank's shard (gather="local");
        with keep=False the returned `rec` is empty and only `pairs` / `ptr` are filled.
        done_mask: bool [I (I - 1) / 2] over itertools.combinations order; pairs marked True are skipped (a resumed job:
        checkpoint.AllPairsCheckpoint.done_mask()).  Every rank must pass the same mask."""
        import torch.distributed as dist
        eng = self.engine
        dev = eng.device
        world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        rank = dist.get_rank() if world > 1 else 0
        counts = [int(c) for c in counts]
        I = len(counts)
        offsets = np.zeros(I + 1, np.int64)
        np.cumsum(counts, out=offsets[1:])
        rows = int(offsets[-1])
        if host is not None:
            if host.dim() != 2 or host.shape[0] != rows:
                raise ValueError("host must be [sum(counts), D]")
            D, dtype = int(host.shape[1]), host.dtype
        if D is None or dtype is None:
            raise ValueError("ranks without host rows must pass D and dtype")
        want_dt = torch.uint8 if norm_type == NORM_HAMMING else torch.float32
        if norm_type not in (NORM_HAMMING, NORM_L2) or dtype != want_dt:
            raise ValueError("NORM_HAMMING needs uint8 rows, NORM_L2 float32 rows")
        sharded = False
        if world > 1 and self.upload != "rank0":
            have = torch.tensor([1 if host is not None else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(have, op=dist.ReduceOp.MIN)
            sharded = bool(have.item())
            if self.upload == "sharded" and not sharded:
                raise ValueError("upload='sharded' needs the host rows on every rank")
        if world > 1 and not sharded and rank == 0 and host is None:
            raise ValueError("rank 0 must hold the host rows")
        mode = mode_of(cross_check, ratio_threshold)
        ratio = float(ratio_threshold or 0.0)
        fmt = self.rec_format
        if fmt is None:
            small = max(counts, default=0) <= 65536
            fmt = (L.REC_Q16T16D16 if norm_type == NORM_HAMMING and D <= 8191 else L.REC_Q16T16D32) if small else L.REC_Q32T32D32
        # Integer-valued (OpenCV SIFT) or genuinely float?  A sample of the host rows decides which encoding to TRY without
        # waiting for the device: non-integer samples go straight to the float path; integer samples take the int8 path, whose
        # kernels still flag any non-integer element anywhere in the store (checked at the end of the call -> float path).
        # Only where every rank sees the rows (same data, same decision); with rank-0 upload the device flag is awaited instead.
        sampled_integer = False
        if norm_type == NORM_L2 and not assume_float and host is not None and rows > 0 and (world == 1 or sharded):
            smp = host[torch.linspace(0, rows - 1, min(rows, 64)).long()].numpy()
            if bool(np.all((smp == np.rint(smp)) & (smp >= 0) & (smp <= 255))):
                sampled_integer = True
            else:
                assume_float = True
        float_path = norm_type == NORM_L2 and assume_float
        G = max(1, self.group_images)
        if world > 1 and sharded:
            G = (G + world - 1) // world * world               # equal upload blocks -> in-place all-gather
        first = max(8, world)                                   # ramp: 8, 16, 32, ... images per group up to G
        if world > 1 and sharded:
            first = (first + world - 1) // world * world
        groups = group_plan(I, I, None) if float_path else group_plan(I, G, min(first, G) if self.ramp else None)
        if done_mask is not None:
            from .checkpoint import pair_index
            done_mask = np.asarray(done_mask, bool)
            if done_mask.shape != (I * (I - 1) // 2,):
                raise ValueError("done_mask must have one entry per pair of itertools.combinations(range(I), 2)")
            groups = [(j0, j1, gp[~done_mask[pair_index(gp[:, 0], gp[:, 1], I)]] if len(gp) else gp) for j0, j1, gp in groups]
        res = self._run(host, offsets, D, dtype, norm_type, groups, world, rank, sharded, float_path, mode, ratio, fmt,
                        consumer, keep, timing, sampled_integer)
        if res == "not-integer":                              # the int8 encoding is not exact for these rows: fp16/tf32 candidates
            return self.match_host(host, counts, norm_type, cross_check, ratio_threshold, D, dtype, consumer, keep, timing,
                                   assume_float=True, done_mask=done_mask)
        return res

    def _run(self, host, offsets, D, dtype, norm_type, groups, world, rank, sharded, float_path, mode, ratio, fmt,
             consumer, keep, timing, sampled_integer=False):
        import torch.distributed as dist
        eng, dev = self.engine, self.engine.device
        S = self._get_streams()
        cs, gs, comp = S["copy"], S["gather"], S["compute"]
        cur = torch.cuda.current_stream(dev)
        ph = _Phases(timing)
        ph.t0 = ph.mark(cur)
        for s in [cs, gs] + comp:
            s.wait_stream(cur)
        if world > 1 and getattr(self, "_board", None) is None:   # collective: every rank gets here on its first call
            self._cpu_group = dist.new_group(backend="gloo")
            self._board = _HostBoard(world, rank, self._cpu_group)
        slices = [partition_pairs(gp, offsets, world) if world > 1 else [gp] for _, _, gp in groups]
        store, ingest_group, events, flag = self._ingest(host, offsets, D, dtype, norm_type, groups, world, rank, sharded, float_path, ph)
        rb = REC_DTYPES[fmt].itemsize
        to_rank0 = world > 1 and self.gather == "rank0"
        owner = rank == 0 or not to_rank0                       # this rank ends up with (its share of) the records
        if events is None:
            # Group 0 now; the other groups' ingest is enqueued AFTER group 0's chunks have been launched, so the GPU is
            # already matching while the host enqueues the rest.  Every ingest collective is still issued before any
            # collective of the return path, on every rank: the order NCCL needs is the same everywhere.
            events = [ingest_group(0)] + [None] * (len(groups) - 1)
        if flag is not None and not sampled_integer:            # integer-valued?  known once the first group is encoded
            events[0].synchronize()
            if int(flag.item()) != 0:
                return "not-integer"

        blocks = []          # owner: (pairs[K, 2], counts int32[K] (host, valid after ev), rec byte offset, rec bytes, ev)
        state = {"hoff": 0, "coff": 0, "consumed": 0, "next_group": 0, "enqueued": -1, "bad": False}
        chunks_left = [(len(sl[rank]) + self.chunk_pairs - 1) // self.chunk_pairs for sl in slices]
        gK = [0] * len(groups)                                   # to_rank0: records this rank produced per group
        gP = [0] * len(groups)
        gev = [[] for _ in groups]                               # events after this rank's copies into the group's send buffers
        gdone = [None] * len(groups)                             # event after group gi's send buffers have been read (gather stream)

        def rec_arena(nbytes_more: int, blk: int):
            """(pinned tensor, byte offset) for the next block of records."""
            if keep:
                off = state["hoff"]
                state["hoff"] = off + nbytes_more
                return self._hbuf("rec", off + nbytes_more), off
            return self._hbuf(f"ring{blk & 1}", nbytes_more), 0

        def counts_of(cn) -> np.ndarray:
            if isinstance(cn, tuple):                            # (pinned arena, byte offset, pairs): valid once the block's event fired
                carena, coff, Pt = cn
                return carena[coff:coff + 4 * Pt].numpy().view(np.int32)
            return cn

        def local_bad() -> bool:
            return flag is not None and int(self._flag_pin[0]) != 0

        def consume_upto(n_blocks: int):
            if consumer is None:
                return
            while state["consumed"] < n_blocks:
                if state["bad"] or local_bad():                  # a non-integer element turned up: this int8 run will be redone on the
                    return                                       # float path (match_host); nothing more of it reaches the consumer
                pr, cn, arena, off, nb, ev = blocks[state["consumed"]]
                ev.synchronize()
                consumer(pr, counts_of(cn), arena[off:off + nb].numpy().view(REC_DTYPES[fmt]))
                state["consumed"] += 1

        def grow_dev(name: str, nbytes: int, used: int) -> torch.Tensor:
            """Device byte buffer that keeps its first `used` bytes when it grows (rare: sized on the first call)."""
            t = self._dev.get(name)
            if t is None or t.numel() < nbytes:
                new = torch.empty(max(int(nbytes), 1 << 20, 0 if t is None else 2 * t.numel()), dtype=torch.uint8, device=dev)
                if t is not None:
                    torch.cuda.synchronize(dev)
                    new[:used].copy_(t[:used])
                    torch.cuda.synchronize(dev)
                self._dev[name] = t = new
            return t

        def finish(gi, slot, chunk, C, optr, rec):
            st = comp[slot]
            with torch.cuda.stream(st):
                ptr_h = optr[:C + 1].cpu().numpy()              # waits for this chunk only; the next one is already enqueued
                K = int(ptr_h[C])
                cnt = np.diff(ptr_h).astype(np.int32)
                if not to_rank0:                                 # straight to this rank's pinned memory
                    consume_upto(len(blocks) - 1)                # the ring half about to be reused has been handed over
                    arena, off = rec_arena(K * rb, len(blocks))
                    e0 = ph.mark(st)
                    if K:
                        arena[off:off + K * rb].copy_(rec[:K * rb], non_blocking=True)
                    ev = torch.cuda.Event(enable_timing=timing)
                    ev.record(st)
                    ph.add("d2h", e0, ev if timing else None)
                    blocks.append((chunk, cnt, arena, off, K * rb, ev))
                else:                                            # into this group's send buffers (device)
                    if gi >= 2 and gdone[gi - 2] is not None:    # same buffers as group gi - 2: its messages must have left
                        st.wait_event(gdone[gi - 2])
                    hdr = grow_dev(f"shdr{gi & 1}", 4 * (gP[gi] + C), 4 * gP[gi])
                    snd = grow_dev(f"srec{gi & 1}", rb * (gK[gi] + K), rb * gK[gi])
                    hdr[4 * gP[gi]:4 * (gP[gi] + C)].copy_(self._dev[f"cnt{slot}"][:C].view(torch.uint8), non_blocking=True)
                    if K:
                        snd[rb * gK[gi]:rb * (gK[gi] + K)].copy_(rec[:K * rb], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(st)
                    gev[gi].append(ev)
                    gK[gi] += K
                    gP[gi] += C
            chunks_left[gi] -= 1
            advance()

        def return_group(gi):
            """to_rank0: this rank's records of group gi -> rank 0 -> pinned host memory (all on the gather stream)."""
            P_all = [len(slices[gi][r]) for r in range(world)]
            if sum(P_all) == 0:
                return
            for ev in gev[gi]:
                gs.wait_event(ev)
            # record counts travel on the HOST (_HostBoard); it doubles as the rendezvous: the NCCL messages below are
            # posted by all ranks at about the same time
            K_all = self._board.exchange(gK[gi] | (int(local_bad()) << 56))      # bit 56: this rank saw a non-integer element
            if K_all is not None:
                state["bad"] |= any(v >> 56 for v in K_all)
                K_all = [v & ((1 << 56) - 1) for v in K_all]
            with torch.cuda.stream(gs):
                e0 = ph.mark(gs)
                hdr, snd = self._dev.get(f"shdr{gi & 1}"), self._dev.get(f"srec{gi & 1}")
                if rank != 0:
                    ops = []
                    if gP[gi]:
                        ops.append(dist.P2POp(dist.isend, hdr[:4 * gP[gi]], 0))
                    if gK[gi]:
                        ops.append(dist.P2POp(dist.isend, snd[:rb * gK[gi]], 0))
                    for req in (dist.batch_isend_irecv(ops) if ops else []):
                        req.wait()
                    ph.add("gather", e0, ph.mark(gs))
                    gdone[gi] = torch.cuda.Event()
                    gdone[gi].record(gs)
                    return
                Pt, Kt = sum(P_all), sum(K_all)
                if self._dev.get("gcnt") is None or self._dev["gcnt"].numel() < 4 * Pt or \
                        self._dev.get("grec") is None or self._dev["grec"].numel() < rb * Kt:
                    gs.synchronize()                              # a buffer about to be replaced may still be read by the last D2H
                gcnt = self._dbuf("gcnt", 4 * Pt, torch.uint8)
                grec = self._dbuf("grec", max(rb * Kt, 1), torch.uint8)
                ops, po, ko = [], 0, 0
                for r in range(world):
                    pb, kb = 4 * P_all[r], rb * K_all[r]
                    if r == 0:
                        if pb:
                            gcnt[po:po + pb].copy_(hdr[:pb], non_blocking=True)
                        if kb:
                            grec[ko:ko + kb].copy_(snd[:kb], non_blocking=True)
                    else:
                        if pb:
                            ops.append(dist.P2POp(dist.irecv, gcnt[po:po + pb], r))
                        if kb:
                            ops.append(dist.P2POp(dist.irecv, grec[ko:ko + kb], r))
                    po += pb
                    ko += kb
                for req in (dist.batch_isend_irecv(ops) if ops else []):
                    req.wait()
                e1 = ph.mark(gs)
                ph.add("gather", e0, e1)
                gdone[gi] = torch.cuda.Event()
                gdone[gi].record(gs)
                consume_upto(len(blocks) - 1)
                arena, off = rec_arena(rb * Kt, len(blocks))
                carena = self._hbuf("cnt", state["coff"] + 4 * Pt)
                coff = state["coff"]
                state["coff"] = coff + 4 * Pt
                carena[coff:coff + 4 * Pt].copy_(gcnt[:4 * Pt], non_blocking=True)
                if Kt:
                    arena[off:off + rb * Kt].copy_(grec[:rb * Kt], non_blocking=True)
                ev = torch.cuda.Event(enable_timing=timing)
                ev.record(gs)
                ph.add("d2h", e1, ev if timing else None)
                pr = np.concatenate([slices[gi][r] for r in range(world)])
                blocks.append((pr, (carena, coff, Pt), arena, off, rb * Kt, ev))

        def advance():
            while state["next_group"] <= state["enqueued"] and chunks_left[state["next_group"]] == 0:
                if to_rank0:
                    return_group(state["next_group"])
                state["next_group"] += 1

        pending = None
        ci = 0
        for gi in range(len(groups)):
            mine = slices[gi][rank]
            if len(mine):
                for st in comp:
                    st.wait_event(events[gi])
            for c0 in range(0, len(mine), self.chunk_pairs):
                slot = ci & 1
                ci += 1
                chunk = mine[c0:c0 + self.chunk_pairs]
                e0 = ph.mark(comp[slot])
                job = self._enqueue_chunk(store, chunk, mode, ratio, fmt, slot, comp[slot])
                ph.add("match", e0, ph.mark(comp[slot]))
                if pending is not None:
                    finish(*pending)
                pending = (gi, slot, chunk, *job)
            state["enqueued"] = gi
            if gi == 0 and ingest_group is not None:
                for g in range(1, len(groups)):
                    events[g] = ingest_group(g)
            advance()
        if pending is not None:
            finish(*pending)
        advance()
        for s in [cs, gs] + comp:
            cur.wait_stream(s)
        bad = None
        if flag is not None:
            bad = flag.clone()
            if world > 1:
                dist.all_reduce(bad, op=dist.ReduceOp.MAX)
        t_end = ph.mark(cur)
        cur.synchronize()                                        # results are host arrays: every copy must have landed
        self.last_phases = ph.summary(t_end)
        if bad is not None and int(bad.item()) != 0:
            return "not-integer"
        if not owner:
            return None
        if state["bad"]:
            return "not-integer"
        consume_upto(len(blocks))
        pairs_out, cnts = [], []
        for pr, cn, *_ in blocks:
            pairs_out.append(pr)
            cnts.append(counts_of(cn))
        pairs_out = np.concatenate(pairs_out) if pairs_out else np.zeros((0, 2), np.int32)
        cnts = np.concatenate(cnts) if cnts else np.zeros(0, np.int32)
        ptr = np.zeros(len(cnts) + 1, np.int64)
        np.cumsum(cnts, out=ptr[1:])
        if keep:                                                 # blocks lie back to back in the arena, in block order
            total = int(ptr[-1])
            assert total * rb == state["hoff"]
            rec = self._hbuf("rec", max(state["hoff"], 1))[:total * rb].numpy().view(REC_DTYPES[fmt])
        else:
            rec = np.zeros(0, REC_DTYPES[fmt])
        return RecCSR(pairs_out, ptr, rec, fmt)
