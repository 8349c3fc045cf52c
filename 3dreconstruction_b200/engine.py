"""Array-first host side of the matcher: owns device buffers (torch = plumbing only) and calls the
C-ABI in include/b200match.h.  Everything that computes runs in libb200match.so on the GPU; a
missing library or device raises — there is no CPU path.
"""
from __future__ import annotations

import ctypes
import os
from typing import NamedTuple, Optional, Sequence

import numpy as np
import torch

from . import _lib as L

NORM_L2, NORM_HAMMING = 4, 6      # cv2.NORM_L2, cv2.NORM_HAMMING
_REC12 = np.dtype([("q", "<i4"), ("t", "<i4"), ("d", "<f4")])     # B200M_REC_Q32T32D32


class UnsupportedNormError(NotImplementedError):
    """A cv2 norm type the reference's matcher accepts (NORM_L1, NORM_L2SQR, NORM_HAMMING2, ...) and this library does not
    implement: raised, never turned into an empty result."""


class MatchArrays(NamedTuple):
    query_idx: np.ndarray    # int32 [K]
    train_idx: np.ndarray    # int32 [K]
    distance: np.ndarray     # float32 [K]   sorted by (distance, query_idx)


class Knn2Arrays(NamedTuple):
    idx1: np.ndarray
    idx2: np.ndarray
    d1: np.ndarray
    d2: np.ndarray


class DeviceDescriptors(NamedTuple):
    """One descriptor store resident in HBM, already in the layout its kernel wants."""
    kind: int
    desc: torch.Tensor                 # uint8 [rows, D] | int8 [rows, Dpad] | float32 [rows, Dpad]
    norm: Optional[torch.Tensor]       # L2_I8: int32 aux block (norms[rows], fold digits, info) | float32 [rows] | None
                                       # HAMMING_TC / HAMMING_FP4: int8 [rows, Dpad] / e2m1 uint8 [rows, Dpad4] +-1 expansion
                                       # of `desc` (which stays the packed bits)
    rows: int
    width: int                         # D as passed to the C-ABI


def mode_of(cross_check: bool, ratio_threshold: Optional[float]) -> int:
    """EnhancedBFMatcher.match's branch (feature_matchers.py:208 vs :225)."""
    if ratio_threshold is not None and not cross_check:
        return L.MODE_RATIO
    return L.MODE_CROSS if cross_check else L.MODE_NN


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


class _DescriptorCache:
    """Prepared descriptor stores keyed on the HOST array they were made from, so that a matcher instance reused over all
    pairs of a run (the reference does exactly that, pipeline.py:99-104) uploads and re-encodes every image once instead of
    once per pair (match_folder visits every image I - 1 times).

    An entry is served only while (1) the array object it was made from is still alive -- the cache holds a weak reference,
    so its memory cannot have been handed to another array -- and (2) a fingerprint of sampled rows is unchanged, which
    catches in-place edits.  The reference never mutates descriptors (SURVEY 8b "borrowed, never mutated"); a caller that
    rewrites single elements in place between calls must call Engine.clear_cache() or construct Engine(cache_entries=0).
    Bounded by entry count and by device bytes (least recently used goes first)."""

    def __init__(self, max_entries: int, max_bytes: int):
        import collections
        self.max_entries, self.max_bytes = int(max_entries), int(max_bytes)
        self.d = collections.OrderedDict()
        self.bytes = 0
        self.hits = self.misses = 0

    @staticmethod
    def _key(arr: np.ndarray, norm_type: int, force_kind):
        return (arr.__array_interface__["data"][0], arr.shape, arr.dtype.str, arr.strides, norm_type, force_kind)

    @staticmethod
    def _fingerprint(arr: np.ndarray) -> int:
        n = arr.shape[0]
        rows = arr if n <= 16 else arr[np.linspace(0, n - 1, 16).astype(np.int64)]
        return hash(rows.tobytes())

    def get(self, arr, norm_type, force_kind):
        if self.max_entries <= 0:
            return None
        e = self.d.get(self._key(arr, norm_type, force_kind))
        if e is not None and e[0]() is not None and e[1] == self._fingerprint(arr):
            self.d.move_to_end(self._key(arr, norm_type, force_kind))
            self.hits += 1
            return e[2]
        self.misses += 1
        return None

    def put(self, arr, norm_type, force_kind, dd) -> None:
        if self.max_entries <= 0:
            return
        import weakref
        try:
            ref = weakref.ref(arr)
        except TypeError:
            return
        nbytes = sum(t.numel() * t.element_size() for t in (dd.desc, dd.norm) if t is not None)
        if nbytes > self.max_bytes:
            return
        key = self._key(arr, norm_type, force_kind)
        old = self.d.pop(key, None)
        if old is not None:
            self.bytes -= old[3]
        self.d[key] = (ref, self._fingerprint(arr), dd, nbytes)
        self.bytes += nbytes
        while len(self.d) > self.max_entries or self.bytes > self.max_bytes:
            _, ev = self.d.popitem(last=False)
            self.bytes -= ev[3]

    def clear(self) -> None:
        self.d.clear()
        self.bytes = 0


class Engine:
    def __init__(self, device: Optional[torch.device | str | int] = None, hamming_path: Optional[str] = None,
                 float_path: Optional[str] = None, cache_entries: int = 64, cache_bytes: int = 4 << 30):
        """hamming_path: "fp4" (tcgen05 kind::mxf4 over bits expanded to e2m1 +-1.0, descriptors up to 64 bytes), "int8"
        (tcgen05 kind::i8 over +-1 int8, same limit), "popc" (the INT-pipe kernel, any width) or "auto" (fp4 where the
        width allows it, else popc); default from B200M_HAMMING_PATH, else auto.  All produce identical results."""
        self.lib = L.load()
        # non-integer float descriptors: "f16" = tcgen05 kind::f16 candidates over a scaled fp16 copy (default), "tf32" =
        # kind::tf32 candidates straight from the fp32 rows; the exact fp32 re-rank makes the results identical.
        self.float_path = float_path or os.environ.get("B200M_FLOAT_PATH", "f16")
        if self.float_path not in ("f16", "tf32"):
            raise ValueError("float_path must be f16 or tf32")
        self.hamming_path = hamming_path or os.environ.get("B200M_HAMMING_PATH", "auto")
        if self.hamming_path not in ("auto", "fp4", "int8", "popc"):
            raise ValueError("hamming_path must be auto, fp4, int8 or popc")
        if not torch.cuda.is_available():
            raise L.B200MatchError(L.ENODEVICE, "Engine", "no CUDA device: the matcher has no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        with torch.cuda.device(self.device):
            L.check(self.lib.b200m_device_info(None, None, None), "b200m_device_info")
        self._ws: Optional[torch.Tensor] = None
        self.cache = _DescriptorCache(cache_entries, cache_bytes)      # host array -> prepared store (match_arrays / knn2_arrays)
        self._out_dev: Optional[torch.Tensor] = None                   # packed result buffer of match_arrays + its pinned mirror
        self._out_pin: Optional[torch.Tensor] = None

    def clear_cache(self) -> None:
        self.cache.clear()

    # ------------------------------------------------------------------ memory helpers
    def _workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
        return self._ws

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def to_device(self, a: np.ndarray) -> torch.Tensor:
        return torch.from_numpy(np.ascontiguousarray(a)).to(self.device, non_blocking=True)

    # ------------------------------------------------------------------ descriptor upload / prep
    def prepare(self, desc, norm_type: int, force_kind: Optional[int] = None) -> DeviceDescriptors:
        """Upload (if needed) and re-encode one descriptor matrix [rows, D]."""
        t = desc if isinstance(desc, torch.Tensor) else self.to_device(desc)
        if t.dim() != 2:
            raise ValueError("descriptors must be [rows, D]")
        t = t.contiguous()
        rows, D = int(t.shape[0]), int(t.shape[1])
        if norm_type == NORM_HAMMING:
            if t.dtype != torch.uint8:
                raise ValueError("NORM_HAMMING needs uint8 descriptors")      # cv2.error in the reference
            want = force_kind
            if want is None:
                want = {"popc": L.KIND_HAMMING, "int8": L.KIND_HAMMING_TC}.get(self.hamming_path, L.KIND_HAMMING_FP4)
            if want == L.KIND_HAMMING:
                return DeviceDescriptors(L.KIND_HAMMING, t, None, rows, D)
            fp4 = want == L.KIND_HAMMING_FP4
            width_fn = self.lib.b200m_hamming_fp4_width if fp4 else self.lib.b200m_hamming_tc_width
            prep_fn = self.lib.b200m_prep_hamming_fp4 if fp4 else self.lib.b200m_prep_hamming_tc
            wpad = ctypes.c_int(0)
            rc = width_fn(D, ctypes.byref(wpad))
            if rc == L.EUNSUPPORTED and force_kind is None and self.hamming_path == "auto":
                return DeviceDescriptors(L.KIND_HAMMING, t, None, rows, D)     # wider than 512 bits: INT-pipe kernel
            L.check(rc, "b200m_hamming_*_width")
            with torch.cuda.device(self.device):
                ex = torch.empty((rows, wpad.value), dtype=torch.uint8 if fp4 else torch.int8, device=self.device)
                L.check(prep_fn(t.data_ptr(), rows, D, ex.data_ptr(), wpad.value, self._stream()), "b200m_prep_hamming_*")
            return DeviceDescriptors(want, t, ex, rows, D)
        if norm_type != NORM_L2:
            raise UnsupportedNormError(f"norm_type {norm_type}: only cv2.NORM_L2 (4) and cv2.NORM_HAMMING (6) are implemented")
        if t.dtype == torch.uint8:
            t = t.to(torch.float32)
        if t.dtype != torch.float32:
            raise ValueError("NORM_L2 needs float32 (or uint8) descriptors")
        with torch.cuda.device(self.device):
            if force_kind not in (L.KIND_L2_F32, L.KIND_L2_F16C):
                dd8, flag = self._try_l2_i8(t)
                if force_kind == L.KIND_L2_I8 or int(flag.item()) == 0:
                    return dd8
            Dpad = (D + 31) // 32 * 32
            if Dpad != D:
                t = torch.nn.functional.pad(t, (0, Dpad - D))
            t = t.contiguous()
            if force_kind == L.KIND_L2_F16C or (force_kind != L.KIND_L2_F32 and self.float_path == "f16"):
                nb = ctypes.c_size_t(0)
                L.check(self.lib.b200m_l2_f16_aux_bytes(rows, Dpad, ctypes.byref(nb)), "b200m_l2_f16_aux_bytes")
                aux = torch.empty((nb.value // 4,), dtype=torch.float32, device=self.device)   # norms | fp16 copy | info
                L.check(self.lib.b200m_prep_l2_f16(t.data_ptr(), rows, Dpad, aux.data_ptr(), self._stream()), "b200m_prep_l2_f16")
                return DeviceDescriptors(L.KIND_L2_F16C, t, aux, rows, Dpad)
            nrm = torch.empty((rows,), dtype=torch.float32, device=self.device)
            L.check(self.lib.b200m_prep_l2_f32(t.data_ptr(), rows, Dpad, nrm.data_ptr(), self._stream()),
                    "b200m_prep_l2_f32")
            return DeviceDescriptors(L.KIND_L2_F32, t, nrm, rows, Dpad)

    def _try_l2_i8(self, t: torch.Tensor):
        """Enqueue the int8 re-encoding of float32 rows `t` (device, contiguous); returns (DeviceDescriptors, flag tensor):
        the encoding is exact iff the flag reads 0 once the stream has run (no synchronisation here)."""
        rows, D = int(t.shape[0]), int(t.shape[1])
        Dpad = (D + 127) // 128 * 128
        with torch.cuda.device(self.device):
            d8 = torch.empty((rows, Dpad), dtype=torch.int8, device=self.device)
            nb = ctypes.c_size_t(0)
            L.check(self.lib.b200m_l2_i8_aux_bytes(rows, ctypes.byref(nb)), "b200m_l2_i8_aux_bytes")
            nrm = torch.empty((nb.value // 4,), dtype=torch.int32, device=self.device)   # norms[rows] + fold digits + info
            flag = torch.zeros((1,), dtype=torch.int32, device=self.device)
            L.check(self.lib.b200m_prep_l2_i8(t.data_ptr(), rows, D, d8.data_ptr(), Dpad, nrm.data_ptr(),
                                              flag.data_ptr(), self._stream()), "b200m_prep_l2_i8")
        return DeviceDescriptors(L.KIND_L2_I8, d8, nrm, rows, Dpad), flag

    # ------------------------------------------------------------------ incremental L2_I8 store (upload in pieces)
    def alloc_l2_i8_store(self, rows: int, D: int):
        """Empty int8 store for `rows` integer-valued float descriptors of width D, to be filled by
        prepare_l2_i8_rows().  Returns (DeviceDescriptors, not_integer flag tensor)."""
        Dpad = (D + 127) // 128 * 128
        with torch.cuda.device(self.device):
            d8 = torch.empty((rows, Dpad), dtype=torch.int8, device=self.device)
            nb = ctypes.c_size_t(0)
            L.check(self.lib.b200m_l2_i8_aux_bytes(rows, ctypes.byref(nb)), "b200m_l2_i8_aux_bytes")
            aux = torch.empty((nb.value // 4,), dtype=torch.int32, device=self.device)
            flag = torch.zeros((1,), dtype=torch.int32, device=self.device)
            L.check(self.lib.b200m_l2_i8_aux_init(aux.data_ptr(), rows, D, self._stream()), "b200m_l2_i8_aux_init")
        return DeviceDescriptors(L.KIND_L2_I8, d8, aux, rows, Dpad), flag

    def prepare_l2_i8_rows(self, dd: DeviceDescriptors, flag: torch.Tensor, src: torch.Tensor, row_begin: int) -> None:
        """Re-encode float32 rows `src` [cnt, D] (device) into rows [row_begin, row_begin + cnt) of `dd` on the current stream."""
        cnt, D = int(src.shape[0]), int(src.shape[1])
        with torch.cuda.device(self.device):
            L.check(self.lib.b200m_prep_l2_i8_rows(src.data_ptr(), row_begin, cnt, dd.rows, D, dd.desc.data_ptr(), dd.width,
                                                   dd.norm.data_ptr(), flag.data_ptr(), self._stream()),
                    "b200m_prep_l2_i8_rows")

    def f32_stats(self, reset: bool = True):
        """(rows re-scanned exhaustively, their distance evaluations) of the L2_F32 path since the last reset."""
        rows, evals = ctypes.c_int64(0), ctypes.c_int64(0)
        with torch.cuda.device(self.device):
            L.check(self.lib.b200m_f32_stats(ctypes.byref(rows), ctypes.byref(evals), int(reset)), "b200m_f32_stats")
        return rows.value, evals.value

    # ------------------------------------------------------------------ pair tables
    @staticmethod
    def make_pairs(spec: Sequence[Sequence[int]]):
        """spec rows: (q_off, n, t_off, m).  Returns (ctypes array, out offsets, total_n, total_m, max_n, max_m)."""
        arr = (L.Pair * max(len(spec), 1))()
        off = 0
        tn = tm = mn = mm = 0
        offs = np.zeros(len(spec) + 1, np.int64)
        for p, (q_off, n, t_off, m) in enumerate(spec):
            arr[p] = L.Pair(int(q_off), int(n), int(t_off), int(m), off)
            off += int(n)
            offs[p + 1] = off
            tn += int(n); tm += int(m); mn = max(mn, int(n)); mm = max(mm, int(m))
        return arr, offs, tn, tm, mn, mm

    # ------------------------------------------------------------------ the hot path (device in, device out)
    def match_device(self, q: DeviceDescriptors, t: DeviceDescriptors, spec, mode: int, ratio: float = 0.0):
        """Enqueue matching for the pairs in `spec`; returns device tensors
        (out_q, out_t, out_d, out_count, out_offsets[np])."""
        if q.kind != t.kind or q.width != t.width:
            raise ValueError("descriptor kind/width mismatch")
        pairs, offs, tn, tm, mn, mm = self.make_pairs(spec)
        P = len(spec)
        nbytes = ctypes.c_size_t(0)
        L.check(self.lib.b200m_workspace_bytes(q.kind, mode, P, tn, tm, mn, mm, q.width, ctypes.byref(nbytes)),
                "b200m_workspace_bytes")
        ws = self._workspace(nbytes.value)
        cap = max(tn, 1)
        out_q = torch.empty(cap, dtype=torch.int32, device=self.device)
        out_t = torch.empty(cap, dtype=torch.int32, device=self.device)
        out_d = torch.empty(cap, dtype=torch.float32, device=self.device)
        out_c = torch.zeros(max(P, 1), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self.lib.b200m_match_pairs(q.kind, q.desc.data_ptr(), _ptr(q.norm), q.rows, t.desc.data_ptr(),
                                            _ptr(t.norm), t.rows, q.width, pairs, P, mode, float(ratio),
                                            out_q.data_ptr(), out_t.data_ptr(), out_d.data_ptr(), out_c.data_ptr(),
                                            ws.data_ptr(), ws.numel(), self._stream())
        L.check(rc, "b200m_match_pairs")
        return out_q, out_t, out_d, out_c, offs

    def knn2_device(self, q: DeviceDescriptors, t: DeviceDescriptors, spec):
        if q.kind != t.kind or q.width != t.width:
            raise ValueError("descriptor kind/width mismatch")
        pairs, offs, tn, tm, mn, mm = self.make_pairs(spec)
        P = len(spec)
        nbytes = ctypes.c_size_t(0)
        L.check(self.lib.b200m_workspace_bytes(q.kind, L.MODE_NN, P, tn, tm, mn, mm, q.width, ctypes.byref(nbytes)),
                "b200m_workspace_bytes")
        ws = self._workspace(nbytes.value)
        cap = max(tn, 1)
        i1 = torch.empty(cap, dtype=torch.int32, device=self.device)
        i2 = torch.empty(cap, dtype=torch.int32, device=self.device)
        d1 = torch.empty(cap, dtype=torch.float32, device=self.device)
        d2 = torch.empty(cap, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self.lib.b200m_knn2_pairs(q.kind, q.desc.data_ptr(), _ptr(q.norm), q.rows, t.desc.data_ptr(),
                                           _ptr(t.norm), t.rows, q.width, pairs, P, i1.data_ptr(), i2.data_ptr(),
                                           d1.data_ptr(), d2.data_ptr(), ws.data_ptr(), ws.numel(), self._stream())
        L.check(rc, "b200m_knn2_pairs")
        return i1, i2, d1, d2, offs

    # ------------------------------------------------------------------ host-array convenience (one pair)
    def _prepare_pair(self, desc1, desc2, norm_type):
        """Both images' stores, from the cache where the host arrays were seen before (see _DescriptorCache)."""
        float_kind = L.KIND_L2_F16C if self.float_path == "f16" else L.KIND_L2_F32
        host = isinstance(desc1, np.ndarray) and isinstance(desc2, np.ndarray)
        if not host:
            a, b = self.prepare(desc1, norm_type), self.prepare(desc2, norm_type)
            if a.kind != b.kind:
                a, b = self.prepare(desc1, norm_type, force_kind=float_kind), self.prepare(desc2, norm_type, force_kind=float_kind)
            return a, b
        out = [self.cache.get(d, norm_type, None) for d in (desc1, desc2)]
        todo = [k for k in (0, 1) if out[k] is None]
        if desc1 is desc2 and len(todo) == 2:
            todo = [0]
        if norm_type == NORM_L2 and desc1.dtype == np.float32 and todo:
            # integer-valued?  Both re-encodings are enqueued first, their flags read back together: one synchronisation
            ups = {k: self.to_device((desc1, desc2)[k]) for k in todo}
            tries = {k: self._try_l2_i8(ups[k]) for k in todo}
            flags = torch.cat([tries[k][1] for k in todo]).cpu().tolist()
            for k, bad in zip(todo, flags):
                out[k] = tries[k][0] if bad == 0 else self.prepare(ups[k], norm_type, force_kind=float_kind)
        else:
            for k in todo:
                out[k] = self.prepare((desc1, desc2)[k], norm_type)
        if desc1 is desc2 and out[1] is None:
            out[1] = out[0]
        for k in todo:
            self.cache.put((desc1, desc2)[k], norm_type, None, out[k])
        a, b = out
        if a.kind != b.kind:      # one side integer-valued, the other not: both go the float way
            both = []
            for d in (desc1, desc2):
                dd = self.cache.get(d, norm_type, float_kind)
                if dd is None:
                    dd = self.prepare(d, norm_type, force_kind=float_kind)
                    self.cache.put(d, norm_type, float_kind, dd)
                both.append(dd)
            a, b = both
        return a, b

    def _match_packed(self, a: DeviceDescriptors, b: DeviceDescriptors, mode: int, ratio: float) -> MatchArrays:
        """One pair -> b200m_match_pairs_rec (12-byte records) into ONE device buffer [records | ptr[2] | count] that comes
        back with ONE device-to-host copy into pinned memory and one synchronisation."""
        if a.kind != b.kind or a.width != b.width:
            raise ValueError("descriptor kind/width mismatch")
        lib = self.lib
        n, m = a.rows, b.rows
        pairs = (L.Pair * 1)(L.Pair(0, n, 0, m, 0))
        nbytes = ctypes.c_size_t(0)
        L.check(lib.b200m_workspace_bytes(a.kind, mode, 1, n, m, n, m, a.width, ctypes.byref(nbytes)), "b200m_workspace_bytes")
        ws = self._workspace(nbytes.value)
        rec_bytes = (max(n, 1) * 12 + 7) // 8 * 8
        total = rec_bytes + 24
        if self._out_dev is None or self._out_dev.numel() < total:
            self._out_dev = torch.empty(max(total, 1 << 18), dtype=torch.uint8, device=self.device)
            self._out_pin = torch.empty(self._out_dev.numel(), dtype=torch.uint8, pin_memory=True)
        dev, pin = self._out_dev, self._out_pin
        base = dev.data_ptr()
        with torch.cuda.device(self.device):
            rc = lib.b200m_match_pairs_rec(a.kind, a.desc.data_ptr(), _ptr(a.norm), a.rows, b.desc.data_ptr(), _ptr(b.norm), b.rows,
                                           a.width, pairs, 1, mode, float(ratio), L.REC_Q32T32D32, base, base + rec_bytes + 16,
                                           base + rec_bytes, ws.data_ptr(), ws.numel(), self._stream())
        L.check(rc, "b200m_match_pairs_rec")
        big = total > (1 << 20)                       # very large images: the count first, then only the records that exist
        lo = rec_bytes if big else 0
        pin[lo:total].copy_(dev[lo:total], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        host = pin.numpy()
        k = int(host[rec_bytes + 8:rec_bytes + 16].view(np.int64)[0]) if n > 0 and m > 0 else 0
        if big and k:
            pin[:k * 12].copy_(dev[:k * 12], non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        rec = host[:k * 12].view(_REC12)
        return MatchArrays(rec["q"].copy(), rec["t"].copy(), rec["d"].copy())

    def match_arrays(self, desc1: np.ndarray, desc2: np.ndarray, norm_type: int = NORM_L2,
                     cross_check: bool = False, ratio_threshold: Optional[float] = None) -> MatchArrays:
        """One image pair, host arrays in, host arrays out (H2D + kernels + D2H).  Arrays seen before come from the
        descriptor cache: no upload, no re-encoding."""
        if desc1.ndim != 2 or desc2.ndim != 2 or desc1.shape[1] != desc2.shape[1]:
            raise ValueError("descriptor width mismatch")
        if desc1.dtype != desc2.dtype:
            raise ValueError("descriptor dtype mismatch")
        a, b = self._prepare_pair(desc1, desc2, norm_type)
        return self._match_packed(a, b, mode_of(cross_check, ratio_threshold), ratio_threshold or 0.0)

    def knn2_arrays(self, desc1: np.ndarray, desc2: np.ndarray, norm_type: int = NORM_L2) -> Knn2Arrays:
        a, b = self._prepare_pair(desc1, desc2, norm_type)
        i1, i2, d1, d2, _ = self.knn2_device(a, b, [(0, a.rows, 0, b.rows)])
        n = a.rows
        return Knn2Arrays(i1[:n].cpu().numpy(), i2[:n].cpu().numpy(), d1[:n].cpu().numpy(), d2[:n].cpu().numpy())


_default_engine: Optional[Engine] = None


def default_engine() -> Engine:
    global _default_engine
    if _default_engine is None:
        _default_engine = Engine()
    return _default_engine
