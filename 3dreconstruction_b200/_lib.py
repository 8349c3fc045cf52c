"""ctypes binding of libb200match.so (include/b200match.h).  Fails loudly: there is no CPU path.

The library is built in-tree by `3dreconstruction_b200/build.py` (explicit nvcc, sm_100a) and
loaded with ctypes — no torch types cross the boundary; torch only supplies device memory and
the stream handle.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("B200MATCH_SO") or os.path.join(HERE, "libb200match.so")   # override: experiment builds only

OK, EINVAL, EWORKSPACE, ECUDA, ENODEVICE, EUNSUPPORTED, ESCRATCH = 0, -1, -2, -3, -4, -5, -6
STATIC_SCRATCH_BYTES = 32 << 20
MODE_RATIO, MODE_CROSS, MODE_NN = 0, 1, 2
REC_Q32T32D32, REC_Q16T16D32, REC_Q16T16D16 = 1, 2, 3
KIND_HAMMING, KIND_L2_I8, KIND_L2_F32, KIND_HAMMING_TC, KIND_HAMMING_FP4, KIND_L2_F16C = 0, 1, 2, 3, 4, 5

EXPORTS = ["b200m_version", "b200m_error_string", "b200m_last_cuda_error", "b200m_device_info",
           "b200m_prep_l2_i8", "b200m_prep_l2_f32", "b200m_workspace_bytes", "b200m_match_pairs",
           "b200m_knn2_pairs", "b200m_launch_count", "b200m_reset_launch_count", "b200m_match_pairs_csr",
           "b200m_timing_enable", "b200m_timing_collect", "b200m_f32_stats", "b200m_l2_i8_aux_bytes",
           "b200m_l2_i8_aux_init", "b200m_prep_l2_i8_rows", "b200m_hamming_tc_width", "b200m_prep_hamming_tc",
           "b200m_gather_correspondences", "b200m_hamming_fp4_width", "b200m_prep_hamming_fp4", "b200m_ransac_homography", "b200m_l2_f16_aux_bytes", "b200m_prep_l2_f16",
           "b200m_l2_i8_aux_rows", "b200m_rec_bytes", "b200m_match_pairs_rec", "b200m_match_pairs_static",
           "b200m_ransac_essential"]


class Pair(ctypes.Structure):
    _fields_ = [("q_off", ctypes.c_int32), ("n", ctypes.c_int32), ("t_off", ctypes.c_int32),
                ("m", ctypes.c_int32), ("out_off", ctypes.c_int64)]


class B200MatchError(RuntimeError):
    def __init__(self, code: int, where: str, detail: str = ""):
        self.code = code
        super().__init__(f"{where}: b200match error {code}" + (f" ({detail})" if detail else ""))


_lib = None


def load() -> ctypes.CDLL:
    """Load the CUDA extension; raise (never fall back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(f"{SO_PATH} is missing: build it with `python 3dreconstruction_b200/build.py` "
                          "(the matcher has no CPU fallback)")
    lib = ctypes.CDLL(SO_PATH)
    vp, i32, i64, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
    lib.b200m_version.restype = i32
    lib.b200m_error_string.restype = ctypes.c_char_p
    lib.b200m_error_string.argtypes = [i32]
    lib.b200m_last_cuda_error.restype = ctypes.c_char_p
    lib.b200m_device_info.argtypes = [ctypes.POINTER(i32)] * 3
    lib.b200m_prep_l2_i8.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp]
    lib.b200m_l2_i8_aux_bytes.argtypes = [i64, ctypes.POINTER(ctypes.c_size_t)]
    lib.b200m_l2_i8_aux_init.argtypes = [vp, i64, i32, vp]
    lib.b200m_prep_l2_i8_rows.argtypes = [vp, i64, i64, i64, i32, vp, i32, vp, vp, vp]
    lib.b200m_prep_l2_f32.argtypes = [vp, i64, i32, vp, vp]
    lib.b200m_hamming_tc_width.argtypes = [i32, ctypes.POINTER(i32)]
    lib.b200m_prep_hamming_tc.argtypes = [vp, i64, i32, vp, i32, vp]
    lib.b200m_hamming_fp4_width.argtypes = [i32, ctypes.POINTER(i32)]
    lib.b200m_prep_hamming_fp4.argtypes = [vp, i64, i32, vp, i32, vp]
    lib.b200m_gather_correspondences.argtypes = [vp, i64, vp, vp, vp, i32, vp, vp, vp, vp]
    lib.b200m_l2_f16_aux_bytes.argtypes = [i64, i32, ctypes.POINTER(ctypes.c_size_t)]
    lib.b200m_prep_l2_f16.argtypes = [vp, i64, i32, vp, vp]
    lib.b200m_ransac_homography.argtypes = [vp, vp, i32, dbl, dbl, i32, vp, vp, vp, vp, vp]
    lib.b200m_ransac_essential.argtypes = [vp, vp, i32, vp, dbl, dbl, i32, vp, vp, vp, vp, vp]
    lib.b200m_workspace_bytes.argtypes = [i32, i32, i32, i64, i64, i32, i32, i32, ctypes.POINTER(ctypes.c_size_t)]
    lib.b200m_match_pairs.argtypes = [i32, vp, vp, i64, vp, vp, i64, i32, ctypes.POINTER(Pair), i32, i32, dbl,
                                      vp, vp, vp, vp, vp, ctypes.c_size_t, vp]
    lib.b200m_knn2_pairs.argtypes = [i32, vp, vp, i64, vp, vp, i64, i32, ctypes.POINTER(Pair), i32,
                                     vp, vp, vp, vp, vp, ctypes.c_size_t, vp]
    lib.b200m_match_pairs_csr.argtypes = [i32, vp, vp, i64, vp, vp, i64, i32, ctypes.POINTER(Pair), i32, i32, dbl,
                                          vp, vp, vp, vp, vp, vp, ctypes.c_size_t, vp]
    lib.b200m_match_pairs_rec.argtypes = [i32, vp, vp, i64, vp, vp, i64, i32, ctypes.POINTER(Pair), i32, i32, dbl,
                                          i32, vp, vp, vp, vp, ctypes.c_size_t, vp]
    lib.b200m_rec_bytes.argtypes = [i32]
    lib.b200m_match_pairs_static.argtypes = [i32, vp, vp, i64, vp, vp, i64, i32, ctypes.POINTER(Pair), i32, i32, dbl,
                                             vp, vp, vp, vp, vp]
    lib.b200m_l2_i8_aux_rows.argtypes = [vp, i64, i64, i64, i32, vp, vp]
    lib.b200m_timing_enable.argtypes = [i32]
    lib.b200m_timing_collect.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i32)]
    lib.b200m_f32_stats.argtypes = [ctypes.POINTER(i64), ctypes.POINTER(i64), i32]
    lib.b200m_launch_count.restype = i64
    lib.b200m_reset_launch_count.restype = None
    _lib = lib
    return lib


def check(rc: int, where: str) -> None:
    if rc != OK:
        lib = load()
        detail = lib.b200m_error_string(rc).decode()
        if rc in (ECUDA, ESCRATCH):
            detail += ": " + lib.b200m_last_cuda_error().decode()
        raise B200MatchError(rc, where, detail)
