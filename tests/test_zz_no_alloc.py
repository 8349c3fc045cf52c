"""b200m_match_pairs_static (include/b200match.h): the matcher with its workspace in a static __device__ array inside the
library, for callers that cannot allocate.  Called through the C-ABI and compared with the existing entry point
b200m_match_pairs (caller-owned workspace) on the same inputs: bit-exact (indices, float32 distance bits, order, counts)
-- both run the same kernels; only where the workspace lives differs -- on calls in a row that reuse the dirty scratch
with different kinds, modes and shapes.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

from conftest import pkg

pytestmark = pytest.mark.gpu
b = pkg()
L = b._lib
synth = b.synth


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return b.Engine()


def _both(eng, a, c, spec, mode, ratio=0.0):
    """(static entry point, existing entry point) results for the pairs in `spec` = [(q_off, n, t_off, m)]."""
    lib = eng.lib
    pairs, offs, tn, tm, mn, mm = eng.make_pairs(spec)
    P, cap = len(spec), max(tn, 1)
    dev = eng.device
    out = []
    for static in (True, False):
        oq = torch.full((cap,), -7, dtype=torch.int32, device=dev)
        ot = torch.full((cap,), -7, dtype=torch.int32, device=dev)
        od = torch.full((cap,), -7.0, dtype=torch.float32, device=dev)
        oc = torch.full((P,), -7, dtype=torch.int32, device=dev)
        args = [a.kind, a.desc.data_ptr(), 0 if a.norm is None else a.norm.data_ptr(), a.rows, c.desc.data_ptr(),
                0 if c.norm is None else c.norm.data_ptr(), c.rows, a.width, pairs, P, mode, float(ratio),
                oq.data_ptr(), ot.data_ptr(), od.data_ptr(), oc.data_ptr()]
        with torch.cuda.device(dev):
            if static:
                rc = lib.b200m_match_pairs_static(*args, eng._stream())
                L.check(rc, "b200m_match_pairs_static")
            else:
                nbytes = ctypes.c_size_t(0)
                L.check(lib.b200m_workspace_bytes(a.kind, mode, P, tn, tm, mn, mm, a.width, ctypes.byref(nbytes)), "ws")
                ws = torch.empty(max(nbytes.value, 1), dtype=torch.uint8, device=dev)
                L.check(lib.b200m_match_pairs(*args, ws.data_ptr(), ws.numel(), eng._stream()), "b200m_match_pairs")
        torch.cuda.synchronize(dev)
        cnt = oc.cpu().numpy()
        res = []
        for p in range(P):
            s, k = int(offs[p]), int(cnt[p])
            res.append((oq[s:s + k].cpu().numpy(), ot[s:s + k].cpu().numpy(), od[s:s + k].cpu().numpy()))
        out.append((cnt, res))
    return out


def _assert_same(got, want, what):
    assert np.array_equal(got[0], want[0]), f"{what}: counts {got[0]} vs {want[0]}"
    assert got[0].sum() > 0, f"{what}: no matches at all -- the comparison would be vacuous"
    for p, (g, w) in enumerate(zip(got[1], want[1])):
        assert np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1]), f"{what}: pair {p} indices"
        assert np.array_equal(g[2].view(np.uint32), w[2].view(np.uint32)), f"{what}: pair {p} distance bits"


def test_static_scratch_equals_caller_workspace_on_calls_in_a_row(eng):
    """Several calls in a row on ONE scratch array: SIFT ratio (int8 tensor path, 3 pairs), then ORB cross-check on both
    Hamming kernels, then float mutual NN, then SIFT again with another shape -- every call leaves the array dirty for the
    next one, every result equals the caller-workspace entry point's."""
    a0, c0 = synth.sift_pair(seed=11, n=900)
    sift = eng.prepare(np.concatenate([a0, c0]), b.NORM_L2)
    spec = [(0, 900, 900, 900), (900, 700, 0, 650), (0, 1, 900, 300)]
    for rep in range(2):                                          # the same call twice in a row
        got, want = _both(eng, sift, sift, spec, L.MODE_RATIO, 0.8)
        _assert_same(got, want, f"sift ratio rep {rep}")
    x, y = synth.binary_pair(seed=12, n=1500, m=1300, nbytes=32)
    for path in ("auto", "popc"):
        e = b.Engine(eng.device, hamming_path=path)
        dx, dy = e.prepare(x, b.NORM_HAMMING), e.prepare(y, b.NORM_HAMMING)
        for mode, ratio in ((L.MODE_CROSS, 0.0), (L.MODE_RATIO, 0.7), (L.MODE_NN, 0.0)):
            got, want = _both(e, dx, dy, [(0, 1500, 0, 1300)], mode, ratio)
            _assert_same(got, want, f"orb {path} mode {mode}")
    sp = synth.superpoint_set(seed=13, images=2, n=1024, d=256, pool=1500)
    f0, f1 = eng.prepare(sp[0], b.NORM_L2), eng.prepare(sp[1], b.NORM_L2)
    got, want = _both(eng, f0, f1, [(0, 1024, 0, 1024)], L.MODE_CROSS)
    _assert_same(got, want, "superpoint mutual")
    got, want = _both(eng, sift, sift, [(0, 513, 900, 257)], L.MODE_NN)
    _assert_same(got, want, "sift nn after the others")


def test_static_scratch_rejects_what_does_not_fit(eng):
    """A batch whose plan exceeds B200M_STATIC_SCRATCH_BYTES is refused with B200M_ESCRATCH and a message; nothing is
    enqueued, and the next call that fits still works."""
    assert L.STATIC_SCRATCH_BYTES == 32 << 20
    a0, c0 = synth.sift_pair(seed=14, n=2048)
    sift = eng.prepare(np.concatenate([a0, c0]), b.NORM_L2)
    spec = [(0, 2048, 2048, 2048)] * 256                          # 256 pairs x 2048 rows: 42 MiB of per-row knn state
    pairs, offs, tn, tm, mn, mm = eng.make_pairs(spec)
    nbytes = ctypes.c_size_t(0)
    L.check(eng.lib.b200m_workspace_bytes(sift.kind, L.MODE_RATIO, len(spec), tn, tm, mn, mm, sift.width, ctypes.byref(nbytes)), "ws")
    assert nbytes.value > L.STATIC_SCRATCH_BYTES
    o = torch.empty(tn, dtype=torch.int32, device=eng.device)
    od = torch.empty(tn, dtype=torch.float32, device=eng.device)
    oc = torch.empty(len(spec), dtype=torch.int32, device=eng.device)
    with torch.cuda.device(eng.device):
        rc = eng.lib.b200m_match_pairs_static(sift.kind, sift.desc.data_ptr(), sift.norm.data_ptr(), sift.rows, sift.desc.data_ptr(),
                                              sift.norm.data_ptr(), sift.rows, sift.width, pairs, len(spec), L.MODE_RATIO, 0.8,
                                              o.data_ptr(), o.data_ptr(), od.data_ptr(), oc.data_ptr(), eng._stream())
    assert rc == L.ESCRATCH
    assert b"static scratch" in eng.lib.b200m_last_cuda_error() and b"scratch" in eng.lib.b200m_error_string(rc)
    with pytest.raises(L.B200MatchError):
        L.check(rc, "b200m_match_pairs_static")
    got, want = _both(eng, sift, sift, [(0, 2048, 2048, 2048)], L.MODE_RATIO, 0.8)
    _assert_same(got, want, "after a refused call")
