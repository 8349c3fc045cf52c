"""GPU parity tests: the CUDA path (through the C-ABI) against the oracle on the same seeded
inputs, against the committed golden vectors produced by the reference's own EnhancedBFMatcher,
and — at BASELINE.json's full sizes — through size-independent properties.

Bar: bit-exact (indices, float32 distance bits, order) for Hamming and integer-valued L2;
relative distance error <= 1e-5 for non-integer float descriptors, near-tie index differences
counted and bounded."""
import numpy as np
import pytest
import torch

from conftest import is_integer_valued, pkg
from oracle import bf_oracle as O

pytestmark = pytest.mark.gpu
b = pkg()
synth = b.synth
RTOL = 1e-5


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return b.Engine()


@pytest.fixture(scope="module", params=["auto", "int8", "popc"])
def heng(request):
    """Hamming engines: "auto" = tensor-core kernel on e2m1 operands (kind::mxf4 GEMM + POPC resolve) up to 64-byte
    descriptors, INT-pipe kernel above; "int8" = the same kernel on int8 operands (kind::i8); "popc" = INT-pipe kernel
    everywhere.  All must be bit-exact against the oracle."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    e = b.Engine(hamming_path=request.param)
    if request.param == "int8":          # wider than 512 bits: the int8 engine has no kernel, the tests fall back explicitly
        prep = e.prepare
        e.prepare = lambda d, nt, force_kind=None: prep(d, nt, pkg("_lib").KIND_HAMMING if force_kind is None and d.shape[1] > 64 else force_kind)
    return e


@pytest.fixture(scope="module", params=["f16", "tf32"])
def feng(request):
    """Float-descriptor engines: "f16" = tcgen05 kind::f16 candidates over the scaled fp16 copy (default), "tf32" =
    kind::tf32 candidates from the fp32 rows.  Both re-rank exactly in fp32."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return b.Engine(float_path=request.param)


def assert_exact(got, want, what=""):
    assert len(got.query_idx) == len(want.query_idx), f"{what}: count {len(got.query_idx)} != {len(want.query_idx)}"
    assert np.array_equal(got.query_idx, want.query_idx), what
    assert np.array_equal(got.train_idx, want.train_idx), what
    assert np.array_equal(got.distance.view(np.uint32), want.distance.view(np.uint32)), what


def _flip_is_a_near_tie(q, g, w, a, c, mode):
    """True iff the difference between the CUDA result `g` and the oracle's `w` for query row q ((train, dist) or None =
    row not in the result) sits where the ORACLE's own decision margin, recomputed in float64, is inside RTOL."""
    a64, c64 = a.astype(np.float64), c.astype(np.float64)
    d = np.sqrt(((a64[q] - c64) ** 2).sum(axis=1))
    if g is not None and w is not None:                       # another train row: the two must be equidistant within RTOL
        return abs(d[g[0]] - d[w[0]]) <= RTOL * max(d[w[0]], 1e-6)
    cc, ratio = mode
    srt = np.sort(d)
    if cc:                                                    # in / out of the mutual set: a tie in either direction
        t = (g or w)[0]
        fwd = len(srt) > 1 and srt[1] - srt[0] <= RTOL * max(srt[0], 1e-6)
        e = np.sort(np.sqrt(((a64 - c64[t]) ** 2).sum(axis=1)))
        bwd = len(e) > 1 and e[1] - e[0] <= RTOL * max(e[0], 1e-6)
        return bool(fwd or bwd)
    if ratio is not None:                                     # in / out of the ratio test: d1 ~ ratio * d2
        return len(srt) > 1 and abs(srt[0] - ratio * srt[1]) <= RTOL * max(ratio * srt[1], 1e-6)
    return False                                              # plain 1-NN keeps every row: membership cannot differ


def assert_close_float(got, want, what="", max_flip_frac=0.01, descs=None, mode=None):
    """Non-integer float descriptors: compare as sets keyed by query; distances within RTOL; differences of index or of
    membership are counted, bounded, and -- with `descs` = (desc1, desc2) and `mode` = (cross_check, ratio) -- each one is
    verified to sit where the oracle's own margin (float64) is inside RTOL: north_star's "near-tie index differences
    counted and reported".  Returns the count."""
    g = {int(q): (int(t), float(d)) for q, t, d in zip(got.query_idx, got.train_idx, got.distance)}
    w = {int(q): (int(t), float(d)) for q, t, d in zip(want.query_idx, want.train_idx, want.distance)}
    flipped = sorted(set(g) ^ set(w))
    for q in set(g) & set(w):
        if g[q][0] != w[q][0]:
            flipped.append(q)
        assert abs(g[q][1] - w[q][1]) <= RTOL * max(abs(w[q][1]), 1e-6) + 1e-7, f"{what}: q={q} {g[q]} vs {w[q]}"
    assert len(flipped) <= max(1, int(max_flip_frac * max(len(w), 1))), f"{what}: {len(flipped)} near-tie flips of {len(w)}"
    if descs is not None:
        for q in flipped:
            assert _flip_is_a_near_tie(q, g.get(q), w.get(q), descs[0], descs[1], mode), \
                f"{what}: q={q} differs ({g.get(q)} vs oracle {w.get(q)}) and the oracle's margin is NOT inside {RTOL}"
    return len(flipped)


MODES = [(False, None), (False, 0.7), (False, 0.75), (True, None)]


# ------------------------------------------------------------------ golden vectors from the reference
def test_golden_vectors(eng, golden):
    for name, g in golden.items():
        if name.startswith("err_"):
            with pytest.raises(Exception):
                eng.match_arrays(g["desc1"], g["desc2"], g["norm"], g["cross_check"], g["ratio"])
            continue
        got = eng.match_arrays(g["desc1"], g["desc2"], g["norm"], g["cross_check"], g["ratio"])
        want = O.Matches(g["q"], g["t"], g["d"])
        if is_integer_valued(g["desc1"]) and is_integer_valued(g["desc2"]):
            assert_exact(got, want, name)
        else:
            assert_close_float(got, want, name)


# ------------------------------------------------------------------ Hamming
@pytest.mark.parametrize("nbytes", [1, 5, 32, 61, 64, 128, 200])
@pytest.mark.parametrize("mode", MODES)
def test_hamming_vs_oracle(heng, nbytes, mode):
    for seed, (n, m) in enumerate([(1, 1), (1, 7), (9, 1), (2, 2), (130, 127), (257, 513), (1000, 777)]):
        a, c = synth.binary_pair(seed=100 + seed, n=max(n, m), nbytes=nbytes)
        a, c = a[:n], c[:m]
        assert_exact(heng.match_arrays(a, c, b.NORM_HAMMING, *mode), O.reference_match(a, c, O.NORM_HAMMING, *mode),
                     f"hamming D={nbytes} n={n} m={m} mode={mode}")


def test_hamming_heavy_ties(heng):
    rng = np.random.default_rng(7)
    for nbytes, hi in [(1, 2), (1, 4), (2, 3), (32, 2)]:
        a = rng.integers(0, hi, size=(300, nbytes), dtype=np.uint8)
        c = np.concatenate([a[:100], a[:100], np.zeros((10, nbytes), np.uint8),
                            rng.integers(0, hi, size=(211, nbytes), dtype=np.uint8)])
        for mode in MODES:
            assert_exact(heng.match_arrays(a, c, b.NORM_HAMMING, *mode),
                         O.reference_match(a, c, O.NORM_HAMMING, *mode), f"ties D={nbytes} hi={hi} {mode}")


def test_hamming_knn2_raw(heng):
    a, c = synth.binary_pair(seed=5, n=700, m=650, nbytes=32)
    got = heng.knn2_arrays(a, c, b.NORM_HAMMING)
    want = O.knn2_c(a, c, O.NORM_HAMMING)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)
    got = heng.knn2_arrays(a, c[:1], b.NORM_HAMMING)          # M == 1: second neighbour absent
    assert np.all(got.idx2 == -1) and np.all(got.d2 == -1)


def test_hamming_config2_full_size(heng):
    """BASELINE config 2: ORB 256-bit, 10k x 10k, cross-check.  Oracle finishes in seconds here."""
    a, c = synth.binary_pair(seed=1, n=10000, nbytes=32)
    got = heng.match_arrays(a, c, b.NORM_HAMMING, True, None)
    assert_exact(got, O.reference_match(a, c, O.NORM_HAMMING, True, None), "config2")
    assert len(got.query_idx) > 5000
    # properties: sorted by (distance, queryIdx); mutual (swap roles -> same pair set)
    key = got.distance.astype(np.float64) * 1e6 + got.query_idx
    assert np.all(np.diff(key) > 0)
    rev = heng.match_arrays(c, a, b.NORM_HAMMING, True, None)
    assert set(zip(got.query_idx.tolist(), got.train_idx.tolist())) == set(zip(rev.train_idx.tolist(), rev.query_idx.tolist()))


def test_hamming_batched_pairs_equal_single(heng):
    """Several pairs with ragged sizes in one C-ABI call == the same pairs one by one."""
    sets = [synth.binary_pair(seed=40 + i, n=n, nbytes=32)[0] for i, n in enumerate([300, 129, 1, 640, 77])]
    store = np.concatenate(sets)
    offs = np.cumsum([0] + [len(s) for s in sets])
    dd = heng.prepare(store, b.NORM_HAMMING)
    spec, pairs = [], [(0, 1), (0, 3), (2, 4), (3, 1), (4, 2), (1, 1)]
    for i, j in pairs:
        spec.append((offs[i], len(sets[i]), offs[j], len(sets[j])))
    for mode in MODES:
        oq, ot, od, oc, ooff = heng.match_device(dd, dd, spec, b.mode_of(*mode), mode[1] or 0.0)
        oq, ot, od, oc = oq.cpu().numpy(), ot.cpu().numpy(), od.cpu().numpy(), oc.cpu().numpy()
        for p, (i, j) in enumerate(pairs):
            want = O.reference_match(sets[i], sets[j], O.NORM_HAMMING, *mode)
            s = slice(int(ooff[p]), int(ooff[p]) + int(oc[p]))
            assert_exact(O.Matches(oq[s], ot[s], od[s]), want, f"batched pair {p} mode {mode}")


def test_hamming_tc_kind_is_used_and_matches_popc(heng):
    """The tensor path must actually be the one that runs under "auto" (no silent detour through the POPC kernel)."""
    a, c = synth.binary_pair(seed=9, n=1500, m=1300, nbytes=32)
    dd = heng.prepare(a, b.NORM_HAMMING)
    L = pkg("_lib")
    assert dd.kind == {"auto": L.KIND_HAMMING_FP4, "int8": L.KIND_HAMMING_TC, "popc": L.KIND_HAMMING}[heng.hamming_path]
    bits = np.unpackbits(a, axis=1, bitorder="little").astype(np.int8) * 2 - 1
    if dd.kind == L.KIND_HAMMING_TC:
        pm1 = dd.norm.cpu().numpy()
        assert pm1.shape == (1500, 256) and np.array_equal(pm1, bits)
        dd61 = heng.prepare(a[:, :29], b.NORM_HAMMING)          # 232 bits -> 256 dims, the last 24 are zero
        p = dd61.norm.cpu().numpy()
        assert p.shape == (1500, 256) and np.array_equal(p[:, :232], bits[:, :232]) and not p[:, 232:].any()
    if dd.kind == L.KIND_HAMMING_FP4:
        nib = dd.norm.cpu().numpy()                              # two e2m1 values per byte, low nibble first
        assert nib.shape == (1500, 128)
        lo, hi = nib & 15, nib >> 4
        dec = np.stack([lo, hi], axis=2).reshape(1500, 256)
        assert np.array_equal(dec, np.where(bits > 0, 0x2, 0xA))
        p = heng.prepare(a[:, :29], b.NORM_HAMMING).norm.cpu().numpy()
        assert p.shape == (1500, 128) and not p[:, 116:].any()


@pytest.mark.parametrize("nbytes", [16, 32, 48, 61, 64])
def test_hamming_tc_column_splits_tails_and_chunk_ties(heng, nbytes):
    """Sizes that exercise one/two CTAs per cluster, ragged last tiles, several column splits; duplicated train rows
    put equal distances into different 16-column chunks, tiles and splits (ties -> lowest train index)."""
    rng = np.random.default_rng(nbytes)
    for n, m in [(128, 256), (129, 257), (255, 4096), (700, 5000), (2100, 1031)]:
        a = rng.integers(0, 256, size=(n, nbytes), dtype=np.uint8)
        c = rng.integers(0, 256, size=(m, nbytes), dtype=np.uint8)
        k = min(n, m) // 3
        c[m - k:] = a[:k]                            # exact copies late in the train set ...
        c[5:5 + k // 2] = a[:k // 2]                 # ... and earlier duplicates of half of them
        c[m // 2] = c[m // 2 + 17] = a[n - 1]
        for mode in MODES:
            assert_exact(heng.match_arrays(a, c, b.NORM_HAMMING, *mode), O.reference_match(a, c, O.NORM_HAMMING, *mode),
                         f"tc D={nbytes} n={n} m={m} mode={mode}")
        got = heng.knn2_arrays(a, c, b.NORM_HAMMING)
        for x, y in zip(got, O.knn2_c(a, c, O.NORM_HAMMING)):
            assert np.array_equal(x, y), f"knn2 D={nbytes} n={n} m={m}"


def test_hamming_akaze_full_size_pair(heng):
    """One pair of BASELINE config 5's shape (AKAZE 16384 x 16384 x 61 bytes, mutual check) against the C oracle."""
    a, c = synth.binary_pair(seed=4, n=16384, nbytes=61)
    got = heng.match_arrays(a, c, b.NORM_HAMMING, True, None)
    assert_exact(got, O.reference_match(a, c, O.NORM_HAMMING, True, None), "akaze full pair")



# ------------------------------------------------------------------ integer-valued L2 (SIFT) — tcgen05 kind::i8
@pytest.mark.parametrize("mode", MODES)
def test_l2_i8_vs_oracle(eng, mode):
    for seed, (n, m, d) in enumerate([(1, 1, 128), (1, 9, 128), (9, 1, 128), (2, 2, 128), (130, 127, 128),
                                      (300, 513, 128), (1000, 777, 128), (200, 300, 64), (150, 150, 8),
                                      (260, 140, 200), (500, 400, 256)]):
        a, c = synth.sift_pair(seed=200 + seed, n=max(n, m), d=d)
        a, c = a[:n], np.ascontiguousarray(c[:m])
        assert_exact(eng.match_arrays(a, c, b.NORM_L2, *mode), O.reference_match(a, c, O.NORM_L2, *mode),
                     f"l2_i8 n={n} m={m} d={d} mode={mode}")


def test_l2_i8_extremes_and_ties(eng):
    rng = np.random.default_rng(9)
    a = rng.choice([0.0, 255.0], size=(200, 128)).astype(np.float32)     # largest possible distances
    c = rng.choice([0.0, 255.0], size=(180, 128)).astype(np.float32)
    c[:50] = a[:50]
    c[50:60] = 0
    for mode in MODES:
        assert_exact(eng.match_arrays(a, c, b.NORM_L2, *mode), O.reference_match(a, c, O.NORM_L2, *mode), f"extreme {mode}")
    a = rng.integers(0, 3, size=(300, 128)).astype(np.float32)           # heavy ties
    c = np.concatenate([a[:100], a[:100], rng.integers(0, 3, size=(150, 128)).astype(np.float32)])
    for mode in MODES:
        assert_exact(eng.match_arrays(a, c, b.NORM_L2, *mode), O.reference_match(a, c, O.NORM_L2, *mode), f"ties {mode}")


def _fold_flag(dd):
    """info[3] of the aux block b200m_prep_l2_i8 wrote: 1 = MODE_RATIO runs the folded-norm kernel."""
    rows = dd.rows
    fold_off = (rows * 4 + 255) // 256 * 256
    info_off = fold_off + (rows * 32 + 255) // 256 * 256
    assert dd.norm.numel() * 4 == info_off + 256
    return int(dd.norm[info_off // 4 + 3].item())


def test_l2_i8_fold_digits_encode_the_norm(eng):
    """sum_k c_k t_k == -(norm >> 1 - bias) for every row, c = {-128 x15, -1} x2 (prep.cu / common.cuh)."""
    a, _ = synth.sift_pair(seed=77, n=3000)
    dd = eng.prepare(a, b.NORM_L2)
    assert dd.kind == b._lib.KIND_L2_I8 and _fold_flag(dd) == 1
    rows = dd.rows
    fold_off = (rows * 4 + 255) // 256 * 256
    info_off = fold_off + (rows * 32 + 255) // 256 * 256
    raw = dd.norm.cpu().numpy()
    norms = raw[:rows].astype(np.int64)
    want = ((a.astype(np.int64) - 128) ** 2).sum(1)
    assert np.array_equal(norms, want)
    digits = raw.view(np.int8)[fold_off:fold_off + rows * 32].reshape(rows, 32).astype(np.int64)
    info = raw[info_off // 4: info_off // 4 + 4]
    c = np.array(([-128] * 15 + [-1]) * 2, np.int64)
    assert info[0] == (norms >> 1).min() and info[1] == (norms >> 1).max()
    assert np.array_equal(digits @ c, -((norms >> 1) - int(info[2])))


@pytest.mark.parametrize("ratio", [0.5, 0.9, 0.99, 1.0, 1.5])
def test_l2_i8_fold_ratio_sweep_with_ties(eng, ratio):
    """Folded-norm kernel + bounds/refine/exhaustive decision ladder of finalize: exact for any ratio, with exact
    duplicates (d = 0), chunk-level ties and odd/even norms mixed."""
    rng = np.random.default_rng(int(ratio * 100))
    a = rng.integers(0, 4, size=(700, 128)).astype(np.float32)
    c = np.concatenate([a[:150], a[100:250], rng.integers(0, 4, size=(1500, 128)).astype(np.float32)])
    assert _fold_flag(eng.prepare(c, b.NORM_L2)) == 1
    assert_exact(eng.match_arrays(a, c, b.NORM_L2, False, ratio), O.reference_match(a, c, O.NORM_L2, False, ratio),
                 f"fold ties ratio={ratio}")
    a, c = synth.sift_pair(seed=31, n=1500, m=1300)
    assert_exact(eng.match_arrays(a, c, b.NORM_L2, False, ratio), O.reference_match(a, c, O.NORM_L2, False, ratio),
                 f"fold sift ratio={ratio}")


def test_l2_i8_fold_disabled_when_norm_spread_too_large(eng):
    """Rows of all 0 (norm 2^21) next to rows of all 128 (norm 0): h spread > digit range -> keyed kernel, same answers."""
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, size=(400, 128)).astype(np.float32)
    c = rng.integers(0, 256, size=(600, 128)).astype(np.float32)
    c[:20] = 0.0
    c[20:40] = 128.0
    assert _fold_flag(eng.prepare(c, b.NORM_L2)) == 0
    for mode in MODES:
        assert_exact(eng.match_arrays(a, c, b.NORM_L2, *mode), O.reference_match(a, c, O.NORM_L2, *mode), f"no-fold {mode}")


def test_l2_i8_fold_small_and_ragged(eng):
    for seed, (n, m) in enumerate([(1, 1), (1, 2), (2, 16), (2, 17), (4, 33), (3, 64), (5, 65), (129, 257), (40, 1000), (700, 255), (256, 512)]):
        a, c = synth.sift_pair(seed=400 + seed, n=max(n, m))
        a, c = a[:n], np.ascontiguousarray(c[:m])
        for ratio in (0.75, 0.95):
            assert_exact(eng.match_arrays(a, c, b.NORM_L2, False, ratio), O.reference_match(a, c, O.NORM_L2, False, ratio),
                         f"fold n={n} m={m} ratio={ratio}")



@pytest.mark.parametrize("D,m", [(128, 20000), (256, 9500)])
def test_l2_i8_fold_long_rows_cross_key_windows(eng, D, m):
    """The fold kernel keeps its running top-2 as packed (value, chunk) words inside a window of 8192 (D = 128) or 4096
    (D = 256) train columns and unpacks at window ends (l2_i8.cu): one column split longer than several windows, best and
    second-best neighbours planted in different windows, exact duplicates across windows (ties -> lowest column)."""
    rng = np.random.default_rng(D + m)
    n = 300                                                   # 3 row tiles of one pair: a single column split
    def rows(k):           # D = 256: values away from 128, or the rows' norms fall below the range the fold digits cover
        x = rng.integers(0, 256 if D == 128 else 48, size=(k, D))
        return np.where(rng.integers(0, 2, size=(k, D)) == 1, 255 - x, x).astype(np.float32) if D == 256 else x.astype(np.float32)
    a, c = rows(n), rows(m)
    win = 8192 if D == 128 else 4096
    for i in range(0, n, 3):                                  # near copies of query rows, one per window, increasingly far
        for w, col in enumerate(range(17 + 5 * i, m, win)):
            c[col] = np.clip(a[i] + rng.integers(-w - 1, w + 2, size=D), 0, 255)
    for i in range(1, n, 7):                                  # exact duplicates of one another in the first and the last window
        c[40 + i] = a[i]
        c[m - 40 - i] = a[i]
    assert _fold_flag(eng.prepare(c, b.NORM_L2)) == 1
    for ratio in (0.75, 1.5):
        assert_exact(eng.match_arrays(a, c, b.NORM_L2, False, ratio), O.reference_match(a, c, O.NORM_L2, False, ratio),
                     f"fold windows D={D} m={m} ratio={ratio}")


def test_l2_i8_knn2_raw(eng):
    a, c = synth.sift_pair(seed=6, n=900, m=800)
    got = eng.knn2_arrays(a, c, b.NORM_L2)
    want = O.knn2_c(a, c, O.NORM_L2)
    assert np.array_equal(got.idx1, want.idx1) and np.array_equal(got.idx2, want.idx2)
    assert np.array_equal(got.d1.view(np.uint32), want.d1.view(np.uint32))
    assert np.array_equal(got.d2.view(np.uint32), want.d2.view(np.uint32))


def test_l2_i8_config1_full_size(eng):
    """BASELINE config 1: SIFT 8192 x 8192 x 128, ratio 0.75."""
    a, c = synth.sift_pair(seed=0, n=8192)
    got = eng.match_arrays(a, c, b.NORM_L2, False, 0.75)
    assert_exact(got, O.reference_match(a, c, O.NORM_L2, False, 0.75), "config1")
    assert 3500 < len(got.query_idx) < 4700
    assert np.all(np.diff(got.distance) >= 0)


def test_l2_uint8_descriptors(eng):
    """cv2 accepts CV_8U with NORM_L2 (integer sum, float sqrt)."""
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, size=(300, 128), dtype=np.uint8)
    c = rng.integers(0, 256, size=(200, 128), dtype=np.uint8)
    got = eng.match_arrays(a, c, b.NORM_L2, False, None)
    want = O.reference_match(a.astype(np.float32), c.astype(np.float32), O.NORM_L2, False, None)
    assert_exact(got, want, "u8 L2")


# ------------------------------------------------------------------ non-integer float L2 (SuperPoint) — TF32 + fp32 re-rank
@pytest.mark.parametrize("mode", [(True, None), (False, 0.8), (False, None)])
def test_l2_f32_vs_oracle(feng, mode):
    s = synth.superpoint_set(seed=21, images=2, n=1500, d=256, pool=2500)
    flips = assert_close_float(feng.match_arrays(s[0], s[1], b.NORM_L2, *mode),
                               O.reference_match(s[0], s[1], O.NORM_L2, *mode), f"f32 {mode}", descs=(s[0], s[1]), mode=mode)
    print(f"f32 mode={mode}: near-tie flips={flips}")
    rng = np.random.default_rng(4)
    a = rng.normal(size=(333, 100)).astype(np.float32) * 3.0          # not unit norm, D not a multiple of 32
    c = (a[rng.permutation(333)[:290]] + rng.normal(size=(290, 100)).astype(np.float32) * 0.05).astype(np.float32)
    assert_close_float(feng.match_arrays(a, c, b.NORM_L2, *mode), O.reference_match(a, c, O.NORM_L2, *mode), f"f32 generic {mode}",
                       descs=(a, c), mode=mode)


def assert_knn2_close(got, want, what=""):
    """knn2 on non-integer floats: distances within RTOL; an index may differ only where the oracle's own
    neighbouring distances are within RTOL of each other."""
    n = len(want.idx1)
    for gi, gd, wi, wd in ((got.idx1, got.d1, want.idx1, want.d1), (got.idx2, got.d2, want.idx2, want.d2)):
        assert np.allclose(gd, wd, rtol=RTOL, atol=1e-7), what
        diff = np.nonzero(gi != wi)[0]
        assert len(diff) <= max(1, n // 100), f"{what}: {len(diff)} index differences"
    return


def test_l2_f32_knn2_raw(feng):
    s = synth.superpoint_set(seed=22, images=2, n=1200, d=256, pool=2000)
    assert_knn2_close(feng.knn2_arrays(s[0], s[1], b.NORM_L2), O.knn2_c(s[0], s[1], O.NORM_L2), "f32 knn2")
    got = feng.knn2_arrays(s[0], s[1][:1], b.NORM_L2)
    assert np.all(got.idx2 == -1) and np.all(got.idx1 == 0)
    d128 = synth.superpoint_set(seed=23, images=2, n=700, d=128, pool=1000)      # DISK / ALIKED width
    assert_knn2_close(feng.knn2_arrays(d128[0], d128[1], b.NORM_L2), O.knn2_c(d128[0], d128[1], O.NORM_L2), "f32 d128")


def test_l2_f32_dense_cluster_takes_exhaustive_fallback(feng):
    """Many train rows within the TF32 error window of each other: the 8-candidate list cannot be proven
    sufficient, so the rows are re-scanned exhaustively in fp32 — results must still match the oracle."""
    rng = np.random.default_rng(11)
    base = rng.normal(size=(1, 64)).astype(np.float32)
    base /= np.linalg.norm(base)
    c = (base + 1e-4 * rng.normal(size=(400, 64))).astype(np.float32)          # 400 near-duplicates
    a = (base + 1e-4 * rng.normal(size=(200, 64))).astype(np.float32)
    feng.f32_stats(reset=True)
    got = feng.knn2_arrays(a, c, b.NORM_L2)
    rows, evals = feng.f32_stats()
    assert rows == 200 and evals == 200 * 400
    want = O.knn2_c(a, c, O.NORM_L2)
    # distances ~1e-3 with gaps ~1e-6: compare distances, and indices where the oracle's gap is resolvable in fp32
    assert np.allclose(got.d1, want.d1, rtol=1e-4) and np.allclose(got.d2, want.d2, rtol=1e-4)
    clear = (want.d2 - want.d1) > 1e-3 * want.d1
    assert np.array_equal(got.idx1[clear], want.idx1[clear])
    # well separated data never needs the fallback
    s = synth.superpoint_set(seed=24, images=2, n=2000, d=256, pool=3000)
    feng.f32_stats(reset=True)
    feng.match_arrays(s[0], s[1], b.NORM_L2, True, None)
    assert feng.f32_stats()[0] == 0


def test_l2_f16c_kind_is_default_and_identical_to_tf32():
    """The fp16 copy is round-to-nearest of src * s with s the power of two that brings max |x| into (1/2, 1]; the two
    candidate kinds return the same bits (both finish with the exact fp32 re-rank), also for rows whose elements span
    the fp16 subnormal range and for large magnitudes."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    L = pkg("_lib")
    e16, e32 = b.Engine(float_path="f16"), b.Engine(float_path="tf32")
    rng = np.random.default_rng(31)
    a = rng.normal(size=(500, 96)).astype(np.float32) * 37.0
    dd = e16.prepare(a, b.NORM_L2)
    assert dd.kind == L.KIND_L2_F16C and e32.prepare(a, b.NORM_L2).kind == L.KIND_L2_F32
    aux = dd.norm.cpu().numpy()
    half_off, Dh = (500 * 4 + 255) // 256 * 256, 128
    info_off = half_off + (500 * Dh * 2 + 255) // 256 * 256
    s = aux[info_off // 4]
    mx = np.abs(a).max()
    assert 0.5 < mx * s <= 1.0 and np.log2(s) == np.round(np.log2(s)) and aux[info_off // 4 + 1] == 1.0 / s
    assert np.allclose(aux[:500], (a.astype(np.float64) ** 2).sum(1), rtol=1e-5)
    h = aux.view(np.uint8)[half_off:half_off + 500 * Dh * 2].view(np.float16).reshape(500, Dh)
    assert np.array_equal(h[:, :96], (a * s).astype(np.float16)) and not h[:, 96:].any()
    # folded-norm rows behind the info words: three fp16 terms that sum to -(s^2 / 2) ||b||^2 (the fp32 value, exactly)
    x_off = info_off + 256
    x = aux.view(np.uint8)[x_off:x_off + 500 * 32].view(np.float16).reshape(500, 16).astype(np.float64)
    assert not x[:, 3:].any()
    assert np.array_equal(x[:, :3].sum(1).astype(np.float32), (np.float32(-0.5) * s * s * aux[:500]).astype(np.float32))
    cases = {"generic": (a, (a[rng.permutation(500)[:420]] + rng.normal(size=(420, 96)) * 0.5).astype(np.float32))}
    wide = (rng.normal(size=(400, 256)) * np.exp(rng.uniform(-14, 0, size=(1, 256)))).astype(np.float32)       # columns down to 1e-6
    cases["wide_range"] = (wide, (wide[rng.permutation(400)[:350]] * (1 + 0.01 * rng.normal(size=(350, 256)))).astype(np.float32))
    big = (rng.normal(size=(300, 128)) * 1e4).astype(np.float32)
    cases["large"] = (big, (big[rng.permutation(300)] + rng.normal(size=(300, 128)) * 50).astype(np.float32))
    sp = synth.superpoint_set(seed=33, images=2, n=1800, d=256, pool=3000)
    cases["superpoint"] = (sp[0], sp[1])
    # two stores with different scales: s_q / s_t = 2^-5 rides in the folded-norm step; 2^-14 is outside its gate (plain kernel)
    cases["scale_ratio_32"] = (sp[0][:900], (sp[1][:1000] * 37.0).astype(np.float32))
    cases["scale_ratio_gate"] = ((sp[0][:700] * 3.0e-4).astype(np.float32), (sp[1][:800] * 4.0).astype(np.float32))
    d192 = synth.superpoint_set(seed=34, images=2, n=1100, d=192, pool=1700)
    cases["d192"] = (d192[0], d192[1])
    for name, (x, y) in cases.items():
        for mode in [(False, 0.8), (True, None), (False, None)]:
            g16, g32 = e16.match_arrays(x, y, b.NORM_L2, *mode), e32.match_arrays(x, y, b.NORM_L2, *mode)
            assert_exact(g16, g32, f"f16 vs tf32 {name} {mode}")
            assert_close_float(g16, O.reference_match(x, y, O.NORM_L2, *mode), f"f16 {name} {mode}", descs=(x, y), mode=mode)
        k16, k32 = e16.knn2_arrays(x, y, b.NORM_L2), e32.knn2_arrays(x, y, b.NORM_L2)
        for u, v in zip(k16, k32):
            assert np.array_equal(u, v), name


def test_l2_f32_config3_shape(eng):
    """BASELINE config 3 shape: SuperPoint 256-d unit-norm, 4096 keypoints, mutual NN (one pair of the 4950)."""
    s = synth.superpoint_set(seed=2, images=2, n=4096, d=256)
    got = eng.match_arrays(s[0], s[1], b.NORM_L2, True, None)
    want = O.reference_match(s[0], s[1], O.NORM_L2, True, None)
    flips = assert_close_float(got, want, "config3 pair", max_flip_frac=0.002, descs=(s[0], s[1]), mode=(True, None))
    assert len(got.query_idx) > 1000 and np.all(np.diff(got.distance) >= 0)
    print("config3 pair: matches", len(got.query_idx), "near-tie flips", flips)


# ------------------------------------------------------------------ plugin level
def test_plugin_match_returns_matchdata(eng):
    a, c = synth.binary_pair(seed=1, n=500, nbytes=32)
    F = b.contracts.FeatureData
    f1 = F(keypoints=list(range(len(a))), descriptors=a, method="ORB")
    f2 = F(keypoints=list(range(len(c))), descriptors=c, method="ORB")
    m = b.B200BFMatcher(norm_type=b.NORM_HAMMING, cross_check=False, ratio_threshold=0.7, engine=eng)
    md = m.match(f1, f2)
    want = O.reference_match(a, c, O.NORM_HAMMING, False, 0.7)
    assert md.method == "BF" and md.score_type == b.contracts.ScoreType.DISTANCE and isinstance(md.matches, list)
    assert [x.queryIdx for x in md.matches] == want.query_idx.tolist()
    assert [x.trainIdx for x in md.matches] == want.train_idx.tolist()
    assert [x.score for x in md.matches] == [float(v) for v in want.distance]
    assert md.matching_time > 0
    # error convention: width mismatch -> printed, empty MatchData (feature_matchers.py:250-252)
    bad = F(keypoints=list(range(4)), descriptors=np.zeros((4, 16), np.uint8), method="ORB")
    assert m.match(f1, bad).matches == []
    # downstream helpers keep working on our MatchData
    assert len(md.filter_by_score(40.0).filtered_matches) <= len(md.matches)


# ------------------------------------------------------------------ all-pairs CSR driver (single GPU)
@pytest.mark.parametrize("kind", ["sift", "orb", "akaze61", "superpoint"])
@pytest.mark.parametrize("mode", [(False, 0.75), (True, None)])
def test_all_pairs_csr_equals_per_pair_oracle(eng, kind, mode):
    ap = pkg("allpairs")
    sizes = [300, 129, 1, 640, 77, 256]
    if kind == "sift":
        imgs = [synth.sift_pair(seed=60 + i, n=max(s, 2))[0][:s] for i, s in enumerate(sizes)]
        # plant overlap so the ratio test has survivors
        imgs[3][:200] = np.clip(imgs[0][:200] + np.random.default_rng(1).integers(-4, 5, size=(200, 128)), 0, 255)
        norm = b.NORM_L2
    elif kind == "superpoint":
        full = synth.superpoint_set(seed=80, images=len(sizes), n=max(sizes), d=256, pool=900, shared=0.6)
        imgs = [np.ascontiguousarray(full[i][:s]) for i, s in enumerate(sizes)]
        norm = b.NORM_L2
    else:
        nb = 32 if kind == "orb" else 61
        imgs = [synth.binary_pair(seed=70 + i, n=max(s, 2), nbytes=nb)[0][:s] for i, s in enumerate(sizes)]
        imgs[3][:200] = imgs[0][:200]
        imgs[3][:200, 0] ^= 1
        norm = b.NORM_HAMMING
    store = ap.build_store(eng, imgs, norm)
    pairs = ap.all_pairs(len(imgs))
    for chunk in (4, 1000):                       # multi-chunk and single-chunk code paths
        res = ap.AllPairsMatcher(eng, chunk_pairs=chunk).match(store, pairs, *mode).to_host()
        assert res.ptr[0] == 0 and len(res.ptr) == len(pairs) + 1 and res.ptr[-1] == len(res.query_idx)
        for p, (i, j) in enumerate(pairs):
            q, t, d = res.pair(p)
            want = O.reference_match(imgs[i], imgs[j], norm, *mode)
            if kind == "superpoint":
                assert_close_float(O.Matches(q, t, d), want, f"{kind} pair {i},{j} chunk={chunk} mode={mode}", 0.02,
                                   descs=(imgs[i], imgs[j]), mode=mode)
            else:
                assert_exact(O.Matches(q, t, d), want, f"{kind} pair {i},{j} chunk={chunk} mode={mode}")


@pytest.mark.parametrize("group", [1, 2, 4, 100])
def test_streaming_all_pairs_equals_per_pair_oracle(eng, group):
    """Upload-overlapped all-pairs path (StreamingAllPairs): same pair set as itertools.combinations, every pair exact;
    ragged image sizes, groups smaller/larger than the set, chunks that straddle groups; twice (arena reuse)."""
    import itertools
    ap = pkg("allpairs")
    sizes = [300, 129, 1, 640, 77, 256, 18]
    imgs = [synth.sift_pair(seed=160 + i, n=max(s, 2))[0][:s] for i, s in enumerate(sizes)]
    imgs[3][:200] = np.clip(imgs[0][:200] + np.random.default_rng(1).integers(-4, 5, size=(200, 128)), 0, 255)
    imgs[5][:60] = imgs[1][:60]
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    sm = ap.StreamingAllPairs(eng, chunk_pairs=5, group_images=group)
    for rep in range(2):
        res = sm.match_host(host, sizes, ratio_threshold=0.8)
        assert sorted(map(tuple, res.pairs)) == list(itertools.combinations(range(len(sizes)), 2))
        assert res.ptr[0] == 0 and res.ptr[-1] == len(res.query_idx)
        for p, (i, j) in enumerate(res.pairs):
            want = O.reference_match(imgs[i], imgs[j], O.NORM_L2, False, 0.8)
            assert_exact(O.Matches(*res.pair(p)), want, f"streaming pair {i},{j} group={group} rep={rep}")


def test_streaming_all_pairs_non_integer_falls_back_to_generic_path(eng):
    ap = pkg("allpairs")
    rng = np.random.default_rng(11)
    sizes = [90, 40, 64]
    imgs = [rng.normal(size=(s, 128)).astype(np.float32) for s in sizes]
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    res = ap.StreamingAllPairs(eng, chunk_pairs=2, group_images=2).match_host(host, sizes, ratio_threshold=0.9)
    for p, (i, j) in enumerate(res.pairs):
        want = O.reference_match(imgs[i], imgs[j], O.NORM_L2, False, 0.9)
        assert_close_float(O.Matches(*res.pair(p)), want, f"streaming fallback pair {i},{j}", 0.05)


def test_streaming_late_low_norm_row_switches_kernels_consistently(eng):
    """a row whose norm falls outside the fold digits' range arrives only with the LAST image group, while
    chunks of earlier groups are in flight.  Every chunk must act on one snapshot of the store's flag (capi.cu): the
    results stay exact whichever kernel each chunk ran."""
    ap = pkg("allpairs")
    sizes = [200, 150, 128, 90, 260, 64]
    imgs = [synth.sift_pair(seed=190 + i, n=max(s, 2))[0][:s] for i, s in enumerate(sizes)]
    imgs[-1][5] = 128.0                         # x - 128 = 0 everywhere: norm 0, far below the digit range
    imgs[-1][6] = 0.0                           # and the other extreme (every element -128)
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    for group, chunk in [(1, 2), (2, 3), (5, 100)]:
        res = ap.StreamingAllPairs(eng, chunk_pairs=chunk, group_images=group).match_host(host, sizes, ratio_threshold=0.85)
        for p, (i, j) in enumerate(res.pairs):
            want = O.reference_match(imgs[i], imgs[j], O.NORM_L2, False, 0.85)
            assert_exact(O.Matches(*res.pair(p)), want, f"late low-norm row: pair {i},{j} group={group} chunk={chunk}")


@pytest.mark.parametrize("kind,fmt", [("orb", None), ("akaze61", None), ("orb", 1), ("sift", 1), ("sift", 2), ("superpoint", None)])
def test_end_to_end_host_path_all_kinds_and_record_formats(eng, kind, fmt):
    """distributed.DistributedAllPairs on one GPU: host rows in, packed records out (6 / 8 / 12 bytes per match), every pair
    against the oracle; with a consumer and keep=False (the streamed mode of the 1000-image workload) the blocks handed
    over are the same records."""
    import itertools
    D = pkg("distributed")
    sizes = [300, 129, 1, 640, 77, 256, 18]
    if kind == "sift":
        imgs = [synth.sift_pair(seed=160 + i, n=max(s, 2))[0][:s] for i, s in enumerate(sizes)]
        norm, mode = b.NORM_L2, (False, 0.8)
    elif kind == "superpoint":
        full = synth.superpoint_set(seed=80, images=len(sizes), n=max(sizes), d=256, pool=900, shared=0.6)
        imgs = [np.ascontiguousarray(full[i][:s]) for i, s in enumerate(sizes)]
        norm, mode = b.NORM_L2, (True, None)
    else:
        nb = 32 if kind == "orb" else 61
        imgs = [synth.binary_pair(seed=70 + i, n=max(s, 2), nbytes=nb)[0][:s] for i, s in enumerate(sizes)]
        imgs[3][:200] = imgs[0][:200]
        norm, mode = b.NORM_HAMMING, (True, None)
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    dap = D.DistributedAllPairs(eng, chunk_pairs=4, group_images=3, rec_format=fmt)
    res = dap.match_host(host, sizes, norm, *mode, timing=True)
    want_fmt = fmt or (3 if norm == b.NORM_HAMMING else 2)
    assert res.rec_format == want_fmt and res.rec.dtype.itemsize == {1: 12, 2: 8, 3: 6}[want_fmt]
    assert sorted(map(tuple, res.pairs)) == list(itertools.combinations(range(len(sizes)), 2))
    assert dap.last_phases["total_ms"] > 0 and dap.last_phases["h2d_ms"] > 0
    onorm = O.NORM_HAMMING if norm == b.NORM_HAMMING else O.NORM_L2
    for p, (i, j) in enumerate(res.pairs):
        want = O.reference_match(imgs[i], imgs[j], onorm, *mode)
        if kind == "superpoint":
            assert_close_float(O.Matches(*res.pair(p)), want, f"e2e {kind} pair {i},{j}", 0.02)
        else:
            assert_exact(O.Matches(*res.pair(p)), want, f"e2e {kind} fmt={fmt} pair {i},{j}")
    got = []
    r2 = dap.match_host(host, sizes, norm, *mode, keep=False,
                        consumer=lambda pr, cn, rec: got.append((pr.copy(), cn.copy(), rec.copy())))
    assert len(r2.rec) == 0 and np.array_equal(r2.ptr, res.ptr) and np.array_equal(r2.pairs, res.pairs)
    assert np.array_equal(np.concatenate([g[0] for g in got]), res.pairs)
    assert np.array_equal(np.concatenate([g[1] for g in got]), np.diff(res.ptr))
    assert np.array_equal(np.concatenate([g[2] for g in got]), res.rec)
    s = res.sorted_by_pair()
    assert [tuple(x) for x in s.pairs] == list(itertools.combinations(range(len(sizes)), 2)) and s.ptr[-1] == res.ptr[-1]


def test_checkpointed_job_interrupted_and_resumed_equals_uninterrupted(eng, tmp_path):
    """an all-pairs job that dies after some blocks and is started again (a new AllPairsCheckpoint on the
    same directory) matches only the unfinished pairs and ends with the same result as one uninterrupted job -- with the
    single-GPU block driver and with the streamed host path (consumer + done_mask); progress.json lists every pair."""
    import itertools
    ap, ckm, D = pkg("allpairs"), pkg("checkpoint"), pkg("distributed")
    sizes = [300, 129, 64, 640, 77, 256, 18]
    I = len(sizes)
    imgs = [synth.binary_pair(seed=370 + i, n=max(s, 2), nbytes=32)[0][:s] for i, s in enumerate(sizes)]
    imgs[3][:200] = imgs[0][:200]
    combos = list(itertools.combinations(range(I), 2))
    want = [O.reference_match(imgs[i], imgs[j], O.NORM_HAMMING, True, None) for i, j in combos]

    def check(c):
        pairs, ptr, rec = c.load()
        assert [tuple(p) for p in pairs] == combos and c.done_mask().all()
        assert sorted(c.progress.completed_pairs) == sorted(ckm.pair_id(p) for p in range(len(combos)))
        for p, w in enumerate(want):
            r = rec[ptr[p]:ptr[p + 1]]
            got = O.Matches(r["q"].astype(np.int32), r["t"].astype(np.int32), r["d"].astype(np.float32))
            assert_exact(got, w, f"checkpointed pair {combos[p]}")

    # (1) block driver over a resident store
    store = ap.build_store(eng, imgs, b.NORM_HAMMING)
    m = ap.AllPairsMatcher(eng, chunk_pairs=3)
    c = ckm.AllPairsCheckpoint(tmp_path / "blocks", I)
    assert ckm.run_checkpointed(m, store, I, c, cross_check=True, block_pairs=5, stop_after_blocks=2) == 10     # "dies" after 2 blocks
    c = ckm.AllPairsCheckpoint(tmp_path / "blocks", I)
    assert c.done_mask().sum() == 10
    assert ckm.run_checkpointed(m, store, I, c, cross_check=True, block_pairs=5) == len(combos) - 10          # the rest only
    assert ckm.run_checkpointed(m, store, I, c, cross_check=True, block_pairs=5) == 0
    check(ckm.AllPairsCheckpoint(tmp_path / "blocks", I))
    # (2) streamed host path: the consumer raises in the third block
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    dap = D.DistributedAllPairs(eng, chunk_pairs=4, group_images=3)
    c = ckm.AllPairsCheckpoint(tmp_path / "stream", I)
    seen = []

    def dying(pr, cn, rec):
        if len(seen) == 2:
            raise KeyboardInterrupt("power cut")
        seen.append(len(pr))
        c.consumer(pr, cn, rec)

    with pytest.raises(KeyboardInterrupt):
        dap.match_host(host, sizes, b.NORM_HAMMING, True, None, consumer=dying, keep=False)
    torch.cuda.synchronize()
    c = ckm.AllPairsCheckpoint(tmp_path / "stream", I)
    assert c.done_mask().sum() == sum(seen) and 0 < sum(seen) < len(combos)
    r = D.DistributedAllPairs(eng, chunk_pairs=4, group_images=3).match_host(host, sizes, b.NORM_HAMMING, True, None,
                                                                             consumer=c.consumer, keep=False, done_mask=c.done_mask())
    assert len(r.pairs) == len(combos) - sum(seen)
    check(ckm.AllPairsCheckpoint(tmp_path / "stream", I))


def test_plugin_raises_for_what_cv2_would_match_but_the_library_does_not(eng):
    """shapes the kernels do not implement must not become an empty MatchData (the reference's cv2 path returns
    real matches for them); genuine input errors still do (feature_matchers.py:250-252)."""
    F = b.contracts.FeatureData
    rng = np.random.default_rng(3)
    wide = rng.integers(0, 256, size=(40, 300)).astype(np.float32)
    f = F(keypoints=list(range(40)), descriptors=wide, method="x")
    m = b.B200BFMatcher(norm_type=b.NORM_L2, ratio_threshold=0.8, engine=eng)
    with pytest.raises(pkg("_lib").B200MatchError):
        m.match(f, f)
    u8 = rng.integers(0, 256, size=(40, 32), dtype=np.uint8)
    fu = F(keypoints=list(range(40)), descriptors=u8, method="x")
    with pytest.raises(pkg("engine").UnsupportedNormError):
        b.B200BFMatcher(norm_type=2, engine=eng).match(fu, fu)            # cv2.NORM_L1
    narrow = F(keypoints=list(range(40)), descriptors=u8[:, :16], method="x")
    md = b.B200BFMatcher(norm_type=b.NORM_HAMMING, engine=eng).match(fu, narrow)   # width mismatch: cv2 fails too -> empty
    assert md.matches == [] and md.method == "BF"


def test_descriptor_cache_serves_only_live_unchanged_arrays():
    """Engine's host-array -> prepared-store cache (the matcher is reused over all pairs of a run, pipeline.py:99-104): a hit
    needs the same live array object and an unchanged fingerprint; results never depend on whether the cache was used."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    e = b.Engine(cache_entries=4)
    a, c = synth.sift_pair(seed=41, n=700)
    want = O.reference_match(a, c, O.NORM_L2, False, 0.8)
    assert_exact(e.match_arrays(a, c, b.NORM_L2, False, 0.8), want, "cold")
    assert (e.cache.hits, e.cache.misses) == (0, 2)
    assert_exact(e.match_arrays(a, c, b.NORM_L2, False, 0.8), want, "cached")
    assert e.cache.hits == 2
    assert_exact(e.match_arrays(c, a, b.NORM_L2, True, None), O.reference_match(c, a, O.NORM_L2, True, None), "cached, swapped, other mode")
    assert e.cache.hits == 4
    a[0] = np.clip(a[0] + 9, 0, 255)                              # in-place edit of a sampled row: the entry must not be served
    assert_exact(e.match_arrays(a, c, b.NORM_L2, False, 0.8), O.reference_match(a, c, O.NORM_L2, False, 0.8), "after an edit")
    assert e.cache.misses == 3
    x, y = synth.binary_pair(seed=42, n=600, nbytes=32)
    assert_exact(e.match_arrays(x, y, b.NORM_HAMMING, True, None), O.reference_match(x, y, O.NORM_HAMMING, True, None), "orb cold")
    x2 = x.copy()
    del x                                                         # the array dies: a new array (even at the same address) is a miss
    h = e.cache.hits
    assert_exact(e.match_arrays(x2, y, b.NORM_HAMMING, True, None), O.reference_match(x2, y, O.NORM_HAMMING, True, None), "orb new object")
    assert e.cache.hits == h + 1                                  # y only
    assert len(e.cache.d) <= 4                                    # bounded (least recently used evicted)
    f = synth.superpoint_set(seed=43, images=2, n=512, d=256, pool=800)
    g1 = e.match_arrays(f[0], f[1], b.NORM_L2, True, None)
    g2 = e.match_arrays(f[0], f[1], b.NORM_L2, True, None)        # float stores from the cache: same bits
    assert_exact(g2, g1, "float cached vs cold")
    mixed = e.match_arrays(a, f[0][:, :128].copy(), b.NORM_L2, False, None)     # one side integer-valued, the other not
    assert len(mixed.query_idx) == len(a)
    e0 = b.Engine(cache_entries=0)
    assert_exact(e0.match_arrays(a, c, b.NORM_L2, False, 0.8), O.reference_match(a, c, O.NORM_L2, False, 0.8), "cache off")
    assert len(e0.cache.d) == 0


def test_late_non_integer_element_reroutes_the_job_and_the_consumer_gets_the_right_answer(eng, tmp_path):
    """Host rows that pass the integer sample of match_host (64 evenly spaced rows) but hold one non-integer element in a
    late image: the int8 kernels flag it, nothing of the int8 run reaches the consumer after that, the job is redone on the
    float path, and what the consumer ends up with is what cv2 gives for these rows."""
    import itertools
    D, ckm = pkg("distributed"), pkg("checkpoint")
    sizes = [300, 200, 256, 180, 220, 150, 260, 190, 240, 210]
    imgs = [synth.sift_pair(seed=460 + i, n=s)[0] for i, s in enumerate(sizes)]
    imgs[4][:100] = imgs[0][:100]
    imgs[9][77, 5] += 0.5                                          # row 2073 of 2206: not one of the sampled rows
    rows = sum(sizes)
    assert 2073 not in torch.linspace(0, rows - 1, 64).long().tolist()
    host = torch.from_numpy(np.concatenate(imgs)).pin_memory()
    ck = ckm.AllPairsCheckpoint(tmp_path, len(sizes))
    calls = []
    r = D.DistributedAllPairs(eng, chunk_pairs=4, group_images=3).match_host(
        host, sizes, b.NORM_L2, False, 0.8, keep=False, consumer=lambda p, c, rec: (calls.append(len(p)), ck.consumer(p, c, rec)))
    combos = list(itertools.combinations(range(len(sizes)), 2))
    assert len(r.pairs) == len(combos)
    pairs, ptr, rec = ck.load()
    assert [tuple(p) for p in pairs] == combos
    for p, (i, j) in enumerate(combos):
        want = O.reference_match(imgs[i], imgs[j], O.NORM_L2, False, 0.8)
        got = O.Matches(rec["q"][ptr[p]:ptr[p + 1]].astype(np.int32), rec["t"][ptr[p]:ptr[p + 1]].astype(np.int32),
                        rec["d"][ptr[p]:ptr[p + 1]].astype(np.float32))
        assert_close_float(got, want, f"rerouted pair {i},{j}", 0.05, descs=(imgs[i], imgs[j]), mode=(False, 0.8))


# ------------------------------------------------------------------ SfM hand-off (feature cache -> all pairs -> correspondences)
@pytest.mark.parametrize("name", ["orb_cross", "orb_ratio07", "sift_ratio075"])
def test_sfm_handoff_equals_reference_converter(eng, name):
    """build_feature_set + match_feature_set (device gather of keypoint coordinates) against what the reference's
    EnhancedBFMatcher + ResultConverter._convert_method produced for every pair (tests/golden/sfm_golden.npz)."""
    from conftest import sfm_golden_sets
    sfm = pkg("sfm")
    g = sfm_golden_sets()[name]
    offs = np.concatenate([[0], np.cumsum(g["count"])])
    F = b.contracts.FeatureData
    feats = {f"img_{i}": F(keypoints=g["xy"][offs[i]:offs[i + 1]], descriptors=g["desc"][offs[i]:offs[i + 1]], method=name)
             for i in range(len(g["count"]))}
    fs = sfm.build_feature_set(eng, feats, g["norm"], {k: (1600, 1200) for k in feats})
    r = sfm.match_feature_set(eng, fs, None, g["cross_check"], g["ratio"])
    assert len(r) == len(g["pairs"])
    for p, (i, j) in enumerate(r.pairs):
        want, got = g["pairs"][(int(i), int(j))], r.pair_dict(p)
        assert np.array_equal(got["query_indices"], want["q"]) and np.array_equal(got["train_indices"], want["t"]), (i, j)
        assert np.array_equal(got["match_scores"], want["s"]) and np.array_equal(got["correspondences"], want["c"]), (i, j)
    blob = r.to_matches_pickle()
    assert all(v["num_matches"] == len(v["correspondences"]) for v in blob["matches_data"].values())
