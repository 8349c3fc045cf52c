"""tools/popc_probe.py — a short program that launches the INT-pipe Hamming kernel (hamming_top2_kernel, csrc/hamming.cu) on
(1) BASELINE configs[1]: one ORB pair 10000 x 10000 x 32 B with cross-check (single fused pass: row top-2 + column minima),
(2) an AKAZE block: all 15 pairs of 6 images x 16384 x 61 B (padded to 64), mutual check,
for an `ncu --set full -k regex:hamming_top2` capture (profiles/r02_hamming_popc_ncu_full.md).  Prints its own CUDA-event
timings and the POPC-roofline fractions."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b = importlib.import_module("3dreconstruction_b200")
ap = importlib.import_module("3dreconstruction_b200.allpairs")
L = b._lib
eng = b.Engine("cuda:0", hamming_path="popc")


def timed(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


a, c = b.synth.binary_pair(seed=1, n=10000, nbytes=32)
da, dc = eng.prepare(a, b.NORM_HAMMING), eng.prepare(c, b.NORM_HAMMING)
ms = timed(lambda: eng.match_device(da, dc, [(0, 10000, 0, 10000)], L.MODE_CROSS), 5)
peak256 = 148 * 16 * 1965e6 / 8 / 1e9
print(f"orb 10k x 10k cross (whole call): {ms:.3f} ms = {0.1 / (ms / 1e3):.1f} Gdist/s ({0.1 / (ms / 1e3) / peak256:.2f} of the POPC roofline {peak256:.0f})")
I, n = 6, 16384
imgs = b.synth.binary_set(seed=4, images=I, n=n, nbytes=61)
store = ap.build_store(eng, [np.ascontiguousarray(x) for x in imgs], b.NORM_HAMMING)
pairs = ap.all_pairs(I)
m = ap.AllPairsMatcher(eng, chunk_pairs=64, pipeline=False)
ms = timed(lambda: m.match(store, pairs, cross_check=True), 3)
gd = len(pairs) * n * n / 1e9 / (ms / 1e3)
peak512 = peak256 / 2
print(f"akaze 6 x 16384 x 64 B, 15 pairs, mutual (whole pass): {ms:.3f} ms = {gd:.1f} Gdist/s ({gd / peak512:.2f} of the POPC roofline {peak512:.0f})")
