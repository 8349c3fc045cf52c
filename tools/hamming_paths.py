"""Time the two Hamming knn paths side by side on one GPU: the INT-pipe POPC kernel (hamming.cu) and the tensor-core
kernel over +-1 expanded bits (hamming_tc.cu).  Prints one JSON line per case.

    python tools/hamming_paths.py [--images 32] [--reps 3]
"""
import argparse
import ctypes
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b = importlib.import_module("3dreconstruction_b200")
ap = importlib.import_module("3dreconstruction_b200.allpairs")
L = b._lib


def timed(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def knn_ms(lib, fn, reps):
    lib.b200m_timing_enable(1)
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    ms, n = ctypes.c_double(0), ctypes.c_int(0)
    lib.b200m_timing_collect(ctypes.byref(ms), ctypes.byref(n))
    lib.b200m_timing_enable(0)
    return ms.value / reps


def main():
    pa = argparse.ArgumentParser()
    pa.add_argument("--images", type=int, default=32)
    pa.add_argument("--reps", type=int, default=3)
    pa.add_argument("--paths", default="popc,int8,fp4")
    pa.add_argument("--cases", default="pair,orb,akaze")
    args = pa.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev); g.manual_seed(0)
    for path in args.paths.split(","):
        eng = b.Engine(dev, hamming_path=path)
        # one ORB pair, 10k x 10k, cross-check and ratio
        a, c = b.synth.binary_pair(seed=1, n=10000, nbytes=32)
        da, dc = eng.prepare(a, b.NORM_HAMMING), eng.prepare(c, b.NORM_HAMMING)
        for mode, name in ((L.MODE_CROSS, "cross"), (L.MODE_RATIO, "ratio0.7")) if "pair" in args.cases else ():
            fn = lambda: eng.match_device(da, dc, [(0, 10000, 0, 10000)], mode, 0.7)   # noqa: E731
            ms = timed(fn, 20)
            k = knn_ms(eng.lib, fn, 20)
            print(json.dumps({"case": f"orb_pair_10000_{name}", "path": path, "ms": ms, "knn_ms": k,
                              "gdist_per_s": 1e8 / 1e9 / (ms / 1e3)}), flush=True)
        # all pairs: ORB 8192 x 32 B and AKAZE 16384 x 64 B
        for name, n, nb, images in (("orb_allpairs", 8192, 32, 2 * args.images), ("akaze_allpairs", 16384, 64, args.images)):
            if name.split("_")[0] not in args.cases:
                continue
            desc = torch.randint(0, 256, (images * n, nb), device=dev, generator=g, dtype=torch.uint8)
            store = ap.build_store(eng, desc, b.NORM_HAMMING, counts=[n] * images)
            pairs = ap.all_pairs(images)
            for pipeline in (True, False):
                m = ap.AllPairsMatcher(eng, chunk_pairs=1024, pipeline=pipeline)
                fn = lambda: m.match(store, pairs, cross_check=True)   # noqa: E731
                ms = timed(fn, args.reps)
                k = knn_ms(eng.lib, fn, args.reps) if not pipeline else None
                gd = len(pairs) * n * n / 1e9 / (ms / 1e3)
                print(json.dumps({"case": f"{name}_{images}x{n}x{nb}B_cross", "path": path, "pipeline": pipeline,
                                  "pairs": len(pairs), "ms": ms, "knn_ms": k, "pairs_per_s": len(pairs) / (ms / 1e3),
                                  "gdist_per_s": gd,
                                  "tensor_top_per_s_both_directions": 2 * 2 * gd * 8 * nb / 1e3 if path != "popc" else None}),
                      flush=True)
            del store, desc


if __name__ == "__main__":
    main()
