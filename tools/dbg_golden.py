import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from conftest import pkg
from oracle import bf_oracle as O
b = pkg(); eng = b.Engine()
z = np.load("/root/repo/tests/golden/bf_golden.npz")
for name in z["_names"]:
    name = str(name)
    if name.startswith("err_"): continue
    norm, cc, ratio = z[f"{name}/params"]
    ratio = None if ratio < 0 else float(ratio)
    a, c = z[f"{name}/desc1"], z[f"{name}/desc2"]
    got = eng.match_arrays(a, c, int(norm), bool(cc), ratio)
    ok = len(got.query_idx) == len(z[f"{name}/q"]) and np.array_equal(got.query_idx, z[f"{name}/q"]) and np.array_equal(got.train_idx, z[f"{name}/t"])
    print(name, a.shape, c.shape, a.dtype, int(norm), bool(cc), ratio, "OK" if ok else "FAIL", len(got.query_idx), len(z[f"{name}/q"]))
    if not ok and a.dtype != np.uint8:
        k = eng.knn2_arrays(a, c, int(norm)); w = O.knn2_c(a, c, int(norm))
        bad = np.nonzero((k.idx1 != w.idx1) | (k.idx2 != w.idx2) | (k.d1 != w.d1) | (k.d2 != w.d2))[0]
        print("  knn2 bad rows:", bad[:20], len(bad))
        for r in bad[:6]:
            print("   row", r, "got", k.idx1[r], k.idx2[r], k.d1[r], k.d2[r], "want", w.idx1[r], w.idx2[r], w.d1[r], w.d2[r])
