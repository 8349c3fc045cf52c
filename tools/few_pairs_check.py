"""tools/few_pairs_check.py — a call with four ragged pairs (2500 / 3 / 1 / 2100 rows) in all three modes, SIFT and ORB,
against the oracle.  This is the shape that exposed the keyed int8 kernel's lost `tc_base += ntiles` (a CTA that drains two
work items of which the first has an odd number of train tiles): every CTA of that call gets more than one item and the
items have 3 tiles.  Written at the end of round 2 when no GPU time was left: run it on a B200 first thing
(`timeout 120 python tools/few_pairs_check.py`), then move it into tests/test_parity_gpu.py."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
b = importlib.import_module("3dreconstruction_b200")
synth = importlib.import_module("3dreconstruction_b200.synth")
from oracle import bf_oracle as O

eng = b.Engine()
sizes = [2500, 3, 1, 2100]
for kind in ("sift", "orb"):
    if kind == "sift":
        sets = [synth.sift_pair(seed=140 + i, n=max(n, 2))[0][:n] for i, n in enumerate(sizes)]
        sets[3][:900] = np.clip(sets[0][:900] + np.random.default_rng(2).integers(-3, 4, size=(900, 128)), 0, 255)
        sets[1][:] = sets[0][7:10]
        norm, onorm = b.NORM_L2, O.NORM_L2
    else:
        sets = [synth.binary_pair(seed=150 + i, n=max(n, 2), nbytes=32)[0][:n] for i, n in enumerate(sizes)]
        sets[3][:900] = sets[0][:900]
        sets[3][:900, 1] ^= 3
        sets[1][:] = sets[0][7:10]
        norm, onorm = b.NORM_HAMMING, O.NORM_HAMMING
    offs = np.cumsum([0] + sizes)
    dd = eng.prepare(np.concatenate(sets), norm)
    pairs = [(0, 3), (1, 0), (2, 3), (3, 0)]
    spec = [(int(offs[i]), sizes[i], int(offs[j]), sizes[j]) for i, j in pairs]
    for mode in [(False, 0.75), (True, None), (False, None)]:
        oq, ot, od, oc, ooff = eng.match_device(dd, dd, spec, b.mode_of(*mode), mode[1] or 0.0)
        oq, ot, od, oc = oq.cpu().numpy(), ot.cpu().numpy(), od.cpu().numpy(), oc.cpu().numpy()
        for p, (i, j) in enumerate(pairs):
            want = O.reference_match(sets[i], sets[j], onorm, *mode)
            sl = slice(int(ooff[p]), int(ooff[p]) + int(oc[p]))
            ok = (np.array_equal(oq[sl], want.query_idx) and np.array_equal(ot[sl], want.train_idx)
                  and np.array_equal(od[sl].view(np.uint32), want.distance.astype(np.float32).view(np.uint32)))
            print(f"{kind} mode {mode} pair {p} ({sizes[i]} x {sizes[j]}): {'identical' if ok else 'MISMATCH'} ({int(oc[p])} matches)", flush=True)
            assert ok
print("ok")
