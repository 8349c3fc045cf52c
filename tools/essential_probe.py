"""tools/essential_probe.py — one launch of ransac_essential_kernel over 148 pairs (one wave), for an ncu capture."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b = importlib.import_module("3dreconstruction_b200")
pose = importlib.import_module("3dreconstruction_b200.pose")
eng = b.Engine("cuda:0")
P = 148
parts, cams = [], []
for p in range(P):
    c, K = b.synth.two_view_pair(seed=30000 + p, n=1200, inlier_frac=0.5, noise=0.7)
    parts.append(c); cams.append([K[0, 0], K[1, 1], K[0, 2], K[1, 2]] * 2)
corr = np.concatenate(parts)
ptr = np.arange(P + 1, dtype=np.int64) * 1200
dc, dp, dk = eng.to_device(corr), eng.to_device(ptr), eng.to_device(np.asarray(cams, np.float64))
for rep in range(2):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    v = pose.estimate_pairs_device(eng, dc, dp, dk, 1.0 / 1920, 0.999, 5000)
    e1.record(); torch.cuda.synchronize()
    its = v.iters.cpu().numpy()
    print(f"launch {rep}: {e0.elapsed_time(e1):.2f} ms, iterations per pair: mean {its.mean():.0f} max {its.max()}, rounds total {int(np.ceil(its / 64).sum())}")
