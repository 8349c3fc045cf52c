"""tools/f32_probe.py — small SuperPoint all-pairs pass (for ncu launch lists of the float path)."""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
b = importlib.import_module("3dreconstruction_b200")
ap = importlib.import_module("3dreconstruction_b200.allpairs")
eng = b.Engine()
I, n, D = int(sys.argv[1]) if len(sys.argv) > 1 else 16, 4096, 256
desc = bench.superpoint_set_gpu(torch, eng.device, I, n, D)
store = ap.build_store(eng, desc.view(I * n, D), b.NORM_L2, counts=[n] * I)
pairs = ap.all_pairs(I)
m = ap.AllPairsMatcher(eng, chunk_pairs=1024)
for _ in range(2):
    r = m.match(store, pairs, cross_check=True)
torch.cuda.synchronize()
print("pairs", len(pairs), "matches", int(r.query_idx.numel()), "fallback", eng.f32_stats())
