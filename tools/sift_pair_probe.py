"""tools/sift_pair_probe.py — BASELINE configs[0]: one SIFT pair, 8192 x 8192 x 128, knn k=2 + ratio 0.75, device-resident
stores (for a launch list under ncu).  Prints the CUDA-event time per call."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
b = importlib.import_module("3dreconstruction_b200")
L = importlib.import_module("3dreconstruction_b200._lib")
synth = importlib.import_module("3dreconstruction_b200.synth")
eng = b.Engine()
a, c = synth.sift_pair(seed=0, n=8192)
da, dc = eng.prepare(a, b.NORM_L2), eng.prepare(c, b.NORM_L2)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
call = lambda: eng.match_device(da, dc, [(0, 8192, 0, 8192)], L.MODE_RATIO, 0.75)
for _ in range(5):
    r = call()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    r = call()
e1.record(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(reps):
    r = call()
host = (time.perf_counter() - t0) / reps * 1e3
torch.cuda.synchronize()
print(f"sift pair 8192 x 8192 ratio 0.75: {e0.elapsed_time(e1) / reps:.4f} ms per call on the stream, {host:.4f} ms of host time per enqueue, {int(r[3][0])} matches")
