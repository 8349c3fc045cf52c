"""tools/orb_pair_probe.py — BASELINE configs[1]: one ORB pair, 10k x 10k x 256 bit, cross-check, device-resident stores
(for a launch list: ncu --metrics gpu__time_duration.sum -k regex:"hamming|finalize|sort_emit|scan" ...).  Prints the
CUDA-event time per call."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
b = importlib.import_module("3dreconstruction_b200")
L = importlib.import_module("3dreconstruction_b200._lib")
eng = b.Engine()
rng = np.random.default_rng(0)
a = rng.integers(0, 256, size=(10000, 32), dtype=np.uint8)
c = rng.integers(0, 256, size=(10000, 32), dtype=np.uint8)
da, dc = eng.prepare(a, b.NORM_HAMMING), eng.prepare(c, b.NORM_HAMMING)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
for _ in range(5):
    r = eng.match_device(da, dc, [(0, 10000, 0, 10000)], L.MODE_CROSS)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    r = eng.match_device(da, dc, [(0, 10000, 0, 10000)], L.MODE_CROSS)
e1.record(); torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(reps):
    r = eng.match_device(da, dc, [(0, 10000, 0, 10000)], L.MODE_CROSS)
host = (time.perf_counter() - t0) / reps * 1e3
torch.cuda.synchronize()
print(f"orb pair 10k x 10k cross-check: {e0.elapsed_time(e1) / reps:.4f} ms per call on the stream, {host:.4f} ms of host time to enqueue one call, {int(r[3][0])} matches")
