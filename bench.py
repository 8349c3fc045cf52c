#!/usr/bin/env python
"""bench.py — headline benchmark of the brute-force matching hot path (BASELINE.json metric:
image pairs/s and Gdist/s, SIFT-L2 / ORB-Hamming, 1/2/4/8 B200 vs CPU cv2).

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W  # reference CPU arm (cv2.BFMatcher)

Workload (config.workload): BASELINE configs[3] — SIFT all-pairs exhaustive matching, 200 images x
8192 keypoints x 128-d (19,900 pairs), knnMatch k=2 + ratio 0.75, sharded over the N ranks by
balanced contiguous pair blocks (strong scaling: the pair list is fixed).  One "step" = one pass over
all pairs.  The int8 store (210 MB) is larger than L2 (126 MB), so no L2 flush is needed between steps.
`also` carries single-pair numbers for configs[0] (SIFT 8k x 8k) and configs[1] (ORB 10k x 10k cross-check).
Prints exactly ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

RATIO = 0.75
HW_INT8_TOPS = 4590.0                 # burst issue-rate peak of tcgen05.mma kind::i8 at 1965 MHz (tools/mma_peak.cu, profiles/r01_mma_peak_microbench.txt)
HW_INT8_TOPS_SUSTAINED = 3730.0       # the same instruction stream back to back for 6 s: the chip settles at its 1000 W cap,
                                      # SM clock ~1540 MHz (profiles/r01_mma_int8_sustained.txt) -- SURVEY 8(d)'s "measured int8 peak"
NCU_DRAM_BYTES_PER_PAIR = (2716.6e6 + 268.5e6) / 2048.0    # ncu --set full of the shipped fold kernel, a 2048-pair launch of this workload (profiles/r02_l2i8_pair_final_ncu_full.md)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--images", type=int, default=200)
    ap.add_argument("--keypoints", type=int, default=8192)
    ap.add_argument("--chunk-pairs", type=int, default=2048)
    ap.add_argument("--cpu-sample-pairs", type=int, default=24)
    ap.add_argument("--no-also", action="store_true")
    ap.add_argument("--group-images", type=int, default=48, help="images per upload group of the e2e arm (StreamingAllPairs)")
    ap.add_argument("--e2e-upload", default="auto", choices=["auto", "sharded", "rank0"],
                    help="N > 1: every rank uploads its block of each image group over its own PCIe link (default) or rank 0 uploads everything")
    ap.add_argument("--e2e-gather", default="rank0", choices=["rank0", "local"],
                    help="N > 1: match records return to rank 0 by NCCL, streamed per image group (default), or stay in every rank's host memory")
    ap.add_argument("--no-parity", action="store_true", help="skip the element-wise check of the e2e result against the resident path")
    ap.add_argument("--workload", default="sift", choices=["sift", "akaze"],
                    help="sift: BASELINE configs[3] (the headline); akaze: configs[4], binary all-pairs with mutual check "
                         "(--images 1000 --keypoints 16384 for the full configuration)")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "sm_max_mhz": 1965.0}, "fallback"


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock + throttle reasons DURING the timed region: an NVML thread sampling every 5 ms (short multi-GPU steps
    end before an nvidia-smi loop prints its first line); nvidia-smi -lms as the fallback when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NVML_REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                    0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.nvml, self.samples, self.stop_flag = None, [], False

    def _nvml_loop(self):
        nv, h = self.nvml
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), int(get_reasons(h))))
            except Exception:
                pass
            time.sleep(0.005)

    def _start_nvml(self) -> bool:
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.nvml = (nv, h)
            self.thread = threading.Thread(target=self._nvml_loop, daemon=True)
            self.thread.start()
            return True
        except Exception:
            self.nvml = None
            return False

    def _stop_nvml(self):
        self.stop_flag = True
        self.thread.join(timeout=1)
        sm = [float(c) for c, _ in self.samples]
        bits = 0
        for _, r in self.samples:
            bits |= r
        busy = [x for x in sm if x > 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
                "reasons": sorted(name for bit, name in self.NVML_REASONS.items() if bits & bit), "source": "nvml, 5 ms"}

    def start(self):
        if self._start_nvml():
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml:
            return self._stop_nvml()
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [x for x in sm if x > 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- synthetic data on the device
def sift_set_gpu(torch, device, images, n, seed=3, d=128, shared=0.3, sigma=6.0):
    """Same recipe as 3dreconstruction_b200.synth.sift_set, generated on the GPU (840 MB of float32)."""
    g = torch.Generator(device=device); g.manual_seed(seed)

    def sift_like(rows):
        x = torch._standard_gamma(torch.full((rows, d), 0.6, device=device), generator=g)
        x = x / (x.norm(dim=1, keepdim=True) + 1e-12)
        x = torch.clamp(x, max=0.2)
        x = x / (x.norm(dim=1, keepdim=True) + 1e-12)
        return torch.clamp(torch.round(512.0 * x), 0, 255)

    pool = sift_like(2 * n)
    k = int(n * shared)
    out = torch.empty((images, n, d), dtype=torch.float32, device=device)
    for i in range(images):
        pick = torch.randperm(2 * n, device=device, generator=g)[:k]
        out[i, :k] = torch.clamp(torch.round(pool[pick] + sigma * torch.randn((k, d), device=device, generator=g)), 0, 255)
        out[i, k:] = sift_like(n - k)
        out[i] = out[i][torch.randperm(n, device=device, generator=g)]
    return out


# --------------------------------------------------------------------------- CPU arm (reference's own engine)
def cpu_match_fn():
    """cv2.BFMatcher driven exactly as EnhancedBFMatcher.match drives it (oracle/bf_oracle.cv2_match);
    falls back to the C restatement (OpenMP) when cv2 is not importable on this box."""
    from oracle import bf_oracle as O
    try:
        import cv2
        threads = os.cpu_count() or 1
        cv2.setNumThreads(threads)
        return (lambda a, b: O.cv2_match(a, b, O.NORM_L2, False, RATIO)), cv2.getNumThreads(), f"cv2.BFMatcher {cv2.__version__}"
    except ImportError:
        O.set_threads(os.cpu_count() or 1)
        return (lambda a, b: O.reference_match(a, b, O.NORM_L2, False, RATIO)), O.num_threads(), "oracle/bf_oracle.c (OpenMP)"


def cpu_time_pairs(host_images, pair_list):
    fn, cores, engine = cpu_match_fn()
    fn(host_images[0][:512], host_images[1][:512])          # warm the thread pool
    t0 = time.perf_counter()
    for i, j in pair_list:
        fn(host_images[i], host_images[j])
    return time.perf_counter() - t0, cores, engine


def cv2_pair_baseline(pairs, norm, cross_check, ratio, what):
    """cpu_baseline of an `also` entry: cv2.BFMatcher on this box's host cores, driven as EnhancedBFMatcher.match drives it
    (oracle/bf_oracle.cv2_match), on `pairs` = [(desc1, desc2)].  Returns (dict for the JSON line, [Matches])."""
    from oracle import bf_oracle as O
    try:
        import cv2
    except ImportError:
        return None, None
    cv2.setNumThreads(os.cpu_count() or 1)
    O.cv2_match(pairs[0][0][:256], pairs[0][1][:256], norm, cross_check, ratio)        # warm the thread pool
    t0 = time.perf_counter()
    res = [O.cv2_match(a, c, norm, cross_check, ratio) for a, c in pairs]
    t = time.perf_counter() - t0
    return {"value": len(pairs) / t, "unit": "pairs/s", "ms_per_pair": 1e3 * t / len(pairs), "cores": cv2.getNumThreads(),
            "kind": "reference", "sample": f"{len(pairs)} pair(s) of {what} on this host, cv2.BFMatcher {cv2.__version__} driven by "
                                           "oracle/bf_oracle.cv2_match (restates EnhancedBFMatcher.match)"}, res


def host_sample_images(args, count):
    synth = importlib.import_module("3dreconstruction_b200.synth")
    return synth.sift_set(seed=3, images=count, n=args.keypoints)


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (cv2.BFMatcher, all host threads) on a
    bounded sample of the same workload per step.  Rank 0 only; other ranks exit 0 without work."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    n_img = 5
    imgs = host_sample_images(args, n_img)
    sample = [(0, 1), (0, 2), (1, 2), (3, 4)]                  # 4 pairs per step  (~2 s of 8-thread cv2)
    for _ in range(args.warmup):
        cpu_time_pairs(imgs, sample[:1])
    t_total, cores, engine = 0.0, 1, ""
    for _ in range(args.steps):
        t, cores, engine = cpu_time_pairs(imgs, sample)
        t_total += t
    pairs_s = len(sample) * args.steps / t_total
    gdist = pairs_s * args.keypoints * args.keypoints / 1e9
    line = {"impl": "reference", "metric": "image pairs/s (SIFT-L2 all-pairs, knn k=2 + ratio 0.75)", "value": pairs_s,
            "unit": "pairs/s", "gdist_per_s": gdist, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "cpu_baseline": {"value": pairs_s, "unit": "pairs/s", "cores": cores, "kind": "port",
                             "sample": f"{len(sample)} pairs/step of the {args.keypoints}x{args.keypoints}x128 workload, {engine} "
                                       "driven by oracle/bf_oracle.cv2_match (restates EnhancedBFMatcher.match)"},
            "e2e": {"value": pairs_s, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def workload_config(args):
    P = args.images * (args.images - 1) // 2
    return {"workload": f"SIFT all-pairs: {args.images} images x {args.keypoints} keypoints x 128-d, {P} pairs, "
                        f"knnMatch k=2 + ratio {RATIO} (BASELINE configs[3])",
            "pairs": P, "l2_flush": "inputs larger than L2 (int8 store > 126 MB)" if args.images * args.keypoints * 128 > 126e6
            else "store fits L2 (reduced size run)", "parallelism": f"pair-list sharding x{args.gpus}"}


# --------------------------------------------------------------------------- the B200 arm
def single_pair_extras(torch, b, eng, pk):
    """configs[0] and configs[1]: one pair, device-resident, kernels only (CUDA events, 20 reps)."""
    L = b._lib
    out = {}

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    a, c = b.synth.sift_pair(seed=0, n=8192)
    da, dc = eng.prepare(a, b.NORM_L2), eng.prepare(c, b.NORM_L2)
    ms = timed(lambda: eng.match_device(da, dc, [(0, 8192, 0, 8192)], L.MODE_RATIO, RATIO))
    gd = 8192 * 8192 / 1e9 / (ms / 1e3)
    out["sift_pair_8192x8192_ratio0.75"] = {"ms_per_pair": ms, "pairs_per_s": 1e3 / ms, "gdist_per_s": gd,
                                            "note": "one pair cannot fill 148 SMs x tensor pipe; latency-bound"}
    a, c = b.synth.binary_pair(seed=1, n=10000, nbytes=32)
    popc_peak = 148 * 16 * pk.get("sm_max_mhz", 1965.0) * 1e6 / 8 / 1e9       # Gdist/s, 256-bit, at max clock
    orb = {}
    for path, e in (("tensor", eng), ("int8", b.Engine(eng.device, hamming_path="int8")), ("popc", b.Engine(eng.device, hamming_path="popc"))):
        da, dc = e.prepare(a, b.NORM_HAMMING), e.prepare(c, b.NORM_HAMMING)
        ms = timed(lambda: e.match_device(da, dc, [(0, 10000, 0, 10000)], L.MODE_CROSS))
        orb[path] = {"ms_per_pair": ms, "pairs_per_s": 1e3 / ms, "gdist_per_s": 1e8 / 1e9 / (ms / 1e3)}
    orb["tensor"]["note"] = ("default path: bits as +-1 int8, tcgen05 kind::i8 in both directions + POPC resolve of the two best "
                             "16-column chunks per row (hamming_tc.cu); one pair = 79 row tiles, launch/tail latency-bound")
    orb["popc"].update({"popc_roofline_gdist_per_s": popc_peak, "frac_of_popc_roofline": orb["popc"]["gdist_per_s"] / popc_peak,
                        "note": "INT-pipe kernel (hamming.cu): single fused pass, row top-2 and column minima from the same distances"})
    cpu, cres = cv2_pair_baseline([(a, c)] * 3, 6, True, None, "10000 x 10000 x 256-bit ORB, crossCheck=True (BASELINE configs[1])")
    if cpu:
        got = eng.match_arrays(a, c, b.NORM_HAMMING, True, None)
        cpu["identical_to_gpu"] = bool(np.array_equal(got.query_idx, cres[0].query_idx) and np.array_equal(got.train_idx, cres[0].train_idx)
                                       and np.array_equal(got.distance, cres[0].distance))
    out["orb_pair_10000x10000_crosscheck"] = {**orb["tensor"], "popc_path": orb["popc"],
                                              "speedup_over_popc_path": orb["popc"]["ms_per_pair"] / orb["tensor"]["ms_per_pair"],
                                              "cpu_baseline": cpu}
    out.update(plugin_call_extras(torch, b, eng))
    out.update(allpairs_extras(torch, b, eng, pk))
    out.update(sfm_extras(torch, b, eng, pk))
    out.update(ransac_extras(torch, b, eng, pk))
    out.update(essential_extras(torch, b, eng, pk))
    return out


def plugin_call_extras(torch, b, eng):
    """configs[0] through the drop-in plugin call a user of the reference makes: B200BFMatcher(norm, cross_check, ratio).match(
    FeatureData, FeatureData) -> MatchData, HOST numpy descriptors in, a Python list of EnhancedDMatch out (wall clock: H2D, int8
    re-encoding, kernels, D2H, the per-match Python objects).  Beside it the reference's way on this host: cv2.BFMatcher driven as
    EnhancedBFMatcher.match drives it (arrays out; the reference then builds the same Python objects on top)."""
    import time
    a, c = b.synth.sift_pair(seed=0, n=8192)
    F = b.contracts.FeatureData
    f1, f2 = F(keypoints=list(range(len(a))), descriptors=a, method="SIFT"), F(keypoints=list(range(len(c))), descriptors=c, method="SIFT")
    m = b.B200BFMatcher(norm_type=b.NORM_L2, cross_check=False, ratio_threshold=RATIO)
    m._engine = eng
    md = m.match(f1, f2)
    torch.cuda.synchronize()
    reps = 10

    def wall(fn, n=reps):
        t0 = time.perf_counter()
        for k in range(n):
            fn(k)
        return (time.perf_counter() - t0) / n * 1e3

    # cached: the same array objects again (a matcher reused over a run sees every image I - 1 times, pipeline.py:99-104)
    ms = wall(lambda k: m.match(f1, f2))
    ms_arr = wall(lambda k: m.match_arrays(a, c))
    # cold: nothing cached -- an engine with the descriptor cache off uploads and re-encodes both images on every call
    # (pageable H2D of 8 MB, int8 re-encoding, kernels, one packed D2H); device blocks are reused between calls
    m0 = b.B200BFMatcher(norm_type=b.NORM_L2, cross_check=False, ratio_threshold=RATIO, engine=b.Engine(eng.device, cache_entries=0))
    m0.match(f1, f2)
    ms_arr_cold = wall(lambda k: m0.match_arrays(a, c))
    ms_cold = wall(lambda k: m0.match(f1, f2))
    # first sight with the cache on: as cold, plus the device allocation of the two stores the cache then keeps (once per image)
    fresh = [(a.copy(), c.copy()) for _ in range(reps)]
    ms_first_sight = wall(lambda k: m.match_arrays(*fresh[k]))
    arr = m.match_arrays(a, c)
    ms_objects = wall(lambda k: importlib.import_module("3dreconstruction_b200.matcher")._materialise(b.contracts.resolve(), arr))
    fn, cores, engine = cpu_match_fn()
    fn(a, c)
    t0 = time.perf_counter()
    want = fn(a, c)
    ms_cpu = (time.perf_counter() - t0) * 1e3
    same = len(md.matches) == len(want.query_idx) and all(x.queryIdx == int(q) and x.trainIdx == int(t) for x, q, t in
                                                          zip(md.matches, want.query_idx, want.train_idx))
    return {"plugin_call_sift_pair_8192x8192": {
        "ms_per_match_call": ms_cold, "ms_per_match_call_cached": ms, "ms_per_match_arrays_call": ms_arr_cold,
        "ms_per_match_arrays_call_cached": ms_arr, "ms_per_match_arrays_first_sight_with_cache": ms_first_sight,
        "ms_python_match_objects": ms_objects, "matches": len(md.matches),
        "h2d_bytes_per_call": int(a.nbytes + c.nbytes), "reference_cpu_ms_per_call": ms_cpu, "reference_engine": engine,
        "reference_threads": cores, "identical_to_reference": bool(same), "speedup": ms_cpu / ms_cold, "speedup_cached": ms_cpu / ms,
        "cache": {"hits": eng.cache.hits, "misses": eng.cache.misses},
        "note": "B200BFMatcher.match(FeatureData, FeatureData) -> MatchData with host numpy descriptors (pageable): the drop-in call. "
                "cold = cache off (H2D + int8 re-encoding of both images, kernels, one packed D2H on every call); cached = the engine's "
                "descriptor cache serves both images (kernels + D2H only), the steady state of a matcher reused over a run; first_sight = "
                "cold plus the one-time device allocation of the stores the cache keeps; match_arrays = the same without the list of EnhancedDMatch "
                "objects, whose construction alone is ms_python_match_objects"}}


def sfm_extras(torch, b, eng, pk):
    """features of every image cached in HBM once, all pairs matched, N x 4 correspondences gathered on
    the device (csrc/adapter.cu) and copied out; beside it the reference's per-match Python assembly
    (result_converters.py:763-766) timed on a sample of the same matches."""
    import time
    import numpy as np
    sfm = importlib.import_module("3dreconstruction_b200.sfm")
    ap = importlib.import_module("3dreconstruction_b200.allpairs")
    I, n = 32, 8192
    desc = sift_set_gpu(torch, eng.device, I, n).view(I * n, 128)
    store = ap.build_store(eng, desc, b.NORM_L2, counts=[n] * I)
    g = torch.Generator(device=eng.device); g.manual_seed(5)
    kp = torch.rand((I * n, 2), device=eng.device, generator=g) * 2000.0
    fs = sfm.FeatureSet([f"img_{i:03d}" for i in range(I)], [(2000, 2000)] * I, store, kp, "SIFT")
    m = ap.AllPairsMatcher(eng, chunk_pairs=2048)
    sfm.match_feature_set(eng, fs, None, False, RATIO, matcher=m)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = sfm.match_feature_set(eng, fs, None, False, RATIO, matcher=m)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    # the gather alone, device-resident CSR
    csr = m.match(store, r.pairs, False, RATIO)
    K = int(csr.query_idx.numel())
    q_off = torch.from_numpy(store.offsets[r.pairs[:, 0]].astype(np.int32)).to(eng.device)
    t_off = torch.from_numpy(store.offsets[r.pairs[:, 1]].astype(np.int32)).to(eng.device)
    corr = torch.empty((max(K, 1), 4), dtype=torch.float32, device=eng.device)
    call = lambda: eng.lib.b200m_gather_correspondences(kp.data_ptr(), I * n, q_off.data_ptr(), t_off.data_ptr(),   # noqa: E731
                                                        csr.ptr.data_ptr(), len(r.pairs), csr.query_idx.data_ptr(),
                                                        csr.train_idx.data_ptr(), corr.data_ptr(), eng._stream())
    call(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        call()
    e1.record(); torch.cuda.synchronize()
    gms = e0.elapsed_time(e1) / 20
    # reference way on the host for the first pairs: one list comprehension per side over Python match objects
    kp_host = kp.cpu().numpy()
    pts = [tuple(map(float, xy)) for xy in kp_host[:2 * n]]
    s = r.pair_slice(0)
    q, t = r.query_idx[s].tolist(), r.train_idx[s].tolist()
    reps = max(1, 20000 // max(len(q), 1))
    t0 = time.perf_counter()
    for _ in range(reps):
        np.hstack([np.array([pts[i] for i in q]), np.array([pts[n + j] for j in t])])
    cpu_us_per_match = (time.perf_counter() - t0) / (reps * max(len(q), 1)) * 1e6
    # N3 on top: the same pass with every pair verified by RANSAC homography before anything leaves the device
    sfm.match_feature_set(eng, fs, None, False, RATIO, matcher=m, ransac=(4.0, 0.99, 2000))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    _, ver = sfm.match_feature_set(eng, fs, None, False, RATIO, matcher=m, ransac=(4.0, 0.99, 2000))
    torch.cuda.synchronize()
    ms_ver = (time.perf_counter() - t0) * 1e3
    return {"sfm_handoff_sift_32x8192": {
        "ms_match_gather_verify_d2h": ms_ver, "ransac_iterations_total": int(ver.iters.sum()),
        "verify_note": "synthetic keypoint coordinates are random, so no pair has a homography: every pair runs RANSAC to its "
                       "iteration cap (2000) -- the worst case for the verification stage",
        "pairs": len(r.pairs), "matches": K, "ms_match_gather_d2h": ms, "pairs_per_s": len(r.pairs) / (ms / 1e3),
        "gather_kernel_ms": gms, "gather_bytes": K * 40, "gather_gbs": K * 40 / (gms / 1e3) / 1e9, "hbm_peak_gbs": pk["hbm_gbs"],
        "reference_python_assembly_us_per_match": cpu_us_per_match,
        "reference_python_assembly_ms_for_these_matches": cpu_us_per_match * K / 1e3,
        "note": "feature cache + all-pairs + device gather of keypoint coordinates into N x 4 correspondences; "
                "the gather is latency-bound at this size (40 algorithmic bytes per match)"}}


def ransac_extras(torch, b, eng, pk):
    """RANSAC homography verification of 2048 pairs in one launch (csrc/ransac.cu), device-resident
    correspondences; beside it cv2.findHomography (what utils.py:141 calls) on a sample of the same pairs on this host."""
    import time
    import numpy as np
    verify = importlib.import_module("3dreconstruction_b200.verify")
    P = 2048
    corr, ptr = b.synth.homography_set(seed=6, pairs=P, n_lo=200, n_hi=4000, inl_lo=0.3, inl_hi=0.9)
    dc, dp = eng.to_device(corr), eng.to_device(ptr)
    run = lambda: verify.verify_pairs_device(eng, dc, dp, 4.0, 0.99, 2000)        # noqa: E731
    for _ in range(3):
        v = run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        v = run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    iters = v.iters.cpu().numpy()
    inl = v.inliers.cpu().numpy()
    mask = v.mask.cpu().numpy()
    out = {"pairs": P, "matches": int(ptr[-1]), "ms_per_launch": ms, "pairs_per_s": P / (ms / 1e3),
           "ransac_iterations_total": int(iters.sum()), "hypotheses_per_s": float(iters.sum()) / (ms / 1e3),
           "point_model_tests_per_s": float((iters.astype(np.float64) * np.diff(ptr)).sum()) / (ms / 1e3),
           "pairs_with_model": int((inl >= 0).sum()),
           "note": "one CTA per pair; cv2's own random sequence replayed (64 tuples per round), fp64 DLT/Jacobi per hypothesis, float32 "
                   "inlier test, inlier refit + Levenberg-Marquardt, mask of the refined H; threshold 4 px, confidence 0.99, maxIters 2000"}
    try:
        import cv2
        sample = list(range(0, P, P // 64))
        same = 0
        t0 = time.perf_counter()
        res = []
        for p in sample:
            c = corr[ptr[p]:ptr[p + 1]]
            res.append(cv2.findHomography(c[:, :2].reshape(-1, 1, 2), c[:, 2:].reshape(-1, 1, 2), cv2.RANSAC, 4.0,
                                          confidence=0.99, maxIters=2000))
        t = time.perf_counter() - t0
        for p, (Hc, mc) in zip(sample, res):
            same += Hc is not None and np.array_equal(mc.ravel(), mask[ptr[p]:ptr[p + 1]])
        out["cpu_cv2_findHomography"] = {"sample_pairs": len(sample), "pairs_per_s": len(sample) / t, "threads": cv2.getNumThreads(),
                                         "masks_identical_to_gpu": int(same), "speedup": out["pairs_per_s"] / (len(sample) / t)}
    except ImportError:
        out["cpu_cv2_findHomography"] = None
    return {"ransac_homography_2048_pairs": out}


def essential_extras(torch, b, eng, pk):
    """SURVEY 8f N3, second half: essential-matrix RANSAC of 1024 pairs in one launch (csrc/essential.cu), device-resident pixel
    correspondences; beside it cv2.findEssentialMat (what essential_estimation.py:118-126 calls) on a sample of the same pairs."""
    import time
    pose = importlib.import_module("3dreconstruction_b200.pose")
    P = 1024
    rng = np.random.default_rng(8)
    parts, cams = [], []
    for p in range(P):
        n = int(rng.integers(150, 2500))
        c, K = b.synth.two_view_pair(seed=20000 + p, n=n, inlier_frac=float(rng.uniform(0.35, 0.9)), noise=0.7)
        parts.append(c)
        cams.append([K[0, 0], K[1, 1], K[0, 2], K[1, 2]] * 2)
    corr = np.concatenate(parts)
    ptr = np.concatenate([[0], np.cumsum([len(x) for x in parts])]).astype(np.int64)
    cams = np.asarray(cams, np.float64)
    thr, prob, iters = 1.0 / 1920, 0.999, 5000            # ~1 pixel; the reference's confidence and iteration cap
    dc, dp, dk = eng.to_device(corr), eng.to_device(ptr), eng.to_device(cams)
    run = lambda: pose.estimate_pairs_device(eng, dc, dp, dk, thr, prob, iters)        # noqa: E731
    for _ in range(2):
        v = run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        v = run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    its, inl, mask = v.iters.cpu().numpy(), v.inliers.cpu().numpy(), v.mask.cpu().numpy()
    out = {"pairs": P, "matches": int(ptr[-1]), "ms_per_launch": ms, "pairs_per_s": P / (ms / 1e3),
           "ransac_iterations_total": int(its.sum()), "hypotheses_per_s": float(its.sum()) / (ms / 1e3),
           "pairs_with_model": int((inl >= 0).sum()),
           "note": "one CTA per pair; cv2's own random sequence and model order replayed; per sample Nister's five-point solver in fp64 "
                   "(null-space basis as cv's SVD completes it, 10 x 20 elimination, Durand-Kerner on the degree-10 polynomial), up to 10 "
                   "models; Sampson error of every point for every model; threshold ~1 px, prob 0.999, maxIters 5000 (the reference's)"}
    try:
        import cv2
        sample = list(range(0, P, P // 32))
        same = 0
        res = []
        t0 = time.perf_counter()
        for p in sample:
            c = corr[ptr[p]:ptr[p + 1]].astype(np.float64)
            K = cams[p]
            m1 = np.stack([(c[:, 0] - K[2]) * (1.0 / K[0]), (c[:, 1] - K[3]) * (1.0 / K[1])], axis=1)
            m2 = np.stack([(c[:, 2] - K[6]) * (1.0 / K[4]), (c[:, 3] - K[7]) * (1.0 / K[5])], axis=1)
            res.append(cv2.findEssentialMat(m1, m2, focal=1.0, pp=(0.0, 0.0), method=cv2.RANSAC, prob=prob, threshold=thr, maxIters=iters))
        t = time.perf_counter() - t0
        for p, (Ec, mc) in zip(sample, res):
            same += Ec is not None and np.array_equal(mc.ravel(), mask[ptr[p]:ptr[p + 1]])
        out["cpu_cv2_findEssentialMat"] = {"sample_pairs": len(sample), "pairs_per_s": len(sample) / t, "threads": cv2.getNumThreads(),
                                           "masks_identical_to_gpu": int(same), "speedup": out["pairs_per_s"] / (len(sample) / t)}
    except ImportError:
        out["cpu_cv2_findEssentialMat"] = None
    return {"ransac_essential_1024_pairs": out}


def superpoint_set_gpu(torch, device, images, n, d=256, seed=2, pool=8192, shared=0.5, sigma=0.05):
    """Same recipe as 3dreconstruction_b200.synth.superpoint_set, generated on the GPU."""
    g = torch.Generator(device=device); g.manual_seed(seed)
    unit = lambda x: x / (x.norm(dim=1, keepdim=True) + 1e-12)   # noqa: E731
    scene = unit(torch.randn((pool, d), device=device, generator=g))
    k = int(n * shared)
    out = torch.empty((images, n, d), dtype=torch.float32, device=device)
    for i in range(images):
        pick = torch.randperm(pool, device=device, generator=g)[:k]
        out[i, :k] = unit(scene[pick] + sigma * torch.randn((k, d), device=device, generator=g))
        out[i, k:] = unit(torch.randn((n - k, d), device=device, generator=g))
        out[i] = out[i][torch.randperm(n, device=device, generator=g)]
    return out


def binary_set_gpu(torch, device, images, n, nbytes, seed=4, shared=0.5, flip=0.08):
    """Same recipe as synth.binary_set (planted correspondences through a shared pool), on the GPU."""
    g = torch.Generator(device=device); g.manual_seed(seed)
    pool = 2 * n
    scene = torch.randint(0, 256, (pool, nbytes), device=device, generator=g, dtype=torch.uint8)
    k = int(n * shared)
    out = torch.empty((images, n, nbytes), dtype=torch.uint8, device=device)
    weights = (2 ** torch.arange(8, device=device, dtype=torch.int32)).view(1, 1, 8)
    for i in range(images):
        pick = torch.randperm(pool, device=device, generator=g)[:k]
        bits = (torch.rand((k, nbytes, 8), device=device, generator=g) < flip).to(torch.int32)
        mask = (bits * weights).sum(dim=2).to(torch.uint8)
        out[i, :k] = scene[pick] ^ mask
        out[i, k:] = torch.randint(0, 256, (n - k, nbytes), device=device, generator=g, dtype=torch.uint8)
        out[i] = out[i][torch.randperm(n, device=device, generator=g)]
    return out


def allpairs_extras(torch, b, eng, pk):
    """configs[2] (SuperPoint 256-d mutual NN, 100 images x 4096, 4950 pairs: the full config) and a
    bounded sample of configs[4] (AKAZE 61-byte rows padded to 64, 16384 keypoints, mutual check)."""
    ap = importlib.import_module("3dreconstruction_b200.allpairs")
    out = {}
    lib = eng.lib
    timed_eng = [eng]                      # the engine timed_pass drives (the Hamming sample runs both kernel paths)

    def timed_pass(store, pairs, reps):
        """(ms per pass with the two-stream chunk pipeline, knn-kernel ms per pass from a second, un-pipelined pass
        whose per-kernel CUDA events bracket nothing but the kernel, matches)."""
        import ctypes
        m = ap.AllPairsMatcher(timed_eng[0], chunk_pairs=1024)
        r = m.match(store, pairs, cross_check=True)           # warm-up (also sizes the buffers)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            r = m.match(store, pairs, cross_check=True)
        e1.record(); torch.cuda.synchronize()
        m2 = ap.AllPairsMatcher(timed_eng[0], chunk_pairs=1024, pipeline=False)
        m2.match(store, pairs, cross_check=True)
        torch.cuda.synchronize()
        lib.b200m_timing_enable(1)
        for _ in range(reps):
            m2.match(store, pairs, cross_check=True)
        torch.cuda.synchronize()
        knn_ms, knn_l = ctypes.c_double(0), ctypes.c_int(0)
        lib.b200m_timing_collect(ctypes.byref(knn_ms), ctypes.byref(knn_l))
        lib.b200m_timing_enable(0)
        return e0.elapsed_time(e1) / reps, knn_ms.value / reps, int(r.query_idx.numel())

    # ---- configs[2]
    I, n, D = 100, 4096, 256
    desc = superpoint_set_gpu(torch, eng.device, I, n, D)
    store = ap.build_store(eng, desc.view(I * n, D), b.NORM_L2, counts=[n] * I)
    del desc
    pairs = ap.all_pairs(I)
    flops = 2.0 * n * n * D * len(pairs) * 2                  # both directions are computed for the mutual check
    sp = {}
    for path in ("f16", "tf32"):
        e = eng if eng.float_path == path else b.Engine(eng.device, float_path=path)
        if path != eng.float_path:
            store = ap.build_store(e, store.dd.desc, b.NORM_L2, counts=[n] * I)
        timed_eng[0] = e
        e.f32_stats(reset=True)
        ms, knn_ms, matches = timed_pass(store, pairs, 2)
        fb_rows, _ = e.f32_stats()
        sp[path] = {"ms_per_pass": ms, "pairs_per_s": len(pairs) / (ms / 1e3),
                    "gdist_per_s_one_direction": len(pairs) * n * n / 1e9 / (ms / 1e3), "matches": matches,
                    "knn_stage_ms": knn_ms, "tflops_achieved_knn_stage": flops / (knn_ms / 1e3) / 1e12,
                    "exhaustive_fallback_rows": fb_rows}
    timed_eng[0] = eng
    peak16, peak32 = pk["bf16_tflops_sustained"], 0.5 * pk["bf16_tflops_sustained"]
    sp["f16"].update({"f16_peak_tflops": peak16, "frac_of_f16_roofline": sp["f16"]["tflops_achieved_knn_stage"] / peak16,
                      "frac_of_tf32_roofline_survey_8d": sp["f16"]["tflops_achieved_knn_stage"] / peak32,
                      "peak_source": "bf16_tflops_sustained of MEASURED_PEAKS.json (kind::f16 runs at the bf16 rate); SURVEY 8d names "
                                     "0.5 x that (TF32) for this config: both fractions given; tools/mma_peak: 2233 / 1117 TFLOP/s burst",
                      "note": "default float path (B200M_KIND_L2_F16C): tcgen05 kind::f16 candidates over a scaled fp16 copy of the rows + "
                              "exact fp32 re-rank; knn stage = candidate + refine + fallback kernels, both directions"})
    sp["tf32"].update({"tf32_peak_tflops": peak32, "frac_of_tf32_roofline": sp["tf32"]["tflops_achieved_knn_stage"] / peak32,
                       "note": "B200M_KIND_L2_F32 (Engine(float_path='tf32')): kind::tf32 candidates from the fp32 rows; same results, bit for bit"})
    # cv2 on the same box for a sample of the same pairs, and the near-tie accounting north_star asks for: rows where the
    # GPU result and cv2 disagree (membership of the mutual set, or the train index)
    sample_pairs = pairs[:: max(1, len(pairs) // 6)][:6]
    host_rows = {int(i): store.dd.desc[i * n:(i + 1) * n, :D].cpu().numpy() for i in np.unique(sample_pairs)}
    cpu, cres = cv2_pair_baseline([(host_rows[int(i)], host_rows[int(j)]) for i, j in sample_pairs], 4, True, None,
                                  "4096 x 4096 x 256 float (BASELINE configs[2]), crossCheck=True")
    if cpu:
        rs = ap.AllPairsMatcher(eng, chunk_pairs=1024).match(store, sample_pairs, cross_check=True).to_host()
        flips, total, max_rel = 0, 0, 0.0
        for p, w in enumerate(cres):
            gq, gt, gd = rs.pair(p)
            g = dict(zip(gq.tolist(), zip(gt.tolist(), gd.tolist())))
            wd = dict(zip(w.query_idx.tolist(), zip(w.train_idx.tolist(), w.distance.tolist())))
            flips += len(set(g) ^ set(wd)) + sum(1 for q in set(g) & set(wd) if g[q][0] != wd[q][0])
            total += len(wd)
            same = [q for q in set(g) & set(wd) if g[q][0] == wd[q][0]]
            if same:
                max_rel = max(max_rel, max(abs(g[q][1] - wd[q][1]) / max(wd[q][1], 1e-12) for q in same))
        cpu.update({"near_tie_flips": flips, "matches_compared": total, "max_rel_distance_error": max_rel,
                    "tolerance": "north_star: relative distance error <= 1e-5, index differences only at near ties (counted here; "
                                 "each one is verified against the oracle's float64 margin in tests/test_parity_gpu.py)"})
    out["superpoint_allpairs_100x4096x256_mutual"] = {
        "pairs": len(pairs), "rows_total": 2 * n * len(pairs), **sp["f16"], "tf32_path": sp["tf32"],
        "speedup_over_tf32_path": sp["tf32"]["ms_per_pass"] / sp["f16"]["ms_per_pass"], "cpu_baseline": cpu}
    del store
    # ---- configs[4], bounded: 48 of the 1000 images (1128 of 499500 pairs), AKAZE 61 bytes padded to 64
    I, n, nb = 48, 16384, 61
    desc = binary_set_gpu(torch, eng.device, I, n, nb)
    desc = torch.nn.functional.pad(desc.view(I * n, nb), (0, 3))
    pairs = ap.all_pairs(I)
    popc_peak = 148 * 16 * pk.get("sm_max_mhz", 1965.0) * 1e6 / 16 / 1e9      # Gdist/s, 512-bit rows, at max clock
    int8_peak = 2.0 * pk["bf16_tflops_sustained"]
    res = {}
    for path in ("tensor", "popc"):
        e = eng if path == "tensor" else b.Engine(eng.device, hamming_path="popc")
        store = ap.build_store(e, desc, b.NORM_HAMMING, counts=[n] * I)
        timed_eng[0] = e
        ms, knn_ms, matches = timed_pass(store, pairs, 2 if path == "tensor" else 1)
        gd = len(pairs) * n * n / 1e9 / (ms / 1e3)
        res[path] = {"ms_per_pass": ms, "pairs_per_s": len(pairs) / (ms / 1e3), "gdist_per_s": gd, "matches": matches,
                     "knn_stage_ms": knn_ms, "full_config_estimate_s_1gpu": 499500 / (len(pairs) / (ms / 1e3))}
        del store
    timed_eng[0] = eng
    tops = 2.0 * 2 * len(pairs) * n * n * 512 / (res["tensor"]["knn_stage_ms"] / 1e3) / 1e12     # both directions, 512 one-bit dims
    fp4_peak = 4.0 * pk["bf16_tflops_sustained"]
    res["tensor"].update({
        "tops_achieved_knn_stage": tops, "fp4_peak_tops": fp4_peak, "frac_of_fp4_roofline": tops / fp4_peak,
        "frac_of_hw_mxf4_sustained_microbench": tops / 7850.0, "frac_of_hw_mxf4_burst_microbench": tops / 8876.0,
        "int8_peak_tops": int8_peak, "frac_of_int8_roofline": tops / int8_peak,
        "peak_source": "4 x bf16_tflops_sustained of MEASURED_PEAKS.json (nominal fp4 : bf16 = 4 : 1); tools/mma_peak.cu kind::mxf4 issue rate "
                       "16384 MAC/clk/SM = 8876 TOP/s burst, 7850 sustained under the power cap (profiles/r01_mma_mxf4_peak.txt); the "
                       "int8 denominators (2 x bf16 sustained) are what the kind::i8 variant of the kernel is held to",
        "note": "default path (hamming_tc.cu, B200M_KIND_HAMMING_FP4): bits as e2m1 +-1.0 nibbles, tcgen05 kind::mxf4 with unit scale factors, "
                "K = 512, forward and backward knn on the tensor cores, exact POPC resolve of two 16-column chunks per row; "
                "knn stage = both directions + resolve"})
    res["popc"].update({
        "popc_roofline_gdist_per_s": popc_peak, "frac_of_popc_roofline": res["popc"]["gdist_per_s"] / popc_peak,
        "note": "INT-pipe kernel (hamming.cu): each distance computed once (fused row top-2 + column minima); POPC roofline at 1965 MHz"})
    hp = desc[:3 * n, :nb].cpu().numpy().reshape(3, n, nb)
    cpu, _ = cv2_pair_baseline([(hp[0], hp[1]), (hp[0], hp[2]), (hp[1], hp[2])], 6, True, None,
                               "16384 x 16384 x 61 B AKAZE (BASELINE configs[4]), crossCheck=True")
    out["akaze_allpairs_sample_48x16384x61B_mutual"] = {
        "pairs": len(pairs), "of_pairs": 499500, "cpu_baseline": cpu, **res["tensor"], "popc_path": res["popc"],
        "speedup_over_popc_path": res["popc"]["ms_per_pass"] / res["tensor"]["ms_per_pass"],
        "matches_d2h_bytes_compact": res["tensor"]["matches"] * 6}
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    b = importlib.import_module("3dreconstruction_b200")
    ap = importlib.import_module("3dreconstruction_b200.allpairs")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    json_out = sys.stdout
    if world > 1:
        # stdout must hold the ONE JSON line: NCCL prints its version banner on file descriptor 1 (at NCCL_DEBUG=VERSION and
        # above, WARN included), so fd 1 is pointed at stderr for the run and the line goes to a saved copy of the real stdout
        sys.stdout.flush()
        json_out = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=device)
    eng = b.Engine(device)
    lib = eng.lib
    pk, pk_src = peaks()
    I, n, D = args.images, args.keypoints, 128
    pairs = ap.all_pairs(I)
    P = len(pairs)

    # ---- inputs: rank 0 synthesises float32 SIFT-like descriptors, prepares the int8 store, broadcasts it
    # every rank holds the host rows (same seed): the e2e arm uploads 1/N of each image group per rank over its own PCIe link
    host_pinned = None
    store = None
    if rank == 0 or args.e2e_upload != "rank0":
        desc = sift_set_gpu(torch, device, I, n)                       # [I, n, 128] float32 on the device
        host_pinned = torch.empty((I * n, D), dtype=torch.float32, pin_memory=True)
        host_pinned.copy_(desc.view(I * n, D))                          # the e2e arm starts from HOST buffers
        if rank == 0:
            store = ap.build_store(eng, desc.view(I * n, D), b.NORM_L2, counts=[n] * I)
        del desc
    if world > 1:
        store = ap.broadcast_store(store, b.NORM_L2, device)
    assert store.dd.kind == b._lib.KIND_L2_I8
    shards = ap.partition_pairs(pairs, store.offsets, world)
    mine = shards[rank]
    matcher = ap.AllPairsMatcher(eng, chunk_pairs=args.chunk_pairs)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step():
        return matcher.match(store, mine, cross_check=False, ratio_threshold=RATIO)

    for _ in range(max(args.warmup, 3)):
        res = step()
    matches_total = int(res.query_idx.numel())
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.b200m_reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        res = step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    launches = torch.tensor([lib.b200m_launch_count()], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    clocks = sampler.stop() if rank == 0 else None
    # ---- the dominant kernel alone: the same K steps once more with the two-stream chunk pipeline OFF, so that the
    # CUDA events the library records around each knn launch (on the launching stream) bracket that kernel only
    import ctypes
    seq = ap.AllPairsMatcher(eng, chunk_pairs=args.chunk_pairs, pipeline=False)
    seq.match(store, mine, cross_check=False, ratio_threshold=RATIO)
    barrier()
    lib.b200m_timing_enable(1)
    ek0, ek1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ek0.record()
    for _ in range(args.steps):
        seq.match(store, mine, cross_check=False, ratio_threshold=RATIO)
    ek1.record()
    barrier()
    knn_ms, knn_launches = ctypes.c_double(0), ctypes.c_int(0)
    lib.b200m_timing_collect(ctypes.byref(knn_ms), ctypes.byref(knn_launches))
    lib.b200m_timing_enable(0)
    seq_ms_total = ek0.elapsed_time(ek1)
    ms_total = float(ms.item())
    value = P * args.steps / (ms_total / 1e3)

    # ---- end to end: HOST float32 descriptors -> H2D -> prep -> all-gather -> match -> gather to rank 0 -> D2H records
    dmod = importlib.import_module("3dreconstruction_b200.distributed")
    streaming = dmod.DistributedAllPairs(eng, chunk_pairs=args.chunk_pairs, group_images=args.group_images,
                                         gather=args.e2e_gather, upload=args.e2e_upload)
    matches_all = torch.tensor([matches_total], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(matches_all, op=dist.ReduceOp.SUM)
    matches_all = int(matches_all.item())
    last = {}

    def e2e_step(timing=False):
        h = streaming.match_host(host_pinned, [n] * I, b.NORM_L2, False, RATIO, D=D, dtype=torch.float32, timing=timing)
        if h is None:
            return 0
        last["h"] = h
        if args.e2e_gather == "rank0" and (int(h.ptr[-1]) != matches_all or len(h.pairs) != P):
            raise RuntimeError(f"e2e path found {int(h.ptr[-1])} matches, device-resident path {matches_all}")
        return h.nbytes

    e2e_step()
    e2e_step()
    barrier()
    e2e_steps = max(1, args.steps)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        d2h = e2e_step()
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = P * e2e_steps / float(t_e2e.item())
    e2e_step(timing=True)                                  # one more step with per-phase CUDA events (outside the timed region)
    barrier()
    phases = dict(streaming.last_phases)
    ph_all = [None] * world
    if world > 1:
        dist.all_gather_object(ph_all, phases)
    else:
        ph_all = [phases]
    # element-wise parity of the sharded end-to-end result with the single-GPU resident path (rank 0 matches ALL pairs once)
    parity = None
    if rank == 0 and args.e2e_gather == "rank0" and not args.no_parity:
        full = matcher.match(store, pairs, cross_check=False, ratio_threshold=RATIO).to_host()
        h = last["h"].sorted_by_pair()
        same = (np.array_equal(h.pairs, pairs) and np.array_equal(h.ptr, full.ptr) and np.array_equal(h.rec["q"], full.query_idx)
                and np.array_equal(h.rec["t"], full.train_idx)
                and np.array_equal(h.rec["d"].view(np.uint32), full.distance.view(np.uint32)))
        if not same:
            raise RuntimeError("end-to-end (sharded) result differs from the single-GPU resident result")
        parity = f"identical to the single-GPU resident result: {len(pairs)} pairs, ptr, queryIdx, trainIdx, distance bits of {int(full.ptr[-1])} matches"
        del full, h

    if rank == 0:
        my_pairs = len(mine)
        ops = 2.0 * n * n * D * my_pairs * args.steps                 # int8 MACs x2, rank 0's share
        achieved = ops / (knn_ms.value / 1e3) / 1e12 if knn_ms.value > 0 else None
        peak = HW_INT8_TOPS_SUSTAINED
        peak_bf16x2 = 2.0 * pk["bf16_tflops_sustained"]
        line = {"metric": "image pairs/s (SIFT-L2 all-pairs, knn k=2 + ratio 0.75)", "value": value, "unit": "pairs/s",
                "gdist_per_s": value * n * n / 1e9, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "int8", "data": "synthetic", "config": workload_config(args),
                "matches_per_step": matches_all,
                "roofline": {"bound": "tensor", "kernel": "l2_i8_fold_kernel (tcgen05 cta_group::2 kind::i8 M256 N256 over a CTA pair; 4 descriptor K-steps + 1 folded-norm K-step, "
                                                        "achieved counts the 4 algorithmic ones)", "achieved": achieved,
                             "peak": peak, "unit": "TOP/s int8 (TFLOP/s-equivalent)", "frac": achieved / peak if achieved else None,
                             "peak_source": "measured int8 tensor peak of this pool's B200, sustained under the 1000 W power cap: tools/mma_peak.cu, "
                                            "back-to-back tcgen05.mma kind::i8 M128 N256 for 6 s (profiles/r01_mma_int8_sustained.txt); SURVEY 8(d) "
                                            "prescribes the measured int8 peak once it exists, 2 x bf16_tflops_sustained before that (kept beside it)",
                             "peak_2x_bf16_sustained": peak_bf16x2,
                             "frac_of_2x_bf16_sustained": achieved / peak_bf16x2 if achieved else None,
                             "peak_2x_bf16_source": f"2 x bf16_tflops_sustained of {pk_src} MEASURED_PEAKS.json",
                             "peak_hw_int8_microbench": HW_INT8_TOPS,
                             "frac_of_hw_int8_microbench": achieved / HW_INT8_TOPS if achieved else None,
                             "peak_hw_source": "tools/mma_peak.cu on this pool's B200: back-to-back tcgen05.mma kind::i8 M128 N256, "
                                               "8192 MAC/clk/SM at 1965 MHz (profiles/r01_mma_peak_microbench.txt); "
                                               "ncu's sm__ops_path_tensor_op_utcimma pct uses the same hardware rate",
                             "kernel_ms_per_step": knn_ms.value / args.steps, "kernel_launches": knn_launches.value,
                             "kernel_share_of_step": knn_ms.value / seq_ms_total,
                             "kernel_timing": "per-launch CUDA events on the launching stream over a second pass of the same K steps with the "
                                              "two-stream chunk pipeline off (sequential step %.1f ms); with the pipeline on, the events of one "
                                              "stream would include waiting for the other stream's kernel" % (seq_ms_total / args.steps),
                             "algorithmic_ops_per_pair": 2.0 * n * n * D,
                             "algorithmic_hbm_bytes_per_pair": 2 * n * (D + 36) + 64 * n,
                             "traffic": NCU_DRAM_BYTES_PER_PAIR * my_pairs * args.steps / max(knn_launches.value, 1),
                             "traffic_source": "ncu --set full of the shipped l2_i8_fold_kernel (CTA pairs), one 2048-pair launch of this workload: dram read "
                                               "2716.6 MB + write 268.5 MB (profiles/r02_l2i8_pair_final_ncu_full.md) = 1.46 MB/pair, scaled to this run's "
                                               "pairs per launch; below the 3.2 MB/pair of algorithmic bytes because the store's rows are re-read from L2 "
                                               "(hit rate 93 %)"},
                "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": int(host_pinned.numel() * 4),
                        "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": 1e3 * float(t_e2e.item()) / e2e_steps,
                        "record_bytes_per_match": 8,
                        "path": ("distributed.DistributedAllPairs.match_host: pinned host float32 descriptors -> H2D in image groups on a copy "
                                 "stream -> b200m_prep_l2_i8_rows -> b200m_match_pairs_rec on the resident groups (two compute streams) -> "
                                 "per-chunk D2H of packed 8-byte records into pinned memory" if world == 1 else
                                 ("every rank: its 1/N block of each image group, pinned host float32 -> H2D on its own PCIe link -> "
                                  "b200m_prep_l2_i8_rows -> ONE in-place NCCL all-gather of the int8 rows per group -> norms / fold digits "
                                  "derived locally (b200m_l2_i8_aux_rows); " if args.e2e_upload != "rank0" else
                                  "rank 0: pinned host float32 descriptors -> H2D in image groups -> b200m_prep_l2_i8_rows -> NCCL broadcast "
                                  "of the int8 rows per group -> aux derived locally; ") +
                                 "every rank: b200m_match_pairs_rec on its slice of the resident groups' pairs; per group, one group behind the "
                                 "matching: NCCL send of the packed records to rank 0 -> D2H into pinned memory" +
                                 ("" if args.e2e_gather == "rank0" else " [gather=local: every rank keeps its records in its own host memory]")),
                        "upload": "n/a" if world == 1 else ("sharded" if args.e2e_upload != "rank0" else "rank0"),
                        "gather": "n/a" if world == 1 else args.e2e_gather, "group_images": args.group_images,
                        "parity_vs_single_gpu": parity,
                        "phases_rank0": ph_all[0],
                        "phases_max_over_ranks": {k: max(p.get(k, 0.0) for p in ph_all) for k in ph_all[0]},
                        "phases_note": "CUDA events per phase over one extra step (sum of the phase's spans on its own stream; phases overlap): "
                                       "front_ms = call start -> first matching kernel, tail_ms = last matching kernel -> last byte in host memory"},
                "gpu_launches": int(launches.item()), "clocks": clocks}
        if world == 1:
            imgs = host_sample_images(args, 4)
            sample = [(0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)] * max(1, args.cpu_sample_pairs // 6)
            t, cores, engine = cpu_time_pairs(imgs, sample)
            line["cpu_baseline"] = {"value": len(sample) / t, "unit": "pairs/s", "cores": cores, "kind": "port",
                                    "sample": f"{len(sample)} pairs of {n}x{n}x128 on this host, {engine} driven by "
                                              "oracle/bf_oracle.cv2_match (restates EnhancedBFMatcher.match)"}
            if not args.no_also:
                line["also"] = single_pair_extras(torch, b, eng, pk)
        print(json.dumps(line), file=json_out, flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_b200_akaze(args):
    """BASELINE configs[4]: AKAZE 61-byte (486-bit) descriptors, all pairs of `--images` images x `--keypoints` keypoints with
    the mutual check, sharded over the ranks.  `value`: device-resident store, every rank matches its balanced block of the
    pair list (results stay on the device and are dropped).  `e2e`: distributed.DistributedAllPairs from pinned host rows
    (sharded upload, one all-gather of the packed bits per image group, e2m1 expansion derived locally), 6-byte match
    records streamed to rank 0 per image group and handed to a consumer there (rank 0 never holds the whole result)."""
    import ctypes
    import torch
    import torch.distributed as dist
    b = importlib.import_module("3dreconstruction_b200")
    ap = importlib.import_module("3dreconstruction_b200.allpairs")
    dmod = importlib.import_module("3dreconstruction_b200.distributed")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    json_out = sys.stdout
    if world > 1:
        sys.stdout.flush()
        json_out = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=device)
    eng = b.Engine(device)
    lib = eng.lib
    pk, pk_src = peaks()
    I, n, nb = args.images, args.keypoints, 61
    pairs = ap.all_pairs(I)
    P = len(pairs)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # every rank synthesises the same set (same seed) on its GPU and keeps a pinned host copy: the e2e arm starts from HOST rows
    desc = binary_set_gpu(torch, device, I, n, nb)                        # [I, n, 61] uint8
    host_pinned = torch.empty((I * n, nb), dtype=torch.uint8, pin_memory=True)
    host_pinned.copy_(desc.view(I * n, nb))
    store = ap.build_store(eng, torch.nn.functional.pad(desc.view(I * n, nb), (0, 3)), b.NORM_HAMMING, counts=[n] * I)
    del desc
    mine = ap.partition_pairs(pairs, store.offsets, world)[rank]
    matcher = ap.AllPairsMatcher(eng, chunk_pairs=args.chunk_pairs)
    SUPER = 8 * args.chunk_pairs                                          # results of this many pairs live on the device at a time

    def step():
        total = 0
        for c0 in range(0, len(mine), SUPER):
            r = matcher.match(store, mine[c0:c0 + SUPER], cross_check=True)
            total += int(r.query_idx.numel())
            del r
        return total

    matcher.match(store, mine[:min(len(mine), 2 * args.chunk_pairs + 1)], cross_check=True)     # bounded warm-up (buffers, attributes)
    for _ in range(max(args.warmup, 0)):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.b200m_reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        matches_mine = step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    launches = torch.tensor([lib.b200m_launch_count()], dtype=torch.int64, device=device)
    matches_all = torch.tensor([matches_mine], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(launches, op=dist.ReduceOp.SUM)
        dist.all_reduce(matches_all, op=dist.ReduceOp.SUM)
    clocks = sampler.stop() if rank == 0 else None
    ms_total, matches_all = float(ms.item()), int(matches_all.item())
    value = P * args.steps / (ms_total / 1e3)
    # the dominant kernel alone on a bounded sample of this rank's pairs (pipeline off: events bracket the kernel only)
    sample = mine[:min(len(mine), 2 * args.chunk_pairs)]
    seq = ap.AllPairsMatcher(eng, chunk_pairs=args.chunk_pairs, pipeline=False)
    seq.match(store, sample, cross_check=True)
    barrier()
    lib.b200m_timing_enable(1)
    seq.match(store, sample, cross_check=True)
    torch.cuda.synchronize()
    knn_ms, knn_l = ctypes.c_double(0), ctypes.c_int(0)
    lib.b200m_timing_collect(ctypes.byref(knn_ms), ctypes.byref(knn_l))
    lib.b200m_timing_enable(0)
    tops = 2.0 * 2 * len(sample) * n * n * 512 / (knn_ms.value / 1e3) / 1e12 if knn_ms.value > 0 else None   # both directions, 512 one-bit dims
    del store, seq
    matcher._bufs.clear()
    torch.cuda.empty_cache()

    # ---- end to end from host rows; rank 0 consumes the records group by group (counts them, keeps three sampled pairs for the oracle)
    dap = dmod.DistributedAllPairs(eng, chunk_pairs=args.chunk_pairs, group_images=args.group_images, gather=args.e2e_gather,
                                   upload=args.e2e_upload)
    seen = {"matches": 0, "pairs": 0, "blocks": 0, "keep": {}}
    want_keys = np.array([0 * I + 1, (I // 3) * I + I // 2, (I - 2) * I + I - 1], np.int64)

    def consumer(pr, cn, rec):
        seen["matches"] += int(cn.sum(dtype=np.int64))
        seen["pairs"] += len(pr)
        seen["blocks"] += 1
        if args.no_parity:
            return
        hit = np.nonzero(np.isin(pr[:, 0].astype(np.int64) * I + pr[:, 1], want_keys))[0]
        if len(hit):
            ptr = np.concatenate([[0], np.cumsum(cn, dtype=np.int64)])
            for k in hit:
                seen["keep"][(int(pr[k, 0]), int(pr[k, 1]))] = rec[ptr[k]:ptr[k + 1]].copy()

    def e2e_step(timing=False):
        seen.update({"matches": 0, "pairs": 0, "blocks": 0})
        r = dap.match_host(host_pinned, [n] * I, b.NORM_HAMMING, True, None, D=nb, dtype=torch.uint8, consumer=consumer, keep=False,
                           timing=timing)
        if rank == 0 and args.e2e_gather == "rank0" and (seen["pairs"] != P or seen["matches"] != matches_all or int(r.ptr[-1]) != matches_all):
            raise RuntimeError(f"e2e path returned {seen['matches']} matches of {seen['pairs']} pairs, resident path {matches_all} of {P}")
        return seen["matches"] * 6 + (0 if r is None else r.ptr.nbytes)

    if args.warmup > 0 or I <= 64:
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        d2h = e2e_step(timing=args.steps == 1)
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = P * args.steps / float(t_e2e.item())
    if args.steps != 1:
        e2e_step(timing=True)
        barrier()
    ph_all = [None] * world
    if world > 1:
        dist.all_gather_object(ph_all, dict(dap.last_phases))
    else:
        ph_all = [dict(dap.last_phases)]
    if rank == 0:
        parity = None
        if not args.no_parity and args.e2e_gather == "rank0":
            from oracle import bf_oracle as O
            O.set_threads(os.cpu_count() or 1)
            hp = host_pinned.numpy().reshape(I, n, nb)
            for (i, j), rec in seen["keep"].items():
                w = O.reference_match(hp[i], hp[j], O.NORM_HAMMING, True, None)
                ok = (np.array_equal(rec["q"].astype(np.int32), w.query_idx) and np.array_equal(rec["t"].astype(np.int32), w.train_idx)
                      and np.array_equal(rec["d"].astype(np.float32), w.distance))
                if not ok:
                    raise RuntimeError(f"e2e records of pair {(i, j)} differ from the oracle")
            parity = f"pairs {sorted(seen['keep'])} of the streamed result identical to the oracle (queryIdx, trainIdx, distance, order)"
        fp4_peak = 4.0 * pk["bf16_tflops_sustained"]
        full = (I, n) == (1000, 16384)
        line = {"metric": "image pairs/s (AKAZE-Hamming all-pairs, mutual cross-check)", "value": value, "unit": "pairs/s",
                "gdist_per_s": value * n * n / 1e9, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "e2m1 (+-1.0 per descriptor bit; fp32 sums of +-1 are exact)", "data": "synthetic",
                "config": {"workload": f"AKAZE binary all-pairs: {I} images x {n} keypoints x 61 B, {P} pairs, mutual cross-check "
                                       f"(BASELINE configs[4]{'' if full else ', REDUCED size'})",
                           "pairs": P, "l2_flush": "inputs larger than L2 (e2m1 store %.2f GB)" % (I * n * 256 / 1e9),
                           "parallelism": f"pair-list sharding x{world}"},
                "matches_per_step": matches_all,
                "roofline": {"bound": "tensor", "kernel": "hamming_tc_kernel<2, CS, FP4> (tcgen05 kind::mxf4.block_scale, forward + backward) + POPC resolve",
                             "achieved": tops, "peak": 7850.0, "unit": "TOP/s (one-bit MACs x 2)", "frac": tops / 7850.0 if tops else None,
                             "peak_source": "tools/mma_peak.cu kind::mxf4 issue rate sustained under the power cap on this pool's B200 "
                                            "(profiles/r01_mma_mxf4_peak.txt; 8876 burst); 4 x bf16_tflops_sustained = %.0f beside it" % fp4_peak,
                             "frac_of_4x_bf16_sustained": tops / fp4_peak if tops else None,
                             "sample": f"{len(sample)} pairs of rank 0's block, pipeline off, library CUDA events around the knn stage",
                             "kernel_ms_sample": knn_ms.value, "kernel_launches_sample": knn_l.value, "traffic": None},
                "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": int(host_pinned.numel()), "d2h_bytes_per_step": int(d2h),
                        "steps": args.steps, "ms_per_step": 1e3 * float(t_e2e.item()) / args.steps, "record_bytes_per_match": 6,
                        "blocks_streamed_to_rank0": seen["blocks"],
                        "upload": "n/a" if world == 1 else ("sharded" if args.e2e_upload != "rank0" else "rank0"),
                        "gather": "n/a" if world == 1 else args.e2e_gather, "group_images": args.group_images,
                        "path": "pinned host uint8 rows -> H2D per image group (every rank its block) -> NCCL all-gather of the packed bits -> "
                                "e2m1 expansion derived locally -> b200m_match_pairs_rec (6-byte records) -> per group NCCL send to rank 0 -> "
                                "D2H into a two-slot pinned ring -> consumer (rank 0 never holds the whole result)",
                        "parity_vs_oracle": parity, "phases_rank0": ph_all[0],
                        "phases_max_over_ranks": {k: max(p.get(k, 0.0) for p in ph_all) for k in ph_all[0]}},
                "gpu_launches": int(launches.item()), "clocks": clocks}
        if world == 1 and not args.no_also:
            try:
                import cv2
                cv2.setNumThreads(os.cpu_count() or 1)
                hp = host_pinned.numpy().reshape(I, n, nb)
                bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)
                bf.match(hp[0][:512], hp[1][:512])
                k = max(1, min(args.cpu_sample_pairs // 6, I - 1))
                t0 = time.perf_counter()
                for j in range(1, k + 1):
                    bf.match(hp[0], hp[j])
                t = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": k / t, "unit": "pairs/s", "cores": cv2.getNumThreads(), "kind": "reference",
                                        "sample": f"{k} pairs of {n}x{n}x61 B on this host, cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match "
                                                  f"{cv2.__version__} (what EnhancedBFMatcher.match calls at feature_matchers.py:227)"}
            except ImportError:
                line["cpu_baseline"] = None
        print(json.dumps(line), file=json_out, flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "akaze":
        run_b200_akaze(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
