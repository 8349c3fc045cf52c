#!/usr/bin/env python
"""tools/scale_akaze.py — bounded sample of BASELINE configs[4] (AKAZE/BRISK binary all-pairs with mutual check,
16k keypoints per image, 61-byte rows padded to 64) sharded over the GPUs of one box, to evidence the scaling of the
Hamming path.  Same plumbing as bench.py's SIFT workload: rank 0 builds the store, NCCL broadcast, balanced
contiguous pair blocks, per-rank CSR, gather to rank 0.

    python tools/scale_akaze.py --images 96                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        tools/scale_akaze.py --images 96

Prints one JSON line (rank 0): pairs/s, Gdist/s, ms per pass (CUDA events, max over ranks), matches."""
import argparse
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (binary_set_gpu)


def main():
    ap_ = argparse.ArgumentParser()
    ap_.add_argument("--images", type=int, default=96)
    ap_.add_argument("--keypoints", type=int, default=16384)
    ap_.add_argument("--reps", type=int, default=2)
    args = ap_.parse_args()
    import torch
    import torch.distributed as dist
    b = importlib.import_module("3dreconstruction_b200")
    ap = importlib.import_module("3dreconstruction_b200.allpairs")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=device)
    eng = b.Engine(device)
    I, n, nb = args.images, args.keypoints, 61
    store = None
    if rank == 0:
        desc = bench.binary_set_gpu(torch, device, I, n, nb)
        desc = torch.nn.functional.pad(desc.view(I * n, nb), (0, 3))
        store = ap.build_store(eng, desc, b.NORM_HAMMING, counts=[n] * I)
    if world > 1:
        store = ap.broadcast_store(store, b.NORM_HAMMING, device)
    pairs = ap.all_pairs(I)
    mine = ap.partition_pairs(pairs, store.offsets, world)[rank]
    m = ap.AllPairsMatcher(eng, chunk_pairs=512)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    r = m.match(store, mine, cross_check=True)
    if world > 1:
        ap.gather_csr(r, device)                     # warm-up: NCCL sets its point-to-point channels up on first use
    barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t_match = t_gather = 0.0
    for _ in range(args.reps):
        e0.record()
        r = m.match(store, mine, cross_check=True)
        e1.record()
        if world > 1:
            ap.gather_csr(r, device)
        e2.record()
        barrier()
        t_match += e0.elapsed_time(e1)
        t_gather += e1.elapsed_time(e2)
    ms = torch.tensor([(t_match + t_gather) / args.reps, t_match / args.reps, t_gather / args.reps], dtype=torch.float64, device=device)
    matches = torch.tensor([int(r.query_idx.numel())], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(matches, op=dist.ReduceOp.SUM)
    if rank == 0:
        msv, ms_match, ms_gather = (float(x) for x in ms.tolist())
        P = len(pairs)
        print(json.dumps({"workload": f"AKAZE all-pairs sample: {I} images x {n} x 61 B (padded 64), mutual check, {P} of 499500 pairs",
                          "n_gpus": world, "ms_per_pass": msv, "ms_match_max_rank": ms_match, "ms_gather_max_rank": ms_gather, "pairs_per_s": P / (msv / 1e3),
                          "gdist_per_s": P * n * n / 1e9 / (msv / 1e3), "matches": int(matches.item()),
                          "includes": "per-rank matching + NCCL gather of the CSRs to rank 0 (store broadcast once, outside)",
                          "full_config_estimate_s": 499500 / (P / (msv / 1e3))}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
