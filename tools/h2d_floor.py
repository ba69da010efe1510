#!/usr/bin/env python
"""Host->device floor of a multi-GPU box: every rank copies the same amount of page-locked host memory to its GPU at the same time
(plain cudaMemcpyAsync through torch, nothing else running), with and without binding the rank to its GPU's NUMA node.  What
bench.py's e2e path can reach at N GPUs is bounded by this; compare `e2e.ms_per_step` of an N-GPU bench line with `ms_per_14GB`.

    torchrun --nproc-per-node N tools/h2d_floor.py [--gb 4] > profiles/h2d_floor_Ngpu.json      (rank 0 prints one JSON line)
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def measure(gb, local, dist, reps=4):
    import torch
    n = int(gb * (1 << 30))
    host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    host.fill_(1)
    dev = torch.empty(n, dtype=torch.uint8, device="cuda")
    times = []
    for _ in range(reps + 1):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t = time.perf_counter()
        dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
        v = torch.tensor([dt], dtype=torch.float64, device="cuda"); dist.all_reduce(v, op=dist.ReduceOp.MAX)
        times.append(float(v.item()))
    del host, dev
    best = min(times[1:])
    return n / best / 1e9, best


if __name__ == "__main__":
    ap = argparse.ArgumentParser(); ap.add_argument("--gb", type=float, default=4.0); a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from bench import numa_bind
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {"n_gpus": world, "gb_per_gpu": a.gb, "cores": os.cpu_count()}
    per, t = measure(a.gb, local, dist)
    out["unbound"] = {"GBps_per_gpu_slowest": round(per, 2), "GBps_aggregate": round(per * world, 1), "ms_per_14GB": round(14.355 / per * 1e3, 1)}
    numa_bind(local)
    per, t = measure(a.gb, local, dist)
    out["numa_bound"] = {"GBps_per_gpu_slowest": round(per, 2), "GBps_aggregate": round(per * world, 1), "ms_per_14GB": round(14.355 / per * 1e3, 1)}
    if rank == 0:
        print(json.dumps(out))
    dist.barrier(); dist.destroy_process_group()
