#!/usr/bin/env python3
"""The box's host<->device copy ceiling for the e2e arm's traffic pattern, without any kernel: every rank copies
blocks of the e2e step's sizes (pinned host memory, cudaMemcpyAsync through torch) - H2D of --h2d-mb and D2H of
--d2h-mb per step, D steps in flight on separate streams - and rank 0 prints the aggregate GB/s. Run under
torchrun with N = 1, 2, 4, 8 (python -m torch.distributed.run --nproc-per-node N tools/pcie_ceiling.py); the result
lines are committed under profiles/ and bench.py prints the matching ceiling beside e2e.pcie_gbs_per_gpu."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h2d-mb", type=float, default=23.15)
    ap.add_argument("--d2h-mb", type=float, default=6.64)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--in-flight", type=int, default=4)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nh, nd = int(a.h2d_mb * 1e6), int(a.d2h_mb * 1e6)
    D = a.in_flight
    h_in = [torch.empty(nh, dtype=torch.uint8).pin_memory() for _ in range(D)]
    h_out = [torch.empty(nd, dtype=torch.uint8).pin_memory() for _ in range(D)]
    d_in = [torch.empty(nh, dtype=torch.uint8, device="cuda") for _ in range(D)]
    d_out = [torch.empty(nd, dtype=torch.uint8, device="cuda") for _ in range(D)]
    streams = [torch.cuda.Stream() for _ in range(D)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(steps, what):
        for j in range(steps):
            k = j % D
            streams[k].synchronize()          # "finish" of the batch that used this slot
            with torch.cuda.stream(streams[k]):
                if what in ("both", "h2d"):
                    d_in[k].copy_(h_in[k], non_blocking=True)
                if what in ("both", "d2h"):
                    h_out[k].copy_(d_out[k], non_blocking=True)
        torch.cuda.synchronize()

    out = {"n_gpus": world, "h2d_mb_per_step": a.h2d_mb, "d2h_mb_per_step": a.d2h_mb, "in_flight": D, "steps": a.steps}
    for what in ("both", "h2d", "d2h"):
        run(20, what)
        barrier()
        t0 = time.perf_counter()
        run(a.steps, what)
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        nbytes = (nh if what != "d2h" else 0) + (nd if what != "h2d" else 0)
        out[what + "_gbs_per_gpu"] = round(nbytes * a.steps / dt / 1e9, 2)
        out[what + "_gbs_aggregate"] = round(nbytes * a.steps * world / dt / 1e9, 2)
        barrier()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
