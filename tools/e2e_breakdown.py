#!/usr/bin/env python3
"""Where the end-to-end step goes on the host: wall clock spent in enqueue (H2D copies are enqueued there), finish
(waiting for kernels + D2H) and read (copying the live rows out of the pinned arena) per batch, D batches in flight.
Run on the GPU box: python tools/e2e_breakdown.py [config] [--steps N]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", nargs="?", default="euroc_sequence")
    ap.add_argument("--steps", type=int, default=40)
    a = ap.parse_args()
    import torch
    import __graft_entry__ as g
    mod = g.build()
    syn = __import__("importlib").import_module("okvis2x_b200.synthetic")
    args = bench.parse_args([])
    ctx = mod.Context(0)
    wl = bench.build_images(syn, a.config, False)
    gc = bench.GpuConfig(args, mod, syn, torch, ctx, wl, 0, 1, False)
    gc.build_world()
    gc.precompute_poses()
    gc.setup()
    F, D = gc.F, gc.D
    gc.run_e2e(0, 2 * D * F)
    t = {"slot_args": 0.0, "enqueue": 0.0, "finish": 0.0, "read": 0.0}
    pending = []
    torch.cuda.synchronize()
    t_all = time.perf_counter()
    for j in range(a.steps):
        b = gc.batches[j % D]
        if len(pending) == D:
            pb = pending.pop(0)
            t0 = time.perf_counter()
            pb.finish()
            t1 = time.perf_counter()
            gc.read(pb)
            t2 = time.perf_counter()
            t["finish"] += t1 - t0
            t["read"] += t2 - t1
        t0 = time.perf_counter()
        sa = gc.slot_args(j * F, False)
        t1 = time.perf_counter()
        gc.enqueue(b, *sa, False, j * F)
        t2 = time.perf_counter()
        t["slot_args"] += t1 - t0
        t["enqueue"] += t2 - t1
        pending.append(b)
    for pb in pending:
        t0 = time.perf_counter()
        pb.finish()
        t1 = time.perf_counter()
        gc.read(pb)
        t["finish"] += t1 - t0
        t["read"] += time.perf_counter() - t1
    total = time.perf_counter() - t_all
    h2d, d2h = gc.bytes_per_frame()
    print(json.dumps({"config": a.config, "ms_per_step": round(total / a.steps * 1e3, 3),
                      "host_ms_per_step": {k: round(v / a.steps * 1e3, 3) for k, v in t.items()},
                      "h2d_mb_per_step": round(h2d * F / 1e6, 2), "d2h_mb_per_step": round(d2h * F / 1e6, 2)}))


if __name__ == "__main__":
    main()
