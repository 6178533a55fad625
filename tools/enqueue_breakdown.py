#!/usr/bin/env python3
"""Host cost of enqueueing one batch, call by call (wall clock around each Python call = ctypes marshalling + the C
side's table uploads and kernel launches), for the throughput batch (F multiframes, images on the device) and for a
single multiframe. Run on the GPU box: python tools/enqueue_breakdown.py [config]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", nargs="?", default="euroc_sequence")
    ap.add_argument("--reps", type=int, default=200)
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
    out = {}
    for label in ("batch", "single"):
        if label == "single":
            gc.F = 1
            gc.setup_single()
        b = gc.batches[0]
        acc = {}
        for i in range(a.reps + 20):
            imgs, poses = gc.slot_args(i * gc.F, label == "batch")
            steps = []
            if gc.conf["landmarks"] or gc.pairs or gc.jobs:
                steps.append(("set_poses", lambda: b.set_poses(poses)))
            steps.append(("enqueue_addrs", lambda: b.enqueue_addrs(imgs, label == "batch")))
            if gc.conf["landmarks"]:
                steps.append(("match_landmarks_enqueue",
                              lambda: b.match_landmarks_enqueue(gc.pool, bench.HAMMING, gc.slices, mod.STAGE_UNINITIALISED)))
            if gc.pairs:
                steps.append(("match_stereo_enqueue", lambda: b.match_stereo_enqueue(gc.pairs, bench.HAMMING)))
            if gc.jobs:
                steps.append(("match_motion_enqueue", lambda: b.match_motion_enqueue(gc.jobs, bench.HAMMING, only_unmatched=True)))
            for name, fn in steps:
                t0 = time.perf_counter()
                fn()
                dt = time.perf_counter() - t0
                if i >= 20:
                    acc.setdefault(name, []).append(dt * 1e6)
            t0 = time.perf_counter()
            b.finish()
            dt = time.perf_counter() - t0
            if i >= 20:
                acc.setdefault("finish (waits for the GPU)", []).append(dt * 1e6)
        t0 = time.perf_counter()
        for _ in range(1000):
            mod.lib().okc_version()
        out[label] = {k: round(float(np.median(v)), 1) for k, v in acc.items()}
        out[label]["enqueue_total_us"] = round(sum(v for k, v in out[label].items() if not k.startswith("finish")), 1)
        out["ctypes_call_overhead_us"] = round((time.perf_counter() - t0) * 1e3, 3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
