"""GPU-box tool: per-kernel device times of ONE stereo frame (batch of 2 images) of the bench workload."""
import os, sys, json, argparse
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, importlib
import __graft_entry__ as g
mod = g.load(); mod.lib()
syn = importlib.import_module("okvis2x_b200.synthetic")
wl = bench.build_workload(syn)
cfg, W, H = wl["cfg"], wl["W"], wl["H"]
det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
ctx = mod.Context(0)
cams = [mod.Camera(ctx, W, H, c["rays"], c["jac"], focal=c["focal"], max_keypoints=bench.MAX_KP) for c in wl["cams"]]
frame0 = [cams[c].detect_describe(wl["images"][c][0], *det, direction=bench.DIRECTION) for c in range(2)]
win_h = bench.make_windows(syn, wl, frame0)
windows = [mod.MapWindow(ctx, rows, owner, proj.shape[1]) for rows, owner, proj in win_h]
gates = [(3.0 + c["fmean"] * 0.06) ** 2 for c in wl["cams"]]
b = mod.Batch(ctx, cams)
b.kernel_timing(True)
acc = {}
for f in range(30):
    b.enqueue([wl["images"][0][f], wl["images"][1][f]], *det, [bench.DIRECTION] * 2, windows=windows,
              proj=[np.ascontiguousarray(win_h[c][2][f]) for c in range(2)], gates=gates, num_slices=cfg["slices"])
    for n, ms in b.kernel_times():
        acc.setdefault(n, []).append(ms)
    b.finish()
print(json.dumps({n: round(float(np.median(v[5:])) * 1e3, 1) for n, v in acc.items()}), "us; sum",
      round(sum(float(np.median(v[5:])) for v in acc.values()) * 1e3, 1))
