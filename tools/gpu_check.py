#!/usr/bin/env python3
"""Quick GPU parity sweep against the oracle (development aid; the real tests live in tests/)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g  # noqa: E402

mod = g.load()
from okvis2x_b200 import synthetic as syn  # noqa: E402
import refbind  # noqa: E402

ref, port = refbind.Ref(), refbind.Port()
ctx = mod.Context(0)
fails = 0


def check(name, ok, extra=""):
    global fails
    print("%-44s %s %s" % (name, "OK  " if ok else "FAIL", extra), flush=True)
    if not ok:
        fails += 1


def kp_close(a, b):
    if len(a) != len(b):
        return False, "n %d vs %d" % (len(a), len(b))
    ok = np.array_equal(a["response"], b["response"]) and np.array_equal(a["octave"], b["octave"])
    dx = float(np.abs(a["x"] - b["x"]).max()) if len(a) else 0.0
    dy = float(np.abs(a["y"] - b["y"]).max()) if len(a) else 0.0
    da = float(np.abs(a["angle"] - b["angle"]).max()) if len(a) else 0.0
    exact = a.tobytes() == b.tobytes()
    return ok and dx <= 1e-4 and dy <= 1e-4 and da <= 1e-4, "n=%d dx=%.2g dy=%.2g dang=%.2g bitexact=%s" % (
        len(a), dx, dy, da, exact)


cases = [("euroc-scene", syn.EUROC, syn.scene_image(752, 480, 1)),
         ("euroc-noise", syn.EUROC, syn.noise_image(752, 480, 2)),
         ("d455-scene", syn.D455, syn.scene_image(848, 480, 8, rects=1500)),
         ("hilti-scene", syn.HILTI, syn.scene_image(720, 540, 3))]
cam_cfg = syn.EUROC["cams"][0]
for name, cfg, img in cases:
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    sc = cam.debug_scores(img)
    check(name + " harris scores", np.array_equal(sc, ref.harris(img)))
    thr = int(cfg["abs_threshold"])
    cand = cam.debug_candidates(img, thr)
    rc = ref.maxima(img, thr)
    rkey = np.lexsort((rc["y"].astype(np.int64) * W + rc["x"], -rc["score"].astype(np.int64)))
    check(name + " nms candidates", len(cand) == len(rc) and cand.tobytes() == rc[rkey].tobytes(), "n=%d" % len(rc))
    ties = len(rc) - len(np.unique(rc["score"]))
    kp = cam.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    rkp = ref.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    ok, extra = kp_close(kp, rkp)
    check(name + " detect keypoints", ok, extra + " score-ties=%d" % ties)
    # plain extraction (no camera maps): the reference's own test configuration
    ext = refbind.RefExtractor(ref)
    kd, dd = cam.describe(img, rkp)
    rkd, rdd = ext.describe(img, rkp)
    ok, extra = kp_close(kd, rkd)
    check(name + " describe plain keypoints", ok, extra)
    check(name + " describe plain descriptors", np.array_equal(dd, rdd),
          "mismatching rows=%d" % (int((dd != rdd).any(1).sum()) if dd.shape == rdd.shape else -1))
    kf, df = cam.detect_describe(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    check(name + " fused detect+describe", kf.tobytes() == kd.tobytes() and np.array_equal(df, dd))
    cam.close()
    # camera-aware extraction
    rays, jac = syn.camera_maps(W, H, cam_cfg["f"], cam_cfg["c"], cam_cfg["d"])
    camA = mod.Camera(ctx, W, H, rays, jac, focal=cam_cfg["f"][0], max_keypoints=4096)
    for dname, direction in (("down", (0.05, 0.99, 0.1)), ("axis", (0.0, 0.0, 1.0))):
        ext.set_camera(rays, jac, np.float32(cam_cfg["f"][0]))
        ext.set_direction(np.asarray(direction, np.float32))
        kd, dd = camA.describe(img, rkp, direction=direction)
        rkd, rdd = ext.describe(img, rkp)
        ok, extra = kp_close(kd, rkd)
        check(name + " describe aware/%s keypoints" % dname, ok, extra)
        check(name + " describe aware/%s descriptors" % dname, np.array_equal(dd, rdd),
              "mismatching rows=%d" % (int((dd != rdd).any(1).sum()) if dd.shape == rdd.shape else -1))
    camA.close()
    ext.close()

# matchers
q, m = syn.sweep_inputs(1500, 40000, seed=5)
t = time.time()
best = ctx.match_batched(q, m, 60)
check("match_batched 1500x40000", np.array_equal(best, port.hamming_best(q, m, 60, nthreads=8)),
      "matches=%d" % int((best != mod.NO_MATCH).sum()))
for pre in (4, 5, 8, 12, 6, 0):
    mod.lib().okc_debug_set_match_prefix(pre)
    check("match_batched prefix=%d" % pre, np.array_equal(ctx.match_batched(q, m, 60), best))
check("match_batched empty map", np.all(ctx.match_batched(q[:5], m[:0], 60) == mod.NO_MATCH))

img = syn.scene_image(752, 480, 1)
cam = mod.Camera(ctx, 752, 480, max_keypoints=4096)
kp, desc = cam.detect_describe(img, 38.0, 0, 150.0, 700)
kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
win = syn.map_window(desc, kp_xy, 2000, 7, 752, 480)
use = (np.arange(len(kp)) % 11 != 0).astype(np.uint8)
sel3d = win["is3d"][win["owner"]]
m3, o3 = win["map"][sel3d], win["owner"][sel3d]
f = 0.5 * (cam_cfg["f"][0] + cam_cfg["f"][1])
tsq = (3.0 + f * 0.06) ** 2
bd, bo, ctrs, errs = ctx.match_to_map(desc, kp_xy, use, m3, o3, win["proj"], tsq, 60.0, 4)
pd, po, pc, pe = port.match_to_map(desc, kp_xy, use, m3, o3, win["proj"], None, tsq, 60.0, 4)
check("match_to_map best", np.array_equal(bd, pd) and np.array_equal(bo, po), "matched=%d" % int((bo >= 0).sum()))
check("match_to_map ctrs/reprErrors", np.array_equal(ctrs, pc) and pe.tobytes() == errs.tobytes(),
      "%s %s" % (ctrs, errs))
cam.close()
print("FAILS:", fails)
sys.exit(1 if fails else 0)
