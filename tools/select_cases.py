"""GPU-box tool: device time of order_candidates and of the two selection kernels (dependency form = default,
one-candidate-at-a-time = A/B baseline) for ONE image of each test case: EuRoC / D455 parameters on a sparse scene, a
dense scene and i.i.d. noise. CUDA events around every launch (okc_debug_batch_kernel_timing). Prints one JSON object."""
import os, sys, json, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")
cases = {"euroc-scene": (syn.EUROC, lambda s: syn.scene_image(752, 480, 1 + s)),
         "euroc-dense": (syn.EUROC, lambda s: syn.dense_image(752, 480, 8 + s)),
         "euroc-noise": (syn.EUROC, lambda s: syn.noise_image(752, 480, 2 + s)),
         "d455-scene": (syn.D455, lambda s: syn.scene_image(848, 480, 8 + s, rects=1500)),
         "d455-dense": (syn.D455, lambda s: syn.dense_image(848, 480, 8 + s)),
         "d455-noise": (syn.D455, lambda s: syn.noise_image(848, 480, 9 + s))}
ctx = mod.Context(0)
out = {}
for name, (cfg, gen) in cases.items():
    img = gen(0)
    h, w = img.shape
    cam = mod.Camera(ctx, w, h, max_keypoints=4096)
    b = mod.Batch(ctx, [cam])
    row = {}
    for mode, label in ((0, "dependency_form_us"), (1, "sequential_us")):
        mod.lib().okc_debug_set_select_wave(mode)
        b.kernel_timing(True)
        acc = {}
        for it in range(12):
            b.enqueue([img], cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"], [(0.0, 1.0, 0.0)])
            b.finish()
            if it >= 2:
                for k, ms in b.kernel_times():
                    acc.setdefault(k, []).append(ms * 1e3)
        b.kernel_timing(False)
        row[label] = round(float(np.median(acc["select_keypoints"])), 1)
        row["order_candidates_us"] = round(float(np.median(acc["order_candidates"])), 1)
        row["harris_nms_us"] = round(float(np.median(acc["harris_nms"])), 1)
    mod.lib().okc_debug_set_select_wave(0)
    kp, _ = b.slot_frame(0)
    c = C = None
    import ctypes
    v = ctypes.c_int(0)
    mod.lib().okc_debug_batch_counter(b.h, 0, 0, ctypes.byref(v))
    row["candidates"], row["keypoints"], row["budget"] = v.value, len(kp), cfg["max_kpt"]
    out[name] = row
    b.close()
    cam.close()
print(json.dumps(out))
