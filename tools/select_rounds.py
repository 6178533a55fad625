"""GPU-box tool: kernel times of the uniformity selection (default dependency-form kernel vs the sequential-visit
baseline) per config image; `rounds` = sweeps of warp 0 (diagnostic counter 5)."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
import importlib, ctypes as C
mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")
ctx = mod.Context(0)
cases = {"euroc-scene": (syn.EUROC, syn.scene_image(752, 480, 1)), "euroc-noise": (syn.EUROC, syn.noise_image(752, 480, 2)),
         "d455-scene": (syn.D455, syn.scene_image(848, 480, 8, rects=1500)), "hilti-scene": (syn.HILTI, syn.scene_image(720, 540, 3)),
         "euroc-dense": (syn.EUROC, syn.dense_image(752, 480, 8)), "d455-dense": (syn.D455, syn.dense_image(848, 480, 8)),
         "hilti-dense": (syn.HILTI, syn.dense_image(720, 540, 8)), "d455-noise": (syn.D455, syn.noise_image(848, 480, 9))}
out = {}
for name, (cfg, img) in cases.items():
    h, w = img.shape
    cam = mod.Camera(ctx, w, h, max_keypoints=4096)
    b = mod.Batch(ctx, [cam])
    res = {}
    for wave in (0, 1):
        mod.lib().okc_debug_set_select_wave(wave)
        b.kernel_timing(True)
        ts = []
        for it in range(6):
            b.enqueue([img], cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"], [(0.0, 1.0, 0.0)])
            ts.append(dict(b.kernel_times()))
            b.finish()
        v = [C.c_int(0) for _ in range(3)]
        for k, which in enumerate((0, 1, 5)):
            mod.lib().okc_debug_batch_counter(b.h, 0, which, C.byref(v[k]))
        res[("wave", "seq")[wave]] = {"select_us": round(1e3 * float(np.median([t["select_keypoints"] for t in ts[2:]])), 1),
                                         "candidates": v[0].value, "kept": v[1].value, "rounds": v[2].value}
        b.kernel_timing(False)
    out[name] = res
    b.close(); cam.close()
mod.lib().okc_debug_set_select_wave(0)
print(json.dumps(out, indent=1))
