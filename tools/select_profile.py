"""GPU-box tool: run the detector a few times on one config image so that ncu can capture the selection kernel.
usage: python tools/select_profile.py [euroc-noise|euroc-scene|d455-noise|d455-scene] [batch]"""
import os, sys, importlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")
name = sys.argv[1] if len(sys.argv) > 1 else "euroc-noise"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cases = {"euroc-scene": (syn.EUROC, lambda s: syn.scene_image(752, 480, 1 + s)),
         "euroc-noise": (syn.EUROC, lambda s: syn.noise_image(752, 480, 2 + s)), "euroc-dense": (syn.EUROC, lambda s: syn.dense_image(752, 480, 8 + s)), "d455-dense": (syn.D455, lambda s: syn.dense_image(848, 480, 8 + s)),
         "d455-scene": (syn.D455, lambda s: syn.scene_image(848, 480, 8 + s, rects=1500)),
         "d455-noise": (syn.D455, lambda s: syn.noise_image(848, 480, 9 + s))}
cfg, gen = cases[name]
imgs = [gen(i) for i in range(B)]
h, w = imgs[0].shape
ctx = mod.Context(0)
cams = [mod.Camera(ctx, w, h, max_keypoints=4096) for _ in range(B)]
b = mod.Batch(ctx, cams)
for it in range(5):
    b.enqueue(imgs, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"], [(0.0, 1.0, 0.0)] * B)
    b.finish()
print("ok")
