import os, sys, importlib
sys.path.insert(0, "/root/repo")
import __graft_entry__ as ge
mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")
ctx = mod.Context(0)
for name, cfg, img in (("euroc-scene", syn.EUROC, syn.scene_image(752, 480, 1)), ("euroc-dense", syn.EUROC, syn.dense_image(752, 480, 8)),
                       ("d455-dense", syn.D455, syn.dense_image(848, 480, 8))):
    h, w = img.shape
    cam = mod.Camera(ctx, w, h, max_keypoints=4096)
    b = mod.Batch(ctx, [cam])
    print(name, flush=True)
    for it in range(3):
        b.enqueue([img], cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"], [(0.0, 1.0, 0.0)])
        b.finish()
    b.close(); cam.close()
