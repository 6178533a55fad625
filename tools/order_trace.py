"""GPU-box tool: in-kernel cycle counts of order_candidates / the selection kernel for ONE image of three cases.
The kernels print them only in a trace build of the library (never the shipped one):
  cd okvis2-x_b200/csrc && nvcc <NVCC_FLAGS of __graft_entry__.py> -DOKC_ORDER_TRACE [-DOKC_SELECT_TRACE] -o ../libokvis_cuda.so <SOURCES>
then `gpurun -- python tools/order_trace.py`, then rebuild without the flags (`__graft_entry__.build(force=True)`)."""
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
