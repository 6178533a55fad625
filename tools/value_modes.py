import sys, time, json, os
sys.path.insert(0, "/root/repo")
import bench, torch
import __graft_entry__ as g
mod = g.build()
syn = __import__("importlib").import_module("okvis2x_b200.synthetic")
name = sys.argv[1]
for graph in (True, False):
    args = bench.parse_args(["--no-graph"] if not graph else [])
    ctx = mod.Context(0)
    wl = bench.build_images(syn, name, False)
    gc = bench.GpuConfig(args, mod, syn, torch, ctx, wl, 0, 1, False)
    gc.build_world(); gc.precompute_poses(); gc.setup()
    F, D = gc.F, gc.D
    n = 60
    def barrier(): torch.cuda.synchronize()
    ms, _ = gc.run_value(n, 5, barrier, lambda x: x)
    # mode B: finish before re-enqueue
    work = [gc.slot_args(j * F, True) for j in range(n)]
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for j, (imgs, poses) in enumerate(work):
            b = gc.batches[j % D]
            if j >= D: b.finish()
            gc.enqueue(b, imgs, poses, True, j * F)
        for b in gc.batches: b.finish()
        torch.cuda.synchronize(); tb = (time.perf_counter() - t0) * 1e3
    print(name, "graph" if graph else "plain", "deep queues ms/step", round(ms / n, 3), "finish-before-reenqueue ms/step", round(tb / n, 3))
    gc.close(); ctx.close()
