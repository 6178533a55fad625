"""GPU-box tool: time the 10k x 1M sweep with each early-out variant of the batched matcher and check they agree.
Usage: python tools/sweep_variants.py [variant ...]   (0 = 2-POPC lower bound, 5 = carry-save prefix, 4/6/8/12 plain)"""
import os, sys, json
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
import importlib

mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")
ctx = mod.Context(0)
variants = [int(a) for a in sys.argv[1:]] or [0, 5]
nq, nm = 10000, 1000000
q, m = syn.sweep_inputs(nq, nm, seed=5)
qs, ms = ctx.descriptor_set(q), ctx.descriptor_set(m)
best = torch.empty(nq, dtype=torch.int64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
cur = torch.cuda.Stream()
torch.cuda.set_stream(cur)
out, ref = {}, None
for v in variants:
    mod.lib().okc_debug_set_match_prefix(v)
    ts = []
    for it in range(7):
        flush.fill_(it)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        ctx.match_sets(qs, ms, 60, best_device=best.data_ptr(), stream=cur.cuda_stream, want_host=False)
        e1.record(cur)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    res = best.cpu().numpy().copy()
    if ref is None:
        ref = res
    t = float(np.median(ts[2:]))
    out[str(v)] = {"ms": round(t, 4), "gpairs_per_s": round(nq * nm / t / 1e6, 1), "agrees": bool(np.array_equal(res, ref))}
mod.lib().okc_debug_set_match_prefix(0)
print(json.dumps(out))
