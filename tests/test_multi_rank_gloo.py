"""CPU-only: the N > 1 host logic (sharding + gather + min) with torch.distributed/gloo, world_size 2 and 3.
The per-rank compute is stood in for by the oracle port (no GPU here); the product's own kernels run the same
rule in tests/test_gpu_match.py::test_match_sets_sharded_min_equals_full and in bench.py --gpus N."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nq, nm, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import __graft_entry__ as g
    g.load()
    import importlib
    syn = importlib.import_module("okvis2x_b200.synthetic")
    sh = importlib.import_module("okvis2x_b200.sharding")
    import refbind
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q, m = syn.sweep_inputs(nq, nm, seed=31)
    m[5] = m[nm - 3] = q[0]            # a tie across shards: the lower global row must win
    lo, hi = sh.map_row_shard(rank, world, nm)
    local = refbind.Port().hamming_best(q, m[lo:hi], 60, index_offset=lo)
    # ordered int64 keys (INT64_MAX = no match), as okc_match_sets_ordered leaves them: one all-reduce(min)
    merged = sh.merge_best(torch.from_numpy(sh.to_ordered(local)), world, dist)
    # frames: every frame of the pool is visited exactly once per sweep over all ranks
    mine = [sh.frame_of(k, rank, world, 12) for k in range(12 // world)]
    allf = [None] * world
    dist.all_gather_object(allf, mine)
    if rank == 0:
        np.save(os.path.join(out_dir, "merged.npy"), sh.from_ordered(merged.numpy()))
        np.save(os.path.join(out_dir, "frames.npy"), np.sort(np.concatenate(allf)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_sweep_and_frames_gloo(world, tmp_path, port, syn, mod):
    nq, nm = 200, 3001   # ragged: the last shard is short
    mp.spawn(_worker, args=(world, _free_port(), nq, nm, str(tmp_path)), nprocs=world, join=True)
    q, m = syn.sweep_inputs(nq, nm, seed=31)
    m[5] = m[nm - 3] = q[0]
    want = port.hamming_best(q, m, 60)
    got = np.load(os.path.join(str(tmp_path), "merged.npy"))
    assert np.array_equal(got, want)
    assert want[0] == 5
    if 12 % world == 0:
        assert np.array_equal(np.load(os.path.join(str(tmp_path), "frames.npy")), np.arange(12))


def test_camera_shards_cover_every_camera_and_split_adjacent_pairs(mod):
    """configs[2]: camera c -> rank c mod N; with N >= 2 the two cameras of an adjacent pair never share a rank, so
    every matchStereo pair is served by the view exchange."""
    import importlib
    sh = importlib.import_module("okvis2x_b200.sharding")
    for world in (1, 2, 4, 5, 8):
        shards = [sh.camera_shard(r, world, 5) for r in range(world)]
        assert sorted(c for s in shards for c in s) == list(range(5))
        if world >= 2:
            for c in range(4):
                assert c % world != (c + 1) % world
    k = np.array([5, 0xFFFFFFFFFFFFFFFF, (60 << 32) | 7], np.uint64)
    assert np.array_equal(sh.from_ordered(sh.to_ordered(k)), k)
    assert sh.to_ordered(k)[1] == sh.ORDERED_NO_MATCH and np.all(np.diff(np.sort(sh.to_ordered(k))) > 0)


def test_shard_bounds(mod):
    import importlib
    sh = importlib.import_module("okvis2x_b200.sharding")
    for nm in (0, 1, 7, 8, 1000000):
        for world in (1, 2, 3, 8):
            spans = [sh.map_row_shard(r, world, nm) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == nm
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all(lo <= hi for lo, hi in spans)
