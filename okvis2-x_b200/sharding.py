"""Host-side sharding rules of the multi-GPU path; one process per GPU, torch.distributed plumbing.

* Frames: independent units, sharded round-robin, no data-path collective.
* Cameras of a multiframe (configs[2]): camera c -> rank c mod N. Detect / describe / back-project / matchToMap are
  per camera; matchStereo needs both cameras of an overlapping pair, so every rank packs the views of its cameras
  (okc_batch_pack_views) and ONE NCCL all-gather exchanges them; the rank that owns camera c matches (c, c + 1).
* Matcher sweep: map rows block-partitioned; every rank computes best[q] = (dist << 32) | GLOBAL row index over its
  rows as an int64 key with INT64_MAX for "none" (okc_match_sets_ordered); ONE all-reduce(min) reproduces the
  reference's single-machine result, first-wins ties included, because the key orders by distance and then by global
  index.
"""
import os

import numpy as np

NO_MATCH_I64 = -1            # 0xFFFFFFFFFFFFFFFF viewed as int64 (OKC_NO_MATCH)
ORDERED_NO_MATCH = (1 << 63) - 1


def frame_of(step, rank, world, n_frames):
    """Frame index processed by `rank` at its local step `step` (round-robin over ranks, cyclic over the pool)."""
    return (rank + step * world) % n_frames


def map_row_shard(rank, world, nm):
    """[lo, hi) of the map rows owned by `rank` (block partition, last shard may be short or empty)."""
    per = (nm + world - 1) // world
    lo = min(nm, rank * per)
    return lo, min(nm, lo + per)


def camera_shard(rank, world, n_cams):
    """Cameras of a multiframe owned by `rank`."""
    return [c for c in range(n_cams) if c % world == rank]


def merge_best(best_local, world, dist=None):
    """Min over the shards of per-shard best keys. best_local: torch int64 tensor [nq] of ORDERED keys (no match =
    INT64_MAX): one all-reduce(min), in place. Works on CUDA tensors with NCCL and on CPU tensors with gloo."""
    if world > 1:
        dist.all_reduce(best_local, op=dist.ReduceOp.MIN)
    return best_local


def to_ordered(keys_u64):
    """uint64 keys with OKC_NO_MATCH -> ordered int64 keys (numpy)."""
    k = np.asarray(keys_u64, np.uint64)
    out = k.view(np.int64).copy()
    out[k == np.uint64(0xFFFFFFFFFFFFFFFF)] = ORDERED_NO_MATCH
    return out


def from_ordered(keys_i64):
    k = np.asarray(keys_i64, np.int64)
    out = k.view(np.uint64).copy()
    out[k == ORDERED_NO_MATCH] = np.uint64(0xFFFFFFFFFFFFFFFF)
    return out


def merge_best_numpy(parts):
    """Same rule on host arrays of uint64 keys (tests, single-process emulation of the ranks)."""
    return np.minimum.reduce([np.asarray(p, np.uint64) for p in parts])


# ---- inputs of the 10k x 1M sweep (BASELINE.json configs[4]) ----
def _flip(rows, p, rng):
    """Every bit of every row flipped with probability p."""
    mask = np.packbits(rng.random((len(rows), 384)) < p, axis=1, bitorder="little")
    return rows ^ mask


def natural_descriptors(mod, ctx):
    """BRISK descriptors of the reference's demo image (tests/golden/brisk_img1_gray_752x480.npz), extracted by this
    pipeline with a low threshold and a large budget: correlated, clustered bit patterns as real frames produce."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    img = np.load(os.path.join(root, "tests", "golden", "brisk_img1_gray_752x480.npz"))["image"]
    cam = mod.Camera(ctx, 752, 480, max_keypoints=8192)
    kp, desc = cam.detect_describe(img, 12.0, 0, 20.0, 8000)
    cam.close()
    return desc


def sweep_inputs(syn, variant, nq, nm, mod=None, ctx=None, seed=5):
    """(queries, map, note). iid: SURVEY 8(d) config 5 (uniform bits + planted matches) - the best case of the
    lower-bound kernel; natural: image descriptors tiled to nm rows with 8 % bit flips, queries = other flips of the
    same descriptors; adversarial: every map row within the threshold of every query."""
    rng = np.random.default_rng(seed)
    if variant == "iid":
        q, m = syn.sweep_inputs(nq, nm, seed=seed)
        return q, m, "i.i.d. uniform bits, 50 % of the queries with a planted row at 0..80 flips (SURVEY 8(d) config 5)"
    if variant == "natural":
        base = natural_descriptors(mod, ctx)
        m = _flip(base[rng.integers(0, len(base), nm)], 0.08, rng)
        q = _flip(base[rng.integers(0, len(base), nq)], 0.08, rng)
        return q, m, ("%d BRISK descriptors of external/brisk/images/img1.ppm (extracted by this pipeline) tiled to the "
                      "map with 8 %% bit flips; queries = other flips of the same descriptors" % len(base))
    if variant == "adversarial":
        base = rng.integers(0, 256, (1, 48), dtype=np.uint8)
        m = _flip(np.repeat(base, nm, 0), 0.03, rng)
        q = _flip(np.repeat(base, nq, 0), 0.03, rng)
        return q, m, "every map row within the threshold of every query (one descriptor, 3 % bit flips on both sides)"
    raise ValueError(variant)


def exact_path_fraction(q, m, threshold, nq=64, nm=20000, seed=1):
    """Fraction of pairs whose 6-word lower bound (popc(parity) + 2 popc(at-least-two), match.cu) is below the
    threshold: an upper bound of the pairs that take the exact 12-word path (the kernel tests against the running
    best, which only falls). Sampled on the host."""
    rng = np.random.default_rng(seed)
    qs = q[rng.integers(0, len(q), nq)][:, :24]
    msub = m[rng.integers(0, len(m), nm)][:, :24]
    hits = total = 0
    for row in qs:
        x = np.unpackbits((msub ^ row[None, :]).reshape(nm, 6, 4), axis=2).reshape(nm, 6, 32)
        n = x.sum(1)
        lb = ((n & 1) + 2 * (n >= 2)).sum(1)
        hits += int((lb < threshold).sum())
        total += nm
    return round(hits / total, 6)


# ---- camera-sharded multiframe: view exchange + cross-camera matchStereo ----
class _DevBuf:
    """A torch uint8 CUDA tensor; .ptr for the C ABI."""

    def __init__(self, torch, nbytes):
        self.t = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        self.ptr = self.t.data_ptr()


class CameraGather:
    """matchStereo over the adjacent camera pairs (c, c + 1) of a camera-sharded multiframe: per batch of F
    multiframes every rank packs the views of its cameras, one all_gather_into_tensor (NCCL over NVLink) brings all
    views to all ranks, and the owner of camera c matches it against camera c + 1 from the gathered buffer - on the
    batch's own stream, so the exchange overlaps the kernels of the other batches in flight."""

    def __init__(self, mod, torch, dist, ctx, gc, hamming):
        self.mod, self.torch, self.dist, self.gc, self.hamming = mod, torch, dist, gc, hamming
        wl = gc.wl
        self.world, self.rank = gc.world, gc.rank
        self.n_cams = wl["n_cams"]
        self.per_rank = (self.n_cams + self.world - 1) // self.world      # padded: NCCL gathers equal blocks
        self.rec = mod.packed_view_bytes(gc.max_kp)
        F = gc.F
        self.send = [_DevBuf(torch, self.rec * self.per_rank * F) for _ in gc.batches]
        self.recv = [_DevBuf(torch, self.rec * self.per_rank * F * self.world) for _ in gc.batches]
        # pairs owned here: camera c (slot j * F + i) against camera c + 1 = record (i, rank1, j1) of the gather
        self.slot0, self.index1, self.cam1 = [], [], []
        for i in range(F):
            for j, c in enumerate(gc.cams_here):
                if c + 1 >= self.n_cams:
                    continue
                r1, j1 = (c + 1) % self.world, (c + 1) // self.world
                self.slot0.append(j * F + i)
                self.index1.append(r1 * self.per_rank * F + j1 * F + i)
                self.cam1.append(c + 1)
        self.f1 = np.full(len(self.slot0), wl["f_mean"])

    def step(self, bi, k0):
        """Exchange + match for the batch `bi`, which holds this rank's cameras of the frames k0 .. k0 + F - 1."""
        gc, torch = self.gc, self.torch
        b = gc.batches[bi]
        b.pack_views(list(range(gc.nc * gc.F)), self.send[bi].ptr)
        with torch.cuda.stream(gc.streams[bi]):
            self.dist.all_gather_into_tensor(self.recv[bi].t, self.send[bi].t, group=getattr(gc, "group", None))
        poses1 = np.stack([gc.pose_rows[gc.frame_of(k0 + s % gc.F), c1] for s, c1 in zip(self.slot0, self.cam1)]) \
            if self.slot0 else np.zeros((0, 24))
        if self.slot0:
            b.match_stereo_packed_enqueue(self.slot0, self.recv[bi].ptr, self.index1, poses1, self.f1, self.hamming)

    def describe(self):
        return {"collective": "one NCCL all_gather_into_tensor per step on the batch's stream",
                "bytes_per_rank_and_step": self.rec * self.per_rank * self.gc.F,
                "pairs_matched_on_this_rank_per_step": len(self.slot0)}
