"""GPU parity: the Hamming matchers through the C ABI against the oracle (compiled brisk::Hamming for the plain
sweep, oracle/port restatement of the Frontend.cpp loops for the gated ones). Integer/index outputs bit-exact;
FP64 outputs (hp_W, reprErrors) bit-exact too because both sides round every operation separately."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_match_batched_small_vs_reference(mod, ctx, ref, port, syn):
    q, m = syn.sweep_inputs(1500, 40000, seed=5)
    best = ctx.match_batched(q, m, 60)
    assert np.array_equal(best, port.hamming_best(q, m, 60, nthreads=8))
    assert np.array_equal(best[:200], ref.hamming_best(q[:200], m, 60))
    assert int((best != mod.NO_MATCH).sum()) > 300


@pytest.mark.parametrize("nq,nm", [(1, 1), (1, 127), (33, 129), (1023, 128), (1025, 5000), (7, 70001), (0, 10),
                                   (5, 0)])
def test_match_batched_ragged_shapes(nq, nm, mod, ctx, port):
    rng = np.random.default_rng(nq * 131 + nm)
    q = rng.integers(0, 256, (nq, 48), dtype=np.uint8)
    m = rng.integers(0, 256, (nm, 48), dtype=np.uint8)
    if nq and nm:
        m[rng.integers(0, nm, max(1, nq // 2))] = q[rng.integers(0, nq, max(1, nq // 2))]
    for thr in (60, 400):  # 400 > 384: every query gets its true nearest row
        got = ctx.match_batched(q, m, thr)
        assert np.array_equal(got, port.hamming_best(q, m, thr)), (nq, nm, thr)


def test_match_batched_ties_first_row_wins(mod, ctx, port):
    rng = np.random.default_rng(7)
    q = rng.integers(0, 256, (64, 48), dtype=np.uint8)
    m = rng.integers(0, 256, (20000, 48), dtype=np.uint8)
    # query i: the same near-duplicate planted at several rows (across stage / chunk / CTA boundaries); the first wins
    base = [3, 127, 128, 129, 4095, 4096, 9999, 19000]
    first = {}
    for i in range(64):
        d = q[i].copy()
        d[i % 48] ^= 1
        rows = [r + 7 * i for r in base[i % 5:]]
        for r in rows:
            m[r] = d
        first[i] = rows[0]
    want = port.hamming_best(q, m, 60)
    got = ctx.match_batched(q, m, 60)
    assert np.array_equal(got, want)
    for i in range(64):
        assert got[i] == (1 << 32) | first[i]
    # threshold semantics: strict '<'
    q2 = q[:1].copy()
    m2 = np.repeat(q2, 3, 0)
    m2[:, 0] ^= 0x0F  # distance 4
    assert np.all(ctx.match_batched(q2, m2, 4) == mod.NO_MATCH)
    assert ctx.match_batched(q2, m2, 5)[0] == (4 << 32) | 0


def test_match_batched_early_out_variants_agree(mod, ctx, syn):
    q, m = syn.sweep_inputs(1100, 30000, seed=9)
    base = ctx.match_batched(q, m, 60)
    try:
        for pre in (12,):
            mod.lib().okc_debug_set_match_prefix(pre)
            assert np.array_equal(ctx.match_batched(q, m, 60), base), pre
        # thresholds at which the lower-bound kernel takes its exact path for most / all pairs
        mod.lib().okc_debug_set_match_prefix(12)
        loose = [ctx.match_batched(q[:300], m[:5000], t) for t in (80, 200, 385)]
        mod.lib().okc_debug_set_match_prefix(0)
        for t, want in zip((80, 200, 385), loose):
            assert np.array_equal(ctx.match_batched(q[:300], m[:5000], t), want), t
    finally:
        mod.lib().okc_debug_set_match_prefix(0)


def test_match_sets_sharded_min_equals_full(mod, ctx, port, syn):
    """Config 5 sharding rule: map rows block-partitioned, per-shard best with the GLOBAL row index, then min."""
    q, m = syn.sweep_inputs(2000, 60000, seed=15)
    full = ctx.match_batched(q, m, 60)
    qs = ctx.descriptor_set(q)
    for shards in (2, 4, 8):
        per = (len(m) + shards - 1) // shards
        parts = []
        for s in range(shards):
            ms = ctx.descriptor_set(m[s * per:(s + 1) * per])
            parts.append(ctx.match_sets(qs, ms, 60, index_offset=s * per))
            ms.close()
        assert np.array_equal(np.minimum.reduce(parts), full), shards
    qs.close()


def test_full_size_sweep_properties(mod, ctx, port, syn):
    """BASELINE config 5 at full size (10k x 1M): checked through size-independent properties + an oracle sample."""
    nq, nm = 10000, 1000000
    q, m = syn.sweep_inputs(nq, nm, seed=5)
    qs, ms = ctx.descriptor_set(q), ctx.descriptor_set(m)
    best = ctx.match_sets(qs, ms, 60)
    # (1) oracle on a sample of queries over the whole map
    sel = np.arange(0, nq, 157)
    assert np.array_equal(best[sel], port.hamming_best(q[sel], m, 60, nthreads=8))
    # (2) every reported match is real: recompute its distance, and it is below the threshold
    hit = best != mod.NO_MATCH
    idx = (best[hit] & np.uint64(0xFFFFFFFF)).astype(np.int64)
    dist = (best[hit] >> np.uint64(32)).astype(np.int64)
    true = np.unpackbits(np.bitwise_xor(q[hit], m[idx]), axis=1).sum(1)
    assert np.array_equal(true, dist) and dist.max() < 60
    assert 0.25 * nq < hit.sum() < 0.5 * nq  # ~half planted, flips U{0..80} -> ~3/8 below 60
    # (3) idempotence and shard-merge invariance at full size
    assert np.array_equal(ctx.match_sets(qs, ms, 60), best)
    half = nm // 2
    a, b = ctx.descriptor_set(m[:half]), ctx.descriptor_set(m[half:])
    merged = np.minimum(ctx.match_sets(qs, a, 60), ctx.match_sets(qs, b, 60, index_offset=half))
    assert np.array_equal(merged, best)
    for s in (qs, ms, a, b):
        s.close()


def _frame(mod, ctx, syn, W=752, H=480, seed=1):
    img = syn.scene_image(W, H, seed)
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    kp, desc = cam.detect_describe(img, 38.0, 0, 150.0, 700)
    cam.close()
    return kp, desc


@pytest.mark.parametrize("slices", [1, 3, 4, 7])
def test_match_to_map_vs_port(slices, mod, ctx, port, syn):
    kp, desc = _frame(mod, ctx, syn)
    kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
    win = syn.map_window(desc, kp_xy, 2000, 7, 752, 480)
    use = (np.arange(len(kp)) % 11 != 0).astype(np.uint8)
    sel = win["is3d"][win["owner"]]
    m3, o3 = win["map"][sel], win["owner"][sel]
    c = syn.EUROC["cams"][0]
    f = 0.5 * (c["f"][0] + c["f"][1])
    for gate in ((3.0 + f * 0.06) ** 2, (3.0 + f * 0.34) ** 2):  # with / without IMU (Frontend.cpp:1764-1765)
        bd, bo, ctrs, errs = ctx.match_to_map(desc, kp_xy, use, m3, o3, win["proj"], gate, 60.0, slices)
        pd, po, pc, pe = port.match_to_map(desc, kp_xy, use, m3, o3, win["proj"], None, gate, 60.0, slices)
        assert np.array_equal(bd, pd) and np.array_equal(bo, po)
        assert np.array_equal(ctrs, pc)
        assert errs.tobytes() == pe.tobytes()  # bit-exact, NaN for empty slices included
        assert (bo >= 0).sum() > 100


def test_match_to_map_edge_cases(mod, ctx, port, syn):
    kp, desc = _frame(mod, ctx, syn)
    kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
    win = syn.map_window(desc, kp_xy, 300, 8, 752, 480)
    none = np.zeros(len(kp), np.uint8)
    bd, bo, ctrs, errs = ctx.match_to_map(desc, kp_xy, none, win["map"], win["owner"], win["proj"], 900.0, 60.0, 4)
    assert np.all(bd == 60.0) and np.all(bo == -1) and np.all(ctrs == 0) and np.all(np.isnan(errs))
    # empty map
    bd, bo, ctrs, errs = ctx.match_to_map(desc, kp_xy, none + 1, win["map"][:0], win["owner"][:0], win["proj"][:0],
                                          900.0, 60.0, 4)
    assert np.all(bo == -1) and np.all(ctrs == 0)
    # more slices than keypoints (segment == 0: the last slice takes everything)
    bd, bo, ctrs, errs = ctx.match_to_map(desc[:3], kp_xy[:3], none[:3] + 1, win["map"], win["owner"], win["proj"],
                                          1e12, 400.0, 8)
    pd, po, pc, pe = port.match_to_map(desc[:3], kp_xy[:3], none[:3] + 1, win["map"], win["owner"], win["proj"], None,
                                       1e12, 400.0, 8)
    assert np.array_equal(bd, pd) and np.array_equal(bo, po) and np.array_equal(ctrs, pc)
    assert errs.tobytes() == pe.tobytes()


def stereo_views(mod, ctx, syn, W=752, H=480, seed=1):
    """An ideal rectified rig (both cameras = EuRoC cam0 intrinsics without distortion, camera 1 displaced by the
    11 cm baseline along the camera x axis) looking at a synthetic stereo pair; e_C = pinhole back-projection."""
    # stereo_pair returns (img, img displaced by +disp px): a feature sits at larger x in the second image, so the
    # second image is the LEFT camera of a real rig (x_right = x_left - d) and the first one the RIGHT camera
    right_cam, left_cam = syn.stereo_pair(W, H, seed)
    c = syn.EUROC["cams"][0]
    T0 = np.asarray(c["T_SC"], np.float64).reshape(4, 4)
    views = []
    for i, img in enumerate((right_cam, left_cam)):   # view 0 = right camera (at +11 cm along x), view 1 = left
        cam = mod.Camera(ctx, W, H, max_keypoints=4096)
        kp, desc = cam.detect_describe(img, 38.0, 0, 150.0, 700)
        cam.close()
        x = (kp["x"].astype(np.float64) - c["c"][0]) / c["f"][0]
        y = (kp["y"].astype(np.float64) - c["c"][1]) / c["f"][1]
        e = np.stack([x, y, np.ones_like(x)], 1)
        e /= np.linalg.norm(e, axis=1, keepdims=True)
        bpv = np.ones(len(kp), np.uint8)
        bpv[::17] = 0
        C_WC = T0[:3, :3]
        r_WC = T0[:3, 3] + (C_WC @ np.asarray([0.11, 0.0, 0.0]) if i == 0 else 0.0)
        views.append(dict(desc=desc, size=kp["size"].astype(np.float64), e_C=e, bp_valid=bpv, C_WC=C_WC,
                          r_WC=r_WC, f=0.5 * (c["f"][0] + c["f"][1])))
    return views


def test_match_stereo_vs_port(mod, ctx, port, syn):
    v0, v1 = stereo_views(mod, ctx, syn)
    for a, b in ((v0, v1), (v1, v0)):
        got = ctx.match_stereo(a, b, 60.0)
        want = port.match_stereo(a, b, 60.0)
        assert np.array_equal(got[0], want[0]), "distances"
        assert np.array_equal(got[1], want[1]), "k1 indices"
        assert got[2].tobytes() == want[2].tobytes(), "hp_W"
        assert np.array_equal(got[3], want[3]), "initialisable"
    assert (got[0] < 60.0).sum() > 50


def test_frame_pipeline_equals_separate_calls(mod, ctx, port, syn):
    """detect -> describe -> matchToMap with everything resident in HBM == the three host-buffer calls == oracle."""
    import torch
    W, H = 752, 480
    c = syn.EUROC["cams"][0]
    rays, jac = syn.camera_maps(W, H, c["f"], c["c"], c["d"])
    cam = mod.Camera(ctx, W, H, rays, jac, focal=c["f"][0], max_keypoints=1024)
    direction = (0.05, 0.99, 0.1)
    img0 = syn.scene_image(W, H, 1)
    kp0, desc0 = cam.detect_describe(img0, 38.0, 0, 150.0, 700, direction=direction)
    kp_xy0 = np.stack([kp0["x"], kp0["y"]], 1).astype(np.float64)
    win = syn.map_window(desc0, kp_xy0, 2000, 7, W, H)
    sel = win["is3d"][win["owner"]]
    # 3-D landmarks only, slots renumbered densely (what Frontend::matchToMap hands to matchToMapByThread)
    slots = np.cumsum(win["is3d"]) - 1
    m3, o3, proj3 = win["map"][sel], slots[win["owner"][sel]].astype(np.int32), win["proj"][win["is3d"]]
    window = mod.MapWindow(ctx, m3, o3, len(proj3))
    gate = (3.0 + 0.5 * (c["f"][0] + c["f"][1]) * 0.06) ** 2
    for seed, on_device, slices in ((1, False, 4), (2, True, 3), (1, True, 7)):
        img = syn.scene_image(W, H, seed)
        d_img = torch.from_numpy(img).cuda()
        torch.cuda.synchronize()
        cam.enqueue_device(d_img.data_ptr(), 38.0, 0, 150.0, 700, direction=direction)
        if on_device:
            d_proj = torch.from_numpy(proj3).cuda()
            torch.cuda.synchronize()
            cam.match_to_map_enqueue(window, None, gate, 60.0, proj_device_ptr=d_proj.data_ptr(), num_slices=slices)
        else:
            cam.match_to_map_enqueue(window, proj3, gate, 60.0, num_slices=slices)
        kp, desc, bd, bo, ctrs, errs = cam.frame_finish(window)
        kp2, desc2 = cam.detect_describe(img, 38.0, 0, 150.0, 700, direction=direction)
        assert kp.tobytes() == kp2.tobytes() and np.array_equal(desc, desc2)
        kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
        use = np.ones(len(kp), np.uint8)
        bd2, bo2, ctrs2, errs2 = ctx.match_to_map(desc, kp_xy, use, m3, o3, proj3, gate, 60.0, slices)
        pd, po, pc, pe = port.match_to_map(desc, kp_xy, use, m3, o3, proj3, None, gate, 60.0, slices)
        for a, b, cc in ((bd, bd2, pd), (bo, bo2, po), (ctrs, ctrs2, pc)):
            assert np.array_equal(a, b) and np.array_equal(a, cc)
        assert errs.tobytes() == errs2.tobytes() == pe.tobytes()
        if seed == 1:
            assert (bo >= 0).sum() > 100
    cam.kernel_timing(True)
    cam.enqueue_device(d_img.data_ptr(), 38.0, 0, 150.0, 700, direction=direction)
    cam.match_to_map_enqueue(window, proj3, gate, 60.0)
    times = cam.kernel_times()
    cam.frame_finish(window)
    cam.kernel_timing(False)
    assert [n for n, _ in times] == ["harris_nms", "order_candidates", "select_keypoints", "border_filter+describe",
                                       "bin_projections", "match_to_map", "match_stats"]
    assert all(t > 0 for _, t in times)
    window.close()
    cam.close()


def _rig_case(mod, ctx, syn, rigname, n_frames):
    """Cameras, images (slot k = camera k % n_cams of frame k // n_cams), detection parameters and slice count of a
    shipped configuration: EuRoC keeps its two calibrations and the sparse stereo sequence; Hilti (5 equidistant
    cameras, 720x540) and D455 (848x480, budget 2500, 3 slices) use dense scenes that saturate their budgets."""
    rig = syn.RIGS[rigname]
    cfg, W, H, n_cams = rig["cfg"], rig["width"], rig["height"], rig["n_cams"]
    cams, focals = [], []
    if rigname == "euroc":
        for c in syn.EUROC["cams"]:
            rays, jac = syn.camera_maps(W, H, c["f"], c["c"], c["d"])
            cams.append(mod.Camera(ctx, W, H, rays, jac, focal=c["f"][0], max_keypoints=rig["max_kp"]))
            focals.append(0.5 * (c["f"][0] + c["f"][1]))
        left, right, shift = syn.stereo_sequence(n_frames, W, H, seed=2)
        imgs = [(left, right)[k % 2][k // 2] for k in range(2 * n_frames)]
    else:
        rays, jac = syn.rig_maps(rig)
        for _ in range(n_cams):
            cams.append(mod.Camera(ctx, W, H, rays, jac, focal=rig["f"][0], max_keypoints=rig["max_kp"]))
            focals.append(0.5 * (rig["f"][0] + rig["f"][1]))
        seq = syn.rig_sequence(n_frames, n_cams, W, H, seed=4, dense=True)
        imgs = [seq[k % n_cams, k // n_cams] for k in range(n_cams * n_frames)]
        shift = np.asarray([(2 * f, f) for f in range(n_frames)], np.int64)
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    return dict(cams=cams, imgs=imgs, shift=shift, det=det, slices=cfg["slices"], n_cams=n_cams, W=W, H=H,
                gates=[(3.0 + f * 0.06) ** 2 for f in focals], landmarks=3 * cfg["max_kpt"] // 2 + 450)


@pytest.mark.parametrize("rigname,on_device", [("euroc", False), ("euroc", True), ("hilti", False), ("hilti", True),
                                               ("d455", False), ("d455", True)])
def test_batch_equals_single_calls(rigname, on_device, mod, ctx, port, syn):
    """A batch of slots (frames of a sequence: cam0, cam1, ..., cam0, ...) == the single-camera entry points, slot by
    slot, with and without the map match; host (pageable / pinned) and device inputs; at the shapes, detection
    parameters and slice counts of the EuRoC, Hilti22 and D455 configurations."""
    import torch
    n_frames = 3 if rigname != "hilti" else 2
    rc = _rig_case(mod, ctx, syn, rigname, n_frames)
    cams, imgs, det, S, nc = rc["cams"], rc["imgs"], rc["det"], rc["slices"], rc["n_cams"]
    ns = len(imgs)
    slot_cams = [cams[k % nc] for k in range(ns)]
    dirs = [(0.05, 0.99, 0.1) if k % 2 == 0 else (-0.02, 0.98, 0.15) for k in range(ns)]
    single = [slot_cams[k].detect_describe(imgs[k], *det, direction=dirs[k]) for k in range(ns)]
    if rigname != "euroc":     # dense scenes: the uniformity radius (Hilti: 50 px) or  (D455: 30 px) binds, not the threshold
        assert min(len(s[0]) for s in single) > (400 if rigname == "hilti" else 1000), [len(s[0]) for s in single]
    wins_h = []
    for cam_i in range(nc):
        kp, desc = single[cam_i]
        kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
        wins_h.append(syn.sequence_window(desc, kp_xy, rc["shift"], rc["landmarks"], 7 + cam_i, rc["W"], rc["H"]))
    wins = [mod.MapWindow(ctx, r, o, p.shape[1]) for r, o, p in wins_h]
    gates = rc["gates"]
    batch = mod.Batch(ctx, slot_cams)
    keep = []
    if on_device:
        d = [torch.from_numpy(i).cuda() for i in imgs]
        dp = [torch.from_numpy(np.ascontiguousarray(wins_h[k % nc][2][k // nc])).cuda() for k in range(ns)]
        torch.cuda.synchronize()
        keep = [d, dp]
        img_arg, proj_arg = [t.data_ptr() for t in d], [t.data_ptr() for t in dp]
    else:
        pinned = torch.from_numpy(np.ascontiguousarray(imgs[1])).pin_memory()   # one pinned source among pageable ones
        keep = [pinned]
        img_arg = [imgs[0], pinned.data_ptr()] + imgs[2:]
        proj_arg = [np.ascontiguousarray(wins_h[k % nc][2][k // nc]) for k in range(ns)]
    # without the map match
    batch.enqueue(img_arg, *det, dirs, images_on_device=on_device)
    batch.finish()
    for k in range(ns):
        kp, desc = batch.slot_frame(k)
        assert kp.tobytes() == single[k][0].tobytes() and np.array_equal(desc, single[k][1]), k
    # with it, twice (idempotent handles)
    for rep in range(2):
        batch.enqueue(img_arg, *det, dirs, images_on_device=on_device,
                      windows=[wins[k % nc] for k in range(ns)], proj=proj_arg, gates=[gates[k % nc] for k in range(ns)],
                      proj_on_device=on_device, num_slices=S)
        batch.finish()
        for k in range(ns):
            kp, desc = batch.slot_frame(k)
            assert kp.tobytes() == single[k][0].tobytes() and np.array_equal(desc, single[k][1]), k
            bd, bo, ctrs, errs = batch.slot_matches(k, len(kp))
            rows, owner, proj = wins_h[k % nc]
            kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
            pd, po, pc, pe = port.match_to_map(desc, kp_xy, np.ones(len(kp), np.uint8), rows, owner, proj[k // nc], None,
                                               gates[k % nc], 60.0, S)
            assert np.array_equal(bd, pd) and np.array_equal(bo, po) and np.array_equal(ctrs[:S], pc), k
            assert errs[:S].tobytes() == pe.tobytes(), k
            if k < nc:
                assert (bo >= 0).sum() > 100
        allr = batch.read_all()
        for k in range(ns):
            n = int(allr["n"][k])
            kp, desc = batch.slot_frame(k)
            bd, bo, ctrs, errs = batch.slot_matches(k, n)
            assert n == len(kp) and allr["kp"][k, :n].tobytes() == kp.tobytes()
            assert np.array_equal(allr["desc"][k, :n], desc) and np.array_equal(allr["dist"][k, :n], bd)
            assert np.array_equal(allr["owner"][k, :n], bo)
            stride = S
            assert np.array_equal(allr["ctrs"].reshape(-1)[k * stride:k * stride + len(ctrs)], ctrs)
            assert allr["errs"].reshape(-1)[k * stride:k * stride + len(errs)].tobytes() == errs.tobytes()
    batch.close()
    for w in wins:
        w.close()
    for c in cams:
        c.close()
    del keep


def _world_rays(v):
    e = v["e_C"] @ np.asarray(v["C_WC"]).T
    return e / np.linalg.norm(e, axis=1, keepdims=True)


@pytest.mark.parametrize("slices,with_previous", [(4, False), (3, True), (1, True)])
def test_match_to_map_uninitialised_vs_port(slices, with_previous, mod, ctx, port, syn):
    """Row a15: landmarks = the other view's keypoints (1-3 descriptor rows each, ray + origin per row)."""
    v0, v1 = stereo_views(mod, ctx, syn)
    rng = np.random.default_rng(42)
    rows, owner, e_rows, r_rows = [], [], [], []
    e0 = _world_rays(v0)
    for j in range(len(v0["desc"])):
        nd = int(rng.integers(1, 4))
        base = np.repeat(v0["desc"][j:j + 1], nd, 0)
        rows.append(syn.flip_bits(base, rng.binomial(384, 0.03, nd), rng))
        owner += [j] * nd
        e_rows += [e0[j]] * nd
        r_rows += [np.asarray(v0["r_WC"])] * nd
    rows = np.concatenate(rows)
    owner = np.asarray(owner, np.int32)
    e_rows, r_rows = np.asarray(e_rows), np.asarray(r_rows)
    kp_e = _world_rays(v1)
    use = (v1["bp_valid"] > 0) & (np.arange(len(kp_e)) % 13 != 0)
    prev = None
    if with_previous:
        # loop-closure mode: some keypoints already carry the landmark they will match again (:1941-1944)
        first = ctx.match_to_map_uninitialised(v1["desc"], kp_e, use, None, rows, owner, e_rows, r_rows, v1["r_WC"],
                                               v1["f"], 60.0, slices)
        prev = np.full(len(kp_e), -1, np.int32)
        hit = np.nonzero(first[1] >= 0)[0]
        prev[hit[::2]] = first[1][hit[::2]]
        prev[hit[1::4]] = 0
    got = ctx.match_to_map_uninitialised(v1["desc"], kp_e, use, prev, rows, owner, e_rows, r_rows, v1["r_WC"], v1["f"],
                                         60.0, slices)
    want = port.match_to_map_uninitialised(v1["desc"], kp_e, use, prev, rows, owner, e_rows, r_rows, len(v0["desc"]),
                                           v1["r_WC"], v1["f"], 60.0, slices)
    assert np.array_equal(got[0], want[0]), "distances"
    assert np.array_equal(got[1], want[1]), "landmark slots"
    assert got[2].tobytes() == want[2].tobytes(), "hps_W"
    assert np.array_equal(got[3], want[3]), "hps written"
    assert np.array_equal(got[4], want[4]), "ctrs"
    assert (want[1] >= 0).sum() > (10 if with_previous else 50)
    assert want[4].sum() >= (want[1] >= 0).sum()


def test_match_motion_stereo_vs_port(mod, ctx, port, syn):
    """Row a16: older keyframe vs the unmatched keypoints of the current frame (packed), per camera."""
    v0, v1 = stereo_views(mod, ctx, syn)
    for a, b in ((v0, v1), (v1, v0)):
        use0 = (np.arange(len(a["desc"])) % 7 != 0).astype(np.uint8)
        keep1 = np.nonzero(np.arange(len(b["desc"])) % 5 != 0)[0]   # the reference packs the unmatched k1 (:2034-2047)
        packed = dict(b, desc=b["desc"][keep1], size=b["size"][keep1], e_C=b["e_C"][keep1], bp_valid=b["bp_valid"][keep1])
        got = ctx.match_motion_stereo(a, use0, packed, 60.0)
        want = port.match_motion_stereo(a, use0, packed, 60.0)
        assert np.array_equal(got[0], want[0]), "distances"
        assert np.array_equal(got[1], want[1]), "k1"
        assert got[2].tobytes() == want[2].tobytes(), "hp_W"
        assert np.array_equal(got[3], want[3]), "initialisable"
        # quality = acos(...): device libm vs glibc, float tolerance of BASELINE.json (1e-4), in practice <= 1 ulp
        assert np.allclose(got[4], want[4], rtol=0, atol=1e-12)
        assert (want[1] >= 0).sum() > 50 and np.all(want[1][use0 == 0] == -1)


@pytest.mark.parametrize("rigname", ["euroc", "hilti", "d455"])
def test_binned_map_match_equals_tile_scan(rigname, mod, ctx, port, syn):
    """Frame-pipeline map match through spatial bins of the projections (default) against the scan of every row and
    against the port: ordinary window, a non-finite projection (bins unusable -> ordered full scan), projections far
    outside the image, a tiny gate (grid coarsened to the cell cap), and one keypoint with more hits than the per-warp
    hit buffer holds (100 copies of its descriptor projected onto it). At the image shape, detection parameters,
    gate and slice count of each shipped configuration (D455: 2500 keypoints against 4200 landmarks, 3 slices)."""
    rig = syn.RIGS[rigname]
    cfg, W, H = rig["cfg"], rig["width"], rig["height"]
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    S = cfg["slices"]
    cam = mod.Camera(ctx, W, H, max_keypoints=rig["max_kp"])
    img = syn.scene_image(W, H, 1) if rigname == "euroc" else syn.dense_image(W, H, 1)
    kp, desc = cam.detect_describe(img, *det)
    kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
    win = syn.map_window(desc, kp_xy, 3 * cfg["max_kpt"] // 2 + 450, 7, W, H)
    rows, owner, proj = win["map"], win["owner"].astype(np.int32), win["proj"].copy()
    gate = (3.0 + 0.5 * (rig["f"][0] + rig["f"][1]) * 0.06) ** 2

    def variants():
        yield "plain", rows, owner, proj, gate
        p = proj.copy()
        p[5] = (np.nan, 10.0)
        p[9] = (np.inf, -np.inf)
        yield "nonfinite", rows, owner, p, gate
        p = proj.copy()
        p[::7] = p[::7] * 50.0 - 9000.0
        yield "far-outside", rows, owner, p, gate
        yield "tiny-gate", rows, owner, proj, 0.25
        # 100 extra rows = descriptor of keypoint 3 with 0..5 flipped bits, all owned by fresh landmarks projected on it
        extra = np.repeat(desc[3:4], 100, 0).copy()
        for j in range(100):
            for bit in range(j % 6):
                extra[j, bit] ^= 1 << (j % 8)
        r2 = np.concatenate([rows, extra])
        L = proj.shape[0]
        o2 = np.concatenate([owner, np.arange(L, L + 100, dtype=np.int32)])
        p2 = np.concatenate([proj, np.repeat(kp_xy[3:4], 100, 0) + 0.5])
        yield "hit-overflow", r2, o2, p2, gate

    try:
        for name, r, o, p, g in variants():
            w = mod.MapWindow(ctx, r, o, p.shape[0])
            res = []
            for binned in (0, 1):
                mod.lib().okc_debug_set_match_binned(binned)
                batch = mod.Batch(ctx, [cam])
                batch.enqueue([img], *det, [(0.0, 1.0, 0.0)], windows=[w], proj=[np.ascontiguousarray(p)],
                              gates=[g], num_slices=S)
                batch.finish()
                k2, d2 = batch.slot_frame(0)
                assert k2.tobytes() == kp.tobytes() and np.array_equal(d2, desc)
                res.append(batch.slot_matches(0, len(kp)))
                batch.close()
            pd, po, pc, pe = port.match_to_map(desc, kp_xy, np.ones(len(kp), np.uint8), r, o, p, None, g, 60.0, S)
            for bd, bo, ctrs, errs in res:
                assert np.array_equal(bd, pd) and np.array_equal(bo, po), name
                assert np.array_equal(ctrs, pc), name
                if name == "nonfinite":   # NaN payloads are not part of the contract
                    assert np.array_equal(errs, pe, equal_nan=True), name
                else:
                    assert errs.tobytes() == pe.tobytes(), name
            w.close()
    finally:
        mod.lib().okc_debug_set_match_binned(1)
        cam.close()


@pytest.mark.parametrize("case", ["identical-descriptors", "threshold-above-384"])
def test_gated_matchers_when_the_candidate_lists_overflow(case, mod, ctx, port, syn):
    """The filter kernels keep 32 candidates per query. Identical descriptors on both sides (every pair at distance 0)
    overflow every list, a threshold above 384 switches the filter off: both make the resolve kernels walk all rows
    themselves - same results as the port, for matchStereo, matchMotionStereo and the uninitialised-landmark matcher."""
    v0, v1 = stereo_views(mod, ctx, syn)
    thr = 60.0
    if case == "identical-descriptors":
        one = v0["desc"][:1]
        v0 = dict(v0, desc=np.repeat(one, len(v0["desc"]), 0))
        v1 = dict(v1, desc=np.repeat(one, len(v1["desc"]), 0))
    else:
        thr = 400.0
    got, want = ctx.match_stereo(v0, v1, thr), port.match_stereo(v0, v1, thr)
    for g, w in zip(got, want):
        assert g.tobytes() == w.tobytes()
    assert (got[1] > 0).sum() > 20          # matches other than the initial k1 = 0 were accepted
    use0 = (np.arange(len(v0["desc"])) % 7 != 0).astype(np.uint8)
    got, want = ctx.match_motion_stereo(v0, use0, v1, thr), port.match_motion_stereo(v0, use0, v1, thr)
    for g, w in zip(got[:4], want[:4]):
        assert g.tobytes() == w.tobytes()
    assert np.allclose(got[4], want[4], rtol=0, atol=1e-12) and (got[1] >= 0).sum() > 20
    # uninitialised landmarks: 2 rows per landmark, landmarks = the other view's keypoints
    e0 = _world_rays(v0)
    rows = np.repeat(v0["desc"], 2, 0)
    owner = np.repeat(np.arange(len(v0["desc"]), dtype=np.int32), 2)
    e_rows, r_rows = np.repeat(e0, 2, 0), np.repeat(np.asarray(v0["r_WC"])[None, :], 2 * len(e0), 0)
    kp_e = _world_rays(v1)
    use = v1["bp_valid"] > 0
    args = (v1["desc"], kp_e, use, None, rows, owner, e_rows, r_rows)
    got = ctx.match_to_map_uninitialised(*args, v1["r_WC"], v1["f"], thr, 4)
    want = port.match_to_map_uninitialised(*args, len(v0["desc"]), v1["r_WC"], v1["f"], thr, 4)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert got[2].tobytes() == want[2].tobytes() and np.array_equal(got[3], want[3]) and np.array_equal(got[4], want[4])
    assert (want[1] >= 0).sum() > 20
