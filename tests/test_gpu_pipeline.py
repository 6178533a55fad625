"""GPU parity of the device-resident frame pipeline (rows a15-a17 inside the pipeline, f1 and f2 as
stages): back-projections after describe, landmark preparation + matchToMapByThread + matchToMapByThreadUnitialised
from a landmark pool kept in HBM, matchStereo between slots, matchMotionStereo against retained keyframes - against the
CPU restatement (oracle/pipeline_port.py -> oracle/port) fed with the very same keypoints / descriptors, and against
the host-buffer entry points. Bar: integer outputs and FP64 results of +, -, *, /, sqrt bit-exact; acos within 1e-12.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

Z0, DISP, STEP = 3.0, 24, (2, 1)


def _rig(mod, syn, ctx, cfgname, n_cams, max_kp):
    if cfgname == "euroc":
        cfg, c = syn.EUROC, syn.EUROC["cams"][0]
        model = mod.Pinhole.make(752, 480, c["f"], c["c"], "radialtangential", c["d"])
        W, H = 752, 480
    elif cfgname == "d455":
        cfg = syn.D455
        W, H = 848, 480
        model = mod.Pinhole.make(W, H, (425.0, 424.2), (423.1, 238.7), "radialtangential", (-0.05, 0.06, 3e-4, -2e-4))
    else:
        cfg = syn.HILTI
        W, H = 720, 540
        model = mod.Pinhole.make(W, H, (351.3, 351.5), (367.7, 253.8), "equidistant", (-0.03696, -0.008917, 0.008912, -0.003785))
    cams = [mod.Camera(ctx, W, H, max_keypoints=max_kp) for _ in range(n_cams)]
    for cam in cams:
        cam.set_geometry(model)
    return cfg, model, cams, W, H


def _poses(mod, r_WC):
    eye = np.eye(3)
    return np.stack([mod.slot_pose(eye, r) for r in r_WC])


def _view(kp, desc, e_C, valid, r_WC, f):
    return dict(desc=desc, size=kp["size"].astype(np.float64), e_C=e_C, bp_valid=valid, C_WC=np.eye(3).reshape(9),
                r_WC=np.asarray(r_WC, np.float64), f=f)


def _check_view(batch, pairs=(), n_jobs=0):
    """okc_batch_view_results (zero-copy arrays over the pinned arena) == the copying readers, live rows only."""
    v = batch.view()
    for c in range(batch.n):
        kp, desc = batch.slot_frame(c)
        n = len(kp)
        assert int(v["n"][c]) == n
        assert v["kp"][c, :n].tobytes() == kp.tobytes() and np.array_equal(v["desc"][c, :n], desc)
        if "dist" in v:
            bd, bo, ctrs, errs = batch.slot_matches(c, n)
            assert np.array_equal(v["dist"][c, :n], bd) and np.array_equal(v["owner"][c, :n], bo)
            assert np.array_equal(v["stats"]["count"][c], ctrs)
            with np.errstate(invalid="ignore", divide="ignore"):
                assert np.array_equal(v["stats"]["sum"][c] / v["stats"]["count"][c].astype(np.float64), errs, equal_nan=True)
        if "u_dist" in v:
            ud, uo, uhp, uhw, uc = batch.slot_uninitialised(c, n)
            assert np.array_equal(v["u_dist"][c, :n], ud) and np.array_equal(v["u_owner"][c, :n], uo)
            assert v["u_hps"][c, :n].tobytes() == uhp.tobytes() and np.array_equal(v["u_written"][c, :n], uhw)
            assert np.array_equal(v["u_ctrs"][c], uc)
    for p, (c0, _) in enumerate(pairs):
        got = batch.pair_stereo(p, int(v["n"][c0]))
        for name, g in zip(("s_dist", "s_k1", "s_hp", "s_init"), got):
            assert v[name][p, :len(g)].tobytes() == g.tobytes(), name
    for j in range(n_jobs):
        got = batch.job_motion(j)
        n0 = int(v["m_n0"][j])
        assert n0 == len(got[0])
        for name, g in zip(("m_dist", "m_k1", "m_hp", "m_init", "m_quality"), got):
            assert v[name][j, :n0].tobytes() == g.tobytes(), name


@pytest.mark.parametrize("cfgname,n_cams,dense", [("euroc", 2, False), ("euroc", 2, True), ("d455", 2, True),
                                                   ("hilti", 3, False)])
def test_frame_pipeline_vs_port(cfgname, n_cams, dense, mod, ctx, port, ref, syn):
    import pipeline_port
    max_kp = 2048 if cfgname == "d455" else 768
    cfg, model, cams, W, H = _rig(mod, syn, ctx, cfgname, n_cams, max_kp)
    tol = 1e-12 if cfgname == "hilti" else 0.0     # equidistant: atan / sqrt of glibc vs the device (last ulp)
    n_frames, n_kf = 7, 3
    imgs = syn.rig_sequence(n_frames, n_cams, W, H, seed=17, step=STEP, disparity=DISP, dense=dense)
    r_all = syn.rig_poses(n_frames, n_cams, model.fu, model.fv, STEP, DISP, Z0)
    f_mean = 0.5 * (model.fu + model.fv)
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    dirs = [(0.0, 1.0, 0.0)] * n_cams
    batch = mod.Batch(ctx, cams)
    cpu = pipeline_port.CpuPipeline(ref, port)
    # ---- keyframes: frames 0 .. n_kf-1, retained on the device; their data also read back to build the world ----
    kfs, views = [], []
    for f in range(n_kf):
        batch.enqueue([imgs[c, f] for c in range(n_cams)], *det, dirs)
        batch.set_poses(_poses(mod, r_all[f]))
        kf = mod.Keyframe(ctx, n_cams, max_kp)
        batch.retain(list(range(n_cams)), kf)
        batch.finish()
        for c in range(n_cams):
            kp, desc = batch.slot_frame(c)
            e_C, valid, nv = batch.slot_back_projections(c, len(kp))
            pe, pvalid, pnv = port.back_project(model, kp)
            assert np.array_equal(valid, pvalid) and nv == pnv
            assert (e_C.tobytes() == pe.tobytes()) if tol == 0.0 else float(np.abs(e_C - pe).max()) <= tol
            views.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid, r_WC=r_all[f, c], frame=f, cam=c))
            assert len(kp) > 50
        kfs.append(kf)
    pool_args = syn.plane_landmarks(views, Z0, model.fu, 2000, seed=5)
    vd, hp_W, quality, off, ov, ok = pool_args
    assert len(hp_W) > 200 and (np.diff(off) > 1).sum() > 50       # real multi-view landmarks
    pool = mod.LandmarkPool(ctx, vd, hp_W, quality, off, ov, ok)
    # eligibility flags of the keyframes' keypoints (estimator state in the reference): a seeded 80 %
    rng = np.random.default_rng(3)
    use0 = {}
    for i, v in enumerate(views):
        use0[(v["frame"], v["cam"])] = (rng.random(len(v["kp"])) < 0.8).astype(np.uint8)
        kfs[v["frame"]].set_view(v["cam"], np.eye(3), v["r_WC"], f_mean, use0[(v["frame"], v["cam"])])
    # ---- the frame under test ----
    f = n_frames - 1
    batch.enqueue([imgs[c, f] for c in range(n_cams)], *det, dirs)
    batch.set_poses(_poses(mod, r_all[f]))
    batch.match_landmarks_enqueue(pool, 60.0, cfg["slices"], mod.STAGE_UNINITIALISED)
    pairs = [(c, c + 1) for c in range(n_cams - 1)]
    batch.match_stereo_enqueue(pairs, 60.0)
    jobs = [(kfs[k], c, c) for k in range(n_kf) for c in range(n_cams)]
    batch.match_motion_enqueue(jobs, 60.0, only_unmatched=True)
    batch.finish()
    cur = []
    n3d_matches = n2d_matches = 0
    for c in range(n_cams):
        kp, desc = batch.slot_frame(c)
        assert len(kp) > 50
        e_C, valid, nv = batch.slot_back_projections(c, len(kp))
        pe, pvalid, _ = port.back_project(model, kp)
        assert np.array_equal(valid, pvalid)
        assert (e_C.tobytes() == pe.tobytes()) if tol == 0.0 else float(np.abs(e_C - pe).max()) <= tol
        pose = (np.eye(3), r_all[f, c], mod.slot_pose(np.eye(3), r_all[f, c])[:12])
        want = cpu.landmark_stage(model, pose, kp, desc, e_C, valid, pool_args, 60.0, cfg["slices"])
        lm = batch.slot_landmarks(c)
        prep = want["prep"]
        assert np.array_equal(lm["visible"], want["visible"]) and lm["nm"] == prep["nm"] > 0
        vis = want["visible"] != 0
        if tol == 0.0:
            assert lm["proj"].tobytes() == want["proj"].tobytes()
        else:
            assert float(np.abs(lm["proj"][vis] - want["proj"][vis]).max()) <= 1e-9
        for k in ("is3d", "rows"):
            assert np.array_equal(lm[k], prep[k]), k
        assert np.array_equal(lm["owner"], prep["owner"]) and np.array_equal(lm["pool"], prep["pool"])
        assert 0 < prep["is3d"].sum() < vis.sum()            # both map matchers have landmarks
        bd, bo, ctrs, errs = batch.slot_matches(c, len(kp))
        wd, wo, wc, we = want["a14"]
        assert np.array_equal(bd, wd) and np.array_equal(bo, wo) and np.array_equal(ctrs, wc)
        if tol == 0.0:
            assert errs.tobytes() == we.tobytes()
        else:
            assert np.allclose(errs, we, rtol=1e-9, atol=0, equal_nan=True)
        ud, uo, uhp, uhw, uc = batch.slot_uninitialised(c, len(kp))
        xd, xo, xhp, xhw, xc = want["a15"]
        assert np.array_equal(ud, xd) and np.array_equal(uo, xo) and np.array_equal(uhw, xhw) and np.array_equal(uc, xc)
        assert uhp[:, :3].tobytes() == xhp[:, :3].tobytes() if tol == 0.0 else np.allclose(uhp[:, :3], xhp[:, :3], rtol=1e-9)
        n3d_matches += int((bo >= 0).sum())
        n2d_matches += int((uo >= 0).sum())
        cur.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid, unmatched=(bo < 0) & (uo < 0)))
    assert n3d_matches > 20 and n2d_matches > 0
    # ---- matchStereo between the cameras ----
    stereo_matches = 0
    for p, (c0, c1) in enumerate(pairs):
        v0 = _view(cur[c0]["kp"], cur[c0]["desc"], cur[c0]["e_C"], cur[c0]["valid"], r_all[f, c0], f_mean)
        v1 = _view(cur[c1]["kp"], cur[c1]["desc"], cur[c1]["e_C"], cur[c1]["valid"], r_all[f, c1], f_mean)
        got = batch.pair_stereo(p, len(cur[c0]["kp"]))
        want = port.match_stereo(v0, v1, 60.0)
        host = ctx.match_stereo(v0, v1, 60.0)          # the host-buffer entry point runs the same kernel
        for g, w, hst in zip(got, want, host):
            assert g.tobytes() == w.tobytes() == hst.tobytes()
        stereo_matches += int((got[0] < 60.0).sum())
    assert stereo_matches > 20
    # ---- matchMotionStereo against the retained keyframes ----
    motion_matches = 0
    for j, (kf, view, slot) in enumerate(jobs):
        k = j // n_cams
        v = views[k * n_cams + view]
        v0 = _view(v["kp"], v["desc"], v["e_C"], v["valid"], v["r_WC"], f_mean)
        v1 = _view(cur[slot]["kp"], cur[slot]["desc"], cur[slot]["e_C"], cur[slot]["valid"], r_all[f, slot], f_mean)
        got = batch.job_motion(j)
        want = cpu.motion(v0, use0[(k, view)], v1, cur[slot]["unmatched"], 60.0)
        assert len(got[0]) == len(v["kp"])
        for name, g, w in zip(("dist", "k1", "hp", "init"), got[:4], want[:4]):
            assert g.tobytes() == w.tobytes(), (j, name)
        assert float(np.abs(got[4] - want[4]).max()) <= 1e-12
        motion_matches += int((got[1] >= 0).sum())
    assert motion_matches > 20
    _check_view(batch, pairs, len(jobs))
    for kf in kfs:
        kf.close()
    pool.close()
    batch.close()
    for cam in cams:
        cam.close()


def test_pipeline_stage_errors(mod, ctx, syn):
    """The stages refuse loudly what they cannot do: no geometry, no poses, slots out of range."""
    cam = mod.Camera(ctx, 320, 240, max_keypoints=256)
    batch = mod.Batch(ctx, [cam, cam])
    img = syn.scene_image(320, 240, 2, rects=100)
    batch.enqueue([img, img], 20.0, 0, 50.0, 200, [(0.0, 1.0, 0.0)] * 2)
    with pytest.raises(mod.OkcError):
        batch.match_stereo_enqueue([(0, 1)])          # no geometry
    batch.finish()
    model = mod.Pinhole.make(320, 240, (200.0, 200.0), (160.0, 120.0), "none", ())
    cam.set_geometry(model)
    batch.enqueue([img, img], 20.0, 0, 50.0, 200, [(0.0, 1.0, 0.0)] * 2)
    with pytest.raises(mod.OkcError):
        batch.match_stereo_enqueue([(0, 1)])          # no poses
    batch.set_poses(np.stack([mod.slot_pose(np.eye(3), (0, 0, 0)), mod.slot_pose(np.eye(3), (0.1, 0, 0))]))
    with pytest.raises(mod.OkcError):
        batch.match_stereo_enqueue([(0, 5)])          # slot out of range
    batch.match_stereo_enqueue([(0, 1)])
    batch.finish()
    kp, desc = batch.slot_frame(0)
    dist, k1, hp, init = batch.pair_stereo(0, len(kp))
    assert len(dist) == len(kp)
    with pytest.raises(mod.OkcError):
        mod.Camera(ctx, 320, 240, max_keypoints=16).set_geometry(mod.Pinhole.make(640, 480, (1, 1), (1, 1), "none", ()))
    batch.close()
    cam.close()


def test_graph_replay_equals_plain_launches(mod, ctx, syn):
    """The launch chains replayed as CUDA graphs (default) against plain launches: five frames of a sequence, each
    with the landmark stage; every output byte equal. The graphs are captured at the second frame, so frames 2-4 are
    replays with new images, poses and extraction directions reaching the kernels through the device tables."""
    cfg, model, cams, W, H = _rig(mod, syn, ctx, "euroc", 2, 768)
    n_frames = 6
    imgs = syn.rig_sequence(n_frames, 2, W, H, seed=31, step=STEP, disparity=DISP, dense=True)
    r_all = syn.rig_poses(n_frames, 2, model.fu, model.fv, STEP, DISP, Z0)
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    # world from frame 0
    b0 = mod.Batch(ctx, cams)
    b0.enqueue([imgs[c, 0] for c in range(2)], *det, [(0.0, 1.0, 0.0)] * 2)
    b0.finish()
    views = []
    for c in range(2):
        kp, desc = b0.slot_frame(c)
        e_C, valid, _ = b0.slot_back_projections(c, len(kp))
        views.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid, r_WC=r_all[0, c]))
    b0.close()
    pool = mod.LandmarkPool(ctx, *syn.plane_landmarks(views, Z0, model.fu, 2000, seed=5))
    results = {}
    for graph in (False, True):
        batch = mod.Batch(ctx, cams)
        batch.set_graph(graph)
        outs = []
        for f in range(1, n_frames):
            d = (0.05 * f, 1.0 - 0.01 * f, 0.1)
            batch.enqueue([imgs[c, f] for c in range(2)], *det, [d, d])
            batch.set_poses(_poses(mod, r_all[f]))
            batch.match_landmarks_enqueue(pool, 60.0, cfg["slices"], mod.STAGE_UNINITIALISED)
            batch.match_stereo_enqueue([(0, 1)], 60.0)
            batch.finish()
            frame = []
            for c in range(2):
                kp, desc = batch.slot_frame(c)
                assert len(kp) > 300
                frame += [kp.tobytes(), desc.tobytes()]
                frame += [x.tobytes() for x in batch.slot_matches(c, len(kp))]
                frame += [x.tobytes() for x in batch.slot_uninitialised(c, len(kp))]
                frame += [x.tobytes() for x in batch.slot_back_projections(c, len(kp))[:2]]
            frame += [x.tobytes() for x in batch.pair_stereo(0, len(batch.slot_frame(0)[0]))]
            _check_view(batch, [(0, 1)])
            outs.append(frame)
        results[graph] = outs
        batch.close()
    assert results[False] == results[True]
    assert len(set(tuple(f) for f in results[True])) == n_frames - 1      # (the frames really differ)
    pool.close()
    for cam in cams:
        cam.close()


def test_large_batch_equals_small_batches(mod, ctx, syn):
    """Four stereo frames in ONE batch of 8 slots against the same frames as four batches of 2 slots: every output
    byte equal. The two shapes take different code paths on purpose - 8 slots: live-row result copies, 48-register
    builds of the one-CTA-per-image kernels, everything on one stream; 2 slots: whole-section copies, 64-register
    builds, landmark preparation and match statistics on the side stream."""
    cfg, model, cams, W, H = _rig(mod, syn, ctx, "euroc", 2, 768)
    n_frames = 6
    imgs = syn.rig_sequence(n_frames, 2, W, H, seed=41, step=STEP, disparity=DISP, dense=True)
    r_all = syn.rig_poses(n_frames, 2, model.fu, model.fv, STEP, DISP, Z0)
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    f_mean = 0.5 * (model.fu + model.fv)
    # world and one retained keyframe from frame 0
    small = mod.Batch(ctx, cams)
    small.enqueue([imgs[c, 0] for c in range(2)], *det, [(0.0, 1.0, 0.0)] * 2)
    small.set_poses(_poses(mod, r_all[0]))
    kf = mod.Keyframe(ctx, 2, 768)
    small.retain([0, 1], kf)
    small.finish()
    views = []
    for c in range(2):
        kp, desc = small.slot_frame(c)
        e_C, valid, _ = small.slot_back_projections(c, len(kp))
        views.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid, r_WC=r_all[0, c]))
        kf.set_view(c, np.eye(3), r_all[0, c], f_mean, (np.arange(len(kp)) % 5 != 0).astype(np.uint8))
    pool = mod.LandmarkPool(ctx, *syn.plane_landmarks(views, Z0, model.fu, 2000, seed=5))
    frames = [2, 3, 4, 5]

    def collect(batch, slots, pair, jobs):
        out = []
        for s in slots:
            kp, desc = batch.slot_frame(s)
            assert len(kp) > 300
            out += [kp.tobytes(), desc.tobytes()]
            out += [x.tobytes() for x in batch.slot_matches(s, len(kp))]
            out += [x.tobytes() for x in batch.slot_uninitialised(s, len(kp))]
            out += [x.tobytes() for x in batch.slot_back_projections(s, len(kp))[:2]]
        out += [x.tobytes() for x in batch.pair_stereo(pair, len(batch.slot_frame(slots[0])[0]))]
        for j in jobs:
            out += [x.tobytes() for x in batch.job_motion(j)]
        return out

    def run(batch, fs):
        """slots: camera-major per frame (frame i: slots 2 i, 2 i + 1)"""
        n = len(fs)
        batch.set_poses(np.concatenate([_poses(mod, r_all[f]) for f in fs]))
        batch.enqueue([imgs[c, f] for f in fs for c in range(2)], *det, [(0.05, 0.99, 0.1)] * (2 * n))
        batch.match_landmarks_enqueue(pool, 60.0, cfg["slices"], mod.STAGE_UNINITIALISED)
        batch.match_stereo_enqueue([(2 * i, 2 * i + 1) for i in range(n)], 60.0)
        batch.match_motion_enqueue([(kf, c, 2 * i + c) for i in range(n) for c in range(2)], 60.0, only_unmatched=True)
        batch.finish()
        return [collect(batch, [2 * i, 2 * i + 1], i, [2 * i, 2 * i + 1]) for i in range(n)]

    want, d2h_small = [], 0
    for rep in range(2):      # (twice: the second pass replays the graphs)
        want, d2h_small = [], 0
        for f in frames:
            want.append(run(small, [f])[0])
            d2h_small += small.transfer_bytes()[1]
    big = mod.Batch(ctx, [cams[s % 2] for s in range(8)])
    for rep in range(2):
        got = run(big, frames)
        assert got == want
        h2d, d2h = big.transfer_bytes()
        assert h2d >= 8 * W * H and 0 < d2h < 0.9 * d2h_small    # live rows only (about 560 of 768 rows per section)
    assert len(set(tuple(x) for x in want)) == len(frames)
    big.close()
    small.close()
    kf.close()
    pool.close()
    for cam in cams:
        cam.close()


def test_per_frame_calls_allocate_nothing_once_warm(mod, ctx, syn):
    """No cudaMalloc / cudaMallocHost / cudaFree on the per-frame path: the library counts every allocation and release
    it makes (okc_debug_alloc_counts); after two warm-up frames with every stage, six more frames - new images, poses,
    pool update, landmark / stereo / motion matchers, retain into a keyframe, finish, views and the copying readers, with graph replay
    and with plain launches, on a 2-slot and an 8-slot batch - leave both counters where they were."""
    import ctypes as C
    cfg, model, cams, W, H = _rig(mod, syn, ctx, "euroc", 2, 768)
    n_frames = 10
    imgs = syn.rig_sequence(n_frames, 2, W, H, seed=43, step=STEP, disparity=DISP, dense=True)
    r_all = syn.rig_poses(n_frames, 2, model.fu, model.fv, STEP, DISP, Z0)
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    f_mean = 0.5 * (model.fu + model.fv)

    def counts():
        a, r = C.c_uint64(), C.c_uint64()
        assert mod.lib().okc_debug_alloc_counts(C.byref(a), C.byref(r)) == 0
        return a.value, r.value

    small = mod.Batch(ctx, cams)
    big = mod.Batch(ctx, [cams[s % 2] for s in range(8)])
    kf, kf2 = mod.Keyframe(ctx, 2, 768), mod.Keyframe(ctx, 2, 768)
    small.enqueue([imgs[c, 0] for c in range(2)], *det, [(0.0, 1.0, 0.0)] * 2)
    small.set_poses(_poses(mod, r_all[0]))
    small.retain([0, 1], kf)
    small.finish()
    views = []
    for c in range(2):
        kp, desc = small.slot_frame(c)
        e_C, valid, _ = small.slot_back_projections(c, len(kp))
        views.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid, r_WC=r_all[0, c]))
        kf.set_view(c, np.eye(3), r_all[0, c], f_mean, (np.arange(len(kp)) % 5 != 0).astype(np.uint8))
    world = syn.plane_landmarks(views, Z0, model.fu, 2000, seed=5)
    pool = mod.LandmarkPool(ctx, *world)
    assert counts()[0] > 20                                    # (set-up does allocate, and is counted)

    def frame(batch, fs, first):
        n = len(fs)
        pool.update(hp_W=world[1], quality=world[2])
        batch.set_poses(np.concatenate([_poses(mod, r_all[f]) for f in fs]))
        batch.enqueue([imgs[c, f] for f in fs for c in range(2)], *det, [(0.05, 0.99, 0.1)] * (2 * n))
        batch.match_landmarks_enqueue(pool, 60.0, cfg["slices"], mod.STAGE_UNINITIALISED)
        batch.match_stereo_enqueue([(2 * i, 2 * i + 1) for i in range(n)], 60.0)
        batch.match_motion_enqueue([(kf, c, 2 * i + c) for i in range(n) for c in range(2)], 60.0, only_unmatched=True)
        batch.retain([0, 1], kf2)
        batch.finish()
        _check_view(batch, [(2 * i, 2 * i + 1) for i in range(n)], 2 * n)
        nk = 0
        for s in range(2 * n):
            kp, _ = batch.slot_frame(s)
            batch.slot_matches(s, len(kp)), batch.slot_uninitialised(s, len(kp)), batch.slot_back_projections(s, len(kp))
            nk += len(kp)
        batch.pair_stereo(0, len(batch.slot_frame(0)[0])), batch.job_motion(0), batch.transfer_bytes()
        assert nk > 300 * 2 * n
        return nk

    for graph in (True, False):
        small.set_graph(graph), big.set_graph(graph)
        for f in (1, 2):                                        # warm: buffers grow to their sizes, graphs are captured
            frame(small, [f], True)
            frame(big, [f, f + 1, f + 2, f + 3], True)
        before = counts()
        for f in (3, 4, 5, 6):
            frame(small, [f], False)
        for f in (3, 4):
            frame(big, [f, f + 1, f + 2, f + 3], False)
        assert counts() == before, (graph, before, counts())
    before = counts()
    for x in (big, small, kf, kf2, pool, *cams):
        x.close()
    after = counts()
    assert after[0] == before[0] and after[1] > before[1] + 20     # (releases are counted too)
