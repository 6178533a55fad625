"""GPU parity of the camera geometry next to the path (SURVEY.md 8(f) rows f2 / f1 first half) against the CPU
restatement in oracle/port/camera.cc. Tolerances: radial-tangential and undistorted cameras use +, -, *, / only ->
bit-exact; the equidistant model calls atan/sqrt, whose device and glibc versions may differ in the last ulp -> 1e-12.
Parity with the reference itself is unpinned (Eigen / okvis_cv cannot be built here)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

EQUI = dict(f=(351.3, 351.5), c=(367.7, 253.8), d=(-0.03696, -0.008917, 0.008912, -0.003785))   # Hilti-like


def _cams(mod, syn):
    e = syn.EUROC["cams"][0]
    return {
        "radtan": (mod.Pinhole.make(752, 480, e["f"], e["c"], "radialtangential", e["d"]), 0.0),
        "equidistant": (mod.Pinhole.make(720, 540, EQUI["f"], EQUI["c"], "equidistant", EQUI["d"]), 1e-12),
        "none": (mod.Pinhole.make(640, 480, (500.0, 501.0), (320.5, 239.5), "none", ()), 0.0),
    }


def _keypoints(mod, w, h, n, seed):
    rng = np.random.default_rng(seed)
    kp = np.zeros(n, mod.KP_DTYPE)
    kp["x"] = rng.uniform(0, w, n).astype(np.float32)
    kp["y"] = rng.uniform(0, h, n).astype(np.float32)
    kp["x"][:4] = (0.0, w - 1.0, 0.0, w - 1.0)          # image corners: strongest distortion
    kp["y"][:4] = (0.0, 0.0, h - 1.0, h - 1.0)
    return kp


@pytest.mark.parametrize("model", ["radtan", "equidistant", "none"])
def test_back_project_vs_port(model, mod, ctx, port, syn):
    cam, tol = _cams(mod, syn)[model]
    kp = _keypoints(mod, cam.width, cam.height, 3000, 3)
    if model == "radtan":
        kp["x"][4], kp["y"][4] = 5000.0, -4000.0         # far outside: Gauss-Newton does not converge in 5 steps
    e, valid, nv = ctx.back_project(cam, kp)
    pe, pvalid, pnv = port.back_project(cam, kp)
    assert np.array_equal(valid, pvalid) and nv == pnv == int(valid.sum())
    if tol == 0.0:
        assert e.tobytes() == pe.tobytes()
    else:
        assert float(np.abs(e - pe).max()) <= tol
    assert np.all(e[:, 2] == 1.0) and valid[:4].all()
    if model == "radtan":
        assert valid[4] == 0
    # empty input
    e0, v0, n0 = ctx.back_project(cam, kp[:0])
    assert len(e0) == 0 and n0 == 0


@pytest.mark.parametrize("model", ["radtan", "equidistant", "none"])
def test_project_landmarks_vs_port(model, mod, ctx, port, syn):
    cam, tol = _cams(mod, syn)[model]
    rng = np.random.default_rng(11)
    n = 4000
    hp = np.concatenate([rng.uniform(-6, 6, (n, 2)), rng.uniform(-2, 12, (n, 1)), np.ones((n, 1))], 1)
    hp[::5] *= rng.uniform(0.2, 3.0, (len(hp[::5]), 1))          # arbitrary positive homogeneous scale
    hp[1::7] *= -1.0                                             # negative scale: projectHomogeneous negates the head
    hp[3] = (0.3, 0.2, 0.0, 1.0)                                 # on the camera plane
    a = 0.3
    C_ = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
    T = np.concatenate([C_.reshape(9), [0.1, -0.05, 0.4]])
    thr = 3.0 + 0.06 * cam.fu
    proj, status, vis = ctx.project_landmarks(cam, T, hp, thr)
    pproj, pstatus, pvis = port.project_landmarks(cam, T, hp, thr)
    assert np.array_equal(status, pstatus) and np.array_equal(vis, pvis)
    if tol == 0.0:
        assert proj.tobytes() == pproj.tobytes()
    else:
        assert float(np.abs(proj - pproj).max()) <= tol * 1e3      # pixels: tolerance scaled by the focal length
    assert set(np.unique(status)) >= {0, 1, 3} and 0 < vis.sum() < n
    assert vis[status == 3].sum() == 0 and vis[status == 0].all()
    # a visible landmark's projection back-projects onto its ray
    ok = np.flatnonzero(status == 0)[:200]
    kp = np.zeros(len(ok), mod.KP_DTYPE)
    kp["x"], kp["y"] = proj[ok, 0].astype(np.float32), proj[ok, 1].astype(np.float32)
    e, valid, _ = ctx.back_project(cam, kp)
    s = np.where(hp[ok, 3:4] < 0, -1.0, 1.0)
    pc = s * ((hp[ok, :3] @ C_.T) + T[9:] * hp[ok, 3:4])
    ray = pc / pc[:, 2:3]
    assert valid.all() and float(np.abs(e - ray).max()) < 5e-3     # float32 pixel rounding of the keypoint


def test_rejects_unknown_distortion(mod, ctx):
    cam = mod.Pinhole.make(640, 480, (500.0, 500.0), (320.0, 240.0), 3, ())
    kp = np.zeros(1, mod.KP_DTYPE)
    with pytest.raises(Exception):
        ctx.back_project(cam, kp)


def test_keyframe_overlap_vs_goldens_and_port(mod, ctx, port):
    """Frontend::doWeNeedANewKeyframe masks: cv2-drawn goldens, then random views against the port."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "overlap_goldens.npz"))
    inter, uni = ctx.keyframe_overlap(g["width"], g["height"], g["kp_offset"], g["keypoints"], g["matched"])
    assert np.array_equal(inter, g["intersection"]) and np.array_equal(uni, g["union"])
    rng = np.random.default_rng(5)
    w, h, off, kps, ms = [], [], [0], [], []
    for v in range(40):
        W, H = int(rng.integers(100, 1500)), int(rng.integers(100, 1100))
        n = int(rng.integers(0, 900))
        kp = np.zeros(n, mod.KP_DTYPE)
        kp["x"] = rng.uniform(-5, W + 5, n).astype(np.float32)     # slightly outside too: clipped discs
        kp["y"] = rng.uniform(-5, H + 5, n).astype(np.float32)
        w.append(W); h.append(H); off.append(off[-1] + n); kps.append(kp)
        ms.append((rng.uniform(0, 1, n) < rng.uniform(0, 1)).astype(np.uint8))
    args = (w, h, off, np.concatenate(kps), np.concatenate(ms))
    gi, gu = ctx.keyframe_overlap(*args)
    pi, pu = port.keyframe_overlap(*args)
    assert np.array_equal(gi, pi) and np.array_equal(gu, pu)
    assert np.all(gi <= gu)


def _landmark_scene(seed, L=3000, frames=6):
    rng = np.random.default_rng(seed)
    V = 2 * frames
    # cameras on a short trajectory looking down +z, small rotations
    C_WC, r_WC = [], []
    for v in range(V):
        a, b = rng.normal(0, 0.15, 2)
        Ry = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
        Rx = np.array([[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]])
        C_WC.append((Ry @ Rx).reshape(9))
        r_WC.append(np.array([0.4 * (v // 2) + 0.11 * (v % 2), rng.normal(0, 0.05), rng.normal(0, 0.3)]))
    kp_per_view = rng.integers(200, 700, V)
    row0 = np.concatenate([[0], np.cumsum(kp_per_view)[:-1]]).astype(np.int64)
    N = int(kp_per_view.sum())
    desc = rng.integers(0, 256, (N, 48), dtype=np.uint8)
    e_C = np.concatenate([rng.uniform(-0.8, 0.8, (N, 2)), np.ones((N, 1))], 1)
    p = np.concatenate([rng.uniform(-4, 6, (L, 1)), rng.uniform(-3, 3, (L, 1)), rng.uniform(0.3, 25, (L, 1))], 1)
    w = rng.uniform(0.5, 2.0, (L, 1))
    hp_W = np.concatenate([p * w, w], 1)
    quality = rng.uniform(0.001, 1.0, L)
    quality[::9] = 1e-4
    visible = (rng.uniform(0, 1, L) < 0.8).astype(np.uint8)
    n_obs = rng.integers(0, 9, L)
    n_obs[:5] = 0
    obs_offset = np.concatenate([[0], np.cumsum(n_obs)]).astype(np.int64)
    obs_view = rng.integers(0, V, int(n_obs.sum())).astype(np.int32)
    obs_kp = (rng.uniform(0, 1, len(obs_view)) * kp_per_view[obs_view]).astype(np.int32)
    views = dict(C_WC=np.array(C_WC), r_WC=np.array(r_WC), row0=row0, desc=desc, e_C=e_C)
    return views, hp_W, quality, visible, obs_offset, obs_view, obs_kp


@pytest.mark.parametrize("loop_closure", [False, True])
def test_prepare_landmarks_vs_port(loop_closure, mod, ctx, port):
    """Frontend::matchToMap landmark preparation (Frontend.cpp:1409-1500): is3d, best-3 descriptor choice, packed pool."""
    views, hp_W, quality, visible, off, ov, ok = _landmark_scene(21)
    r1 = np.array([1.0, 0.02, -0.1])
    fsum = 458.65 + 457.30
    got = ctx.prepare_landmarks(views, hp_W, quality, visible, off, ov, ok, r1, fsum, loop_closure)
    want = port.prepare_landmarks(views, hp_W, quality, visible, off, ov, ok, r1, fsum, loop_closure)
    assert got["nm"] == want["nm"] > 0
    for k in ("is3d", "rows", "owner", "filled", "row_obs", "pool"):
        assert np.array_equal(got[k], want[k]), k
    for k in ("e_W", "r_W"):
        assert got[k].tobytes() == want[k].tobytes(), k
    rows = got["rows"]
    assert np.all(rows[visible == 0] == 0) and np.all(rows[visible != 0] >= 1) and rows.max() == 3
    assert 0 < got["is3d"].sum() < visible.sum()
    assert np.all(np.diff(got["owner"]) >= 0)                       # landmarks ascending = std::map order
    assert (got["filled"] == 0).sum() > 0                           # landmarks without a usable observation keep 1 row
    if not loop_closure:
        # every kept row satisfies the two pruning rules it was selected under
        l = got["owner"][got["filled"] != 0]
        assert len(l) and len(l) < got["nm"] + 1
    # the pool feeds the map matcher as is
    e = np.linalg.norm(got["e_W"][got["filled"] != 0], axis=1)
    assert float(np.abs(e - 1.0).max()) < 1e-12


def test_prepare_landmarks_rejects_bad_observations(mod, ctx):
    views, hp_W, quality, visible, off, ov, ok = _landmark_scene(3, L=50)
    bad = ok.copy()
    bad[0] = 10 ** 6
    with pytest.raises(Exception):
        ctx.prepare_landmarks(views, hp_W, quality, visible, off, ov, bad, np.zeros(3), 900.0)
    empty = ctx.prepare_landmarks(views, hp_W[:0], quality[:0], visible[:0], off[:1], ov[:0], ok[:0], np.zeros(3), 900.0)
    assert empty["nm"] == 0
