"""Seeded synthetic inputs of the shapes BASELINE.json names.

Used by tests/, bench.py and __graft_entry__.smoke(); everything is a pure function of the seed so that the
CUDA path and the oracle see identical bytes. Camera-awareness maps follow the construction of
okvis_cv/include/okvis/cameras/implementation/PinholeCamera.hpp:172-200 (rays = normalised back-projection,
Jacobian = d(project)/d(point) at the ray) for a pinhole + radial-tangential camera; they are an INPUT of the
path (to both the product and the oracle), not part of it.
"""
import numpy as np

# config/euroc/okvis2.yaml:3-33 and its frontend_parameters block
EUROC = dict(
    width=752, height=480,
    cams=[dict(f=(458.654880721, 457.296696463), c=(367.215803962, 248.37534061),
               d=(-0.28340811217, 0.0739590738929, 0.000193595028569, 1.76187114545e-05),
               T_SC=[0.0148655429818, -0.999880929698, 0.00414029679422, -0.0216401454975,
                     0.999557249008, 0.0149672133247, 0.025715529948, -0.064676986768,
                     -0.0257744366974, 0.00375618835797, 0.999660727178, 0.00981073058949, 0, 0, 0, 1]),
          dict(f=(457.587426604, 456.13442556), c=(379.99944652, 255.238185386),
               d=(-0.283683654496, 0.0745128430929, -0.000104738949098, -3.55590700274e-05),
               T_SC=[0.0125552670891, -0.999755099723, 0.0182237714554, -0.0198435579556,
                     0.999598781151, 0.0130119051815, 0.0251588363115, 0.0453689425024,
                     -0.0253898008918, 0.0179005838253, 0.999517347078, 0.00786212447038, 0, 0, 0, 1])],
    radius=38.0, abs_threshold=150.0, max_kpt=700, octaves=0, matching_threshold=60.0, slices=4)
D455 = dict(width=848, height=480, radius=30.0, abs_threshold=5.0, max_kpt=2500, octaves=0, matching_threshold=60.0,
            slices=3)
HILTI = dict(width=720, height=540, radius=50.0, abs_threshold=20.0, max_kpt=700, octaves=0, matching_threshold=60.0,
             slices=4)


def scene_image(width, height, seed, rects=400, noise=3):
    """Random grey rectangles on a mid-grey canvas, 3x3 box blur, +-noise uniform noise (config 1 generator)."""
    rng = np.random.default_rng(seed)
    img = np.full((height, width), 128.0, np.float32)
    for _ in range(rects):
        x0, y0 = int(rng.integers(0, width)), int(rng.integers(0, height))
        w, h = (int(v) for v in rng.integers(8, 120, 2))
        img[y0:y0 + h, x0:x0 + w] = float(rng.integers(0, 256))
    p = np.pad(img, 1, mode="edge")
    img = sum(p[dy:dy + height, dx:dx + width] for dy in range(3) for dx in range(3)) / 9.0
    img = img + rng.integers(-noise, noise + 1, (height, width))
    return np.clip(img, 0, 255).astype(np.uint8)


def dense_image(width, height, seed, rects=6000, size=(4, 24), noise=3):
    """Many small rectangles: corner-dense texture that saturates the keypoint budgets (EuRoC: the 700 cap binds with
    ~7 500 NMS candidates; 848x480 at radius 30: ~1 500 keypoints, the most the uniformity stamps leave room for)."""
    rng = np.random.default_rng(seed)
    img = np.full((height, width), 128.0, np.float32)
    for _ in range(rects):
        x0, y0 = int(rng.integers(0, width)), int(rng.integers(0, height))
        w, h = (int(v) for v in rng.integers(size[0], size[1], 2))
        img[y0:y0 + h, x0:x0 + w] = float(rng.integers(0, 256))
    p = np.pad(img, 1, mode="edge")
    img = sum(p[dy:dy + height, dx:dx + width] for dy in range(3) for dx in range(3)) / 9.0
    img = img + rng.integers(-noise, noise + 1, (height, width))
    return np.clip(img, 0, 255).astype(np.uint8)


def noise_image(width, height, seed):
    """i.i.d. uniform u8: the reference's own test input (okvis_cv/test/TestFrame.cpp:62-65), worst case for NMS."""
    return np.random.default_rng(seed).integers(0, 256, (height, width), dtype=np.uint8)


def stereo_pair(width, height, seed, rects=400):
    """Left image + right image = left shifted by a per-row-band constant integer disparity of 8..40 px."""
    rng = np.random.default_rng(seed + 1000)
    wide = scene_image(width + 64, height, seed, rects)
    left = np.ascontiguousarray(wide[:, 48:48 + width])
    right = np.empty_like(left)
    band = 0
    while band < height:
        hgt = int(rng.integers(24, 96))
        disp = int(rng.integers(8, 41))
        right[band:band + hgt] = wide[band:band + hgt, 48 - disp:48 - disp + width]
        band += hgt
    return left, np.ascontiguousarray(right)


def _radtan_distort(x, y, d):
    k1, k2, p1, p2 = d
    mx2, my2, mxy = x * x, y * y, x * y
    rho2 = mx2 + my2
    rad = k1 * rho2 + k2 * rho2 * rho2
    xd = x + x * rad + 2.0 * p1 * mxy + p2 * (rho2 + 2.0 * mx2)
    yd = y + y * rad + 2.0 * p2 * mxy + p1 * (rho2 + 2.0 * my2)
    # Jacobian of the distortion
    j00 = 1 + rad + k1 * 2.0 * mx2 + k2 * rho2 * 4 * mx2 + 2.0 * p1 * y + 6 * p2 * x
    j01 = k1 * 2.0 * x * y + k2 * 4 * rho2 * x * y + p1 * 2.0 * x + 2.0 * p2 * y
    j10 = j01
    j11 = 1 + rad + k1 * 2.0 * my2 + k2 * rho2 * 4 * my2 + 6 * p1 * y + 2.0 * p2 * x
    return xd, yd, j00, j01, j10, j11


def camera_maps(width, height, f, c, d):
    """(rays float32 HxWx3, jacobians float32 HxWx6) for a pinhole + radtan camera (Appendix E of SURVEY.md)."""
    fu, fv = f
    cu, cv = c
    u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    xd, yd = (u - cu) / fu, (v - cv) / fv
    x, y = xd.copy(), yd.copy()
    for _ in range(5):  # Gauss-Newton undistortion
        xx, yy, j00, j01, j10, j11 = _radtan_distort(x, y, d)
        ex, ey = xd - xx, yd - yy
        # solve (J^T J) dx = J^T e
        a, b, cc = j00 * j00 + j10 * j10, j00 * j01 + j10 * j11, j01 * j01 + j11 * j11
        gx, gy = j00 * ex + j10 * ey, j01 * ex + j11 * ey
        det = a * cc - b * b
        x = x + (cc * gx - b * gy) / det
        y = y + (-b * gx + a * gy) / det
    xx, yy, j00, j01, j10, j11 = _radtan_distort(x, y, d)
    ok = (xd - xx) ** 2 + (yd - yy) ** 2 < 1e-4
    n = np.sqrt(x * x + y * y + 1.0)
    rays = np.stack([x / n, y / n, 1.0 / n], -1)
    rays[~ok] = 0.0
    # project(ray): point Jacobian (PinholeCamera.hpp impl:336-346) at the normalised ray
    X, Y, Z = rays[..., 0], rays[..., 1], np.where(ok, rays[..., 2], 1.0)
    px, py = X / Z, Y / Z
    _, _, d00, d01, d10, d11 = _radtan_distort(px, py, d)
    rz, rz2 = 1.0 / Z, 1.0 / (Z * Z)
    J = np.stack([fu * d00 * rz, fu * d01 * rz, -fu * (X * d00 + Y * d01) * rz2,
                  fv * d10 * rz, fv * d11 * rz, -fv * (X * d10 + Y * d11) * rz2], -1)
    J[~ok] = 0.0
    return np.ascontiguousarray(rays, np.float32), np.ascontiguousarray(J, np.float32)


def random_descriptors(n, seed):
    return np.random.default_rng(seed).integers(0, 256, (n, 48), dtype=np.uint8)


def flip_bits(desc, nflips, rng):
    """Copy of descriptor rows with nflips[i] distinct random bits flipped in row i."""
    out = desc.copy()
    for i in range(len(out)):
        k = int(nflips[i])
        if k <= 0:
            continue
        bits = rng.choice(384, size=k, replace=False)
        np.bitwise_xor.at(out[i], bits // 8, (1 << (bits % 8)).astype(np.uint8))
    return out


def sweep_inputs(nq, nm, seed=5, planted_frac=0.5, max_flips=80):
    """Config 5: i.i.d. descriptors; for planted_frac of the queries one map row = query with U{0..max_flips} flips."""
    rng = np.random.default_rng(seed)
    q = rng.integers(0, 256, (nq, 48), dtype=np.uint8)
    m = rng.integers(0, 256, (nm, 48), dtype=np.uint8)
    planted = rng.random(nq) < planted_frac
    idx = rng.integers(0, nm, nq)
    flips = rng.integers(0, max_flips + 1, nq)
    sel = np.nonzero(planted)[0]
    m[idx[sel]] = flip_bits(q[sel], flips[sel], rng)
    return q, m


def map_window(desc, kp_xy, num_landmarks, seed, width, height, is3d_frac=0.7):
    """Config 2: a keyframe-window of landmarks built from frame descriptors.

    Returns dict(map rows [M,48], owner [M] (landmark slot, ascending), proj [L,2], is3d [L]).
    Landmark l copies descriptor (l mod K) with Binomial(384, 0.05) flips, 1..3 rows each; its projection is the
    keypoint position + N(0, 5 px) for the first K landmarks and uniform in the image for the rest.
    """
    rng = np.random.default_rng(seed)
    K = len(desc)
    rows, owner = [], []
    proj = np.zeros((num_landmarks, 2), np.float64)
    for l in range(num_landmarks):
        nd = int(rng.integers(1, 4))
        if K > 0:
            src = l % K
            base = np.repeat(desc[src:src + 1], nd, 0)
            if l < K:
                r = flip_bits(base, rng.binomial(384, 0.05, nd), rng)
                proj[l] = kp_xy[src] + rng.normal(0.0, 5.0, 2)
            else:
                r = rng.integers(0, 256, (nd, 48), dtype=np.uint8)
                proj[l] = rng.random(2) * (width, height)
        else:
            r = rng.integers(0, 256, (nd, 48), dtype=np.uint8)
            proj[l] = rng.random(2) * (width, height)
        rows.append(r)
        owner += [l] * nd
    return dict(map=np.concatenate(rows, 0) if rows else np.zeros((0, 48), np.uint8),
                owner=np.asarray(owner, np.int32), proj=proj, is3d=rng.random(num_landmarks) < is3d_frac)


def stereo_sequence(n_frames, width=752, height=480, seed=2, rects=900, step=(2, 1)):
    """Config 2: a stereo sequence cut from one master canvas translated by `step` px per frame.

    Returns (left [n,H,W] u8, right [n,H,W] u8, shift [n,2] = (dx, dy) of frame f relative to frame 0). The right
    image is the left view displaced by a per-row-band integer disparity of 8..40 px (constant over the sequence).
    """
    mw, mh = width + step[0] * (n_frames - 1) + 64, height + step[1] * (n_frames - 1)
    master = scene_image(mw, mh, seed, rects=rects)
    rng = np.random.default_rng(seed + 1000)
    disp = np.zeros(mh, np.int64)
    band = 0
    while band < mh:
        hgt = int(rng.integers(24, 96))
        disp[band:band + hgt] = int(rng.integers(8, 41))
        band += hgt
    cols = np.arange(mw)[None, :] - disp[:, None]
    master_r = np.take_along_axis(master, np.clip(cols, 0, mw - 1), axis=1)
    left = np.empty((n_frames, height, width), np.uint8)
    right = np.empty_like(left)
    shift = np.zeros((n_frames, 2), np.int64)
    for f in range(n_frames):
        x0, y0 = 48 + step[0] * f, step[1] * f
        left[f] = master[y0:y0 + height, x0:x0 + width]
        right[f] = master_r[y0:y0 + height, x0:x0 + width]
        shift[f] = (step[0] * f, step[1] * f)
    return left, right, shift


def sequence_window(desc0, kp_xy0, shift, num_landmarks, seed, width, height):
    """Keyframe window for stereo_sequence: landmarks from frame 0's descriptors (map_window), 3-D ones only with
    densely renumbered slots, and per-frame projections = frame-0 projection moved with the canvas."""
    win = map_window(desc0, kp_xy0, num_landmarks, seed, width, height)
    sel = win["is3d"][win["owner"]]
    slots = np.cumsum(win["is3d"]) - 1
    rows = np.ascontiguousarray(win["map"][sel])
    owner = slots[win["owner"][sel]].astype(np.int32)
    proj0 = win["proj"][win["is3d"]]
    proj = proj0[None, :, :] - shift[:, None, :].astype(np.float64)
    return rows, owner, np.ascontiguousarray(proj)


# ------------------------------------------------------------------------------------------------------------
# A geometrically consistent world for the stages after describe (landmark preparation, the two map matchers,
# matchMotionStereo, matchStereo): a fronto-parallel textured plane at depth Z0 seen by a rig of cameras that look
# along +z (C_WC = identity) and translate parallel to the plane. Then every camera image is the master canvas cut
# at an integer offset: `step` px per frame, `disparity` px per camera.
# ------------------------------------------------------------------------------------------------------------
def rig_sequence(n_frames, n_cams, width, height, seed, step=(2, 1), disparity=24, dense=False, rects=900):
    """images[c][f] (u8 H x W): camera c of frame f = master[f * step_y :, 48 + f * step_x - c * disparity :]."""
    mw = width + step[0] * (n_frames - 1) + disparity * (n_cams - 1) + 64
    mh = height + step[1] * (n_frames - 1)
    area = (mw * mh) / float(752 * 480)   # rectangle counts are quoted for a 752 x 480 canvas
    master = dense_image(mw, mh, seed, rects=int(6000 * area)) if dense else scene_image(mw, mh, seed, rects=int(rects * area))
    x_base = 16 + disparity * (n_cams - 1)
    out = np.empty((n_cams, n_frames, height, width), np.uint8)
    for c in range(n_cams):
        for f in range(n_frames):
            x0, y0 = x_base + step[0] * f - disparity * c, step[1] * f
            out[c, f] = master[y0:y0 + height, x0:x0 + width]
    return out


def rig_poses(n_frames, n_cams, fx, fy, step=(2, 1), disparity=24, Z0=3.0):
    """r_WC[f][c] of the rig above (C_WC = I): content moves by -step px per frame and +disparity px per camera."""
    tx, ty, B = step[0] * Z0 / fx, step[1] * Z0 / fy, disparity * Z0 / fx
    r = np.zeros((n_frames, n_cams, 3))
    for f in range(n_frames):
        for c in range(n_cams):
            r[f, c] = (f * tx - c * B, f * ty, 0.0)
    return r


def hamming_rows(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def plane_landmarks(views, Z0, fx, max_landmarks, seed, quality=(0.002, 1.0), uninitialised=0.3, keep=0.6):
    """Landmarks of a keyframe window from its views' keypoints.

    views: list (oldest first) of dict(kp, desc, e_C, valid, r_WC) with C_WC = I. A keypoint's world point is its ray
    cut with the plane z = Z0; keypoints of different views within one pixel footprint and with Hamming distance
    < 60 are observations of the same landmark; a fraction `uninitialised` of them gets a quality of 1e-6 and so
    is not 3D yet. Returns the arguments of okc_landmark_pool_create /
    okc_prepare_landmarks: (views dict, hp_W, quality, obs_offset, obs_view, obs_kp); observations newest first,
    as the reference walks them (Frontend.cpp:1420-1424)."""
    rng = np.random.default_rng(seed)
    cell = 1.5 * Z0 / fx
    grid = {}
    pos, first_desc, obs = [], [], []
    for v, view in enumerate(views):
        e, ok = view["e_C"], view["valid"]
        for k in range(len(view["kp"])):
            if not ok[k]:
                continue
            P = view["r_WC"] + e[k] * (Z0 / e[k][2])
            gx, gy = int(np.floor(P[0] / cell)), int(np.floor(P[1] / cell))
            hit = -1
            for yy in (gy - 1, gy, gy + 1):
                for xx in (gx - 1, gx, gx + 1):
                    for l in grid.get((xx, yy), ()):
                        if abs(pos[l][0] - P[0]) < cell and abs(pos[l][1] - P[1]) < cell and \
                                hamming_rows(first_desc[l], view["desc"][k]) < 60:
                            hit = l
                            break
            if hit < 0:
                # (only a fraction `keep` of the scene points are landmarks: the rest stays for matchMotionStereo)
                if len(pos) >= max_landmarks or ((gx * 73856093) ^ (gy * 19349663)) % 1000 >= int(keep * 1000):
                    continue
                hit = len(pos)
                pos.append(P)
                first_desc.append(view["desc"][k])
                obs.append([])
                grid.setdefault((gx, gy), []).append(hit)
            if all(o[0] != v for o in obs[hit]):
                obs[hit].append((v, k))
    L = len(pos)
    w = rng.uniform(0.5, 2.0, (L, 1))
    hp_W = np.concatenate([np.asarray(pos).reshape(L, 3) * w, w], 1)
    q = rng.uniform(quality[0], quality[1], L)
    # a landmark of (near) zero quality fails the reference's is3d test (Frontend.cpp:1431-1438): it is matched by
    # matchToMapByThreadUnitialised instead of matchToMapByThread
    q[rng.random(L) < uninitialised] = 1e-6
    obs_offset = np.zeros(L + 1, np.int64)
    obs_view, obs_kp = [], []
    for l in range(L):
        for v, k in reversed(obs[l]):
            obs_view.append(v)
            obs_kp.append(k)
        obs_offset[l + 1] = len(obs_view)
    n_per = [len(v["kp"]) for v in views]
    row0 = np.concatenate([[0], np.cumsum(n_per)[:-1]]).astype(np.int64)
    eye = np.eye(3).reshape(9)
    vd = dict(C_WC=np.array([eye for _ in views]), r_WC=np.array([v["r_WC"] for v in views]), row0=row0,
              desc=np.concatenate([v["desc"] for v in views]).reshape(-1, 48),
              e_C=np.concatenate([v["e_C"] for v in views]).reshape(-1, 3))
    return vd, hp_W, q, obs_offset, np.asarray(obs_view, np.int32), np.asarray(obs_kp, np.int32)


def _equidistant_distort(x, y, k):
    k1, k2, k3, k4 = k
    r = np.sqrt(x * x + y * y)
    th = np.arctan(r)
    th2 = th * th
    poly = 1.0 + th2 * (k1 + th2 * (k2 + th2 * (k3 + th2 * k4)))
    thd = th * poly
    safe = np.where(r > 1e-8, r, 1.0)
    s = np.where(r > 1e-8, thd / safe, 1.0)
    # d(thd)/d(th) and ds/dr for the Jacobian
    dthd = 1.0 + th2 * (3 * k1 + th2 * (5 * k2 + th2 * (7 * k3 + th2 * 9 * k4)))
    ds = np.where(r > 1e-8, (dthd / (1.0 + r * r) * safe - thd) / (safe * safe), 0.0)
    j00 = s + ds * x * x / safe
    j01 = ds * x * y / safe
    j11 = s + ds * y * y / safe
    return s * x, s * y, j00, j01, j01, j11


def camera_maps_equidistant(width, height, f, c, k):
    """(rays float32 HxWx3, jacobians float32 HxWx6) for a pinhole + equidistant camera (Hilti22 config): the same
    construction as camera_maps (PinholeCamera.hpp:172-200), an INPUT of the path to both the product and the oracle."""
    fu, fv = f
    cu, cv = c
    u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    xd, yd = (u - cu) / fu, (v - cv) / fv
    # undistort: theta_d = |(xd, yd)|, Newton on theta
    rd = np.sqrt(xd * xd + yd * yd)
    th = rd.copy()
    k1, k2, k3, k4 = k
    for _ in range(10):
        th2 = th * th
        fth = th * (1.0 + th2 * (k1 + th2 * (k2 + th2 * (k3 + th2 * k4)))) - rd
        dfth = 1.0 + th2 * (3 * k1 + th2 * (5 * k2 + th2 * (7 * k3 + th2 * 9 * k4)))
        th = th - fth / dfth
    ok = (th > -1e-9) & (th < 1.5)
    th = np.clip(th, 0.0, 1.5)
    scale = np.where(rd > 1e-8, np.tan(th) / np.where(rd > 1e-8, rd, 1.0), 1.0)
    x, y = xd * scale, yd * scale
    n = np.sqrt(x * x + y * y + 1.0)
    rays = np.stack([x / n, y / n, 1.0 / n], -1)
    rays[~ok] = 0.0
    X, Y, Z = rays[..., 0], rays[..., 1], np.where(ok, rays[..., 2], 1.0)
    _, _, d00, d01, d10, d11 = _equidistant_distort(X / Z, Y / Z, k)
    rz, rz2 = 1.0 / Z, 1.0 / (Z * Z)
    J = np.stack([fu * d00 * rz, fu * d01 * rz, -fu * (X * d00 + Y * d01) * rz2,
                  fv * d10 * rz, fv * d11 * rz, -fv * (X * d10 + Y * d11) * rz2], -1)
    J[~ok] = 0.0
    return np.ascontiguousarray(rays, np.float32), np.ascontiguousarray(J, np.float32)


# frontend configurations of the shipped okvis2.yaml files with one camera model per rig:
# every camera of a synthetic rig uses the first camera's intrinsics so that the constant disparity between its images
# is geometrically consistent (rig_sequence / rig_poses)
RIGS = dict(
    euroc=dict(cfg=EUROC, width=752, height=480, n_cams=2, distortion="radialtangential",
               f=EUROC["cams"][0]["f"], c=EUROC["cams"][0]["c"], k=EUROC["cams"][0]["d"], max_kp=768),
    # config/hilti22/okvis2.yaml:3-74 (cam0)
    hilti=dict(cfg=HILTI, width=720, height=540, n_cams=5, distortion="equidistant",
               f=(351.31400364193297, 351.4911744656785), c=(367.8522793375995, 253.8402144980996),
               k=(-0.03696737352869157, -0.008917880497032812, 0.008912969593422046, -0.0037685977496087313), max_kp=768),
    # config/rsD455/okvis2.yaml (848x480 as BASELINE.json names the shape; intrinsics scaled from the 640x480 file)
    d455=dict(cfg=D455, width=848, height=480, n_cams=2, distortion="radialtangential",
              f=(425.0, 424.2), c=(423.1, 238.7), k=(-0.05, 0.06, 3e-4, -2e-4), max_kp=2560),
)


def rig_maps(rig):
    if rig["distortion"] == "equidistant":
        return camera_maps_equidistant(rig["width"], rig["height"], rig["f"], rig["c"], rig["k"])
    return camera_maps(rig["width"], rig["height"], rig["f"], rig["c"], rig["k"])
