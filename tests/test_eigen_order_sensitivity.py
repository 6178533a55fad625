"""CPU-only: how much do the matcher results depend on the order in which Eigen evaluates its small fixed-size
expressions? (SURVEY.md hard part H5: "count and report as 0")

okvis_frontend cannot be compiled here (no Eigen), so the FP64 gates of the matchers are restated in oracle/port with ONE
evaluation order: 3-vector dot = (x0 y0 + x1 y1) + x2 y2, normalized() = v / sqrt(v.v), 2x2 inverse = adjugate * (1 / det),
3x3 * 3x1 rows = (c0 x + c1 y) + c2 z. The sensitivity builds (oracle/Makefile `variants`, PORT_EIGEN_VARIANT bits)
restate the same loops with the OTHER plausible orders; this test runs every matcher on the workloads of the parity
tests and of bench.py with each build and counts what changes: accept / reject decisions and matched indices must not
change at all (the gates are far from their razor edges on these inputs); the triangulated points may move in the
last bits. The counts are quoted in DESIGN.md section 5.
"""
import json
import os

import numpy as np
import pytest

VARIANTS = {1: "dot3 = x0y0 + (x1y1 + x2y2)", 2: "normalized() = v * (1 / |v|)", 4: "2x2 inverse by element-wise division",
            8: "3x3 * 3x1 rows = c0x + (c1y + c2z)", 15: "all four"}


def _rot(a, b, c):
    ca, sa, cb, sb, cc, sc = np.cos(a), np.sin(a), np.cos(b), np.sin(b), np.cos(c), np.sin(c)
    Rx = np.array([[1, 0, 0], [0, ca, -sa], [0, sa, ca]])
    Ry = np.array([[cb, 0, sb], [0, 1, 0], [-sb, 0, cb]])
    Rz = np.array([[cc, -sc, 0], [sc, cc, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _world(mod, syn, ref, port, rigname, dense, n_frames=6, n_kf=3):
    """The synthetic world of bench.py (plane at Z0 seen by a translating rig), turned by a fixed rotation so that no
    pose is the identity (an identity rotation hides the order of the 3x3 * 3x1 products)."""
    import pipeline_port
    import refbind
    rig = syn.RIGS[rigname]
    W, H, n_cams = rig["width"], rig["height"], min(rig["n_cams"], 2)
    model = mod.Pinhole.make(W, H, rig["f"], rig["c"], rig["distortion"], rig["k"])
    cfg = rig["cfg"]
    det = (cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    imgs = syn.rig_sequence(n_frames, n_cams, W, H, seed=23, dense=dense)
    r_plain = syn.rig_poses(n_frames, n_cams, rig["f"][0], rig["f"][1])
    R = _rot(0.31, -0.47, 0.23)          # world <- plain frame
    cpu = pipeline_port.CpuPipeline(ref, port)
    ext = refbind.RefExtractor(ref, True, False)
    rays, jac = syn.rig_maps(rig)
    ext.set_camera(rays, jac, np.float32(rig["f"][0]))
    ext.set_direction(np.asarray((0.0, 0.0, 1.0), np.float32))
    views_plain, frames = [], {}
    for f in list(range(n_kf)) + [n_frames - 1]:
        for c in range(n_cams):
            kp, desc = cpu.detect_describe(ext, imgs[c, f], det)
            e_C, valid, _ = port.back_project(model, kp)
            v = dict(kp=kp, desc=desc, e_C=e_C, valid=valid, r_WC=r_plain[f, c])
            frames[(f, c)] = v
            if f < n_kf:
                views_plain.append(v)
    ext.close()
    vd, hp_W, quality, off, ov, ok = syn.plane_landmarks(views_plain, 3.0, rig["f"][0], 2000, seed=5)
    # rotate the world: poses C_WC = R, centres R r, landmarks R p
    vd = dict(vd, C_WC=np.array([R.reshape(9) for _ in range(len(vd["C_WC"]))]), r_WC=vd["r_WC"] @ R.T)
    hp = hp_W.copy()
    hp[:, :3] = hp_W[:, :3] @ R.T
    return dict(model=model, cfg=cfg, R=R, frames=frames, n_cams=n_cams, n_kf=n_kf, cur=n_frames - 1,
                pool=(vd, hp, quality, off, ov, ok), f_mean=0.5 * (rig["f"][0] + rig["f"][1]), mod=mod)


def _run_all(port, w):
    """Every matcher on the world with one build of the port; returns {name: arrays}."""
    import pipeline_port
    cpu = pipeline_port.CpuPipeline(w["ref"], port)
    R, mod = w["R"], w["mod"]
    out = {}
    cur = [w["frames"][(w["cur"], c)] for c in range(w["n_cams"])]
    sv = lambda v: dict(desc=v["desc"], size=v["kp"]["size"].astype(np.float64), e_C=v["e_C"], bp_valid=v["valid"],
                        C_WC=R.reshape(9), r_WC=R @ v["r_WC"], f=w["f_mean"])
    for c, v in enumerate(cur):
        r = R @ v["r_WC"]
        pose = (R, r, mod.slot_pose(R, r)[:12])
        lm = cpu.landmark_stage(w["model"], pose, v["kp"], v["desc"], v["e_C"], v["valid"], w["pool"], 60.0,
                                w["cfg"]["slices"])
        out["prep%d" % c] = dict(is3d=lm["prep"]["is3d"], rows=lm["prep"]["rows"], owner=lm["prep"]["owner"],
                                 visible=lm["visible"], pool=lm["prep"]["pool"], e_W=lm["prep"]["e_W"])
        out["a14_%d" % c] = dict(dist=lm["a14"][0], owner=lm["a14"][1], ctrs=lm["a14"][2])
        out["a15_%d" % c] = dict(dist=lm["a15"][0], owner=lm["a15"][1], written=lm["a15"][3], ctrs=lm["a15"][4],
                                 hps=lm["a15"][2][:, :3])
        unmatched = (lm["a14"][1] < 0) & (lm["a15"][1] < 0)
        for k in range(w["n_kf"]):
            kv = w["frames"][(k, c)]
            d, k1, hp, init, q = cpu.motion(sv(kv), np.ones(len(kv["kp"]), np.uint8), sv(v), unmatched, 60.0)
            out["a16_%d_%d" % (k, c)] = dict(dist=d, k1=k1, init=init, hps=hp[:, :3], quality=q)
    if w["n_cams"] > 1:
        d, k1, hp, init = port.match_stereo(sv(cur[0]), sv(cur[1]), 60.0)
        out["a17"] = dict(dist=d, k1=k1, init=init, hps=hp[:, :3])
    return out


DECISIONS = ("dist", "owner", "k1", "init", "written", "ctrs", "is3d", "rows", "visible", "pool")


@pytest.mark.parametrize("rigname,dense", [("euroc", True), ("euroc", False), ("d455", True)])
def test_matcher_decisions_do_not_depend_on_the_eigen_evaluation_order(rigname, dense, mod, syn, ref, port, tmp_path):
    import refbind
    w = _world(mod, syn, ref, port, rigname, dense)
    w["ref"] = ref
    base = _run_all(port, w)
    n_matches = sum(int((v["owner"] >= 0).sum()) for k, v in base.items() if k.startswith(("a14", "a15"))) + \
        sum(int((v["k1"] >= 0).sum()) for k, v in base.items() if k.startswith("a16")) + \
        int((base["a17"]["dist"] < 60.0).sum())
    assert n_matches > 200      # the gates are exercised: hundreds of accepted matches
    report = {}
    for variant, what in VARIANTS.items():
        other = _run_all(refbind.Port(variant), w)
        compared, flips, moved, max_rel = _compare(base, other)
        report[str(variant)] = dict(order=what, decisions_compared=compared, decision_flips=flips,
                                    fp64_values_changed=moved, max_relative_change=max_rel)
        assert flips == 0, (variant, what, flips)
        assert max_rel < 1e-6, (variant, max_rel)       # rounding-level changes only, amplified by the triangulation
    out = os.environ.get("OKC_SENSITIVITY_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(json.dumps({"rig": rigname, "dense": dense, "matches": n_matches, "variants": report}) + "\n")


def _test_rig_views(syn, ref, seed=1, W=752, H=480):
    """The inputs of tests/test_gpu_match.py::stereo_views (ideal rectified rig with the EuRoC cam0 extrinsics), with
    the keypoints of the compiled reference instead of the product's."""
    import refbind
    right_cam, left_cam = syn.stereo_pair(W, H, seed)
    c = syn.EUROC["cams"][0]
    T0 = np.asarray(c["T_SC"], np.float64).reshape(4, 4)
    ext = refbind.RefExtractor(ref, True, False)
    views = []
    for i, img in enumerate((right_cam, left_cam)):
        kp, desc = ext.describe(img, ref.detect(img, 38.0, 0, 150.0, 700))
        x = (kp["x"].astype(np.float64) - c["c"][0]) / c["f"][0]
        y = (kp["y"].astype(np.float64) - c["c"][1]) / c["f"][1]
        e = np.stack([x, y, np.ones_like(x)], 1)
        e /= np.linalg.norm(e, axis=1, keepdims=True)
        bpv = np.ones(len(kp), np.uint8)
        bpv[::17] = 0
        C_WC = T0[:3, :3]
        r_WC = T0[:3, 3] + (C_WC @ np.asarray([0.11, 0.0, 0.0]) if i == 0 else 0.0)
        views.append(dict(desc=desc, size=kp["size"].astype(np.float64), e_C=e, bp_valid=bpv, C_WC=C_WC, r_WC=r_WC,
                          f=0.5 * (c["f"][0] + c["f"][1])))
    ext.close()
    return views


def _matcher_suite(port, v0, v1, syn):
    """The calls of test_match_stereo_vs_port, test_match_motion_stereo_vs_port and
    test_match_to_map_uninitialised_vs_port on one build of the port."""
    out = {}
    for tag, (a, b) in (("01", (v0, v1)), ("10", (v1, v0))):
        d, k1, hp, init = port.match_stereo(a, b, 60.0)
        out["a17_" + tag] = dict(dist=d, k1=k1, init=init, hps=hp[:, :3] / hp[:, 3:4])
        use0 = (np.arange(len(a["desc"])) % 7 != 0).astype(np.uint8)
        keep1 = np.nonzero(np.arange(len(b["desc"])) % 5 != 0)[0]
        packed = dict(b, desc=b["desc"][keep1], size=b["size"][keep1], e_C=b["e_C"][keep1], bp_valid=b["bp_valid"][keep1])
        d, k1, hp, init, q = port.match_motion_stereo(a, use0, packed, 60.0)
        out["a16_" + tag] = dict(dist=d, k1=k1, init=init, hps=hp[:, :3], quality=q)
    rng = np.random.default_rng(42)
    world = lambda v: (lambda e: e / np.linalg.norm(e, axis=1, keepdims=True))(v["e_C"] @ np.asarray(v["C_WC"]).T)
    rows, owner, e_rows, r_rows = [], [], [], []
    e0 = world(v0)
    for j in range(len(v0["desc"])):
        nd = int(rng.integers(1, 4))
        rows.append(syn.flip_bits(np.repeat(v0["desc"][j:j + 1], nd, 0), rng.binomial(384, 0.03, nd), rng))
        owner += [j] * nd
        e_rows += [e0[j]] * nd
        r_rows += [np.asarray(v0["r_WC"])] * nd
    kp_e = world(v1)
    use = (v1["bp_valid"] > 0) & (np.arange(len(kp_e)) % 13 != 0)
    for slices in (1, 4):
        r = port.match_to_map_uninitialised(v1["desc"], kp_e, use, None, np.concatenate(rows), np.asarray(owner, np.int32),
                                            np.asarray(e_rows), np.asarray(r_rows), len(v0["desc"]), v1["r_WC"], v1["f"],
                                            60.0, slices)
        out["a15_s%d" % slices] = dict(dist=r[0], owner=r[1], hps=r[2][:, :3], written=r[3], ctrs=r[4])
    return out


def _compare(base, other):
    flips = moved = compared = 0
    max_rel = 0.0
    for name, arrs in base.items():
        for key, a in arrs.items():
            b = other[name][key]
            if key in DECISIONS:
                compared += a.size
                flips += int(np.count_nonzero(a != b))
            else:
                ok = np.isfinite(a) & np.isfinite(b)
                moved += int(np.count_nonzero(a[ok] != b[ok])) + int(np.count_nonzero(np.isfinite(a) != np.isfinite(b)))
                den = np.maximum(np.abs(a[ok]), 1e-300)
                max_rel = max(max_rel, float(np.max(np.abs(a[ok] - b[ok]) / den, initial=0.0)))
    return compared, flips, moved, max_rel


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_matcher_test_inputs_do_not_depend_on_the_eigen_evaluation_order(seed, syn, ref, port):
    import refbind
    v0, v1 = _test_rig_views(syn, ref, seed)
    base = _matcher_suite(port, v0, v1, syn)
    assert sum(int((v["k1"] >= 0).sum()) for k, v in base.items() if "k1" in v) > 200
    report = {}
    for variant, what in VARIANTS.items():
        compared, flips, moved, max_rel = _compare(base, _matcher_suite(refbind.Port(variant), v0, v1, syn))
        report[str(variant)] = dict(order=what, decisions_compared=compared, decision_flips=flips,
                                    fp64_values_changed=moved, max_relative_change=max_rel)
        assert flips == 0, (variant, what, flips)
        assert max_rel < 1e-6, (variant, max_rel)
    out = os.environ.get("OKC_SENSITIVITY_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(json.dumps({"rig": "test_gpu_match inputs", "seed": seed, "variants": report}) + "\n")
