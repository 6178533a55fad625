"""GPU parity: detect / describe through the C ABI against the compiled reference (oracle/_ref) on the same seeded
inputs, and against the committed goldens. Bar (BASELINE.json): keypoint sets, order, responses and descriptors
bit-exact; float sub-pixel positions and angles within 1e-4."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, kp_assert_close

pytestmark = pytest.mark.gpu

TOL = 1e-4


def _cases(syn):
    return {
        "euroc-scene": (syn.EUROC, syn.scene_image(752, 480, 1)),
        "euroc-noise": (syn.EUROC, syn.noise_image(752, 480, 2)),
        "d455-scene": (syn.D455, syn.scene_image(848, 480, 8, rects=1500)),
        "hilti-scene": (syn.HILTI, syn.scene_image(720, 540, 3)),
        "odd-size": (syn.HILTI, syn.scene_image(333, 211, 4, rects=150)),
    }


CASE_NAMES = ["euroc-scene", "euroc-noise", "d455-scene", "hilti-scene", "odd-size"]


@pytest.fixture(scope="module")
def cases(syn):
    return _cases(syn)


@pytest.mark.parametrize("name", CASE_NAMES)
def test_harris_scores_bit_exact(name, cases, mod, ctx, ref):
    cfg, img = cases[name]
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    assert np.array_equal(cam.debug_scores(img), ref.harris(img))
    cam.close()


@pytest.mark.parametrize("name", CASE_NAMES)
def test_nms_candidate_set(name, cases, mod, ctx, ref):
    cfg, img = cases[name]
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    thr = int(cfg["abs_threshold"])
    cand = cam.debug_candidates(img, thr)
    rc = ref.maxima(img, thr)
    assert len(cand) == len(rc)
    key = lambda c: np.lexsort((c["y"].astype(np.int64) * W + c["x"], -c["score"].astype(np.int64)))
    assert cand[key(cand)].tobytes() == rc[key(rc)].tobytes()
    # the device order is the reference's std::sort order (score desc; tie order of libstdc++ introsort)
    assert cand.tobytes() == ref.std_sort(rc).tobytes()
    cam.close()


def test_large_image_takes_the_strided_gather(mod, ctx, ref, syn):
    """1280 x 800: more bitmap words than order_candidates keeps in registers (16 per thread), so the candidate
    gather runs in its strided two-pass form; candidates in std::sort order, keypoints and descriptors against the
    compiled reference."""
    import refbind
    cfg, img = syn.EUROC, syn.scene_image(1280, 800, 6, rects=1200)
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    thr = int(cfg["abs_threshold"])
    cand, rc = cam.debug_candidates(img, thr), ref.maxima(img, thr)
    assert len(rc) > 1500 and cand.tobytes() == ref.std_sort(rc).tobytes()
    rkp = ref.detect(img, cfg["radius"], 0, cfg["abs_threshold"], 2000)
    kp, desc = cam.detect_describe(img, cfg["radius"], 0, cfg["abs_threshold"], 2000)
    ext = refbind.RefExtractor(ref)
    rkd, rdd = ext.describe(img, rkp)
    ext.close()
    assert len(rkd) > 300
    kp_assert_close(kp, rkd, TOL)
    assert np.array_equal(desc, rdd)
    cam.close()


@pytest.mark.parametrize("name", CASE_NAMES)
def test_detect_keypoints(name, cases, mod, ctx, ref):
    cfg, img = cases[name]
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    kp = cam.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    rkp = ref.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    assert len(rkp) > 50
    kp_assert_close(kp, rkp, TOL)
    cam.close()


@pytest.mark.parametrize("name", CASE_NAMES)
@pytest.mark.parametrize("mode", ["plain", "aware-down", "aware-axis", "plain-norot"])
def test_describe(name, mode, cases, mod, ctx, ref, syn):
    import refbind
    cfg, img = cases[name]
    H, W = img.shape
    rkp = ref.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    rot = mode != "plain-norot"
    ext = refbind.RefExtractor(ref, rotation_invariant=rot)
    if mode.startswith("aware"):
        c = syn.EUROC["cams"][0]
        rays, jac = syn.camera_maps(W, H, c["f"], c["c"], c["d"])
        direction = (0.05, 0.99, 0.1) if mode == "aware-down" else (0.0, 0.0, 1.0)
        cam = mod.Camera(ctx, W, H, rays, jac, focal=c["f"][0], max_keypoints=4096)
        ext.set_camera(rays, jac, np.float32(c["f"][0]))
        ext.set_direction(np.asarray(direction, np.float32))
    else:
        direction = (0.0, 1.0, 0.0)
        cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    kd, dd = cam.describe(img, rkp, rotation_invariant=rot, direction=direction)
    rkd, rdd = ext.describe(img, rkp)
    kp_assert_close(kd, rkd, TOL)
    assert dd.shape == rdd.shape
    bad = int((dd != rdd).any(1).sum())
    assert bad == 0, "%d of %d descriptor rows differ" % (bad, len(dd))
    cam.close()
    ext.close()


@pytest.mark.parametrize("name", CASE_NAMES)
def test_fused_detect_describe_equals_two_calls(name, cases, mod, ctx):
    cfg, img = cases[name]
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    kp = cam.detect(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    kd, dd = cam.describe(img, kp)
    kf, df = cam.detect_describe(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    assert kf.tobytes() == kd.tobytes() and np.array_equal(df, dd)
    # determinism / idempotence: the same call twice gives the same bytes
    kf2, df2 = cam.detect_describe(img, cfg["radius"], 0, cfg["abs_threshold"], cfg["max_kpt"])
    assert kf2.tobytes() == kf.tobytes() and np.array_equal(df2, df)
    cam.close()


@pytest.mark.parametrize("name", ["scene320", "noise256", "scene400"])
def test_committed_goldens(name, mod, ctx, syn):
    """Fixtures generated from the compiled reference in the build container (tests/golden/make_golden.py)."""
    src = open(os.path.join(GOLDEN, "make_golden.py")).read().split("def main():")[0].replace("g.build()", "")
    ns = {"__file__": os.path.join(GOLDEN, "make_golden.py")}
    exec(compile(src, "make_golden_defs", "exec"), ns)
    W, H, kind, seed, rects, radius, thr, maxk = ns["CASES"][name]
    img = ns["case_image"](name)
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    assert np.array_equal(cam.debug_scores(img)[H // 2], g["scores_row"])
    kp = cam.detect(img, radius, 0, thr, maxk)
    kp_assert_close(kp, g["kp_detect"], TOL)
    kd, dd = cam.describe(img, g["kp_detect"])
    kp_assert_close(kd, g["kp_plain"], TOL)
    assert np.array_equal(dd, g["desc_plain"])
    cam.close()
    rays, jac, focal = ns["case_maps"](name)
    camA = mod.Camera(ctx, W, H, rays, jac, focal=float(focal), max_keypoints=4096)
    ka, da = camA.describe(img, g["kp_detect"], direction=(0.05, 0.99, 0.1))
    kp_assert_close(ka, g["kp_aware"], TOL)
    assert np.array_equal(da, g["desc_aware"])
    camA.close()
    best = ctx.match_batched(g["desc_plain"], g["desc_aware"], 60)
    assert np.array_equal(best, g["best_plain_vs_aware"])


def test_edge_cases(mod, ctx, ref, syn):
    # flat image: no candidates, no keypoints, nothing to describe
    flat = np.full((120, 160), 77, np.uint8)
    cam = mod.Camera(ctx, 160, 120, max_keypoints=64)
    assert len(cam.detect(flat, 38.0, 0, 150.0, 700)) == 0
    kp, desc = cam.detect_describe(flat, 38.0, 0, 150.0, 700)
    assert len(kp) == 0 and desc.shape == (0, 48)
    # keypoints all inside the border band are filtered out by describe (RoiPredicate)
    img = syn.scene_image(160, 120, 5, rects=80)
    rkp = ref.detect(img, 10.0, 0, 20.0, 700)
    kd, dd = cam_describe_ref_equal(mod, ctx, ref, img, rkp)
    # max_num_keypoints budget is honoured exactly like the reference (stop at kept == maxNumKpt)
    cam2 = mod.Camera(ctx, 160, 120, max_keypoints=4096)
    for budget in (1, 7, 50):
        kp_assert_close(cam2.detect(img, 10.0, 0, 20.0, budget), ref.detect(img, 10.0, 0, 20.0, budget), TOL)
    # capacity overflow is reported, not truncated silently
    with pytest.raises(mod.OkcError):
        cam.detect(img, 2.0, 0, 1.0, 100000)
    # more octaves than the image can carry is refused loudly - every time, and the camera stays usable
    # (a refused pyramid must not be left half-built behind the handle)
    for _ in range(3):
        with pytest.raises(mod.OkcError):
            cam2.detect(img, 10.0, 6, 20.0, 700)
    kp_assert_close(cam2.detect(img, 10.0, 1, 20.0, 700), ref.detect(img, 10.0, 1, 20.0, 700), TOL)
    kp_assert_close(cam2.detect(img, 10.0, 0, 20.0, 700), ref.detect(img, 10.0, 0, 20.0, 700), TOL)
    cam.close()
    cam2.close()


@pytest.mark.parametrize("variant", [0, 1])
def test_candidate_overflow_is_reported_and_survivable(variant, mod, ctx, ref, syn):
    """More than 131 072 NMS maxima (a flat image with threshold 0: every pixel is a plateau maximum,
    harris-score-calculator.cc:90-110 keeps ties) is OKC_ERR_CAPACITY from both selection kernels - no stale or
    unwritten candidate keys are consumed - and the same handle then detects normally."""
    flat = np.full((480, 752), 90, np.uint8)
    img = syn.scene_image(752, 480, 31)
    cam = mod.Camera(ctx, 752, 480, max_keypoints=1024)
    try:
        mod.lib().okc_debug_set_select_wave(variant)
        for _ in range(2):
            with pytest.raises(mod.OkcError) as e:
                cam.detect(flat, 38.0, 0, 0.0, 700)
            assert e.value.code == 3  # OKC_ERR_CAPACITY
            with pytest.raises(mod.OkcError):
                cam.detect_describe(flat, 38.0, 0, 0.0, 700)
        kp_assert_close(cam.detect(img, 38.0, 0, 150.0, 700), ref.detect(img, 38.0, 0, 150.0, 700), TOL)
    finally:
        mod.lib().okc_debug_set_select_wave(0)
        cam.close()


def cam_describe_ref_equal(mod, ctx, ref, img, rkp):
    import refbind
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    ext = refbind.RefExtractor(ref)
    kd, dd = cam.describe(img, rkp)
    rkd, rdd = ext.describe(img, rkp)
    kp_assert_close(kd, rkd, TOL)
    assert np.array_equal(dd, rdd)
    cam.close()
    ext.close()
    return kd, dd


def _ref_order(ref, scores):
    import refbind
    n = len(scores)
    pts = np.zeros(n, refbind.PWS_DTYPE)
    pts["score"] = scores
    pts["x"] = np.arange(n) % 65536
    pts["y"] = np.arange(n) // 65536
    out = ref.std_sort(pts)
    return out["y"].astype(np.int64) * 65536 + out["x"]


def test_std_sort_reproduction(mod, ctx, ref):
    """The device reproduces libstdc++ std::sort (tie order included) on tie-heavy, sorted, constant and
    heapsort-triggering inputs, in shared memory (n <= 8192) and in the global-memory path."""
    cam = mod.Camera(ctx, 752, 480, max_keypoints=64)
    rng = np.random.default_rng(0)
    cases = []
    for trial in range(40):
        n = int(rng.integers(0, 9000))
        cases.append(rng.integers(0, int(rng.choice([2, 5, 50, 1000, 10 ** 6])), n))
    for n in (1, 2, 16, 17, 18, 33, 8191, 8192, 8193, 30000, 131072):
        cases.append(rng.integers(-50, 50, n))
        cases.append(rng.integers(-2 ** 31, 2 ** 31 - 1, n))
    cases += [np.arange(5000), np.arange(5000)[::-1].copy(), np.zeros(4000, np.int64)]
    for n in (2000, 4096, 20000):  # median-of-3 killers: reach the depth limit -> heapsort fallback
        k = n // 2
        kill = np.zeros(n, np.int64)
        i = np.arange(k)
        kill[:k] = np.where(i % 2 == 0, i + 1, k + i + 1)
        kill[k:] = 2 * (i + 1)
        cases += [kill, -kill]
    for sc in cases:
        got = cam.debug_std_sort(sc).astype(np.int64)
        assert np.array_equal(got, _ref_order(ref, sc)), "n=%d" % len(sc)
    cam.close()


@pytest.mark.parametrize("octaves", [1, 2, 3])
@pytest.mark.parametrize("name", ["euroc-scene", "euroc-noise", "d455-scene", "odd-size"])
def test_scale_space_detect(name, octaves, cases, mod, ctx, ref):
    """Row a6: 2*octaves layers (two-third / half sampling with the reference's region-dependent rounding), per-layer
    Harris + 2-D NMS, 3-D NMS against bilinear scores of the neighbouring layers, per-layer uniformity budgets."""
    import refbind
    cfg, img = cases[name]
    H, W = img.shape
    cam = mod.Camera(ctx, W, H, max_keypoints=4096)
    kp = cam.detect(img, cfg["radius"], octaves, cfg["abs_threshold"], cfg["max_kpt"])
    rkp = ref.detect(img, cfg["radius"], octaves, cfg["abs_threshold"], cfg["max_kpt"])
    assert len(rkp) > 30 and len(np.unique(rkp["octave"])) >= min(octaves, 2)
    kp_assert_close(kp, rkp, TOL)
    # fused detect + describe on the base image (scale_invariant = false ignores the keypoint size)
    kf, df = cam.detect_describe(img, cfg["radius"], octaves, cfg["abs_threshold"], cfg["max_kpt"])
    ext = refbind.RefExtractor(ref)
    rkd, rdd = ext.describe(img, rkp)
    ext.close()
    kp_assert_close(kf, rkd, TOL)
    assert np.array_equal(df, rdd)
    cam.close()


@pytest.mark.parametrize("name", CASE_NAMES)
def test_selection_kernels_agree(name, cases, mod, ctx):
    """The wave form of the uniformity selection (many final candidates decided and stamped per round) against the
    sequential-visit stamping kernel, which follows the reference loop literally: same keypoints, same order, binding
    and non-binding keypoint caps, small radius (dense acceptance) included."""
    cfg, img = cases[name]
    h, w = img.shape
    cam = mod.Camera(ctx, w, h, max_keypoints=4096)
    try:
        for radius, thr, cap in [(cfg["radius"], cfg["abs_threshold"], cfg["max_kpt"]), (cfg["radius"], 5.0, 50),
                                 (20.0, cfg["abs_threshold"], 4000), (cfg["radius"], 1.0, 0)]:
            mod.lib().okc_debug_set_select_wave(1)
            want = cam.detect(img, radius, 0, thr, cap)
            for variant in (0,):
                mod.lib().okc_debug_set_select_wave(variant)
                got = cam.detect(img, radius, 0, thr, cap)
                assert len(got) == len(want), (variant, radius, thr, cap, len(got), len(want))
                assert got.tobytes() == want.tobytes(), (variant, radius, thr, cap)
    finally:
        mod.lib().okc_debug_set_select_wave(0)
        cam.close()


@pytest.mark.parametrize("shape", [(480, 752), (41, 131), (540, 720), (101, 250)])
def test_harris_rows_per_warp_do_not_change_the_result(shape, mod, ctx, syn):
    """The launcher picks how many image rows a Harris warp walks (8..96, by batch size); scores, NMS candidates and
    keypoints must not depend on it (halo rows, chunk ends that fall outside the image, 3-row loop trips)."""
    h, w = shape
    img = syn.scene_image(w, h, 21)
    cam = mod.Camera(ctx, w, h, max_keypoints=2048)
    try:
        mod.lib().okc_debug_set_harris_rows(0)
        scores = cam.debug_scores(img)
        cands = cam.debug_candidates(img, 50)
        kps = cam.detect(img, 30.0, 0, 50.0, 1500)
        assert len(cands) > 0
        for rows in (1, 2, 3, 5, 8, 12, 16, 24, 32, 48, 64, 96, 1000):
            mod.lib().okc_debug_set_harris_rows(rows)
            assert np.array_equal(cam.debug_scores(img), scores), rows
            assert cam.debug_candidates(img, 50).tobytes() == cands.tobytes(), rows
            assert cam.detect(img, 30.0, 0, 50.0, 1500).tobytes() == kps.tobytes(), rows
    finally:
        mod.lib().okc_debug_set_harris_rows(0)
        cam.close()
