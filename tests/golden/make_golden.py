#!/usr/bin/env python3
"""Generates the committed golden fixtures (run in the build container, where /root/reference exists).

  python tests/golden/make_golden.py

Sources of truth:
  * oracle/_ref/libbrisk_ref.so = the reference's own external/brisk sources compiled unchanged (oracle/Makefile)
    -> detect / describe / Hamming goldens.
  * the Python cv2 4.13.0 wheel of this image -> goldens for the three OpenCV-defined pieces of arithmetic that the
    oracle's cv shim restates (cv::eigen 2x2, 2x3*3x2 Mat product, cv::normalize(Vec3f)); see SURVEY.md 8(c).
Outputs are small .npz files next to this script; inputs are regenerated from seeds by okvis2x_b200.synthetic.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g  # noqa: E402

g.build()
from okvis2x_b200 import synthetic as syn  # noqa: E402
import refbind  # noqa: E402

CASES = {
    # name: (W, H, image kind, seed, rects, radius, abs_threshold, max_kpt)
    "scene320": (320, 240, "scene", 11, 120, 38.0, 150.0, 700),
    "noise256": (256, 192, "noise", 12, 0, 30.0, 5.0, 2500),
    "scene400": (400, 300, "scene", 13, 300, 50.0, 20.0, 700),
}
CAM = syn.EUROC["cams"][0]


def case_image(name):
    W, H, kind, seed, rects, *_ = CASES[name]
    return syn.scene_image(W, H, seed, rects=rects) if kind == "scene" else syn.noise_image(W, H, seed)


def case_maps(name):
    """Camera-awareness maps for the case: EuRoC cam0 intrinsics scaled to the case's width."""
    W, H = CASES[name][:2]
    s = W / 752.0
    f = (CAM["f"][0] * s, CAM["f"][1] * s)
    c = (CAM["c"][0] * s, CAM["c"][1] * s)
    rays, jac = syn.camera_maps(W, H, f, c, CAM["d"])
    return rays, jac, np.float32(f[0])


def main():
    import cv2
    ref = refbind.Ref()
    for name, (W, H, kind, seed, rects, radius, thr, maxk) in CASES.items():
        img = case_image(name)
        scores = ref.harris(img)
        cand = ref.maxima(img, int(thr))
        kp = ref.detect(img, radius, 0, thr, maxk)
        ext = refbind.RefExtractor(ref)
        kpp, dp = ext.describe(img, kp)
        rays, jac, focal = case_maps(name)
        ext.set_camera(rays, jac, focal)
        ext.set_direction(np.asarray((0.05, 0.99, 0.1), np.float32))
        kpa, da = ext.describe(img, kp)
        ext.close()
        best = ref.hamming_best(dp, da, 60) if len(dp) and len(da) else np.zeros(0, np.uint64)
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            image_sha256=np.frombuffer(hashlib.sha256(img.tobytes()).digest(), np.uint8),
                            scores_sha256=np.frombuffer(hashlib.sha256(scores.tobytes()).digest(), np.uint8),
                            scores_row=scores[H // 2].copy(), cand=cand, kp_detect=kp, kp_plain=kpp, desc_plain=dp,
                            kp_aware=kpa, desc_aware=da, best_plain_vs_aware=best)
        print(name, "cand", len(cand), "kp", len(kp), "described", len(kpp), len(kpa),
              "score ties among candidates:", len(cand) - len(np.unique(cand["score"])))

    # OpenCV arithmetic goldens (cv2 wheel)
    rng = np.random.default_rng(99)
    A = rng.normal(0, 1, (256, 2, 2)).astype(np.float32)
    A[:32, 0, 1] = 0.0  # the |p| <= eps branch
    A[32:64] *= 1e-3
    ev = np.stack([cv2.eigen(a)[1].reshape(2) for a in A]).astype(np.float32)
    J = rng.normal(0, 300, (256, 2, 3)).astype(np.float32)
    E = rng.normal(0, 0.003, (256, 3, 2)).astype(np.float32)
    P = np.stack([cv2.gemm(j, e, 1.0, None, 0.0) for j, e in zip(J, E)]).astype(np.float32)
    V = rng.normal(0, 1, (256, 3)).astype(np.float32)
    V[0] = 0
    # cv::normalize(Vec3f) is the matx.hpp template (v * (1. / norm(v)), product in double), NOT the array function
    # cv2.normalize exposes (that one scales in float and differs in the last ulp), so only the norm value is pinned
    # by cv2; the scaling step is restated from matx.hpp in the test.
    nrm = np.asarray([cv2.norm(v.reshape(1, 3), cv2.NORM_L2) for v in V], np.float64)
    np.savez_compressed(os.path.join(HERE, "opencv_arith.npz"), eig_in=A, eig_out=ev, gemm_a=J, gemm_b=E, gemm_out=P,
                        norm_in=V, norm_val=nrm, cv2_version=np.asarray(cv2.__version__))
    print("opencv_arith ok")


if __name__ == "__main__":
    main()
