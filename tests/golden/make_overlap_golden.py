"""Goldens for the keyframe-overlap masks (Frontend::doWeNeedANewKeyframe, Frontend.cpp:1186-1295), drawn with the real
cv2.circle / bitwise_and / bitwise_or / countNonZero of OpenCV 4.13 in this container.
Run from the repo root: python tests/golden/make_overlap_golden.py"""
import os
import numpy as np
import cv2

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
               ("class_id", "<i4")])
rng = np.random.default_rng(42)
views = [(752, 480, 700), (752, 480, 40), (720, 540, 350), (1388, 700, 1000), (640, 480, 0), (333, 211, 90)]
width, height, offs, kps, matched, inter, uni = [], [], [0], [], [], [], []
for w, h, n in views:
    kp = np.zeros(n, KP)
    kp["x"] = rng.uniform(0, w, n).astype(np.float32)
    kp["y"] = rng.uniform(0, h, n).astype(np.float32)
    if n >= 8:   # centres exactly on .5 cells (round-half-even) and on the borders
        kp["x"][:8] = (5.0, 15.0, 25.0, 0.0, w - 0.01, 45.0, 4.99, 5.01)
        kp["y"][:8] = (5.0, 25.0, 15.0, 0.0, h - 0.01, 35.0, 4.99, 5.01)
    m = (rng.uniform(0, 1, n) < 0.4).astype(np.uint8)
    rows, cols = h // 10, w // 10
    det = np.zeros((rows, cols), np.uint8)
    mat = np.zeros((rows, cols), np.uint8)
    radius = int(float(min(rows, cols)) * 0.09)
    for k in range(n):
        # keypoint.pt * 0.1 (Point2f * double -> float), then cv::Point(Point2f) = cvRound
        c = (int(np.rint(np.float32(np.float64(kp["x"][k]) * 0.1))), int(np.rint(np.float32(np.float64(kp["y"][k]) * 0.1))))
        cv2.circle(det, c, radius, 255, cv2.FILLED)
        if m[k]:
            cv2.circle(mat, c, radius, 255, cv2.FILLED)
    inter.append(cv2.countNonZero(cv2.bitwise_and(mat, det)))
    uni.append(cv2.countNonZero(cv2.bitwise_or(mat, det)))
    width.append(w); height.append(h); offs.append(offs[-1] + n); kps.append(kp); matched.append(m)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "overlap_goldens.npz"),
                    width=np.array(width, np.int32), height=np.array(height, np.int32), kp_offset=np.array(offs, np.int64),
                    keypoints=np.concatenate(kps), matched=np.concatenate(matched), intersection=np.array(inter, np.int32),
                    union=np.array(uni, np.int32), opencv=cv2.__version__)
print(inter, uni)
