#!/usr/bin/env python3
"""Generates tests/golden/brisk_img1_gray_752x480.npz (run in the build container, where /root/reference exists): the
reference's own demo image external/brisk/images/img1.ppm, converted to grey the way brisk/src/demo-free.cc:101-118
does (cv::cvtColor BGR2GRAY, Python cv2 4.13 here) and cropped to 752 x 480 at (24, 80). A natural-texture input for
the matcher sweep's "natural descriptors" variant (bench.py) and the detect / describe parity tests."""
import os
import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
img = cv2.imread("/root/reference/external/brisk/images/img1.ppm")
g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
np.savez_compressed(os.path.join(HERE, "brisk_img1_gray_752x480.npz"), image=np.ascontiguousarray(g[80:560, 24:776]))
