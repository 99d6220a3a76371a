#!/usr/bin/env python
"""Golden vectors for Draw::visualize* (src/core/draw.cpp) from OpenCV itself (cv2 4.13 in this container; third party, not part of
/root/reference):
  * hsv2bgr : cv2.cvtColor(COLOR_HSV2BGR) for every H in [0, 180], V in [0, 255] at S = 255 (all that visualizeDepth produces) and
              at S = 128 (pins the general formula);
  * cvt_*   : cv2.convertScaleAbs(x, a, b) = saturate_cast<uchar>(|x a + b|) for the (a, b) pairs of draw.cpp on inputs where
              x a + b >= 0, where it is the same float multiply-add + cvRound + saturation as Mat::convertTo(CV_8UC1, a, b)
              (Mat::convertTo itself has no Python binding).
Writes tests/golden/draw_golden.npz.  usage: python tests/golden/make_draw.py"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
out = {}
H, V = np.meshgrid(np.arange(181, dtype=np.uint8), np.arange(256, dtype=np.uint8), indexing="ij")
for S in (255, 128):
    hsv = np.stack([H, np.full_like(H, S), V], -1)
    out["hsv2bgr_s%d" % S] = cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR)
rng = np.random.default_rng(20251104)
for name, a, b, lo, hi in (("gray", 255.0, 0.0, 0.0, 1.2), ("depth_raw", 100.0, 0.0, 0.5, 2.5), ("sigma", -500.0, 255.0, 0.0, 0.51),
                           ("age", 10.0, 0.0, 0.0, 30.0)):
    x = np.concatenate([rng.uniform(lo, hi, 4096), np.linspace(lo, hi, 1024), (np.arange(512) + 0.5) / max(abs(a), 1.0)]).astype(np.float32)
    x = x[(x * np.float32(a) + np.float32(b)) >= 0].reshape(1, -1)
    out["cvt_%s_x" % name] = x
    out["cvt_%s_y" % name] = cv2.convertScaleAbs(x, alpha=a, beta=b)
    out["cvt_%s_ab" % name] = np.array([a, b], np.float32)
out["cv2_version"] = np.array(cv2.__version__)
np.savez_compressed(os.path.join(HERE, "draw_golden.npz"), **out)
print({k: (v.shape, v.dtype) for k, v in out.items()})
