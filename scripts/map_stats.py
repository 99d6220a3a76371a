#!/usr/bin/env python
"""Search-length statistics of Mapper::update's epipolar search on the bench workload (debug build
libdvo_cuda_stats.so, -DDVO_MAP_STATS).  usage: map_stats.py [frames] [streams]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "direct-visual-odometry_b200")); sys.path.insert(0, ROOT)
import dvo_b200
import bench

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 120
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
lib = dvo_b200.load_library(os.path.join(ROOT, "direct-visual-odometry_b200", "libdvo_cuda_stats.so"))
lib.dvo_debug_map_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
dv = dvo_b200.Dvo(lib=lib, n_streams=S, width=640, height=480, levels=3, culls=2, K=bench.default_K(), max_frames=2, max_keyframes=frames + 4)
dv.vo_reserve(frames + 2)
fb = S * 640 * 480 * 4
dev = dv.dev_alloc(fb)
d0, s0 = bench.depth_prior(S, seed=4321)
st = (C.c_ulonglong * 8)(); hist = (C.c_uint * 128)()
for i in range(frames):
    dv.synth_render_dev(dev, None, i, 1000, bench.TRANS_STEP, bench.ROT_STEP_DEG)
    dv.vo_push_dev(dev, d0 if i == 0 else None, s0 if i == 0 else None)
    if i % 20 == 19 or i == frames - 1:
        lib.dvo_debug_map_stats(st, hist, 1)
        h = np.array(hist[:40], np.float64)
        print("frames %3d-%3d: steps/warp %.1f  lane efficiency %.3f  warps %d  active lanes/warp %.1f  distinct keyframes/warp %.2f  hist(1..12) %s" % (
            i - 19, i, st[0] / max(st[2], 1), st[0] / max(st[1], 1), st[2], st[4] / max(st[2], 1), st[3] / max(st[2], 1),
            np.round(h[1:13] / max(h.sum(), 1), 3).tolist()))
dv.close()
