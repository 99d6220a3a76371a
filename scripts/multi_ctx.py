#!/usr/bin/env python
"""Does splitting the 256 streams of a GPU over several dvo contexts (one CUDA stream each) help?  Kernels of different contexts
can overlap, so one context's launch tails (k_track waits for its slowest stream) fill with another context's work.
usage: multi_ctx.py [frames] [total streams]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "direct-visual-odometry_b200")); sys.path.insert(0, ROOT)
import dvo_b200
import bench

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 100
S_total = int(sys.argv[2]) if len(sys.argv) > 2 else 256
lib = dvo_b200.load_library()
n_res = 48
for n_ctx in (1, 2, 4):
    S = S_total // n_ctx
    fb = S * 640 * 480 * 4
    ctxs = []
    for k in range(n_ctx):
        dv = dvo_b200.Dvo(lib=lib, n_streams=S, width=640, height=480, levels=3, culls=2, K=bench.default_K(), max_frames=2, max_keyframes=frames + 8)
        dv.vo_reserve(frames + 6)
        dev = dv.dev_alloc(fb * n_res)
        for i in range(n_res):
            dv.synth_render_dev(dev + i * fb, None, i, 1000 + k * S, bench.TRANS_STEP, bench.ROT_STEP_DEG)
        d0, s0 = bench.depth_prior(S, seed=4321 + k)
        ctxs.append((dv, dev, d0, s0))

    def fptr(dev, i):
        per = 2 * (n_res - 1)
        j = i % per
        return dev + (j if j < n_res else per - j) * fb

    for dv, dev, d0, s0 in ctxs:
        dv.vo_push_dev(dev, d0, s0)
    for f in range(1, 4):
        for dv, dev, _, _ in ctxs:
            dv.vo_push_dev(fptr(dev, f))
    for dv, *_ in ctxs:
        dv.sync()
    t0 = time.perf_counter()
    for f in range(4, 4 + frames):
        for dv, dev, _, _ in ctxs:
            dv._ck(lib.dvo_vo_push_dev(dv.ctx, fptr(dev, f), None, None, None))
    for dv, *_ in ctxs:
        dv.sync()
    dt = time.perf_counter() - t0
    print("%d context(s) x %d streams: %.0f frames/s (%.3f ms per %d-frame step)" % (n_ctx, S, S_total * frames / dt, 1e3 * dt / frames, S_total), flush=True)
    for dv, dev, _, _ in ctxs:
        dv.dev_free(dev); dv.close()
