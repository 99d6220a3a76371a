#!/usr/bin/env python
"""Where k_track spends a frame (stream 0, ns from %globaltimer): pixel loop / reduction / serial Gauss-Newton
tail, on the bench workload.  Needs the experiment build variants/libdvo_trkprof.so (-DDVO_TRK_PROF).
usage: trk_phases.py [frames] [streams]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "direct-visual-odometry_b200")); sys.path.insert(0, ROOT)
import dvo_b200
import bench

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 60
S = int(sys.argv[2]) if len(sys.argv) > 2 else 256
lib = dvo_b200.load_library(os.environ.get("DVO_LIB") or os.path.join(ROOT, "direct-visual-odometry_b200", "variants", "libdvo_trkprof.so"))
lib.dvo_debug_trk_prof.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
dv = dvo_b200.Dvo(lib=lib, n_streams=S, width=640, height=480, levels=3, culls=2, K=bench.default_K(), max_frames=2, max_keyframes=frames + 4)
dv.vo_reserve(frames + 2)
dev = dv.dev_alloc(S * 640 * 480 * 4)
d0, s0 = bench.depth_prior(S, seed=4321)
st = (C.c_ulonglong * 8)()
import numpy as np
lib.dvo_debug_trk_streams.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
ns = np.zeros(S, np.uint32)
prev = None
for i in range(frames):
    dv.synth_render_dev(dev, None, i, 1000, bench.TRANS_STEP, bench.ROT_STEP_DEG)
    dv.vo_push_dev(dev, d0 if i == 0 else None, s0 if i == 0 else None)
    it = dv.vo_counters()["gn_iters"].astype(np.int64)
    if prev is not None and i % 10 == 9:
        lib.dvo_debug_trk_streams(dv.ctx, ns.ctypes.data_as(C.c_void_p), S)
        d = it - prev
        us = ns / 1e3
        print("frame %3d: iterations/stream mean %.2f max %d hist %s | stream time in k_track: mean %.0f us  p50 %.0f  p90 %.0f  max %.0f" % (
            i, d.mean(), d.max(), np.bincount(d, minlength=8)[:12].tolist(), us.mean(), np.percentile(us, 50), np.percentile(us, 90), us.max()))
    prev = it
    if i % 20 == 19:
        lib.dvo_debug_trk_prof(dv.ctx, st, 1)
        n = 20.0
        print("frames %3d-%3d (per frame, stream 0): loop %.1f us  reduce %.1f us  tail %.1f us (solve %.1f  concatenate %.1f  exp+rest %.1f)  "
              "iterations %.2f  kernel %.1f us" % (i - 19, i, st[0] / n / 1e3, st[1] / n / 1e3, st[2] / n / 1e3, st[5] / n / 1e3, st[6] / n / 1e3,
                                                  st[7] / n / 1e3, st[3] / n, st[4] / n / 1e3))
dv.close()
