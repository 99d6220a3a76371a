#!/usr/bin/env python
"""Why the steady-state epipolar search cannot share memory traffic between neighbouring pixels: runs bench.py's
sequence 0 through the oracle (CPU, ~10 s) and prints, at frames 40 / 150 / 300 / 459, how the AGE of the keyframe's pixels (= which
born keyframe a pixel's search reads, mapper.cpp:99-101) is distributed over Mapper::update's ROI."""
import sys, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/direct-visual-odometry_b200')
import numpy as np
import bench
from dvo_b200 import synth
from oracle.oracle import Oracle
orc = Oracle(fast=True)
n_ring = 48
g, d, poses, K = synth.sequence(n_ring, 640, 480, K=bench.default_K(), seed=1000, trans_step=bench.TRANS_STEP, rot_step_deg=bench.ROT_STEP_DEG)
d0, s0 = synth.random_depth_prior((120, 160), seed=4321)
vo = orc.vo_create(K)
t0 = time.time()
for i in range(460):
    f = bench.ring_index(i, n_ring)
    vo.odometrize(g[f], d0 if i == 0 else None, s0 if i == 0 else None)
    if i in (40, 150, 300, 459):
        fr = vo.ref_frame()
        a = fr.image(2, "age"); dep = fr.image(2, "depth")
        roi = a[12:108, 16:144]
        same_x = (roi[:, 1:] == roi[:, :-1]).mean()
        same_y = (roi[1:, :] == roi[:-1, :]).mean()
        vals, cnt = np.unique(roi, return_counts=True)
        order = np.argsort(-cnt)
        print("frame %d: keyframes %d, distinct ages in ROI %d, same age as x-neighbour %.2f, y-neighbour %.2f; top ages %s; age>0 frac %.2f (%.0fs)"
              % (i, vo.history_size(), len(vals), same_x, same_y, [(int(vals[k]), int(cnt[k])) for k in order[:8]], (roi > 0).mean(), time.time()-t0), flush=True)
        # run lengths along rows
        runs = []
        for r in roi:
            ch = np.flatnonzero(np.diff(r) != 0)
            runs.append(np.diff(np.concatenate([[-1], ch, [len(r)-1]])))
        runs = np.concatenate(runs)
        print("   mean run length of equal age along a row: %.1f, median %.0f; pixels in runs >= 8: %.2f" % (runs.mean(), np.median(runs), runs[runs >= 8].sum() / runs.sum()))
