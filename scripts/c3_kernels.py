#!/usr/bin/env python
"""Per-kernel device times of bench.py's C3 configuration (ONE synthetic 1920x1080 stream, full VO loop) at three points of the
500-frame run, with and without the per-kernel event records:

    python scripts/c3_kernels.py [culls ...]        # default: 2 0

Prints, per window, the plain ms per frame (50 frames) and {kernel: (ms per frame, launches)} (the next 50 frames)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "direct-visual-odometry_b200"))
import bench  # noqa: E402
import dvo_b200  # noqa: E402


def main():
    lib = dvo_b200.load_library()
    W, H, n_res = 1920, 1080, 40
    K = np.array([[0.82 * W, 0, 0.5 * W - 0.5], [0, 0.82 * W, 0.5 * H - 0.5], [0, 0, 1]], np.float32)
    for culls in [int(a) for a in sys.argv[1:]] or [2, 0]:
        sc = 3 if culls else 12
        roi = dict(roi_x0=20 * sc, roi_x1=140 * sc, roi_y0=20 * sc, roi_y1=100 * sc,
                   map_roi_x0=16 * sc, map_roi_x1=144 * sc, map_roi_y0=12 * sc, map_roi_y1=108 * sc)
        dv = dvo_b200.Dvo(lib=lib, n_streams=1, width=W, height=H, levels=3, culls=culls, K=K, max_frames=2,
                          max_keyframes=520, rng_table=(1 << 24) if culls else (1 << 28), **roi)
        dv.vo_reserve(504)
        fb = W * H * 4
        dev = dv.dev_alloc(fb * n_res); dd = dv.dev_alloc(fb)
        for i in range(n_res):
            dv.synth_render_dev(dev + i * fb, dd if i == 0 else None, i, 1234, 0.002, 0.07)
        d0 = np.empty((1, H, W), np.float32)
        dv.copy_to_host(d0, dd)
        d0 = np.ascontiguousarray(d0[0, ::1 << culls, ::1 << culls])
        dv.vo_push_dev(dev, d0, np.full_like(d0, 0.1))
        f = 1
        for lo in (4, 200, 400):
            while f < lo:
                dv.vo_push_dev(dev + bench.ring_index(f, n_res) * fb); f += 1
            dv.timer_start()
            for _ in range(50):
                dv.vo_push_dev(dev + bench.ring_index(f, n_res) * fb); f += 1
            ms = dv.timer_stop()
            dv.kernel_timing(True)
            for _ in range(50):
                dv.vo_push_dev(dev + bench.ring_index(f, n_res) * fb); f += 1
            kt = dv.kernel_timing(False)
            c1 = dv.vo_counters()
            print("C3 culls=%d frames %d..%d: plain %.3f ms/frame; keyframes so far %d; per kernel (ms/frame, launches): %s"
                  % (culls, lo, lo + 99, ms / 50, int(c1["n_propagate"].sum()),
                     {k: (round(v[0] / 50, 4), v[1]) for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])}), flush=True)
        dv.dev_free(dev); dv.dev_free(dd)
        dv.close()


if __name__ == "__main__":
    main()
