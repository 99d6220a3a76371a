#!/usr/bin/env python
"""Small hot-path runs for compute-sanitizer (scripts/sanitize.sh): every hand-rolled synchronisation of the library on inputs small
enough for a 50-100x slowdown.
  A  4 streams x 160x120: k_track takes the cooperative B > 1 split (inter-CTA barrier: atomicAdd arrive / gen publish / spin),
     propagate + update + regularize as separate kernels, 10 frames with keyframe switches
  B  48 streams x 160x120: one CTA per stream, propagate in k_track's epilogue, k_map_update as a programmatic dependent
     (ticket + ready list), scan by the last phase-A CTA, commit by the last regularize CTA, 8 frames, 3+ keyframes per stream
  C  1 stream 320x240 at culls = 0, ROI = whole frame: 1 200 phase-A CTAs into the last-CTA scan, phase B ranks, mt19937 replay;
     granular entry points (Mapper::update with a 4-keyframe history, propagate with collisions, regularize)
Results are checked for finiteness only - parity is pytest's job; this run is about hazards."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "direct-visual-odometry_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import dvo_b200  # noqa: E402

W, H = 640, 480


def K_of(w, h):
    return np.array([[0.82 * w, 0, 0.5 * w - 0.5], [0, 0.82 * w, 0.5 * h - 0.5], [0, 0, 1]], np.float32)


def vo_case(S, n_frames, tag):
    dv = dvo_b200.Dvo(n_streams=S, width=W, height=H, levels=3, culls=2, K=K_of(W, H), max_frames=2, max_keyframes=n_frames + 2)
    nb = S * H * W * 4
    g = dv.dev_alloc(nb)
    rng = np.random.default_rng(7)
    d0 = np.maximum(rng.normal(1.5, 0.5, (S, 120, 160)), 0.5).astype(np.float32)
    for i in range(n_frames):
        dv.synth_render_dev(g, None, i, 500, 0.012, 0.3)
        T = dv.vo_push_dev(g, d0 if i == 0 else None, np.full_like(d0, 0.5) if i == 0 else None, want_pose=True)
        assert np.isfinite(T).all()
    st = dv.vo_state()
    assert dv.vo_error() == 0
    print("%s: %d streams x %d frames, keyframes per stream %d..%d, launches %d" % (tag, S, n_frames, st["n_keyframes"].min(),
                                                                                   st["n_keyframes"].max(), dv.launch_count()))
    dv.dev_free(g)
    dv.close()


def granular_case():
    w, h = 320, 240
    dv = dvo_b200.Dvo(n_streams=1, width=w, height=h, levels=1, culls=0, K=K_of(w, h), max_frames=8, max_keyframes=8,
                      map_roi_x0=0, map_roi_x1=w - 1, map_roi_y0=0, map_roi_y1=h - 1, rng_table=1 << 12)
    rng = np.random.default_rng(3)
    xs = np.arange(w + 40, dtype=np.float32)[None]; ys = np.arange(h, dtype=np.float32)[:, None]
    tex = (0.5 + 0.2 * np.sin(0.11 * xs + 0.05 * ys) + 0.2 * np.sin(0.31 * xs - 0.17 * ys)).astype(np.float32)
    slots = []
    for k in range(4):
        s = dv.frame_create(np.ascontiguousarray(tex[:, 4 * k:4 * k + w]))
        xi = np.array([-0.02 * k, 0, 0, 0, 0, 0], np.float32)
        dv.frame_set_xi(s, xi, xi)
        slots.append(s)
    obj = dv.frame_create(np.ascontiguousarray(tex[:, 20:20 + w]))
    rel = np.array([-0.1, 0, 0, 0, 0, 0], np.float32)
    dv.frame_set_xi(obj, rel, rel)
    depth = np.maximum(rng.normal(1.5, 0.5, (h, w)), 0.5).astype(np.float32)
    age = rng.integers(0, 4, (h, w)).astype(np.float32)
    dv.frame_set_depth(slots[3], depth, np.full_like(depth, 0.5), age)
    vu, mask = dv.map_update(obj, slots[3], slots[::-1])          # history_slots[age]
    dv.map_regularize(slots[3])
    nw = dv.frame_create(np.ascontiguousarray(tex[:, 24:24 + w]))
    dv.map_propagate(slots[3], nw, np.array([0.01, -0.01, 0.3, 0, 0, 0.02], np.float32))   # forward motion: many collisions
    d = dv.frame_download(nw, 0, "depth")
    assert np.isfinite(d).all()
    print("C: Mapper::update valid %d, resets %d (mt19937 draws %d), propagate + regularize ok" % (int(vu[0]), int((mask & 8).astype(bool).sum()),
                                                                                                 int(dv.rng_position()[0])))
    dv.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "ABC"
    if "A" in which:
        vo_case(4, 10, "A")
    if "B" in which:
        vo_case(48, 8, "B")
    if "C" in which:
        granular_case()
