"""GPU parity at BASELINE.json's FULL sizes:
  C3  1920x1080 synthetic sequence, full VO loop — Frame(gray, K, 3, 2) with the reference's ROIs scaled x3 (480x270 finest
      level), and Frame(gray, K, 3, 0) (1920x1080 finest level);
  C4  3840x2160 mapping-only stress: Mapper::update on dense hypotheses (depth = max(randn(1.5, 0.5), 0.5), sigma = 0.5,
      frame.hpp:18-21) with rel_xi = (-0.1, 0, 0, 0, 0, 0) (test/update.cpp:58), ROI = whole frame, then Mapper::regularize.
The oracle runs the same inputs on the host (a few tens of seconds at 4K); masks / ages / counts bit-exact, maps bit-identical
(teacher-forced poses, see tests/test_gpu_vo.py)."""
import numpy as np
import pytest

from conftest import make_dvo
from test_gpu_vo import _run_pair

pytestmark = pytest.mark.gpu

# Per-iteration updates at these sizes: the tolerance stays north_star's 1e-5.  Where the two free-running trackers differ by more
# (observed once: 1.03e-5 on a 0.022 update at 480x270), the test solves the reference's stacked system (optimize.cpp:87-97) in fp64
# at EACH side's own pose (exact_solve, tests/test_gpu_vo.py): both results must be within 1e-5 of their exact solution - the
# difference is the poses' accumulated difference from the earlier iterations of the frame, not either solver.


def K_for(w, h):
    return np.array([[0.82 * w, 0, 0.5 * w - 0.5], [0, 0.82 * w, 0.5 * h - 0.5], [0, 0, 1]], np.float32)


def render_sequence(n, w, h, seed=1234, trans_step=0.006, rot_step_deg=0.2):
    """n frames + frame-0 ground-truth depth from the device ray-caster (csrc/dvo_synth.cuh): the numpy renderer needs ~10 s per
    1080p frame."""
    K = K_for(w, h)
    dv = make_dvo(width=w, height=h, levels=1, culls=0, K=K, max_frames=2, max_keyframes=2)
    gd = dv.dev_alloc(w * h * 4); dd = dv.dev_alloc(w * h * 4)
    g = np.empty((n, h, w), np.float32); d0 = np.empty((1, h, w), np.float32)
    for i in range(n):
        dv.synth_render_dev(gd, dd if i == 0 else None, i, seed, trans_step, rot_step_deg)
        dv.copy_to_host(g[i:i + 1], gd)
        if i == 0:
            dv.copy_to_host(d0, dd)
    dv.dev_free(gd); dv.dev_free(dd)
    dv.close()
    return g, d0[0], K


def test_c3_1080p_vo_reference_defaults_scaled(orc):
    """levels = 3, culls = 2 -> 480x270 finest level; optimize.cpp:33-36 / mapper.cpp:90 ROIs scaled x3."""
    g, d, K = render_sequence(8, 1920, 1080, trans_step=0.002, rot_step_deg=0.07)  # same pixel motion as 0.006 m / 0.2 deg at 640x480
    d0 = orc.cull_image(d, 2)
    s0 = np.full_like(d0, 0.1)
    roi = dict(roi_x0=60, roi_x1=420, roi_y0=60, roi_y1=300, map_roi_x0=48, map_roi_x1=432, map_roi_y0=36, map_roi_y1=324)
    n_kf, n_upd, _ = _run_pair(orc, g, K, d0, s0, levels=3, culls=2, exact_solve=True, **roi)
    assert n_kf + n_upd == 8 and n_upd >= 2


def test_c3_1080p_vo_full_resolution(orc):
    """levels = 3, culls = 0: the finest level is the 1920x1080 frame itself (2 M pixels per Gauss-Newton iteration)."""
    g, d, K = render_sequence(3, 1920, 1080, seed=99)
    s0 = np.full_like(d, 0.1)
    roi = dict(roi_x0=240, roi_x1=1680, roi_y0=240, roi_y1=1200, map_roi_x0=192, map_roi_x1=1728, map_roi_y0=144, map_roi_y1=1296)
    n_kf, n_upd, _ = _run_pair(orc, g, K, d, s0, levels=3, culls=0, exact_solve=True, **roi)
    assert n_kf + n_upd == 3


def plane_views(w, h, shift_px):
    """Two views of a fronto-parallel textured plane: the second is the first shifted along x (a pure x translation of a plane at
    constant depth).  Band-limited texture, seed 4321."""
    rng = np.random.default_rng(4321)
    x = np.arange(w + abs(int(shift_px)) + 1, dtype=np.float32)[None, :]
    y = np.arange(h, dtype=np.float32)[:, None]
    t = np.zeros((h, x.shape[1]), np.float32)
    for _ in range(8):
        fx, fy = rng.uniform(0.004, 0.08, 2).astype(np.float32)
        t += np.float32(rng.uniform(0.3, 1.0)) * np.sin(fx * x + fy * y + np.float32(rng.uniform(0, 6.28)))
    t = (0.5 + 0.45 * t / np.abs(t).max()).astype(np.float32)
    s = abs(int(shift_px))
    return np.ascontiguousarray(t[:, :w]), np.ascontiguousarray(t[:, s:s + w])


def test_c4_4k_mapping_update_regularize(orc):
    W, H = 3840, 2160
    K = K_for(W, H)
    Z, tx = 1.5, 0.1
    g_ref, g_obj = plane_views(W, H, K[0, 0] * tx / Z)
    rng = np.random.default_rng(4321)
    depth = np.maximum(rng.normal(1.5, 0.5, (H, W)), 0.5).astype(np.float32)   # Scene ctor prior, frame.hpp:18-21
    sigma = np.full((H, W), 0.5, np.float32)
    age = np.zeros((H, W), np.float32)
    rel = np.array([-tx, 0, 0, 0, 0, 0], np.float32)                            # test/update.cpp:58
    roi = dict(map_roi_x0=0, map_roi_x1=W - 1, map_roi_y0=0, map_roi_y1=H - 1)
    # ---- CUDA path
    dv = make_dvo(width=W, height=H, levels=1, culls=0, K=K, max_frames=3, rng_table=1 << 24, **roi)  # one mt19937 draw per reset pixel
    ref = dv.frame_create(g_ref); obj = dv.frame_create(g_obj)
    zero = np.zeros(6, np.float32)
    dv.frame_set_xi(ref, zero, zero)
    dv.frame_set_xi(obj, rel, rel)
    dv.frame_set_depth(ref, depth, sigma, age)
    vu, _ = dv.map_update(obj, ref, [ref], want_mask=False)
    got_u = [dv.frame_download(ref, 0, k)[0].copy() for k in ("depth", "sigma", "age")]
    dv.map_regularize(ref)
    got_r = dv.frame_download(ref, 0, "depth")[0].copy()
    dv.close()
    # ---- oracle on the host
    gx = orc.gradiate(g_ref, True); gy = orc.gradiate(g_ref, False)
    od, os_, oa, _, valid, rc = orc.mapper_update(depth, sigma, age, g_obj, [g_ref], [gx], [gy], np.zeros((1, 6), np.float32), rel, rel, K,
                                                  params=orc.params(**roi))
    assert rc == 0
    assert int(vu[0]) == valid and valid > 10000, (int(vu[0]), valid)
    assert got_u[2].tobytes() == oa.tobytes()                                    # reset mask (age) bit-exact
    assert got_u[0].tobytes() == od.tobytes() and got_u[1].tobytes() == os_.tobytes()
    assert got_r.tobytes() == orc.regularize(od, os_).tobytes()


def test_c3_500_frames_free_running_drift(orc):
    """BASELINE configs[2] as written: 500 frames of the synthetic 1920x1080 sequence, Frame(gray, K, 3, 2), FREE-RUNNING on both sides
    (no teacher forcing): how far the GPU trajectory drifts from the oracle's.  The numbers go to BASELINE.md; asserted is what
    holds by construction - identical behaviour until the first schedule mismatch, finite poses throughout."""
    W, H, n = 1920, 1080, 500
    K = K_for(W, H)
    roi = dict(roi_x0=60, roi_x1=420, roi_y0=60, roi_y1=300, map_roi_x0=48, map_roi_x1=432, map_roi_y0=36, map_roi_y1=324)
    dv = make_dvo(width=W, height=H, levels=3, culls=2, K=K, max_frames=2, max_keyframes=n + 2, **roi)
    vo = orc.vo_create(K, 3, 2, orc.params(**roi))
    gd = dv.dev_alloc(W * H * 4); dd = dv.dev_alloc(W * H * 4)
    g = np.empty((1, H, W), np.float32); d = np.empty((1, H, W), np.float32)
    first_sched = first_kf = None
    max_rel = 0.0; path = 0.0; prev = np.zeros(3)
    for i in range(n):
        dv.synth_render_dev(gd, dd if i == 0 else None, i, 1234, 0.002, 0.07)
        dv.copy_to_host(g, gd)
        if i == 0:
            dv.copy_to_host(d, dd)
            d0 = orc.cull_image(d[0], 2); s0 = np.full_like(d0, 0.5)
        To = vo.odometrize(g[0], d0 if i == 0 else None, s0 if i == 0 else None)
        Tg = dv.vo_push(g, d0 if i == 0 else None, s0 if i == 0 else None)[0]
        assert np.isfinite(Tg).all()
        st = dv.vo_state()
        if i:
            lg = [(a["level"], a["iteration"], a["valid_pixels"]) for a in dv.vo_last_log()[0]]
            lo = [(a["level"], a["iteration"], a["valid_pixels"]) for a in vo.last_log()]
            if first_sched is None and lg != lo:
                first_sched = i
            if first_sched is None:
                assert np.abs(Tg - To).max() <= 1e-4, i
        if first_kf is None and bool(st["was_keyframe"][0]) != vo.last_was_keyframe():
            first_kf = i
        path += float(np.linalg.norm(To[:3, 3] - prev)); prev = To[:3, 3].copy()
        max_rel = max(max_rel, float(np.linalg.norm(Tg[:3, 3] - To[:3, 3])) / max(1.0, float(np.linalg.norm(To[:3, 3]))))
    print("FREE-RUN c3_ %s" % dict(frames=n, first_schedule_mismatch=first_sched, first_keyframe_mismatch=first_kf,
                                   max_rel_translation_err=max_rel, path_length=path, final_translation_norm=float(np.linalg.norm(To[:3, 3]))))
    assert dv.vo_error() == 0
    dv.dev_free(gd); dv.dev_free(dd)
    dv.close()
