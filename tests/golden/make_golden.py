#!/usr/bin/env python
"""Generates tests/golden/ref_golden.npz by running THE REFERENCE'S OWN CODE — its unmodified
translation units under /root/reference compiled over the OpenCV-API shim (oracle/_ref, see
oracle/Makefile and oracle/ref_capi.cpp) — on small seeded inputs.  The reference ships no golden
vectors of its own (its test/ programs assert nothing), so these are the committed
"outputs of the reference itself run here" that the oracle and the CUDA path are checked against on
machines where /root/reference does not exist (the GPU box).

    python tests/golden/make_golden.py        # needs /root/reference; rewrites ref_golden.npz

Inputs are stored compactly and exactly: gray as uint8 (the loader's gray = u8/255,
src/core/loader.cpp:61), depth as uint16 millimetres (KinectLoader's depth/5000-style integer
source), sigma as integer hundredths, age as uint8.  A W x H top-level image is turned into the
4W x 4H input of Frame(gray, K, 3, 2) by pixel replication, which Convert::cullImage (top-left
sample of every 4x4 block, convert.cpp:7-20) inverts exactly.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "direct-visual-odometry_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from dvo_b200 import synth  # noqa: E402
from oracle.ref import Ref  # noqa: E402

W, H = 160, 120
K_TOP = np.array([[0.82 * W, 0, 0.5 * W - 0.5], [0, 0.82 * W, 0.5 * H - 0.5], [0, 0, 1]], np.float32)


def k_full():
    K = (K_TOP * np.float32(4.0)).astype(np.float32)
    K[2, 2] = 1.0
    return K


def up4(a):
    return np.ascontiguousarray(np.repeat(np.repeat(a, 4, axis=0), 4, axis=1))


def gray_of(u8):
    return (u8.astype(np.float32) / np.float32(255.0)).astype(np.float32)


def depth_of(mm):
    return (mm.astype(np.float32) / np.float32(1000.0)).astype(np.float32)


def sigma_of(c):
    return (c.astype(np.float32) / np.float32(100.0)).astype(np.float32)


def render(n, seed, trans_step, rot_step, invalid_border=False):
    g, d, poses, _ = synth.sequence(n, W, H, K=K_TOP, seed=seed, trans_step=trans_step, rot_step_deg=rot_step)
    u8 = np.clip(np.rint(g * 255.0), 0, 255).astype(np.uint8)
    mm = np.clip(np.rint(d * 1000.0), 0, 65535).astype(np.uint16)
    return u8, mm, poses


def logs_to_arrays(logs):
    meta = np.array([[L["level"], L["iteration"], L["valid_pixels"]] for L in logs], np.int32).reshape(-1, 3)
    res = np.array([L["residual"] for L in logs], np.float32)
    upd = np.array([L["update"] for L in logs], np.float32).reshape(-1, 6)
    return meta, res, upd


def main():
    ref = Ref()
    out = {"K_top": K_TOP, "K_full": k_full()}
    Kf = k_full()

    # ---- A. Tracker::track + Track::optimize + Transform::warpImage (main.cpp geometry) --------
    u8, mm, _ = render(3, 1234, 0.006, 0.25)
    invalid = np.zeros((H, W), bool)
    invalid[:, :3] = True; invalid[40:48, 100:118] = True       # remap border / holes -> INVALID (-2.0)
    sig = np.full((H, W), 30, np.uint8)
    out.update(track_ref_u8=u8[0], track_obj_u8=u8[2], track_depth_mm=mm[0], track_sigma_c=sig, track_invalid=invalid)
    g_ref = gray_of(u8[0]); g_ref[invalid] = -2.0
    g_obj = gray_of(u8[2])
    d_ref, s_ref = depth_of(mm[0]), sigma_of(sig)
    f_ref = ref.frame_create(up4(g_ref), Kf, 3, 2, up4(d_ref), up4(s_ref))
    f_obj = ref.frame_create(up4(g_obj), Kf, 3, 2)
    xi, logs = ref.track(f_obj, f_ref)
    meta, res, upd = logs_to_arrays(logs)
    out.update(track_xi=xi, track_log_meta=meta, track_log_residual=res, track_log_update=upd)
    out["track_warped_top"] = ref.warp_image(xi, f_ref.image(2, "gray"), f_ref.image(2, "depth"), f_ref.K(2))
    out["track_gradx_top_valid"] = np.packbits(f_ref.image(2, "gradx") > -2.0)
    out["track_grady_top_valid"] = np.packbits(f_ref.image(2, "grady") > -2.0)
    gn_xi = np.array([[0, 0, 0, 0, 0, 0], [0.004, -0.002, 0.003, 0.001, -0.002, 0.0015]], np.float32)
    gn = []
    for lv in range(3):
        for x in gn_xi:
            u, r, v = ref.optimize(f_obj.image(lv, "gray"), f_ref.image(lv, "gray"), f_ref.image(lv, "depth"), f_ref.image(lv, "sigma"),
                                   f_ref.K(lv), x, lv)
            gn.append(np.concatenate([[lv], x, u, [r, v]]).astype(np.float64))
    out["gn_cases"] = np.array(gn)  # level, xi[6], update[6], residual, valid

    # ---- B. Implement::propagate / regularize (test/propagate.cpp, test/regularize.cpp recipes) --
    rng = np.random.default_rng(31)
    p_mm = mm[1].copy(); p_mm[40:44, 50:90] = 0
    p_sig = rng.integers(2, 60, size=(H, W)).astype(np.uint8)
    p_age = rng.integers(0, 4, size=(H, W)).astype(np.uint8)
    p_xi = np.array([[0.01, -0.01, 0, 0, 0, 0], [0.03, 0.02, -0.15, 0.02, -0.03, 0.05]], np.float32)
    out.update(prop_depth_mm=p_mm, prop_sigma_c=p_sig, prop_age=p_age, prop_xi=p_xi)
    for i, x in enumerate(p_xi):
        d, s, a = ref.propagate(depth_of(p_mm), sigma_of(p_sig), p_age.astype(np.float32), x, K_TOP)
        out["prop%d_depth" % i], out["prop%d_sigma" % i], out["prop%d_age" % i] = d, s, a.astype(np.uint8)
        assert np.array_equal(a, a.astype(np.uint8).astype(np.float32))
    r_mm = np.clip(np.rint(rng.normal(1.5, 0.5, size=(50, 50)) * 1000), 100, 65535).astype(np.uint16)
    r_sig = np.clip(np.rint(np.abs(rng.normal(0.3, 0.2, size=(50, 50))) * 100), 1, 255).astype(np.uint8)
    r_mm[25, 25], r_sig[25, 25] = 3000, 10
    r_mm[3, 3:9] = 7500
    out.update(reg_depth_mm=r_mm, reg_sigma_c=r_sig, reg_out=ref.regularize(depth_of(r_mm), sigma_of(r_sig)))

    # ---- C. Mapper::update with a two-keyframe history (wide baseline so the gate accepts) -------
    u8w, mmw, poses = render(4, 12, 0.2, 0.2)
    xis = [np.asarray(ref.se3_log(np.linalg.inv(p.astype(np.float64)).astype(np.float32))) for p in poses]
    m_mm = np.clip(np.rint(mmw[1] * rng.uniform(0.9, 1.1, size=(H, W))), 1, 65535).astype(np.uint16)
    m_sig = rng.integers(5, 40, size=(H, W)).astype(np.uint8)
    m_age = rng.integers(0, 2, size=(H, W)).astype(np.uint8)
    rel = ref.se3_concatenate(xis[3], -xis[1])
    f_old, f_kf, f_o = (ref.frame_create(up4(gray_of(u8w[i])), Kf, 3, 2) for i in (0, 1, 3))
    f_kf.set_depth_sigma_age(depth_of(m_mm), sigma_of(m_sig), m_age.astype(np.float32))
    f_old.set_xi(xi=xis[0]); f_kf.set_xi(xi=xis[1]); f_o.set_xi(xi=xis[3], rel_xi=rel, ref_frame=f_kf)
    ref.rng_reset()
    valid = ref.mapper_update([f_kf, f_old], f_o)
    out.update(upd_u8=np.stack([u8w[0], u8w[1], u8w[3]]), upd_depth_mm=m_mm, upd_sigma_c=m_sig, upd_age=m_age,
               upd_xi=np.stack([xis[0], xis[1], xis[3]]).astype(np.float32), upd_rel=rel, upd_valid=np.int32(valid),
               upd_out_depth=f_kf.image(2, "depth"), upd_out_sigma=f_kf.image(2, "sigma"),
               upd_out_age=f_kf.image(2, "age").astype(np.uint8))

    # ---- D. VisualOdometry::odometrize over a fast-motion sequence ------------------------------
    u8v, _, _ = render(9, 5, 0.08, 0.3)
    d0, _ = synth.random_depth_prior((H, W), seed=4321)
    ref.rng_reset()
    vo = ref.vo_create(Kf)
    T, kf, vu, nh, n_it, metas, ress, upds, xi_l, rel_l = [], [], [], [], [], [], [], [], [], []
    for i in range(len(u8v)):
        T.append(vo.odometrize(up4(gray_of(u8v[i])), d0 if i == 0 else None))
        kf.append(vo.last_was_keyframe()); vu.append(vo.last_valid_update()); nh.append(vo.history_size())
        m, r, u = logs_to_arrays(vo.last_log())
        n_it.append(len(r)); metas.append(m); ress.append(r); upds.append(u)
        # the frame's relative pose = Tracker::track's fold xi <- concatenate(xi, update) (tracker.cpp:46-51),
        # recomputed with the reference's own se3::concatenate from the exactly-logged updates
        x = np.zeros(6, np.float32)
        for uu in u:
            nx = ref.se3_concatenate(x, uu)
            if not np.isnan(nx).any():
                x = nx
        rel_l.append(x)
        if i > 0 and vo.last_was_keyframe():
            assert np.array_equal(vo.ref_frame().xi()[1], x)
    fr = vo.ref_frame()
    out.update(vo_u8=u8v, vo_init_depth=d0, vo_T=np.stack(T), vo_was_keyframe=np.array(kf, np.uint8), vo_valid_update=np.array(vu, np.int32),
               vo_history_size=np.array(nh, np.int32), vo_n_iters=np.array(n_it, np.int32), vo_log_meta=np.concatenate(metas),
               vo_log_residual=np.concatenate(ress), vo_log_update=np.concatenate(upds),
               vo_final_depth=fr.image(2, "depth"), vo_final_sigma=fr.image(2, "sigma"), vo_final_age=fr.image(2, "age").astype(np.uint8),
               vo_final_kf_xi=fr.xi()[0], vo_rel_xi=np.stack(rel_l))
    path = os.path.join(HERE, "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote %s: %d arrays, %.0f KiB" % (path, len(out), os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
