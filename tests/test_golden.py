"""Golden vectors: outputs of THE REFERENCE'S OWN CODE (oracle/_ref, generated here by
tests/golden/make_golden.py and committed as tests/golden/ref_golden.npz) against
  * the plain-C oracle  — CPU tests, bit-exact;
  * the CUDA path through the C ABI — `gpu` tests: masks / counts / ages / scatter winners bit-exact,
    per-iteration pose updates within 1e-5 (se3 norm), poses within 1e-4, depth and sigma maps
    within 1e-4 relative (north_star tolerances; in practice the maps are bit-identical).
Nothing here reads /root/reference: these run on the GPU box.
"""
import os

import numpy as np
import pytest

from helpers import se3_norm

HERE = os.path.dirname(os.path.abspath(__file__))
UPD_TOL, POSE_TOL, REL_TOL = 1e-5, 1e-4, 1e-4


@pytest.fixture(scope="module")
def G():
    return dict(np.load(os.path.join(HERE, "golden", "ref_golden.npz")))


def up4(a):
    return np.ascontiguousarray(np.repeat(np.repeat(a, 4, axis=0), 4, axis=1))


def gray_of(u8):
    return (u8.astype(np.float32) / np.float32(255.0)).astype(np.float32)


def depth_of(mm):
    return (mm.astype(np.float32) / np.float32(1000.0)).astype(np.float32)


def sigma_of(c):
    return (c.astype(np.float32) / np.float32(100.0)).astype(np.float32)


def track_inputs(G):
    g_ref = gray_of(G["track_ref_u8"]); g_ref[G["track_invalid"]] = -2.0
    return g_ref, gray_of(G["track_obj_u8"]), depth_of(G["track_depth_mm"]), sigma_of(G["track_sigma_c"])


def maps_close(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return a.tobytes() == b.tobytes() or bool(np.isclose(a, b, rtol=REL_TOL, atol=0).all())


# ================================================================================================
# CPU: oracle == golden, bit for bit
# ================================================================================================
def test_oracle_track_golden(orc, G):
    g_ref, g_obj, d, s = track_inputs(G)
    Kf = G["K_full"]
    f_ref = orc.frame_create(up4(g_ref), Kf, 3, 2, up4(d), up4(s))
    f_obj = orc.frame_create(up4(g_obj), Kf, 3, 2)
    assert np.array_equal(f_ref.K(2), G["K_top"])
    xi, logs = orc.track(f_obj, f_ref)
    assert np.array_equal(xi, G["track_xi"])
    assert np.array_equal(np.array([[L["level"], L["iteration"], L["valid_pixels"]] for L in logs], np.int32), G["track_log_meta"])
    assert np.array_equal(np.array([L["residual"] for L in logs], np.float32), G["track_log_residual"])
    assert np.array_equal(np.array([L["update"] for L in logs], np.float32), G["track_log_update"])
    assert np.array_equal(orc.warp_image(xi, f_ref.image(2, "gray"), f_ref.image(2, "depth"), f_ref.K(2)), G["track_warped_top"])
    assert np.array_equal(np.packbits(f_ref.image(2, "gradx") > -2.0), G["track_gradx_top_valid"])
    assert np.array_equal(np.packbits(f_ref.image(2, "grady") > -2.0), G["track_grady_top_valid"])
    for row in G["gn_cases"]:
        lv = int(row[0]); x = row[1:7].astype(np.float32)
        im = {k: f_ref.image(lv, k) for k in ("gray", "depth", "sigma", "gradx", "grady")}
        warped = orc.warp_image(x, im["gray"], im["depth"], f_ref.K(lv))
        u, r, v, _ = orc.optimize(f_obj.image(lv, "gray"), warped, im["depth"], im["sigma"], im["gradx"], im["grady"], f_ref.K(lv), x, lv)
        assert v == int(row[14]) and np.float32(r) == np.float32(row[13]) and np.array_equal(u, row[7:13].astype(np.float32))


def test_oracle_map_golden(orc, G):
    K = G["K_top"]
    for i, x in enumerate(G["prop_xi"]):
        d, s, a = orc.propagate(depth_of(G["prop_depth_mm"]), sigma_of(G["prop_sigma_c"]), G["prop_age"].astype(np.float32), x, K)
        assert np.array_equal(d, G["prop%d_depth" % i]) and np.array_equal(s, G["prop%d_sigma" % i])
        assert np.array_equal(a, G["prop%d_age" % i].astype(np.float32))
    assert np.array_equal(orc.regularize(depth_of(G["reg_depth_mm"]), sigma_of(G["reg_sigma_c"])), G["reg_out"])
    grays = [gray_of(u) for u in G["upd_u8"]]  # old keyframe, current keyframe, object frame
    born = [grays[1], grays[0]]
    xis = G["upd_xi"]
    out = orc.mapper_update(depth_of(G["upd_depth_mm"]), sigma_of(G["upd_sigma_c"]), G["upd_age"].astype(np.float32), grays[2], born,
                            [orc.gradiate(b, True) for b in born], [orc.gradiate(b, False) for b in born], np.stack([xis[1], xis[0]]),
                            xis[2], G["upd_rel"], K)
    od, os_, oa, mask, valid, rc = out
    assert rc == 0 and valid == int(G["upd_valid"]) and valid > 100
    assert np.array_equal(od, G["upd_out_depth"]) and np.array_equal(os_, G["upd_out_sigma"])
    assert np.array_equal(oa, G["upd_out_age"].astype(np.float32))


def test_oracle_vo_golden(orc, G):
    vo = orc.vo_create(G["K_full"], 3, 2)
    d0 = G["vo_init_depth"]; s0 = np.full_like(d0, 0.5)
    pos = 0
    for i, u8 in enumerate(G["vo_u8"]):
        T = vo.odometrize(up4(gray_of(u8)), d0 if i == 0 else None, s0 if i == 0 else None)
        assert np.array_equal(T, G["vo_T"][i]), i
        assert vo.last_was_keyframe() == bool(G["vo_was_keyframe"][i]) and vo.history_size() == int(G["vo_history_size"][i])
        n = int(G["vo_n_iters"][i])
        logs = vo.last_log() if i else []
        assert len(logs) == n
        for k, L in enumerate(logs):
            assert np.array_equal([L["level"], L["iteration"], L["valid_pixels"]], G["vo_log_meta"][pos + k])
            assert L["residual"] == G["vo_log_residual"][pos + k] and np.array_equal(L["update"], G["vo_log_update"][pos + k])
        pos += n
        if i and not vo.last_was_keyframe():
            assert vo.last_valid_update() == int(G["vo_valid_update"][i])
        if i:
            assert np.array_equal(vo.last_frame().xi()[1], G["vo_rel_xi"][i])
    fr = vo.ref_frame()
    assert np.array_equal(fr.image(2, "depth"), G["vo_final_depth"]) and np.array_equal(fr.image(2, "sigma"), G["vo_final_sigma"])
    assert np.array_equal(fr.image(2, "age"), G["vo_final_age"].astype(np.float32))
    assert np.array_equal(fr.xi()[0], G["vo_final_kf_xi"])


# ================================================================================================
# GPU: CUDA path (through the C ABI) against the reference's golden outputs
# ================================================================================================
@pytest.mark.gpu
def test_gpu_track_golden(G):
    from conftest import make_dvo
    g_ref, g_obj, d, s = track_inputs(G)
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=G["K_full"])
    rs = dv.frame_create(up4(g_ref)); dv.frame_set_depth(rs, up4(d), up4(s), None, full_res=True)
    os_ = dv.frame_create(up4(g_obj))
    assert np.array_equal(dv.level_K(2), G["K_top"])
    # validity masks of the gradients bit-exact
    assert np.array_equal(np.packbits(dv.frame_download(rs, 2, "gradx")[0] > -2.0), G["track_gradx_top_valid"])
    assert np.array_equal(np.packbits(dv.frame_download(rs, 2, "grady")[0] > -2.0), G["track_grady_top_valid"])
    # Track::optimize single steps: valid count exact, residual and update within tolerance
    for row in G["gn_cases"]:
        lv = int(row[0])
        upd, res, valid, _ = dv.gn_step(os_, rs, lv, row[1:7].astype(np.float32))
        assert int(valid[0]) == int(row[14])
        assert se3_norm(upd[0], row[7:13]) <= UPD_TOL
        assert abs(float(res[0]) - row[13]) <= 1e-6 * max(1.0, abs(row[13]))
    # Tracker::track: same schedule, same valid counts, updates within 1e-5, pose within 1e-4
    xi, logs = dv.track(os_, rs)
    meta = np.array([[L["level"], L["iteration"], L["valid_pixels"]] for L in logs[0]], np.int32)
    assert np.array_equal(meta, G["track_log_meta"])
    for L, u, r in zip(logs[0], G["track_log_update"], G["track_log_residual"]):
        assert se3_norm(L["update"], u) <= UPD_TOL
        assert abs(L["residual"] - r) <= 1e-6 * max(1.0, abs(r))
    assert se3_norm(xi[0], G["track_xi"]) <= POSE_TOL
    # Transform::warpImage at the reference's final pose: INVALID mask exact, values bit-exact
    w = dv.warp_image(rs, 2, G["track_xi"])[0]
    assert np.array_equal(w <= -2.0, G["track_warped_top"] <= -2.0)
    assert w.tobytes() == G["track_warped_top"].tobytes()
    dv.close()


@pytest.mark.gpu
def test_gpu_map_golden(G):
    from conftest import make_dvo
    K = G["K_top"]
    z = np.zeros((120, 160), np.float32)
    dv = make_dvo(width=160, height=120, levels=3, culls=0, K=K)
    a = dv.frame_create(z); b = dv.frame_create(z)
    for i, x in enumerate(G["prop_xi"]):
        dv.frame_set_depth(a, depth_of(G["prop_depth_mm"]), sigma_of(G["prop_sigma_c"]), G["prop_age"].astype(np.float32))
        dv.map_propagate(a, b, x)
        assert np.array_equal(dv.frame_download(b, 2, "age")[0], G["prop%d_age" % i].astype(np.float32))  # scatter winners
        assert maps_close(dv.frame_download(b, 2, "depth")[0], G["prop%d_depth" % i])
        assert maps_close(dv.frame_download(b, 2, "sigma")[0], G["prop%d_sigma" % i])
    dv.close()
    dv = make_dvo(width=50, height=50, levels=1, culls=0)
    s = dv.frame_create(np.zeros((50, 50), np.float32))
    dv.frame_set_depth(s, depth_of(G["reg_depth_mm"]), sigma_of(G["reg_sigma_c"]), None)
    dv.map_regularize(s)
    assert dv.frame_download(s, 0, "depth")[0].tobytes() == G["reg_out"].tobytes()
    dv.close()
    # Mapper::update, two-keyframe history
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=G["K_full"])
    slots = [dv.frame_create(up4(gray_of(u))) for u in G["upd_u8"]]  # old kf, current kf, object
    xis = G["upd_xi"]
    dv.frame_set_xi(slots[0], xis[0], np.zeros(6, np.float32))
    dv.frame_set_xi(slots[1], xis[1], np.zeros(6, np.float32))
    dv.frame_set_xi(slots[2], xis[2], G["upd_rel"])
    dv.frame_set_depth(slots[1], depth_of(G["upd_depth_mm"]), sigma_of(G["upd_sigma_c"]), G["upd_age"].astype(np.float32))
    vu, mask = dv.map_update(slots[2], slots[1], [slots[1], slots[0]])
    assert int(vu[0]) == int(G["upd_valid"])
    assert np.array_equal(dv.frame_download(slots[1], 2, "age")[0], G["upd_out_age"].astype(np.float32))  # reset mask
    assert maps_close(dv.frame_download(slots[1], 2, "depth")[0], G["upd_out_depth"])
    assert maps_close(dv.frame_download(slots[1], 2, "sigma")[0], G["upd_out_sigma"])
    dv.close()


@pytest.mark.gpu
def test_gpu_vo_golden(G):
    """Full VO against the reference's recorded run.  Teacher-forced (tests/test_gpu_vo.py explains
    why): every frame the GPU tracker runs freely from the same keyframe state and its iterations
    are compared; mapping then proceeds from the reference's own relative pose."""
    from conftest import make_dvo
    n = len(G["vo_u8"])
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=G["K_full"], max_frames=2, max_keyframes=n + 2)
    d0 = G["vo_init_depth"]; s0 = np.full_like(d0, 0.5)
    pos = 0
    for i, u8 in enumerate(G["vo_u8"]):
        if i:
            dv.vo_force_pose(G["vo_rel_xi"][i])
        T = dv.vo_push(up4(gray_of(u8)), d0 if i == 0 else None, s0 if i == 0 else None)
        st = dv.vo_state()
        assert dv.vo_error() == 0
        k = int(G["vo_n_iters"][i])
        if i:
            logs = dv.vo_last_log()[0]
            meta = np.array([[L["level"], L["iteration"], L["valid_pixels"]] for L in logs], np.int32).reshape(-1, 3)
            assert np.array_equal(meta, G["vo_log_meta"][pos:pos + k]), i
            for L, u in zip(logs, G["vo_log_update"][pos:pos + k]):
                assert se3_norm(L["update"], u) <= UPD_TOL
            assert se3_norm(logs[-1]["xi_after"], G["vo_rel_xi"][i]) <= POSE_TOL
        pos += k
        assert bool(st["was_keyframe"][0]) == bool(G["vo_was_keyframe"][i])
        assert int(st["n_keyframes"][0]) == int(G["vo_history_size"][i])
        if i and not G["vo_was_keyframe"][i]:
            assert int(st["valid_updates"][0]) == int(G["vo_valid_update"][i])
        tn = float(np.linalg.norm(G["vo_T"][i][:3, 3]))
        assert float(np.linalg.norm(T[0][:3, 3] - G["vo_T"][i][:3, 3])) <= POSE_TOL * max(1.0, tn)
        assert np.abs(T[0] - G["vo_T"][i]).max() <= POSE_TOL
    slot = dv.vo_ref_slot(0)
    assert np.array_equal(dv.frame_download(slot, 2, "age")[0], G["vo_final_age"].astype(np.float32))
    assert maps_close(dv.frame_download(slot, 2, "depth")[0], G["vo_final_depth"])
    assert maps_close(dv.frame_download(slot, 2, "sigma")[0], G["vo_final_sigma"])
    dv.close()
