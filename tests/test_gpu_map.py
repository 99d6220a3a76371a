"""GPU parity: Map::Implement::{regularize, propagate, update} + Mapper::update
(src/map/implement.cpp:156-256, src/map/mapper.cpp:76-137) against the CPU oracle, through the
C ABI.  Masks / ages / scatter winners bit-exact; depth and sigma bit-exact as well (same fp32
operation order, no FMA) — the 1e-4 relative tolerance of north_star is the fallback stated in the
assertion helper."""
import numpy as np
import pytest

from conftest import make_dvo
from helpers import punch_invalid

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4  # north_star: inverse-depth and variance maps within 1e-4 relative


def assert_maps(a, b, what):
    a = np.asarray(a); b = np.asarray(b)
    if a.tobytes() == b.tobytes():
        return
    bad = ~np.isclose(a, b, rtol=REL_TOL, atol=0)
    assert not bad.any(), "%s: %d px beyond %g relative (max abs %g)" % (what, int(bad.sum()), REL_TOL, float(np.abs(a - b).max()))


# ------------------------------------------------------------------------------------------------
# regularize — test/regularize.cpp:26-38 recipe (50x50 randn with one confident spike) and more
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("w,h,seed", [(50, 50, 0), (160, 120, 1), (257, 131, 2), (1, 1, 3), (7, 1, 4), (1, 9, 5)])
def test_regularize_bit_exact(orc, w, h, seed):
    rng = np.random.default_rng(seed)
    depth = rng.normal(1.5, 0.5, (h, w)).astype(np.float32)
    sigma = np.abs(rng.normal(0.3, 0.2, (h, w))).astype(np.float32) + np.float32(1e-3)
    if w > 30:
        depth[h // 2, w // 2] = 3.0; sigma[h // 2, w // 2] = 0.1     # test/regularize.cpp:30-31
        depth[3, 3:9] = 7.5                                          # exercises cv::min(., 6.0) :178
    dv = make_dvo(width=w, height=h, levels=1, culls=0)
    s = dv.frame_create(np.zeros((h, w), np.float32))
    dv.frame_set_depth(s, depth, sigma, None, full_res=False)
    dv.map_regularize(s)
    out = dv.frame_download(s, 0, "depth")[0]
    ref = orc.regularize(depth, sigma)
    assert out.tobytes() == ref.tobytes()
    assert dv.frame_download(s, 0, "sigma")[0].tobytes() == sigma.tobytes()  # sigma untouched (:174)
    dv.close()


def test_regularize_recull(orc):
    """Frame::updateDepth (frame.cpp:56-61): coarser depth levels re-culled from the regularised map."""
    w, h, L = 160, 120, 3
    rng = np.random.default_rng(9)
    depth = rng.uniform(0.3, 7, (h, w)).astype(np.float32)
    sigma = rng.uniform(0.01, 0.6, (h, w)).astype(np.float32)
    dv = make_dvo(width=w, height=h, levels=L, culls=0)
    s = dv.frame_create(np.zeros((h, w), np.float32))
    dv.frame_set_depth(s, depth, sigma, None, full_res=False)
    dv.map_regularize(s)
    ref = orc.regularize(depth, sigma)
    for l in range(L):
        t = L - 1 - l
        assert dv.frame_download(s, l, "depth")[0].tobytes() == orc.cull_image(ref, t).tobytes()
        assert dv.frame_download(s, l, "sigma")[0].tobytes() == orc.cull_image(sigma, t).tobytes()
    dv.close()


# ------------------------------------------------------------------------------------------------
# propagate — test/propagate.cpp:51-62 recipe (xi = (0.01,-0.01,0,0,0,0) repeated) and 6-DoF
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("xi", [(0.01, -0.01, 0, 0, 0, 0), (0.03, 0.02, -0.15, 0.02, -0.03, 0.05),
                                (0, 0, 0.4, 0, 0, 0), (0, 0, -2.5, 0, 0, 0), (0, 0, 0, 0, 0, 0)])
def test_propagate_bit_exact(orc, xi):
    from dvo_b200 import synth
    w, h = 160, 120
    _, d, _, K = synth.sequence(1, w, h, seed=3)
    depth = d[0].copy(); depth[10:20, 30:60] = 0.0; depth[50, 50] = 1e-7; depth[60, 60] = 0.005
    rng = np.random.default_rng(1)
    sigma = rng.uniform(0.05, 0.5, (h, w)).astype(np.float32)
    age = rng.integers(0, 5, (h, w)).astype(np.float32)
    xi = np.array(xi, np.float32)
    dv = make_dvo(width=w, height=h, levels=3, culls=0, K=K)
    z = np.zeros((h, w), np.float32)
    a = dv.frame_create(z); b = dv.frame_create(z)
    cur_d, cur_s, cur_a = depth, sigma, age
    src, dst = a, b
    Kt = dv.level_K(2)
    for rep in range(3):  # repeated propagate like test/propagate.cpp:57-62
        dv.frame_set_depth(src, cur_d, cur_s, cur_a, full_res=False)
        dv.map_propagate(src, dst, xi)
        od, os_, oa = orc.propagate(cur_d, cur_s, cur_a, xi, Kt)
        gd = dv.frame_download(dst, 2, "depth")[0]; gs = dv.frame_download(dst, 2, "sigma")[0]
        ga = dv.frame_download(dst, 2, "age")[0]
        assert ga.tobytes() == oa.tobytes(), "age / scatter winners differ (rep %d)" % rep
        assert gd.tobytes() == od.tobytes() and gs.tobytes() == os_.tobytes()
        for l in range(2):  # Frame::updateDepthSigmaAge re-cull (frame.cpp:47-54)
            assert dv.frame_download(dst, l, "depth")[0].tobytes() == orc.cull_image(od, 2 - l).tobytes()
        cur_d, cur_s, cur_a = od, os_, oa
    dv.close()


# ------------------------------------------------------------------------------------------------
# Mapper::update
# ------------------------------------------------------------------------------------------------
def _update_case(orc, seed, n_kf, prior, baseline, w=640, h=480, levels=3, culls=2, invalid=True, **params):
    """n_kf keyframes (ages 0..n_kf-1 scattered over the map) + one object frame."""
    from dvo_b200 import synth
    g, d, poses, K = synth.sequence(n_kf + 1, w, h, seed=seed, trans_step=baseline, rot_step_deg=0.2)
    if invalid:
        g = np.stack([punch_invalid(x, seed=seed + i) for i, x in enumerate(g)])
    dv = make_dvo(width=w, height=h, levels=levels, culls=culls, K=K, **params)
    top = levels - 1
    tw, th = dv.level_size(top)
    Kt = dv.level_K(top)
    rng = np.random.default_rng(seed)
    slots = [dv.frame_create(x) for x in g]
    # camera-from-world twists: xi_i = log(inv(T_wc_i)) (world frame = first camera)
    xis = [orc.se3_log(np.linalg.inv(p.astype(np.float64)).astype(np.float32)) for p in poses]
    for s, x in zip(slots, xis):
        dv.frame_set_xi(s, x, np.zeros(6, np.float32))
    ref_i = n_kf - 1                      # newest keyframe
    obj_i = n_kf
    rel = orc.se3_concatenate(xis[obj_i], -xis[ref_i])  # obj-from-ref
    dv.frame_set_xi(slots[obj_i], xis[obj_i], rel)
    gt = orc.cull_image(d[ref_i], culls)
    if prior == "gt":
        depth = (gt * rng.uniform(0.9, 1.1, gt.shape)).astype(np.float32)
        sigma = rng.uniform(0.05, 0.4, gt.shape).astype(np.float32)
    else:
        depth, sigma = synth.random_depth_prior(gt.shape, seed=seed)
    age = rng.integers(0, n_kf, gt.shape).astype(np.float32)
    dv.frame_set_depth(slots[ref_i], depth, sigma, age, full_res=False)
    hist = [slots[ref_i - a] for a in range(n_kf)]   # index = age
    of = [orc.frame_create(x, K, levels, culls) for x in g]
    born = [of[ref_i - a] for a in range(n_kf)]
    o = orc.mapper_update(depth, sigma, age, of[obj_i].image(top, "gray"),
                          [b.image(top, "gray") for b in born], [b.image(top, "gradx") for b in born],
                          [b.image(top, "grady") for b in born], np.stack([xis[ref_i - a] for a in range(n_kf)]),
                          xis[obj_i], rel, Kt, orc.params(**{k: v for k, v in params.items() if k.startswith(("map_", "accept_"))}))
    return dv, slots, ref_i, obj_i, hist, o, (depth, sigma, age)


@pytest.mark.parametrize("seed,n_kf,prior,baseline", [
    (11, 1, "gt", 0.02), (12, 3, "gt", 0.03), (13, 1, "random", 0.02), (14, 4, "random", 0.05), (15, 2, "gt", 0.08)])
def test_mapper_update_bit_exact(orc, seed, n_kf, prior, baseline):
    dv, slots, ref_i, obj_i, hist, o, _ = _update_case(orc, seed, n_kf, prior, baseline)
    od, os_, oa, omask, ovu, rc = o
    assert rc == 0
    vu, mask = dv.map_update(slots[obj_i], slots[ref_i], hist)
    np.testing.assert_array_equal(mask[0], omask)          # candidate / accepted / fused / reset masks
    assert vu[0] == ovu
    gd = dv.frame_download(slots[ref_i], 2, "depth")[0]; gs = dv.frame_download(slots[ref_i], 2, "sigma")[0]
    ga = dv.frame_download(slots[ref_i], 2, "age")[0]
    assert ga.tobytes() == oa.tobytes()
    assert_maps(gd, od, "depth"); assert_maps(gs, os_, "sigma")
    n_reset = int((omask & 8 > 0).sum())
    assert int(dv.rng_position()[0]) == n_reset          # mt19937 draws consumed (gaussian.cpp:22)
    print("candidates %d accepted %d fused %d reset %d" % tuple(int((omask & b > 0).sum()) for b in (1, 2, 4, 8)))
    for l in range(2):  # ref->updateDepthSigma re-cull (mapper.cpp:135)
        assert dv.frame_download(slots[ref_i], l, "depth")[0].tobytes() == orc.cull_image(gd, 2 - l).tobytes()
        assert dv.frame_download(slots[ref_i], l, "sigma")[0].tobytes() == orc.cull_image(gs, 2 - l).tobytes()
    dv.close()


def test_mapper_update_rng_is_cumulative(orc):
    """Two consecutive updates draw consecutive mt19937 outputs in row-major reject order (H6)."""
    dv, slots, ref_i, obj_i, hist, o, (depth, sigma, age) = _update_case(orc, 21, 1, "random", 0.1, accept_sigma_hi=5.0)
    vu, mask = dv.map_update(slots[obj_i], slots[ref_i], hist)
    n1 = int((mask[0] & 8 > 0).sum())
    assert n1 > 0, "case must produce resets"
    np.testing.assert_array_equal(mask[0], o[3])
    d1 = dv.frame_download(slots[ref_i], 2, "depth")[0]
    raw = orc.rng_u32(n1)
    expect = np.minimum(orc.dist_from_u32(raw), np.float32(4.0))
    np.testing.assert_array_equal(d1[mask[0] & 8 > 0], expect)     # row-major rank -> draw index
    # second update continues the stream
    vu2, mask2 = dv.map_update(slots[obj_i], slots[ref_i], hist)
    n2 = int((mask2[0] & 8 > 0).sum())
    assert int(dv.rng_position()[0]) == n1 + n2
    if n2:
        d2 = dv.frame_download(slots[ref_i], 2, "depth")[0]
        raw = orc.rng_u32(n1 + n2)[n1:]
        np.testing.assert_array_equal(d2[mask2[0] & 8 > 0], np.minimum(orc.dist_from_u32(raw), np.float32(4.0)))
    dv.close()


def test_mapper_update_history_overflow(orc):
    """age >= history size: FrameHistory::operator[] -> vector::at throws (frame.hpp:176)."""
    import dvo_b200
    dv, slots, ref_i, obj_i, hist, o, (depth, sigma, age) = _update_case(orc, 31, 1, "gt", 0.02)
    age2 = age.copy(); age2[60, 80] = 3.0
    dv.frame_set_depth(slots[ref_i], depth, sigma, age2, full_res=False)
    with pytest.raises(dvo_b200.DvoError):
        dv.map_update(slots[obj_i], slots[ref_i], hist)
    dv.close()
