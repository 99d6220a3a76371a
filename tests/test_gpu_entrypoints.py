"""GPU parity of the granular entry points added for the C++ drop-in layer:
 * dvo_map_update_points  = Map::Implement::update (src/map/implement.cpp:182-214) per point,
   the call test/update.cpp:86-95 makes for every pixel
 * dvo_se3_{exp,log,concatenate} = math::se3 (src/math/se3.cpp:76-131), test/se3.cpp:18-20 recipe
 * dvo_vo_submit / dvo_vo_collect = pipelined dvo_vo_push (same results)
Integer/mask results bit-exact; float results bit-exact as well (same operation order)."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_dvo

pytestmark = pytest.mark.gpu


def test_update_points_matches_oracle(orc, seq160):
    """test/update.cpp recipe: xi = (-0.1, 0, 0, 0, 0, 0)-style baseline, prior randn(1.7, 0.5) clipped, sigma 0.5."""
    g, d, xis, K = seq160
    L, culls = 1, 2
    dv = make_dvo(width=640, height=480, levels=L, culls=culls, K=K)
    obj = dv.frame_create(g[3]); born = dv.frame_create(g[0])
    fo = orc.frame_create(g[3], K, L, culls); fb = orc.frame_create(g[0], K, L, culls)
    og, bg = fo.image(0, "gray"), fb.image(0, "gray")
    gx, gy = fb.image(0, "gradx"), fb.image(0, "grady")
    Kl = fb.K(0)
    h, w = og.shape
    rng = np.random.default_rng(7)
    n = 600
    px = rng.integers(0, w, n).astype(np.int32); py = rng.integers(0, h, n).astype(np.int32)
    depth = np.clip(rng.normal(1.7, 0.5, n), 1.0, 4.0).astype(np.float32)
    sigma = np.full(n, 0.5, np.float32); sigma[::7] = 0.05
    rel = np.array([-0.03, 0.004, 0.002, 0.002, -0.003, 0.001], np.float32)
    od, os_ = dv.map_update_points(obj, born, rel, px, py, depth, sigma)
    n_ok = 0
    for i in range(n):
        rd, rs = orc.map_update_pixel(og, bg, gx, gy, rel, Kl, int(px[i]), int(py[i]), float(depth[i]), float(sigma[i]))
        assert np.float32(rd).tobytes() == od[0, i].tobytes(), (i, rd, od[0, i])
        assert np.float32(rs).tobytes() == os_[0, i].tobytes(), (i, rs, os_[0, i])
        n_ok += rd > 0
    assert n_ok > n // 10  # the recipe does produce matches
    dv.close()


SE3_TOL = 1e-5  # north_star: pose updates within 1e-5 in se3 norm (device libm vs glibc differ by an ulp in acos/sin)


def test_se3_ops_match_oracle(orc):
    rng = np.random.default_rng(3)
    xi = np.concatenate([rng.uniform(1, 3, (8, 6)), rng.normal(0, 0.05, (24, 6)), np.zeros((1, 6))]).astype(np.float32)  # test/se3.cpp:18
    dv = make_dvo(width=64, height=48, levels=1, culls=0)
    T = dv.se3_exp(xi)
    for i in range(len(xi)):
        assert np.abs(T[i] - orc.se3_exp(xi[i])).max() <= SE3_TOL, i
    back = dv.se3_log(T)
    for i in range(len(xi)):
        assert np.linalg.norm(back[i] - orc.se3_log(T[i]).reshape(6)) <= SE3_TOL, i
    assert np.abs(back[8:] - xi[8:]).max() < 1e-4  # exp/log round trip for small twists (test/se3.cpp:19-20)
    b = rng.normal(0, 0.1, xi.shape).astype(np.float32)
    cc = dv.se3_concatenate(xi, b)
    for i in range(len(xi)):
        # large rotations (|omega| ~ 3.5 rad): theta = acos(.) is ill-conditioned near pi -> compare through exp
        ref = orc.se3_concatenate(xi[i], b[i]).reshape(6)
        if i >= 8:
            assert np.linalg.norm(cc[i] - ref) <= SE3_TOL, i
        else:
            assert np.abs(orc.se3_exp(cc[i]) - orc.se3_exp(ref)).max() <= 1e-4, i
    dv.close()


def test_submit_collect_equals_push(seq160):
    """The pipelined submission produces the poses and maps of the synchronous dvo_vo_push."""
    g, d, xis, K = seq160
    S = 2
    frames = [np.stack([g[i], g[i][:, ::-1].copy()]) for i in range(len(g))]
    rng = np.random.default_rng(11)
    d0 = np.maximum(rng.normal(1.5, 0.5, (S, 120, 160)), 0.5).astype(np.float32)
    s0 = np.full_like(d0, 0.5)

    def run(pipelined):
        dv = make_dvo(n_streams=S, width=640, height=480, levels=3, culls=2, K=K)
        lib = dv.lib
        poses = []
        dv.vo_push(frames[0], d0, s0)
        if not pipelined:
            for f in frames[1:]:
                poses.append(dv.vo_push(f).copy())
        else:
            nb = frames[0].nbytes
            hbuf = [lib.dvo_host_alloc(nb) for _ in frames[1:]]
            tbuf = [lib.dvo_host_alloc(S * 64) for _ in frames[1:]]
            for hp, f in zip(hbuf, frames[1:]):
                C.memmove(hp, f.ctypes.data, nb)
            inflight = 0
            for k in range(len(hbuf)):
                dv._ck(lib.dvo_vo_submit(dv.ctx, hbuf[k], tbuf[k])); inflight += 1
                if inflight == 2:
                    dv._ck(lib.dvo_vo_collect(dv.ctx)); inflight -= 1
            while inflight:
                dv._ck(lib.dvo_vo_collect(dv.ctx)); inflight -= 1
            assert lib.dvo_vo_collect(dv.ctx) != 0  # nothing in flight -> DVO_ERR_STATE
            for tp in tbuf:
                poses.append(np.ctypeslib.as_array(C.cast(tp, C.POINTER(C.c_float)), shape=(S, 4, 4)).copy())
            for p in hbuf + tbuf:
                lib.dvo_host_free(p)
        slot = dv.vo_ref_slot(0)
        maps = [dv.frame_download(slot, 2, k).copy() for k in ("depth", "sigma", "age")]
        dv.close()
        return poses, maps

    pa, ma = run(False)
    pb, mb = run(True)
    assert len(pa) == len(pb)
    for a, b in zip(pa, pb):
        assert a.tobytes() == b.tobytes()
    for a, b in zip(ma, mb):
        assert a.tobytes() == b.tobytes()


def test_error_behaviour_of_the_ingest_entry_points():
    """Misuse returns a negative dvo_status with a message (the reference would abort() or read out of bounds): nothing is computed."""
    import ctypes as C
    import dvo_b200
    dv = make_dvo(width=64, height=48, levels=2, culls=1)
    lib = dv.lib
    slot = C.c_int(-1)
    z8 = np.zeros((1, 48, 64, 4), np.uint8)
    # Kinect frame before dvo_kinect_configure: call-sequence error
    rc = lib.dvo_frame_create_kinect(dv.ctx, C.byref(slot), z8.ctypes.data_as(C.c_void_p), z8.ctypes.data_as(C.c_void_p))
    assert rc == -6 and b"dvo_kinect_configure" in lib.dvo_last_error(dv.ctx)
    # channels other than 3 / 4, invalid_code outside -1..255, one map without the other
    assert lib.dvo_frame_create_bgr8(dv.ctx, C.byref(slot), z8.ctypes.data_as(C.c_void_p), 2) < 0
    assert lib.dvo_frame_create_gray8(dv.ctx, C.byref(slot), z8.ctypes.data_as(C.c_void_p), 256) < 0
    m = np.zeros((48, 64), np.float32)
    assert lib.dvo_set_undistort_map(dv.ctx, m.ctypes.data_as(C.POINTER(C.c_float)), None) < 0
    # three frames in flight
    with pytest.raises(dvo_b200.DvoError):
        hp = lib.dvo_host_alloc(64 * 48 * 4)
        try:
            dv._ck(lib.dvo_vo_submit(dv.ctx, hp, None)); dv._ck(lib.dvo_vo_submit(dv.ctx, hp, None)); dv._ck(lib.dvo_vo_submit(dv.ctx, hp, None))
        finally:
            while lib.dvo_vo_collect(dv.ctx) == 0:
                pass
            lib.dvo_host_free(hp)
    assert slot.value == -1   # no slot was ever handed out
    dv.close()
