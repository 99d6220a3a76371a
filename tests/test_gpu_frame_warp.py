"""GPU parity: Frame pyramid/gradients (frame.hpp:91-117, convert.cpp:7-75) and
Transform::warpImage (transform.cpp:35-51) against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

from conftest import make_dvo
from helpers import punch_invalid

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("levels,culls,w,h", [(3, 2, 640, 480), (4, 1, 512, 424), (1, 0, 96, 64), (2, 0, 101, 67)])
def test_pyramid_and_gradients_bit_exact(orc, levels, culls, w, h):
    from dvo_b200 import synth
    g, d, _, K = synth.sequence(1, w, h, seed=7)
    gray = punch_invalid(g[0], seed=3)
    depth = d[0].copy(); depth[::7, ::5] = 0.0
    sigma = np.where(depth > 0, 0.1, 1.0).astype(np.float32)
    dv = make_dvo(width=w, height=h, levels=levels, culls=culls, K=K)
    of = orc.frame_create(gray, K, levels, culls, depth, sigma)
    slot = dv.frame_create(gray)
    dv.frame_set_depth(slot, depth, sigma, None, full_res=True)
    for l in range(levels):
        assert dv.level_size(l) == of.level_size(l)
        np.testing.assert_array_equal(dv.level_K(l), of.K(l))
        for kind in ("gray", "gradx", "grady", "depth", "sigma"):
            a = dv.frame_download(slot, l, kind)[0]
            b = of.image(l, kind)
            assert a.tobytes() == b.tobytes(), (l, kind)
    # top-level sized update (Frame::updateDepthSigmaAge) + re-cull
    tw, th = dv.level_size(levels - 1)
    rng = np.random.default_rng(5)
    d2 = rng.uniform(0.3, 4, (th, tw)).astype(np.float32); s2 = rng.uniform(0.01, 1, (th, tw)).astype(np.float32)
    a2 = rng.integers(0, 4, (th, tw)).astype(np.float32)
    dv.frame_set_depth(slot, d2, s2, a2, full_res=False)
    of.set_depth_sigma_age(d2, s2, a2)
    for l in range(levels):
        for kind in ("depth", "sigma"):
            assert dv.frame_download(slot, l, kind)[0].tobytes() == of.image(l, kind).tobytes()
    assert dv.frame_download(slot, levels - 1, "age")[0].tobytes() == of.image(levels - 1, "age").tobytes()
    dv.close()


@pytest.mark.parametrize("xi", [
    (0, 0, 0, 0, 0, 0), (0.02, -0.01, 0.03, 0, 0, 0), (0.01, 0.02, -0.015, 0.01, -0.02, 0.015),
    (-0.1, 0, 0, 0, 0, 0), (0.3, -0.2, 0.1, 0.2, 0.1, -0.3)])
def test_warp_image_matches_oracle(orc, xi):
    from dvo_b200 import synth
    w, h, levels, culls = 512, 424, 4, 1
    g, d, _, K = synth.sequence(1, w, h, seed=11)
    gray = punch_invalid(g[0], seed=1)
    depth = d[0].copy(); depth[100:140, 200:260] = 0.0
    sigma = np.full_like(depth, 0.1)
    dv = make_dvo(width=w, height=h, levels=levels, culls=culls, K=K)
    of = orc.frame_create(gray, K, levels, culls, depth, sigma)
    slot = dv.frame_create(gray)
    dv.frame_set_depth(slot, depth, sigma, None, full_res=True)
    xi = np.array(xi, np.float32)
    for l in range(levels):
        a = dv.warp_image(slot, l, xi)[0]
        b = orc.warp_image(xi, of.image(l, "gray"), of.image(l, "depth"), of.K(l))
        # validity mask bit-exact; values bit-exact as well (same fp32 operation order, no FMA)
        np.testing.assert_array_equal(a <= -2.0, b <= -2.0)
        assert a.tobytes() == b.tobytes(), "level %d: %d pixels differ, max %g" % (l, int((a != b).sum()), float(np.abs(a - b).max()))
    dv.close()


def test_streams_are_independent(orc):
    """n_streams=3 with different inputs == three single-stream contexts."""
    from dvo_b200 import synth
    w, h = 160, 120
    gs = []; ds = []
    for s in range(3):
        g, d, _, K = synth.sequence(1, w, h, seed=20 + s)
        gs.append(punch_invalid(g[0], seed=s)); ds.append(d[0])
    gs = np.stack(gs); ds = np.stack(ds); sg = np.full_like(ds, 0.2)
    dv3 = make_dvo(n_streams=3, width=w, height=h, levels=3, culls=0, K=K)
    slot3 = dv3.frame_create(gs); dv3.frame_set_depth(slot3, ds, sg, None, full_res=True)
    xi = np.array([[0.01, 0, 0.02, 0.01, 0, 0], [0, 0.02, 0, 0, 0.02, 0], [-0.02, 0.01, 0.01, 0, 0, 0.03]], np.float32)
    out3 = dv3.warp_image(slot3, 2, xi)
    for s in range(3):
        dv1 = make_dvo(n_streams=1, width=w, height=h, levels=3, culls=0, K=K)
        s1 = dv1.frame_create(gs[s]); dv1.frame_set_depth(s1, ds[s], sg[s], None, full_res=True)
        assert dv1.warp_image(s1, 2, xi[s])[0].tobytes() == out3[s].tobytes()
        dv1.close()
    dv3.close()


def test_mask_compaction_bit_exact():
    dv = make_dvo(width=64, height=48, levels=1, culls=0)
    rng = np.random.default_rng(0)
    for n, p in [(1, 1.0), (255, 0.5), (256, 0.0), (257, 0.3), (19200, 0.37), (54272, 0.9), (1 << 20, 0.01)]:
        m = (rng.random(n) < p).astype(np.uint8)
        idx = dv.mask_compact(m)
        np.testing.assert_array_equal(idx, np.flatnonzero(m).astype(np.int32))
    dv.close()
