"""SURVEY §8 f-4, headless outputs: Draw::visualize* (src/core/draw.cpp:7-100) as BGR bytes, Convert::inversePose
(src/core/convert.cpp:31-39), the trajectory CSV and the PNG writer.

  * CPU: the oracle's restatement of OpenCV's convertTo(CV_8UC1, a, b) and 8-bit COLOR_HSV2BGR against vectors OpenCV itself
    produced (tests/golden/draw_golden.npz, made by tests/golden/make_draw.py with cv2 4.13), exhaustively over the (H, V)
    pairs visualizeDepth can emit;
  * `gpu`: dvo_visualize through the C ABI against the oracle, byte for byte, on every level of a frame with invalid pixels.
"""
import os
import struct
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def G():
    return dict(np.load(os.path.join(HERE, "golden", "draw_golden.npz")))


def test_oracle_hsv2bgr_matches_opencv(orc, G):
    for S in (255, 128):
        ref = G["hsv2bgr_s%d" % S]
        for h in range(181):
            for v in range(0, 256, 1 if S == 255 else 5):
                assert orc.hsv2bgr(h, S, v) == tuple(int(x) for x in ref[h, v]), (h, S, v)


def test_oracle_convert_matches_opencv(orc, G):
    for name in ("gray", "depth_raw", "sigma", "age"):
        x = G["cvt_%s_x" % name][0]; y = G["cvt_%s_y" % name][0]; a, b = (float(t) for t in G["cvt_%s_ab" % name])
        got = np.array([orc.convert_8u(float(t), a, b) for t in x], np.uint8)
        assert np.array_equal(got, y), name


def test_oracle_visualize_rules(orc):
    inv = np.float32(-2.0)
    g = np.array([[0.0, 0.5, 1.0, inv, 1.5]], np.float32)
    out = orc.visualize("gray", g)[0]
    assert out[:3].tolist() == [[0, 0, 0], [128, 128, 128], [255, 255, 255]]     # cvRound(127.5) = 128 (half to even)
    assert out[3].tolist() == [255, 0, 0] and out[4].tolist() == [255, 255, 255]  # invalid -> blue; saturation
    d = np.array([[0.0, 5e-7, 0.70, 1.0, 3.5, 9.0]], np.float32)
    out = orc.visualize("depth", d)[0]
    assert out[0].tolist() == [0, 0, 0] and out[1].tolist() == [0, 0, 0]          # < EPSILON -> black
    assert out[2].tolist() == [0, 0, 255]                                         # H = 0: red
    assert out[3].tolist() == list(orc.hsv2bgr(21, 255, 255))                     # (1.0 - 0.7) * 70 = 21.0000..: (uchar) truncates
    assert out[4].tolist() == out[5].tolist() == list(orc.hsv2bgr(180, 255, 255))  # clamp at 180 (wraps to H = 0)
    s = np.array([[0.0, 0.1, 0.5, 0.9, 0.0, 0.2]], np.float32)
    out2 = orc.visualize("depth_sigma", d, s)[0]
    assert out2[2].tolist() == [0, 0, 5] and out2[5].tolist() == [0, 0, 155]      # V = (uchar)(-500 * min(sigma, 0.5) + 255)
    assert out2[3].tolist() == list(orc.hsv2bgr(21, 255, 5))                      # sigma clamped at 0.5 -> V = 5
    assert orc.visualize("depth_raw", d)[0][:, 0].tolist() == [50, 50, 70, 100, 250, 250]
    assert orc.visualize("sigma", s)[0][:, 0].tolist() == [255, 205, 5, 0, 255, 155]
    gr = np.array([[inv, 0.25, -0.5, 0.0]], np.float32)
    assert orc.visualize("gradient", gr)[0].tolist() == [[255, 0, 0], [0, 63, 0], [0, 0, 127], [0, 0, 0]]
    assert orc.visualize("age", np.array([[0, 3, 25.5, 26]], np.float32))[0][:, 0].tolist() == [0, 30, 255, 255]


def _png_decode(path):
    """Minimal PNG reader (8-bit RGB, filter 0) for the writer's round trip."""
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + data) & 0xffffffff
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", data[:10])
            assert (depth, ctype) == (8, 2)
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, 3)


def test_host_helpers_need_no_gpu(tmp_path):
    """inversePose, the trajectory CSV and the PNG writer are host code of the C-ABI library: no device, no context."""
    import dvo_b200
    lib = dvo_b200.load_library()
    api = dvo_b200.Dvo.__new__(dvo_b200.Dvo); api.lib = lib; api._ck = lambda rc: (_ for _ in ()).throw(RuntimeError(rc)) if rc else None
    T = np.arange(16, dtype=np.float32).reshape(4, 4) + 1
    inv = api.inverse_pose(T)
    assert np.array_equal(inv[:3, :3], T[:3, :3].T) and np.array_equal(inv[:3, 3], -T[:3, 3]) and not inv[3].any()  # convert.cpp:31-39
    p = str(tmp_path / "traj.csv")
    api.write_trajectory_csv(p, np.stack([T, inv]), first_frame=7)
    api.write_trajectory_csv(p, T[None], first_frame=9, append=True)
    rows = np.loadtxt(p, delimiter=",")
    assert rows[:, 0].tolist() == [7, 8, 9] and np.array_equal(rows[:, 1:].astype(np.float32), np.stack([T, inv, T]).reshape(3, 16))
    rng = np.random.default_rng(5)
    for shape in ((3, 5, 3), (120, 160, 3), (300, 400, 3)):  # the last spans several 65535-byte stored blocks
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        q = str(tmp_path / "x.png")
        api.write_png(q, img)
        assert np.array_equal(_png_decode(q), img[:, :, ::-1])


def test_oracle_inverse_pose(orc):
    T = np.arange(16, dtype=np.float32).reshape(4, 4) * 0.5
    inv = np.asarray(orc.inverse_pose(T)).reshape(4, 4)
    assert np.array_equal(inv[:3, :3], T[:3, :3].T) and np.array_equal(inv[:3, 3], -T[:3, 3]) and not inv[3].any()


@pytest.mark.gpu
def test_gpu_visualize_matches_oracle(orc, tmp_path):
    from conftest import make_dvo
    from dvo_b200 import synth
    rng = np.random.default_rng(11)
    W, H = 160, 120
    K = np.array([[140, 0, 79.5], [0, 140, 59.5], [0, 0, 1]], np.float32)
    gray = rng.uniform(0, 1, (H, W)).astype(np.float32)
    gray[rng.uniform(size=(H, W)) < 0.05] = -2.0                       # invalid pixels (undistortion border)
    gray[0, :4] = [0.5, 127.5 / 255, 1.0, 1.3]
    depth = np.maximum(rng.normal(1.5, 0.8, (H, W)), 0).astype(np.float32)
    depth[rng.uniform(size=(H, W)) < 0.05] = 0.0
    sigma = rng.uniform(0, 0.7, (H, W)).astype(np.float32)
    age = rng.integers(0, 40, (H, W)).astype(np.float32)
    dv = make_dvo(width=W, height=H, levels=3, culls=0, K=K)
    slot = dv.frame_create(gray)
    dv.frame_set_depth(slot, depth, sigma, age)
    seen = 0
    for lv in range(3):
        img = {k: dv.frame_download(slot, lv, k)[0] for k in ("gray", "depth", "sigma", "gradx", "grady")}
        pairs = [("gray", "gray", img["gray"], None), ("depth", "depth", img["depth"], None), ("depth_sigma", "depth_sigma", img["depth"], img["sigma"]),
                 ("depth_raw", "depth_raw", img["depth"], None), ("sigma", "sigma", img["sigma"], None), ("gradx", "gradient", img["gradx"], None),
                 ("grady", "gradient", img["grady"], None)]
        if lv == 2:
            pairs.append(("age", "age", dv.frame_download(slot, lv, "age")[0], None))
        for kind, okind, src, aux in pairs:
            got = dv.visualize(slot, lv, kind)[0]
            want = orc.visualize(okind, src, aux)
            assert got.shape == want.shape and np.array_equal(got, want), (lv, kind, int((got != want).any(-1).sum()))
            assert np.array_equal(dv.visualize_image(kind, src, aux), want), (lv, kind, "raw image form")
            seen += 1
    assert seen == 22
    vis = dv.visualize(slot, 2, "depth_sigma")[0]
    assert (vis.reshape(-1, 3).max(0) > 0).all() and (vis[depth < 1e-6] == 0).all()
    p = str(tmp_path / "depth.png")
    dv.write_png(p, vis)
    assert np.array_equal(_png_decode(p), vis[:, :, ::-1])
    with pytest.raises(Exception):
        dv.visualize(slot, 0, "age")                                   # age exists on the finest level only
    dv.close()
