"""Kinect ingest: KinectLoader::getMappedImages (src/core/loader.cpp:90-147) + Transform::mapDepthtoGray
(src/core/transform.cpp:53-78) + Frame(gray, depth, sigma, K, 4, 1) (test/track.cpp:24-25) on the device, from the raw 8-bit BGRA and
16-bit depth images.  Expected values: the OpenCV calls done by cv2 4.13, the per-pixel registration restated in numpy in the
reference's fp32 operation order (the same restatement that prepared tests/golden/bundled_golden.npz), the bilinear tap by the
oracle's Convert::getSubpixel.  Calibration = external/camera-calibration/data/kinectv2_00/config.yaml."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

K1 = np.array([[1.0581375e+03, 0, 9.4575586e+02], [0, 1.0610057e+03, 5.1298480e+02], [0, 0, 1]], np.float32)   # intrinsic1 (colour)
D1 = np.array([0.038666584, -0.03998715, -0.009501372, -0.001274522, -0.009535416], np.float32)
K2 = np.array([[363.38058, 0, 256.42273], [0, 364.06924, 202.84938], [0, 0, 1]], np.float32)                   # intrinsic2 (depth)
D2 = np.array([0.086782634, -0.30465686, -0.003397709, 0.001191156, 0.14221883], np.float32)
ROT = np.array([[0.9999723, -0.005255498, -0.005274575], [0.005228787, 0.9999735, -0.005065211],
                [0.005301055, 0.005037491, 0.99997324]], np.float32)
TRA = np.array([-0.05776357, -0.012686561, -0.003236694], np.float32)
CW, CH, DW, DH = 1920, 1080, 512, 424


def inv_T():
    T = np.eye(4, dtype=np.float32)          # ExtrinsicParams::invT (params_struct.hpp:38-46): [R^T | -t]
    T[:3, :3] = ROT.T
    T[:3, 3] = -TRA
    return T


def synthetic_pair(seed):
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:CH, 0:CW].astype(np.float32)
    base = 127 + 80 * np.sin(0.013 * xs + 0.021 * ys) + 40 * np.sin(0.11 * xs - 0.07 * ys)
    bgra = np.stack([np.clip(base + rng.normal(0, 6, base.shape), 0, 255) for _ in range(3)] + [np.full(base.shape, 255.0)], axis=-1).astype(np.uint8)
    yd, xd = np.mgrid[0:DH, 0:DW].astype(np.float32)
    depth_m = 1.2 + 0.6 * np.sin(0.01 * xd) * np.cos(0.013 * yd) + 0.002 * rng.normal(0, 1, (DH, DW))
    d16 = np.clip(depth_m * 5000.0, 0, 65535).astype(np.uint16)
    d16[rng.random((DH, DW)) < 0.05] = 0           # missing measurements
    d16[100:140, 200:260] = 0
    d16[:, :12] = 600                              # 0.12 m: projects far outside the colour image
    return bgra, d16


def expected_frame(orc, bgra, d16, maps, culls):
    """(gray, depth, sigma) at the finest level of Frame(., levels, culls)."""
    (cmx, cmy), (dmx, dmy) = maps
    gray8 = cv2.cvtColor(bgra, cv2.COLOR_BGR2GRAY)                                                    # loader.cpp:143
    g = (gray8.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)                        # :144
    d = (d16.astype(np.float32) * np.float32(1.0 / 5000.0)).astype(np.float32)                         # :145
    g_u = cv2.remap(g, cmx, cmy, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=-2.0)  # :110-111
    d_u = cv2.remap(d, dmx, dmy, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=0)     # :112-113
    st = 1 << culls
    dd = np.ascontiguousarray(d_u[::st, ::st])
    h, w = dd.shape
    ys, xs = np.mgrid[0:h, 0:w]
    X = (xs * st).astype(np.float32); Y = (ys * st).astype(np.float32)
    valid = ~(np.abs(dd) < np.float32(1e-6))
    fx, fy, cx, cy = K2[0, 0], K2[1, 1], K2[0, 2], K2[1, 2]
    T = inv_T()
    with np.errstate(all="ignore"):
        Xc = (dd * (X - cx)) / fx; Yc = (dd * (Y - cy)) / fy; Zc = dd
        tx = ((T[0, 0] * Xc + T[0, 1] * Yc) + T[0, 2] * Zc) + T[0, 3]
        ty = ((T[1, 0] * Xc + T[1, 1] * Yc) + T[1, 2] * Zc) + T[1, 3]
        tz = ((T[2, 0] * Xc + T[2, 1] * Yc) + T[2, 2] * Zc) + T[2, 3]
        u = (tx * K1[0, 0]) / tz + K1[0, 2]
        v = (ty * K1[1, 1]) / tz + K1[1, 2]
    gray = np.full((h, w), -2.0, np.float32)
    sigma = np.ones((h, w), np.float32)
    for yy, xx in zip(*np.nonzero(valid)):
        gray[yy, xx] = orc.get_subpixel(g_u, float(u[yy, xx]), float(v[yy, xx]), dense=False)
    sigma[valid] = np.float32(0.1)
    return gray, dd, sigma


@pytest.mark.gpu
@pytest.mark.parametrize("levels,culls", [(4, 1), (3, 2)])
def test_kinect_frame_matches_loader_restatement(orc, levels, culls):
    from conftest import make_dvo
    cmaps = cv2.initUndistortRectifyMap(K1, D1, None, K1, (CW, CH), cv2.CV_32FC1)   # KinectLoader::createUndistortMap, loader.cpp:77-88
    dmaps = cv2.initUndistortRectifyMap(K2, D2, None, K2, (DW, DH), cv2.CV_32FC1)
    S = 2
    pairs = [synthetic_pair(s) for s in range(S)]
    dv = make_dvo(width=DW, height=DH, levels=levels, culls=culls, K=K2, n_streams=S)
    dv.kinect_configure((CW, CH), 4, K1, inv_T(), cmaps, dmaps)
    slot = dv.frame_create_kinect(np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs]))
    top = levels - 1
    got = {k: dv.frame_download(slot, top, k) for k in ("gray", "depth", "sigma", "age")}
    for s in range(S):
        gray, depth, sigma = expected_frame(orc, pairs[s][0], pairs[s][1], (cmaps, dmaps), culls)
        assert (gray > -2.0).sum() > 0.5 * gray.size and (gray <= -2.0).sum() > 100      # both branches exercised
        assert got["depth"][s].tobytes() == depth.tobytes()
        assert got["sigma"][s].tobytes() == sigma.tobytes()
        assert np.array_equal(got["gray"][s] <= -2.0, gray <= -2.0)                       # validity mask
        assert got["gray"][s].tobytes() == gray.tobytes()
        assert not got["age"][s].any()
    # the registered frame is a normal frame: gradients come out as Convert::gradiate of the registered gray
    gx = dv.frame_download(slot, top, "gradx")[0]
    g0, _, _ = expected_frame(orc, pairs[0][0], pairs[0][1], (cmaps, dmaps), culls)
    assert gx.tobytes() == orc.gradiate(g0, True).tobytes()
    dv.close()
