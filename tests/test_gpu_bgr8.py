"""BGR ingest: the whole of Loader::getNormalizedUndistortedImages (src/core/loader.cpp:32-42,55-63) on the device —
cvtColor(BGR2GRAY), convertTo(CV_32FC1, 1/255), remap(INTER_NEAREST, BORDER_CONSTANT, INVALID) through the maps of
cv::initUndistortRectifyMap — fused with Convert::cullImage.  cv2 4.13 is the reference implementation of the three OpenCV calls;
the CPU tests pin their arithmetic, the GPU tests require the device output to equal the cv2 pipeline bit for bit."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

K_UND = np.array([[780, 0, 378], [0, 796, 220], [0, 0, 1]], np.float32)         # loader.cpp:17
D_UND = np.array([-0.0462, 0.152, -0.00429, 0.0117, -0.0725], np.float32)       # loader.cpp:18


def cv2_pipeline(bgr, mx, my, culls):
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)                                                   # loader.cpp:60
    f = (gray.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)                     # :61 (fp32 multiply, test_gpu_gray8)
    if mx is not None:
        f = cv2.remap(f, mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=-2.0)  # :39-40
    return np.ascontiguousarray(f[::1 << culls, ::1 << culls])                                     # Convert::cullImage


def test_bgr2gray_fixed_point_matches_opencv():
    rng = np.random.default_rng(0)
    for ch in (3, 4):
        img = rng.integers(0, 256, (257, 131, ch), dtype=np.uint8)
        b, g, r = (img[..., i].astype(np.int64) for i in range(3))
        mine = ((b * 3735 + g * 19235 + r * 9798 + (1 << 14)) >> 15).astype(np.uint8)
        assert np.array_equal(mine, cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))


def test_remap_nearest_is_round_half_even_with_constant_border():
    rng = np.random.default_rng(1)
    mx, my = cv2.initUndistortRectifyMap(K_UND, D_UND, None, K_UND, (640, 480), cv2.CV_32FC1)
    mx = mx.copy(); my = my.copy()
    mx[5, :40] = np.arange(40) + 0.5          # exact ties: cvRound rounds half to even
    my[7, :40] = np.arange(40) + 0.5
    mx[9, 0] = 1e9; my[9, 1] = -1e9; mx[9, 2] = np.nan
    src = rng.random((480, 640)).astype(np.float32)
    ref = cv2.remap(src, mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=-2.0)
    with np.errstate(invalid="ignore"):
        sx = np.clip(np.rint(np.nan_to_num(mx, nan=-1e9)), -32768, 32767).astype(np.int64)
        sy = np.clip(np.rint(np.nan_to_num(my, nan=-1e9)), -32768, 32767).astype(np.int64)
    ok = (sx >= 0) & (sx < 640) & (sy >= 0) & (sy < 480)
    mine = np.full((480, 640), -2.0, np.float32)
    mine[ok] = src[sy[ok], sx[ok]]
    assert np.array_equal(mine, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("ch,culls,levels", [(3, 2, 3), (4, 2, 3), (3, 0, 2), (3, 1, 4)])
def test_frame_from_bgr8_equals_cv2_pipeline(ch, culls, levels):
    from conftest import make_dvo
    rng = np.random.default_rng(7)
    S = 2
    bgr = rng.integers(0, 256, (S, 480, 640, ch), dtype=np.uint8)
    mx, my = cv2.initUndistortRectifyMap(K_UND, D_UND, None, K_UND, (640, 480), cv2.CV_32FC1)
    mx = mx.copy(); my = my.copy()
    mx[8, 16:48] = np.arange(32) + 0.5; mx[12, 0] = 1e9; my[12, 4] = np.nan     # ties, far outside, NaN (rows / cols that survive culling)
    dv = make_dvo(width=640, height=480, levels=levels, culls=culls, n_streams=S)
    top = levels - 1
    for maps in ((mx, my), (None, None)):
        dv.set_undistort_map(*maps)
        slot = dv.frame_create_bgr8(bgr)
        got = dv.frame_download(slot, top, "gray")
        for s in range(S):
            want = cv2_pipeline(bgr[s], maps[0], maps[1], culls)
            assert got[s].tobytes() == want.tobytes(), (ch, culls, s, maps[0] is None)
        dv.frame_release(slot)
    dv.close()


@pytest.mark.gpu
def test_vo_from_bgr8_equals_float_frames(seq160):
    """dvo_vo_push_bgr8 through the undistortion map == dvo_vo_push on the frames the reference's loader would have produced."""
    from conftest import make_dvo
    from dvo_b200 import synth
    grays, _, _, K = seq160
    n = len(grays)
    codes = np.clip(np.rint(grays * 255.0), 0, 255).astype(np.uint8)
    bgr = np.stack([codes, codes, codes], axis=-1)                      # B = G = R -> gray code = the same value (weights sum to 2^15)
    assert np.array_equal(cv2.cvtColor(bgr[0], cv2.COLOR_BGR2GRAY), codes[0])
    mx, my = cv2.initUndistortRectifyMap(K_UND, D_UND, None, K_UND, (640, 480), cv2.CV_32FC1)
    d0, s0 = synth.random_depth_prior((120, 160), seed=11)
    runs = []
    for mode in ("float", "bgr8"):
        dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2, max_keyframes=n + 2)
        if mode == "bgr8":
            dv.set_undistort_map(mx, my)
        Ts = []
        for i in range(n):
            a = (d0, s0) if i == 0 else (None, None)
            if mode == "float":
                f = (codes[i].astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)
                f = cv2.remap(f, mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=-2.0)
                Ts.append(dv.vo_push(f, *a)[0].copy())
            else:
                Ts.append(dv.vo_push_bgr8(bgr[i], *a)[0].copy())
        assert dv.vo_error() == 0
        slot = dv.vo_ref_slot(0)
        runs.append((np.stack(Ts), dv.frame_download(slot, 2, "depth")[0].copy(), dv.frame_download(slot, 2, "sigma")[0].copy()))
        dv.close()
    assert runs[1][0].tobytes() == runs[0][0].tobytes()
    assert runs[1][1].tobytes() == runs[0][1].tobytes() and runs[1][2].tobytes() == runs[0][2].tobytes()
