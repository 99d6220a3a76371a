"""8-bit ingest: Loader::getNormalizedImages' convertTo(CV_32FC1, 1.0/255.0) (src/core/loader.cpp:59-61) on the
device, fused with Convert::cullImage.  The CPU test pins the arithmetic against cv2 (OpenCV multiplies 8-bit input in fp32);
the GPU tests require the 8-bit entry points to be bit-identical to feeding the converted float frames."""
import ctypes as C

import numpy as np
import pytest


def to_float(codes, invalid_code=-1):
    g = (codes.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)
    if invalid_code >= 0:
        g[codes == invalid_code] = -2.0
    return g


def test_convert_to_matches_opencv():
    cv2 = pytest.importorskip("cv2")
    u = np.arange(256, dtype=np.uint8).reshape(1, 256)
    # NORM_MINMAX with alpha = 1, beta = 0 on data spanning 0..255 is src.convertTo(dst, CV_32F, 1.0 / 255.0, 0)
    ref = cv2.normalize(u, None, alpha=1.0, beta=0.0, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    assert np.array_equal(ref, to_float(u))
    assert (ref != (u.astype(np.float64) * (1.0 / 255.0)).astype(np.float32)).sum() > 0  # the fp64 product is NOT what OpenCV does


@pytest.mark.gpu
@pytest.mark.parametrize("levels,culls,w,h", [(3, 2, 640, 480), (4, 1, 512, 424), (2, 0, 101, 67)])
def test_frame_from_gray8_bit_exact(levels, culls, w, h):
    from conftest import make_dvo
    rng = np.random.default_rng(3)
    codes = rng.integers(0, 256, (h, w), dtype=np.uint8)
    codes[:, :5] = 255  # "border"
    for inv in (-1, 255, 0):
        dv = make_dvo(width=w, height=h, levels=levels, culls=culls)
        a = dv.frame_create_gray8(codes, inv)
        b = dv.frame_create(to_float(codes, inv))
        for lv in range(levels):
            for kind in ("gray", "gradx", "grady"):
                assert dv.frame_download(a, lv, kind)[0].tobytes() == dv.frame_download(b, lv, kind)[0].tobytes(), (inv, lv, kind)
        dv.close()


@pytest.mark.gpu
def test_vo_from_gray8_bit_exact(seq160):
    """dvo_vo_push_gray8 and the pipelined dvo_vo_submit_gray8 == dvo_vo_push on the converted float frames."""
    from conftest import make_dvo
    from dvo_b200 import synth
    grays, _, _, K = seq160
    codes = np.clip(np.rint(grays * 255.0), 0, 255).astype(np.uint8)
    n = len(codes)
    d0, s0 = synth.random_depth_prior((120, 160), seed=11)
    runs = []
    for mode in ("float", "gray8", "submit8"):
        dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2, max_keyframes=n + 2)
        lib = dv.lib
        Ts = []
        if mode == "submit8":
            Ts.append(dv.vo_push_gray8(codes[0], -1, d0, s0).copy())
            hp = [lib.dvo_host_alloc(codes[0].nbytes) for _ in range(2)]
            tp = [lib.dvo_host_alloc(64) for _ in range(2)]
            hb = [np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=codes[0].shape) for p in hp]
            tb = [np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(4, 4)) for p in tp]
            pend = []
            for i in range(1, n):
                k = i & 1
                hb[k][...] = codes[i]
                dv._ck(lib.dvo_vo_submit_gray8(dv.ctx, hp[k], -1, tp[k]))
                pend.append(k)
                if len(pend) == 2:
                    dv._ck(lib.dvo_vo_collect(dv.ctx)); Ts.append(tb[pend.pop(0)].copy())
            while pend:
                dv._ck(lib.dvo_vo_collect(dv.ctx)); Ts.append(tb[pend.pop(0)].copy())
            for p in hp + tp:
                lib.dvo_host_free(p)
        else:
            for i in range(n):
                a = (d0, s0) if i == 0 else (None, None)
                T = dv.vo_push(to_float(codes[i]), *a) if mode == "float" else dv.vo_push_gray8(codes[i], -1, *a)
                Ts.append(T[0].copy())
        assert dv.vo_error() == 0
        slot = dv.vo_ref_slot(0)
        runs.append((np.stack([np.asarray(t).reshape(4, 4) for t in Ts]), dv.frame_download(slot, 2, "depth")[0].copy(),
                     dv.frame_download(slot, 2, "sigma")[0].copy()))
        dv.close()
    for r in runs[1:]:
        assert r[0].tobytes() == runs[0][0].tobytes()
        assert r[1].tobytes() == runs[0][1].tobytes() and r[2].tobytes() == runs[0][2].tobytes()
