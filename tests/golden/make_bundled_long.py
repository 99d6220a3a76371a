#!/usr/bin/env python
"""Generates tests/golden/bundled_long.npz: BASELINE.json configs[1] — main.cpp's loop (main.cpp:16-18,22-49) — on the two
bundled inputs the config names, run by THE REFERENCE'S OWN CODE (oracle/_ref: its unmodified translation units over the
OpenCV-API shim, strict IEEE build):

    python tests/golden/make_bundled_long.py      # needs /root/reference (data/ + src/); rewrites the .npz

  k2_*  `main ../data/kinectv2_00/info.txt` — all 56 lines.  Core::Loader keeps the FIRST token of every info.txt line
        (include/core/loader.hpp:38-45), i.e. the 1920x1080 4-channel colour PNG; cvtColor(BGR2GRAY) + convertTo(1/255)
        (loader.cpp:55-63) and the hard-coded 640x480 undistortion map (loader.cpp:15-30) then crop its top-left corner:
        cv::remap's destination has the MAP's size.  K = logicool_00/config.yaml's (main.cpp:22), Frame(gray, K, 3, 2).
  l0_*  default `main` on data/logicool0 — the first N_L0 = 150 of 501 frames (640x480 BGR).

Inputs are prepared with cv2 4.13 exactly as the reference's loader does; only the pixels Frame(., 3, 2) samples are stored
(160x120 8-bit codes + the constant INVALID border mask of the undistortion).  Outputs per frame: pose, relative pose,
keyframe decision, history size, valid_update, Gauss-Newton log; keyframe maps at checkpoints (every 25 frames and the last).
The frame-0 depth prior stands in for Scene's cv::randn draw."""
import os
import sys
import time

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("DVO_REFERENCE", "/root/reference")
for p in (ROOT, os.path.join(ROOT, "direct-visual-odometry_b200"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from dvo_b200 import synth  # noqa: E402
from make_bundled import f32, logs_to_arrays, read_list, up, yaml_all  # noqa: E402
from oracle.ref import Ref  # noqa: E402

N_L0 = 150      # frames of data/logicool0 (501 bundled)
CHECK_EVERY = 25


def loader_codes(path, mx, my):
    """Loader::getNormalizedUndistortedImages up to the 8-bit code: imread(UNCHANGED), cvtColor(BGR2GRAY), remap NEAREST with
    border INVALID (NEAREST moves whole pixels, so remapping the codes and converting afterwards is the same image), then the
    pixels cullImage(gray, 2) keeps.  Returns int16 codes, -1 = INVALID."""
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    code = cv2.COLOR_BGRA2GRAY if img.shape[2] == 4 else cv2.COLOR_BGR2GRAY   # cvtColor(COLOR_BGR2GRAY) accepts 3 or 4 channels
    g8 = cv2.cvtColor(img, code)
    und = cv2.remap(g8.astype(np.int16), mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=-1)
    return np.ascontiguousarray(und[::4, ::4])


def run(ref, paths, K, tag, out):
    Ku = np.array([[780, 0, 378], [0, 796, 220], [0, 0, 1]], np.float32)      # loader.cpp:17
    Du = np.array([-0.0462, 0.152, -0.00429, 0.0117, -0.0725], np.float32)    # loader.cpp:18
    mx, my = cv2.initUndistortRectifyMap(Ku, Du, None, Ku, (640, 480), cv2.CV_32FC1)
    codes = np.stack([loader_codes(p, mx, my) for p in paths]).astype(np.int16)
    invalid = codes[0] < 0
    assert all(np.array_equal(c < 0, invalid) for c in codes), "the INVALID border comes from the map alone"
    gray = np.where(codes < 0, np.float32(-2.0), (codes.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)).astype(np.float32)
    d0, s0 = synth.random_depth_prior((120, 160), seed=4321)
    ref.rng_reset()
    vo = ref.vo_create(K)
    n = len(paths)
    T, rel, kf, nkf, vu, metas, ck_i, ck_d, ck_s, ck_a = [], [], [], [], [], [], [], [], [], []
    t0 = time.time()
    for i in range(n):
        T.append(vo.odometrize(up(gray[i], 4), d0 if i == 0 else None, s0 if i == 0 else None))
        kf.append(vo.last_was_keyframe()); nkf.append(vo.history_size()); vu.append(vo.last_valid_update() if i and not kf[-1] else 0)
        logs = vo.last_log() if i > 0 else []
        metas.append(logs)
        x = np.zeros(6, np.float32)   # Tracker::track's fold (tracker.cpp:46-51) with the reference's own se3::concatenate
        for L in logs:
            nx = ref.se3_concatenate(x, L["update"])
            if not np.isnan(nx).any():
                x = nx
        rel.append(x)
        if i > 0 and vo.last_was_keyframe():
            assert np.array_equal(vo.ref_frame().xi()[1], x)
        if i % CHECK_EVERY == 0 or i == n - 1:
            fr = vo.ref_frame()
            a = fr.image(2, "age")
            assert a.max() < 256 and np.array_equal(a, np.rint(a))
            ck_i.append(i); ck_d.append(fr.image(2, "depth")); ck_s.append(fr.image(2, "sigma")); ck_a.append(a.astype(np.uint8))
    n_it = np.array([len(m) for m in metas], np.int32)
    meta, res, upd = logs_to_arrays([L for m in metas for L in m])
    c8 = np.where(codes < 0, 0, codes).astype(np.uint8)
    out.update({tag + "K": f32(K), tag + "codes": c8, tag + "invalid": invalid, tag + "init_depth": d0,
                tag + "T": np.stack(T).astype(np.float32), tag + "rel_xi": np.stack(rel).astype(np.float32),
                tag + "keyframe": np.array(kf, np.uint8), tag + "n_keyframes": np.array(nkf, np.int32),
                tag + "valid_update": np.array(vu, np.int32), tag + "n_iters": n_it, tag + "log_meta": meta,
                tag + "log_residual": res, tag + "log_update": upd, tag + "ck_frame": np.array(ck_i, np.int32),
                tag + "ck_depth": np.stack(ck_d), tag + "ck_sigma": np.stack(ck_s), tag + "ck_age": np.stack(ck_a)})
    vu = np.array(vu)
    print("%s %d frames in %.0f s: %d keyframes, %d GN iterations, update frames with valid_update > 0: %d (max %d, sum %d), "
          "max age %d" % (tag, n, time.time() - t0, nkf[-1], int(n_it.sum()), int((vu > 0).sum()), int(vu.max()), int(vu.sum()),
                          int(ck_a[-1].max())))


def main():
    ref = Ref()
    out = {}
    cfg = yaml_all(os.path.join(REF, "external/camera-calibration/data/logicool_00/config.yaml"))
    K = f32(cfg["intrinsic"])                                                  # main.cpp:22,33: loader.Rgb().K()
    k2 = [r[0] for r in read_list(os.path.join(REF, "data/kinectv2_00/info.txt"))]
    assert len(k2) == 56
    run(ref, k2, K, "k2_", out)
    l0 = [r[0] for r in read_list(os.path.join(REF, "data/logicool0/info.txt"))][:N_L0]
    run(ref, l0, K, "l0_", out)
    path = os.path.join(HERE, "bundled_long.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%.0f KB)" % (path, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
