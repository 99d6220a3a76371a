#!/usr/bin/env python
"""Generates tests/golden/bundled_golden.npz: BASELINE.json configs[0] and configs[1] on the reference's
BUNDLED data, with outputs produced by THE REFERENCE'S OWN CODE (oracle/_ref: its unmodified translation
units over the OpenCV-API shim).

    python tests/golden/make_bundled.py       # needs /root/reference (data/ + src/); rewrites the .npz

Inputs are prepared exactly as the reference's loaders would (OpenCV calls done with cv2 4.13, the
per-pixel registration restated below):
  C1  test/track.cpp:12-25 — data/KINECT_50MM entries 9 and 10 through KinectLoader::getMappedImages
      (src/core/loader.cpp:90-147: cvtColor BGR2GRAY, /255, depth/5000, undistort both with INTER_NEAREST,
      then Transform::mapDepthtoGray, src/core/transform.cpp:53-78) -> 512x424 gray / depth / sigma;
      Frame(., 4, 1); Tracker::track.
  C2  main.cpp:22-49 — data/logicool0 through Loader::getNormalizedUndistortedImages
      (src/core/loader.cpp:15-42: hard-coded K, D; remap INTER_NEAREST, border -2.0) -> 640x480 gray;
      VisualOdometry(K).odometrize per frame, K from external/camera-calibration/data/logicool_00/config.yaml.
Only the pixels Frame(.) samples are stored (cullImage keeps the top-left pixel of every 2^culls block,
convert.cpp:7-20): C1 at 256x212 (float32, registered gray is bilinear), C2 at 160x120 as int16 (u8 code or -1
for the INVALID border, gray = (float)code * (float)(1.0/255.0) exactly as convertTo(CV_32FC1, 1.0/255.0))."""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("DVO_REFERENCE", "/root/reference")
for p in (ROOT, os.path.join(ROOT, "direct-visual-odometry_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from dvo_b200 import synth  # noqa: E402
from oracle.ref import Ref  # noqa: E402

N_C2 = 14  # frames of data/logicool0 (501 bundled)


def yaml_mats(path):
    fs = cv2.FileStorage(path, cv2.FILE_STORAGE_READ)
    out = {}
    for k in ("intrinsic", "distortion", "rgb_intrinsic", "rgb_distortion", "depth_intrinsic", "depth_distortion", "rotation", "translation",
              "extrinsic"):
        n = fs.getNode(k)
        if not n.empty() and n.isMap():
            out[k] = n.mat()
    fs.release()
    return out


def read_list(info):
    d = os.path.dirname(info)
    rows = [ln.split() for ln in open(info).read().splitlines() if ln.strip()]
    return [[os.path.join(d, t) for t in r] for r in rows]


def normalized_gray(path):
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)          # loader.cpp:60 / :143 (4-channel input uses B, G, R)
    # convertTo(CV_32FC1, 1.0/255.0) on 8-bit input multiplies in fp32 by (float)(1.0/255.0) (cvt_32f; checked against cv2 4.13:
    # the double-precision product differs in the last bit for 126 of the 256 codes)
    return gray, (gray.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)


def undistort_nearest(img, K, D, size, border):
    mx, my = cv2.initUndistortRectifyMap(K, D, None, K, size, cv2.CV_32FC1)
    return cv2.remap(img, mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_CONSTANT, borderValue=border)


def up(a, k):
    return np.ascontiguousarray(np.repeat(np.repeat(a, k, axis=0), k, axis=1))


def logs_to_arrays(logs):
    meta = np.array([[L["level"], L["iteration"], L["valid_pixels"]] for L in logs], np.int32).reshape(-1, 3)
    res = np.array([L["residual"] for L in logs], np.float32)
    upd = np.array([L["update"] for L in logs], np.float32).reshape(-1, 6)
    return meta, res, upd


def f32(x):
    return np.asarray(x, np.float32)


def map_depth_to_gray(ref, depth, gray, rgb_K, depth_K, invT):
    """Transform::mapDepthtoGray (src/core/transform.cpp:53-78) in the reference's fp32 operation order:
    backProject (d*(x-cx))/fx, transform R*p + t left to right, project (X*fx)/Z + cx (transform.cpp:7-33);
    the bilinear tap is the reference's own Convert::getSubpixel (called through oracle/_ref)."""
    h, w = depth.shape
    ys, xs = np.mgrid[0:h, 0:w]
    xs = xs.astype(np.float32); ys = ys.astype(np.float32)
    valid = ~(np.abs(depth) < np.float32(1e-6))                      # math::isEpsilon
    fx, fy, cx, cy = (f32(depth_K[0, 0]), f32(depth_K[1, 1]), f32(depth_K[0, 2]), f32(depth_K[1, 2]))
    X = (depth * (xs - cx)) / fx
    Y = (depth * (ys - cy)) / fy
    Z = depth
    R = f32(invT[:3, :3]); t = f32(invT[:3, 3])
    tx = ((R[0, 0] * X + R[0, 1] * Y) + R[0, 2] * Z) + t[0]
    ty = ((R[1, 0] * X + R[1, 1] * Y) + R[1, 2] * Z) + t[1]
    tz = ((R[2, 0] * X + R[2, 1] * Y) + R[2, 2] * Z) + t[2]
    gfx, gfy, gcx, gcy = (f32(rgb_K[0, 0]), f32(rgb_K[1, 1]), f32(rgb_K[0, 2]), f32(rgb_K[1, 2]))
    with np.errstate(all="ignore"):
        u = (tx * gfx) / tz + gcx
        v = (ty * gfy) / tz + gcy
    mapped = np.full((h, w), -2.0, np.float32)
    sigma = np.ones((h, w), np.float32)
    pts = np.stack([u[valid], v[valid]], axis=1).astype(np.float32)
    mapped[valid] = ref.get_subpixel_many(gray, pts, dense=False)
    sigma[valid] = np.float32(0.1)
    return mapped, sigma


def kinect_mapped(ref, cfg, rgb_path, depth_path):
    K1, D1 = f32(cfg["intrinsic1"]), f32(cfg["distortion1"])
    K2, D2 = f32(cfg["intrinsic2"]), f32(cfg["distortion2"])
    _, g = normalized_gray(rgb_path)
    dimg = cv2.imread(depth_path, cv2.IMREAD_UNCHANGED)
    d = (dimg.astype(np.float32) * np.float32(1.0 / 5000.0)).astype(np.float32)   # loader.cpp:145 (fp32 multiply, as above)
    g_u = undistort_nearest(g, K1, D1, (1920, 1080), -2.0)
    d_u = undistort_nearest(d, K2, D2, (512, 424), 0)
    invT = np.eye(4, dtype=np.float32)
    invT[:3, :3] = f32(cfg["rotation"]).T                                    # params_struct.hpp:38-46 (sic: -t, not -R^T t)
    invT[:3, 3] = -f32(cfg["translation"]).reshape(3)
    mapped, sigma = map_depth_to_gray(ref, d_u, g_u, K1, K2, invT)
    return mapped, d_u, sigma, K2


def yaml_all(path):
    fs = cv2.FileStorage(path, cv2.FILE_STORAGE_READ)
    out = {}
    for k in ("intrinsic", "distortion", "intrinsic1", "distortion1", "intrinsic2", "distortion2", "rotation", "translation"):
        n = fs.getNode(k)
        if not n.empty():
            out[k] = n.mat()
    fs.release()
    return out


def main():
    ref = Ref()
    out = {}
    # ---- C1: test/track.cpp on data/KINECT_50MM entries 9 -> 10 --------------------------------------
    cfg = yaml_all(os.path.join(REF, "external/camera-calibration/data/kinectv2_00/config.yaml"))
    lst = read_list(os.path.join(REF, "data/KINECT_50MM/info.txt"))
    frames = [kinect_mapped(ref, cfg, *lst[i]) for i in (9, 10)]
    K2 = frames[0][3]
    cull = lambda a: np.ascontiguousarray(a[::2, ::2])                       # what Frame(., 4, 1) samples
    g1, d1, s1 = (cull(a) for a in frames[0][:3])
    g2, d2, s2 = (cull(a) for a in frames[1][:3])
    out.update(c1_K=K2, c1_gray=np.stack([g1, g2]), c1_depth=np.stack([d1, d2]), c1_sigma10=np.rint(np.stack([s1, s2]) * 10).astype(np.uint8))
    pre = ref.frame_create(up(g1, 2), K2, 4, 1, up(d1, 2), up(s1, 2))        # pre_frame, test/track.cpp:24
    cur = ref.frame_create(up(g2, 2), K2, 4, 1, up(d2, 2), up(s2, 2))        # cur_frame, :25
    xi, logs = ref.track(cur, pre)                                           # Tracker::track(obj = cur, ref = pre)
    meta, res, upd = logs_to_arrays(logs)
    out.update(c1_xi=xi, c1_log_meta=meta, c1_log_residual=res, c1_log_update=upd)
    # pose BEFORE every logged iteration (tracker.cpp:46-51: xi <- concatenate(xi, update) unless NaN), with the reference's
    # own se3::concatenate: the reference's tracker DIVERGES on this pair (first update 1.4 m; README.md:4 says as much), so the
    # CUDA path is compared iteration by iteration from these poses, not by a free run of a chaotic sequence
    x = np.zeros(6, np.float32); before = []
    for L in logs:
        before.append(x.copy())
        nx = ref.se3_concatenate(x, L["update"])
        if not np.isnan(nx).any():
            x = nx
    assert np.array_equal(x, xi)
    out.update(c1_log_xi_before=np.stack(before).astype(np.float32))
    for lv in range(4):                                                      # first iteration of every level from xi = 0
        u, r, v = ref.optimize(cur.image(lv, "gray"), pre.image(lv, "gray"), pre.image(lv, "depth"), pre.image(lv, "sigma"), pre.K(lv),
                               np.zeros(6, np.float32), lv)
        out["c1_gn%d" % lv] = np.concatenate([u, [r, v]]).astype(np.float64)
    print("C1: %d iterations, xi = %s" % (len(logs), np.round(xi, 5)))
    # ---- C2: main.cpp on data/logicool0 -------------------------------------------------------------
    cfg2 = yaml_all(os.path.join(REF, "external/camera-calibration/data/logicool_00/config.yaml"))
    K = f32(cfg2["intrinsic"])
    Ku = np.array([[780, 0, 378], [0, 796, 220], [0, 0, 1]], np.float32)      # loader.cpp:17 (hard-coded, not the YAML's)
    Du = np.array([-0.0462, 0.152, -0.00429, 0.0117, -0.0725], np.float32)    # loader.cpp:18
    lst2 = read_list(os.path.join(REF, "data/logicool0/info.txt"))[:N_C2]
    codes = []
    for (path,) in [r[:1] for r in lst2]:
        u8, _ = normalized_gray(path)
        und = undistort_nearest(u8.astype(np.int16), Ku, Du, (640, 480), -1)  # NEAREST picks whole pixels: remap the codes
        codes.append(np.ascontiguousarray(und[::4, ::4]))                     # what Frame(., 3, 2) samples
    codes = np.stack(codes).astype(np.int16)
    gray = np.where(codes < 0, np.float32(-2.0), (codes.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)).astype(np.float32)
    d0, s0 = synth.random_depth_prior((120, 160), seed=4321)                  # stands in for Scene's cv::randn prior
    ref.rng_reset()
    vo = ref.vo_create(K)
    T, rel, kf, nkf, vu, metas = [], [], [], [], [], []
    for i in range(N_C2):
        T.append(vo.odometrize(up(gray[i], 4), d0 if i == 0 else None, s0 if i == 0 else None))
        kf.append(vo.last_was_keyframe()); nkf.append(vo.history_size()); vu.append(vo.last_valid_update())
        logs = vo.last_log() if i > 0 else []
        metas.append(logs)
        # the frame's relative pose = Tracker::track's fold xi <- concatenate(xi, update) (tracker.cpp:46-51),
        # recomputed with the reference's own se3::concatenate from the exactly-logged updates
        x = np.zeros(6, np.float32)
        for L in logs:
            nx = ref.se3_concatenate(x, L["update"])
            if not np.isnan(nx).any():
                x = nx
        rel.append(x)
        if i > 0 and vo.last_was_keyframe():
            assert np.array_equal(vo.ref_frame().xi()[1], x)
    out.update(c2_K=K, c2_codes=codes, c2_init_depth=d0, c2_T=np.stack(T).astype(np.float32), c2_keyframe=np.array(kf, np.uint8),
               c2_n_keyframes=np.array(nkf, np.int32), c2_valid_update=np.array(vu, np.int32), c2_rel_xi=np.stack(rel).astype(np.float32))
    n_it = np.array([len(m) for m in metas], np.int32)
    flat = [L for m in metas for L in m]
    meta, res, upd = logs_to_arrays(flat)
    out.update(c2_n_iters=n_it, c2_log_meta=meta, c2_log_residual=res, c2_log_update=upd)
    kfm = vo.ref_frame()
    out.update(c2_kf_depth=kfm.image(2, "depth"), c2_kf_sigma=kfm.image(2, "sigma"), c2_kf_age=kfm.image(2, "age").astype(np.uint8))
    print("C2: %d frames, %d keyframes, iterations per frame %s" % (N_C2, nkf[-1], n_it.tolist()))
    path = os.path.join(HERE, "bundled_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%.0f KB)" % (path, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
