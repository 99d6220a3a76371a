"""Shared helpers for the parity tests (oracle <-> CUDA through the C ABI)."""
import numpy as np


def se3_norm(a, b):
    return float(np.linalg.norm(np.asarray(a, np.float64) - np.asarray(b, np.float64)))


def punch_invalid(gray, seed=0, n_rect=6, frac_px=0.01):
    """Mark rectangles and scattered pixels INVALID (-2.0), like the undistortion border / missing
    Kinect depth of the reference's loaders (src/core/loader.cpp:39-40, transform.cpp:60)."""
    rng = np.random.default_rng(seed)
    g = gray.copy()
    h, w = g.shape[-2:]
    for _ in range(n_rect):
        x0 = int(rng.integers(0, w - 8)); y0 = int(rng.integers(0, h - 8))
        ww = int(rng.integers(4, max(5, w // 8))); hh = int(rng.integers(4, max(5, h // 8)))
        g[..., y0:y0 + hh, x0:x0 + ww] = -2.0
    m = rng.random(g.shape) < frac_px
    g[m] = -2.0
    g[..., :, :3] = -2.0  # border strip as produced by cv::remap(BORDER_CONSTANT, INVALID)
    return g


def oracle_frame_pair(orc, gray_ref, gray_obj, K, levels, culls, depth=None, sigma=None, full_res=True):
    """(ref, obj) oracle frames; depth/sigma for ref either full-res (6-arg ctor) or top-level sized."""
    if full_res:
        ref = orc.frame_create(gray_ref, K, levels, culls, depth, sigma)
    else:
        ref = orc.frame_create(gray_ref, K, levels, culls)
        ref.set_depth_sigma_age(depth, sigma, None)
    obj = orc.frame_create(gray_obj, K, levels, culls)
    return ref, obj


def oracle_gn_inputs(orc, ref, obj, level, xi):
    g = {k: ref.image(level, k) for k in ("gray", "depth", "sigma", "gradx", "grady")}
    K = ref.K(level)
    warped = orc.warp_image(xi, g["gray"], g["depth"], K)
    return obj.image(level, "gray"), warped, g["depth"], g["sigma"], g["gradx"], g["grady"], K
