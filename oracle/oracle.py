"""ctypes binding of the CPU oracle (oracle/libdvo_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
F32P = C.POINTER(C.c_float)
U8P = C.POINTER(C.c_uint8)


class OrcParams(C.Structure):
    _fields_ = [
        ("min_residual", C.c_float), ("min_update", C.c_float), ("max_iterations", C.c_int),
        ("roi_level", C.c_int), ("roi_x0", C.c_int), ("roi_x1", C.c_int), ("roi_y0", C.c_int), ("roi_y1", C.c_int),
        ("min_depth", C.c_double), ("sigma_lo", C.c_float), ("sigma_hi", C.c_float),
        ("step_default", C.c_float), ("step_level1", C.c_float), ("step_level2", C.c_float),
        ("map_roi_x0", C.c_int), ("map_roi_x1", C.c_int), ("map_roi_y0", C.c_int), ("map_roi_y1", C.c_int),
        ("min_movement", C.c_float), ("max_forward", C.c_int),
        ("accept_depth_lo", C.c_float), ("accept_depth_hi", C.c_float), ("accept_sigma_hi", C.c_float),
        ("huber_delta", C.c_float), ("min_gradient", C.c_float),
    ]


class OrcIterLog(C.Structure):
    _fields_ = [("level", C.c_int), ("iteration", C.c_int), ("valid_pixels", C.c_int), ("residual", C.c_float),
                ("update", C.c_float * 6), ("xi_before", C.c_float * 6), ("xi_after", C.c_float * 6)]


class OrcRng(C.Structure):
    _fields_ = [("mt", C.c_uint32 * 624), ("idx", C.c_int), ("draws", C.c_uint64)]


def build(fast=False, out_dir=None):
    """Compile the C restatement (gcc).  fast=True uses the reference's release flags."""
    name = "libdvo_oracle_fast.so" if fast else "libdvo_oracle.so"
    out = os.path.join(out_dir or HERE, name)
    src = os.path.join(HERE, "dvo_oracle.c")
    if os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(os.path.join(HERE, "dvo_oracle.h"))):
        return out
    flags = ["-Ofast", "-march=native"] if fast else ["-O2", "-ffp-contract=off", "-fno-fast-math"]
    subprocess.check_call(["gcc", "-std=c11", *flags, "-fPIC", "-shared", "-o", out, src, "-lm"])
    return out


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def _p(a):
    return a.ctypes.data_as(F32P)


class Oracle:
    def __init__(self, path=None, fast=False):
        self.lib = C.CDLL(path or build(fast))
        L = self.lib
        L.orc_norm.restype = C.c_double
        L.orc_rng_dist.restype = C.c_float
        L.orc_dist_from_u32.restype = C.c_float
        L.orc_dist_from_u32.argtypes = [C.c_uint32]
        L.orc_rng_next_u32.restype = C.c_uint32
        L.orc_get_pixel.restype = C.c_float
        L.orc_get_subpixel.restype = C.c_float
        L.orc_get_subpixel.argtypes = [F32P, C.c_int, C.c_int, C.c_float, C.c_float]
        L.orc_get_subpixel_dense.restype = C.c_float
        L.orc_get_subpixel_dense.argtypes = [F32P, C.c_int, C.c_int, C.c_float, C.c_float]
        L.orc_warp_point.argtypes = [F32P, F32P, C.c_float, C.c_float, C.c_float, F32P]
        L.orc_frame_create.restype = C.c_void_p
        L.orc_frame_create.argtypes = [F32P, F32P, F32P, C.c_int, C.c_int, F32P, C.c_int, C.c_int]
        L.orc_frame_destroy.argtypes = [C.c_void_p]
        L.orc_frame_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_frame_image.restype = F32P
        L.orc_frame_image.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_frame_K.argtypes = [C.c_void_p, C.c_int, F32P]
        L.orc_frame_set_depth_sigma_age.argtypes = [C.c_void_p, F32P, F32P, F32P]
        L.orc_frame_get_xi.argtypes = [C.c_void_p, F32P, F32P]
        L.orc_track.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(OrcParams), F32P, C.POINTER(OrcIterLog), C.c_int,
                                C.c_double, C.c_float]
        L.orc_optimize.argtypes = [F32P, F32P, F32P, F32P, F32P, F32P, F32P, F32P, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(OrcParams), F32P, F32P, C.POINTER(C.c_int), U8P]
        L.orc_map_update_pixel.argtypes = [F32P, F32P, F32P, F32P, C.c_int, C.c_int, F32P, F32P, C.c_int, C.c_int,
                                           C.c_float, C.c_float, F32P, F32P]
        L.orc_gaussian_update.argtypes = [F32P, F32P, C.c_float, C.c_float, C.POINTER(OrcRng)]
        L.orc_gaussian_fuse.argtypes = [F32P, F32P, C.c_float, C.c_float]
        L.orc_vo_create.restype = C.c_void_p
        L.orc_vo_create.argtypes = [F32P, C.c_int, C.c_int, C.POINTER(OrcParams)]
        L.orc_vo_destroy.argtypes = [C.c_void_p]
        L.orc_vo_odometrize.argtypes = [C.c_void_p, F32P, C.c_int, C.c_int, F32P, F32P, F32P]
        for n in ("orc_vo_ref_frame", "orc_vo_last_frame"):
            getattr(L, n).restype = C.c_void_p
            getattr(L, n).argtypes = [C.c_void_p]
        for n in ("orc_vo_history_size", "orc_vo_last_was_keyframe", "orc_vo_last_valid_update"):
            getattr(L, n).argtypes = [C.c_void_p]
        L.orc_vo_rng_draws.restype = C.c_uint64
        L.orc_vo_rng_draws.argtypes = [C.c_void_p]
        L.orc_vo_last_update_mask.restype = U8P
        L.orc_vo_last_update_mask.argtypes = [C.c_void_p]
        L.orc_vo_last_log.argtypes = [C.c_void_p, C.POINTER(OrcIterLog), C.c_int]
        L.orc_mapper_update.argtypes = [F32P, F32P, F32P, C.c_int, C.c_int, F32P, C.POINTER(F32P), C.POINTER(F32P),
                                        C.POINTER(F32P), F32P, C.c_int, F32P, F32P, F32P, C.POINTER(OrcParams),
                                        C.POINTER(OrcRng), U8P, C.POINTER(C.c_int)]

    # ---- params
    def params(self, **kw):
        p = OrcParams()
        self.lib.orc_params_default(C.byref(p))
        for k, v in kw.items():
            setattr(p, k, v)
        return p

    # ---- math
    def rodrigues(self, w):
        w = _f32(w); R = np.empty(9, np.float32)
        self.lib.orc_rodrigues(_p(w), _p(R))
        return R.reshape(3, 3)

    def so3_log(self, R):
        R = _f32(R).reshape(-1); w = np.empty(3, np.float32)
        self.lib.orc_so3_log(_p(R), _p(w))
        return w

    def se3_exp(self, xi):
        xi = _f32(xi); T = np.empty(16, np.float32)
        self.lib.orc_se3_exp(_p(xi), _p(T))
        return T.reshape(4, 4)

    def se3_log(self, T):
        T = _f32(T).reshape(-1); xi = np.empty(6, np.float32)
        self.lib.orc_se3_log(_p(T), _p(xi))
        return xi

    def se3_concatenate(self, a, b):
        a = _f32(a); b = _f32(b); o = np.empty(6, np.float32)
        self.lib.orc_se3_concatenate(_p(a), _p(b), _p(o))
        return o

    def norm(self, v):
        v = _f32(v).reshape(-1)
        return self.lib.orc_norm(_p(v), C.c_int(v.size))

    def inverse_pose(self, T):
        T = _f32(T).reshape(-1); o = np.empty(16, np.float32)
        self.lib.orc_inverse_pose(_p(T), _p(o))
        return o.reshape(4, 4)

    # ---- rng / gaussian
    def rng(self, seed=5489):
        r = OrcRng()
        self.lib.orc_rng_seed(C.byref(r), C.c_uint32(seed))
        return r

    def rng_u32(self, n, seed=5489):
        r = self.rng(seed)
        return np.array([self.lib.orc_rng_next_u32(C.byref(r)) for _ in range(n)], dtype=np.uint32)

    def dist_from_u32(self, raw):
        return np.array([self.lib.orc_dist_from_u32(int(x)) for x in np.asarray(raw).reshape(-1)], np.float32)

    def gaussian_fuse(self, depth, sigma, d, s):
        a = np.array([depth], np.float32); b = np.array([sigma], np.float32)
        ok = self.lib.orc_gaussian_fuse(_p(a), _p(b), C.c_float(d), C.c_float(s))
        return float(a[0]), float(b[0]), bool(ok)

    def gaussian_update(self, depth, sigma, d, s, rng):
        a = np.array([depth], np.float32); b = np.array([sigma], np.float32)
        ok = self.lib.orc_gaussian_update(_p(a), _p(b), C.c_float(d), C.c_float(s), C.byref(rng))
        return float(a[0]), float(b[0]), bool(ok)

    # ---- core
    def warp_point(self, T, K, x, y, d):
        T = _f32(T).reshape(-1); K = _f32(K).reshape(-1); o = np.empty(2, np.float32)
        self.lib.orc_warp_point(_p(T), _p(K), x, y, d, _p(o))
        return o

    def get_subpixel(self, img, px, py, dense=False):
        img = _f32(img); h, w = img.shape
        f = self.lib.orc_get_subpixel_dense if dense else self.lib.orc_get_subpixel
        return f(_p(img), w, h, C.c_float(px), C.c_float(py))

    def cull_image(self, img, times):
        img = _f32(img); h, w = img.shape
        out = np.empty((h >> times, w >> times), np.float32)
        self.lib.orc_cull_image(_p(img), C.c_int(w), C.c_int(h), C.c_int(times), _p(out))
        return out

    def cull_intrinsic(self, K, times):
        K = _f32(K).reshape(-1); o = np.empty(9, np.float32)
        self.lib.orc_cull_intrinsic(_p(K), C.c_int(times), _p(o))
        return o.reshape(3, 3)

    def gradiate(self, gray, is_x):
        gray = _f32(gray); h, w = gray.shape; out = np.empty_like(gray)
        self.lib.orc_gradiate(_p(gray), C.c_int(w), C.c_int(h), C.c_int(1 if is_x else 0), _p(out))
        return out

    def warp_image(self, xi, gray, depth, K):
        xi = _f32(xi); gray = _f32(gray); depth = _f32(depth); K = _f32(K).reshape(-1)
        h, w = gray.shape; out = np.empty_like(gray)
        self.lib.orc_warp_image(_p(xi), _p(gray), _p(depth), _p(K), C.c_int(w), C.c_int(h), _p(out))
        return out

    # ---- track
    def solve_svd(self, A, b):
        A = _f32(A); b = _f32(b).reshape(-1); m, n = A.shape; x = np.empty(n, np.float32)
        self.lib.orc_solve_svd(_p(A), _p(b), C.c_int(m), C.c_int(n), _p(x))
        return x

    def optimize(self, obj_gray, warped_gray, ref_depth, ref_sigma, gx, gy, K, xi, level_index, params=None):
        arrs = [_f32(a) for a in (obj_gray, warped_gray, ref_depth, ref_sigma, gx, gy)]
        h, w = arrs[0].shape
        K = _f32(K).reshape(-1); xi = _f32(xi)
        P = params or self.params()
        upd = np.empty(6, np.float32); res = np.empty(1, np.float32); valid = C.c_int(0)
        mask = np.empty((h, w), np.uint8)
        self.lib.orc_optimize(*[_p(a) for a in arrs], _p(K), _p(xi), C.c_int(w), C.c_int(h), C.c_int(level_index),
                              C.byref(P), _p(upd), _p(res), C.byref(valid), mask.ctypes.data_as(U8P))
        return upd, float(res[0]), valid.value, mask

    def optimize_system(self, obj_gray, warped_gray, ref_depth, ref_sigma, gx, gy, K, xi, level_index, params=None):
        """Track::optimize plus the stacked system it hands to cv::solve: (upd, residual, valid, A[valid, 6], b[valid]) with
        A x = b the reference's `cv::solve(A, -B, x, DECOMP_SVD)` and upd = -x (optimize.cpp:87-98)."""
        arrs = [_f32(a) for a in (obj_gray, warped_gray, ref_depth, ref_sigma, gx, gy)]
        h, w = arrs[0].shape
        K = _f32(K).reshape(-1); xi = _f32(xi)
        P = params or self.params()
        upd = np.empty(6, np.float32); res = np.empty(1, np.float32); valid = C.c_int(0)
        A = np.empty((h * w, 6), np.float32); B = np.empty(h * w, np.float32)
        self.lib.orc_optimize_system.argtypes = None
        self.lib.orc_optimize_system(*[_p(a) for a in arrs], _p(K), _p(xi), C.c_int(w), C.c_int(h), C.c_int(level_index),
                                     C.byref(P), _p(upd), _p(res), C.byref(valid), None, _p(A), _p(B))
        return upd, float(res[0]), valid.value, A[:valid.value].copy(), B[:valid.value].copy()

    # ---- frames
    def frame_create(self, gray, K, levels, culls, depth=None, sigma=None):
        gray = _f32(gray); h, w = gray.shape; K = _f32(K).reshape(-1)
        d = _f32(depth) if depth is not None else None
        s = _f32(sigma) if sigma is not None else None
        f = self.lib.orc_frame_create(_p(gray), _p(d) if d is not None else None, _p(s) if s is not None else None,
                                      w, h, _p(K), levels, culls)
        return OrcFrame(self, f, owned=True)

    def track(self, obj, ref, params=None, stop_update=None, stop_residual=None, log_cap=256):
        P = params or self.params()
        if stop_update is None:
            stop_update = float(np.float32(P.min_update))
        if stop_residual is None:
            stop_residual = P.min_residual
        xi = np.empty(6, np.float32)
        log = (OrcIterLog * log_cap)()
        n = self.lib.orc_track(obj.h, ref.h, C.byref(P), _p(xi), log, log_cap, C.c_double(stop_update),
                               C.c_float(stop_residual))
        out = []
        for i in range(min(n, log_cap)):
            L = log[i]
            out.append(dict(level=L.level, iteration=L.iteration, valid_pixels=L.valid_pixels, residual=L.residual,
                            update=np.array(L.update[:], np.float32), xi_before=np.array(L.xi_before[:], np.float32),
                            xi_after=np.array(L.xi_after[:], np.float32)))
        return xi, out

    # ---- map
    def map_update_pixel(self, obj_gray, ref_gray, gx, gy, rel_xi, K, px, py, depth, sigma):
        arrs = [_f32(a) for a in (obj_gray, ref_gray, gx, gy)]
        h, w = arrs[0].shape
        rel_xi = _f32(rel_xi); K = _f32(K).reshape(-1)
        nd = np.empty(1, np.float32); ns = np.empty(1, np.float32)
        self.lib.orc_map_update_pixel(*[_p(a) for a in arrs], w, h, _p(rel_xi), _p(K), int(px), int(py),
                                      C.c_float(depth), C.c_float(sigma), _p(nd), _p(ns))
        return float(nd[0]), float(ns[0])

    def propagate(self, depth, sigma, age, xi, K):
        depth = _f32(depth); sigma = _f32(sigma); age = _f32(age); xi = _f32(xi); K = _f32(K).reshape(-1)
        h, w = depth.shape
        od = np.empty_like(depth); os_ = np.empty_like(depth); oa = np.empty_like(depth)
        self.lib.orc_propagate(_p(depth), _p(sigma), _p(age), _p(xi), _p(K), C.c_int(w), C.c_int(h), _p(od), _p(os_), _p(oa))
        return od, os_, oa

    def regularize(self, depth, sigma):
        depth = _f32(depth); sigma = _f32(sigma); h, w = depth.shape; out = np.empty_like(depth)
        self.lib.orc_regularize(_p(depth), _p(sigma), C.c_int(w), C.c_int(h), _p(out))
        return out

    VIS = {"gray": 0, "depth": 1, "depth_sigma": 2, "depth_raw": 3, "sigma": 4, "gradient": 5, "age": 6}

    def visualize(self, kind, src, aux=None):
        """Draw::visualize* (draw.cpp:7-100): float image -> (h, w, 3) BGR bytes."""
        src = _f32(src); out = np.empty(src.shape + (3,), np.uint8)
        aux = _f32(aux) if aux is not None else None
        self.lib.orc_visualize(C.c_int(self.VIS[kind]), _p(src), _p(aux) if aux is not None else None, C.c_int(src.size),
                               out.ctypes.data_as(C.c_void_p))
        return out

    def hsv2bgr(self, h, s, v):
        out = (C.c_uint8 * 3)()
        self.lib.orc_hsv2bgr_8u(C.c_uint8(h), C.c_uint8(s), C.c_uint8(v), out)
        return tuple(out)

    def convert_8u(self, v, a, b):
        self.lib.orc_convert_8u.restype = C.c_uint8
        return int(self.lib.orc_convert_8u(C.c_float(v), C.c_float(a), C.c_float(b)))

    def mapper_update(self, depth, sigma, age, obj_gray, born_gray, born_gx, born_gy, hist_xi, obj_xi, rel_xi, K,
                      params=None, rng=None):
        """In-place on copies; returns (depth, sigma, age, mask, valid_update, rc)."""
        depth = _f32(depth).copy(); sigma = _f32(sigma).copy(); age = _f32(age).copy()
        obj_gray = _f32(obj_gray); h, w = depth.shape
        n = len(born_gray)
        keep = [[_f32(a) for a in lst] for lst in (born_gray, born_gx, born_gy)]
        ptrs = [(F32P * n)(*[_p(a) for a in lst]) for lst in keep]
        hist_xi = _f32(hist_xi).reshape(n, 6); obj_xi = _f32(obj_xi); rel_xi = _f32(rel_xi); K = _f32(K).reshape(-1)
        P = params or self.params()
        rng = rng if rng is not None else self.rng()
        mask = np.zeros((h, w), np.uint8); vu = C.c_int(0)
        rc = self.lib.orc_mapper_update(_p(depth), _p(sigma), _p(age), w, h, _p(obj_gray), ptrs[0], ptrs[1], ptrs[2],
                                        _p(hist_xi), n, _p(obj_xi), _p(rel_xi), _p(K), C.byref(P), C.byref(rng),
                                        mask.ctypes.data_as(U8P), C.byref(vu))
        return depth, sigma, age, mask, vu.value, rc

    # ---- VO
    def vo_create(self, K, levels=3, culls=2, params=None):
        K = _f32(K).reshape(-1)
        P = params or self.params()
        return OrcVO(self, self.lib.orc_vo_create(_p(K), levels, culls, C.byref(P)))


class OrcFrame:
    def __init__(self, orc, handle, owned):
        self.orc = orc; self.h = handle; self.owned = owned

    def __del__(self):
        if self.owned and self.h:
            self.orc.lib.orc_frame_destroy(self.h)
            self.h = None

    def level_size(self, level):
        w = C.c_int(); h = C.c_int()
        if self.orc.lib.orc_frame_level_size(self.h, level, C.byref(w), C.byref(h)):
            raise IndexError(level)
        return w.value, h.value

    def image(self, level, which):
        idx = {"gray": 0, "depth": 1, "sigma": 2, "gradx": 3, "grady": 4, "age": 5}[which]
        w, h = self.level_size(level)
        p = self.orc.lib.orc_frame_image(self.h, level, idx)
        return np.ctypeslib.as_array(p, shape=(h, w)).copy()

    def K(self, level):
        o = np.empty(9, np.float32)
        self.orc.lib.orc_frame_K(self.h, level, _p(o))
        return o.reshape(3, 3)

    def set_depth_sigma_age(self, depth=None, sigma=None, age=None):
        d = _f32(depth) if depth is not None else None
        s = _f32(sigma) if sigma is not None else None
        a = _f32(age) if age is not None else None
        self.orc.lib.orc_frame_set_depth_sigma_age(self.h, _p(d) if d is not None else None,
                                                   _p(s) if s is not None else None, _p(a) if a is not None else None)

    def xi(self):
        a = np.empty(6, np.float32); b = np.empty(6, np.float32)
        self.orc.lib.orc_frame_get_xi(self.h, _p(a), _p(b))
        return a, b


class OrcVO:
    def __init__(self, orc, handle):
        self.orc = orc; self.h = handle

    def __del__(self):
        if self.h:
            self.orc.lib.orc_vo_destroy(self.h)
            self.h = None

    def odometrize(self, gray, init_depth=None, init_sigma=None):
        gray = _f32(gray); h, w = gray.shape
        d = _f32(init_depth) if init_depth is not None else None
        s = _f32(init_sigma) if init_sigma is not None else None
        T = np.empty(16, np.float32)
        rc = self.orc.lib.orc_vo_odometrize(self.h, _p(gray), w, h, _p(d) if d is not None else None,
                                            _p(s) if s is not None else None, _p(T))
        if rc:
            raise RuntimeError("orc_vo_odometrize rc=%d" % rc)
        return T.reshape(4, 4)

    def ref_frame(self):
        return OrcFrame(self.orc, self.orc.lib.orc_vo_ref_frame(self.h), owned=False)

    def last_frame(self):
        return OrcFrame(self.orc, self.orc.lib.orc_vo_last_frame(self.h), owned=False)

    def history_size(self):
        return self.orc.lib.orc_vo_history_size(self.h)

    def last_was_keyframe(self):
        return bool(self.orc.lib.orc_vo_last_was_keyframe(self.h))

    def last_valid_update(self):
        return self.orc.lib.orc_vo_last_valid_update(self.h)

    def rng_draws(self):
        return int(self.orc.lib.orc_vo_rng_draws(self.h))

    def last_log(self):
        log = (OrcIterLog * 256)()
        n = self.orc.lib.orc_vo_last_log(self.h, log, 256)
        return [dict(level=L.level, iteration=L.iteration, valid_pixels=L.valid_pixels, residual=L.residual,
                     update=np.array(L.update[:], np.float32), xi_before=np.array(L.xi_before[:], np.float32),
                     xi_after=np.array(L.xi_after[:], np.float32)) for L in log[:min(n, 256)]]

    def last_update_mask(self):
        f = self.ref_frame()
        lv = 0
        w = h = None
        while True:
            try:
                w, h = f.level_size(lv)
                lv += 1
            except IndexError:
                break
        p = self.orc.lib.orc_vo_last_update_mask(self.h)
        if not p:
            return None
        return np.ctypeslib.as_array(p, shape=(h, w)).copy()
