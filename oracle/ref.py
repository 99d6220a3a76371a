"""ctypes binding of oracle/_ref/libdvo_ref.so — the REFERENCE'S OWN translation units compiled over
the OpenCV-API shim (oracle/Makefile, oracle/ref_capi.cpp).

TEST INFRASTRUCTURE ONLY.  Used to pin oracle/dvo_oracle.c against the reference code itself, to
generate tests/golden/*, and (the -Ofast/OpenMP build) as bench.py's reference arm.  The method
names mirror oracle.oracle.Oracle so the same test bodies drive both.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("DVO_REFERENCE_ROOT", "/root/reference")
F32P = C.POINTER(C.c_float)


class RefIterLog(C.Structure):
    _fields_ = [("level", C.c_int), ("iteration", C.c_int), ("valid_pixels", C.c_int), ("residual", C.c_float),
                ("update", C.c_float * 6)]


def lib_path(fast=False):
    return os.path.join(HERE, "_ref", "libdvo_ref_fast.so" if fast else "libdvo_ref.so")


def available(fast=False):
    return os.path.exists(lib_path(fast))


def build(fast=False):
    """(Re)build from /root/reference when it is present; otherwise use the prebuilt file."""
    if os.path.isdir(os.path.join(REF_ROOT, "src")):
        subprocess.check_call(["make", "-s", "-C", HERE, "REF=" + REF_ROOT,
                               "_ref/libdvo_ref_fast.so" if fast else "_ref/libdvo_ref.so"])
    if not available(fast):
        raise FileNotFoundError(lib_path(fast))
    return lib_path(fast)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(F32P) if a is not None else None


def _logs(log, n):
    return [dict(level=L.level, iteration=L.iteration, valid_pixels=L.valid_pixels, residual=L.residual,
                 update=np.array(L.update[:], np.float32)) for L in log[:n]]


class Ref:
    def __init__(self, path=None, fast=False):
        self.lib = L = C.CDLL(path or build(fast))
        L.ref_norm.restype = C.c_double
        L.ref_norm.argtypes = [F32P, C.c_int]
        for n in ("ref_get_subpixel", "ref_get_subpixel_dense"):
            getattr(L, n).restype = C.c_float
            getattr(L, n).argtypes = [F32P, C.c_int, C.c_int, C.c_float, C.c_float]
        L.ref_get_pixel.restype = C.c_float
        L.ref_get_pixel.argtypes = [F32P, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_get_subpixel_many.argtypes = [F32P, C.c_int, C.c_int, F32P, C.c_int, C.c_int, F32P]
        L.ref_warp_point.argtypes = [F32P, F32P, C.c_float, C.c_float, C.c_float, F32P]
        L.ref_gaussian_update.argtypes = [F32P, F32P, C.c_float, C.c_float]
        L.ref_gaussian_fuse.argtypes = [F32P, F32P, C.c_float, C.c_float]
        L.ref_optimize.argtypes = [F32P, F32P, F32P, F32P, F32P, F32P, C.c_int, C.c_int, C.c_int, F32P, F32P, C.POINTER(C.c_int)]
        L.ref_frame_create.restype = C.c_void_p
        L.ref_frame_create.argtypes = [F32P, F32P, F32P, C.c_int, C.c_int, F32P, C.c_int, C.c_int]
        L.ref_frame_destroy.argtypes = [C.c_void_p]
        L.ref_frame_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_frame_image.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P]
        L.ref_frame_K.argtypes = [C.c_void_p, C.c_int, F32P]
        L.ref_frame_set_depth_sigma_age.argtypes = [C.c_void_p, F32P, F32P, F32P]
        L.ref_frame_get_xi.argtypes = [C.c_void_p, F32P, F32P]
        L.ref_frame_set_xi.argtypes = [C.c_void_p, F32P, F32P, C.c_void_p]
        L.ref_track.argtypes = [C.c_void_p, C.c_void_p, F32P, C.POINTER(RefIterLog), C.c_int]
        L.ref_map_update_pixel.argtypes = [F32P, F32P, F32P, F32P, C.c_int, C.c_int, F32P, F32P, C.c_int, C.c_int, C.c_float,
                                           C.c_float, F32P, F32P]
        L.ref_mapper_update.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_void_p]
        L.ref_vo_create.restype = C.c_void_p
        L.ref_vo_create.argtypes = [F32P]
        L.ref_vo_destroy.argtypes = [C.c_void_p]
        L.ref_vo_odometrize.argtypes = [C.c_void_p, F32P, C.c_int, C.c_int, F32P, F32P]
        for n in ("ref_vo_history_size", "ref_vo_last_was_keyframe", "ref_vo_last_valid_update"):
            getattr(L, n).argtypes = [C.c_void_p]
        L.ref_vo_last_log.argtypes = [C.c_void_p, C.POINTER(RefIterLog), C.c_int]
        L.ref_vo_history_frame.restype = C.c_void_p
        L.ref_vo_history_frame.argtypes = [C.c_void_p, C.c_int]

    # ---- math
    def _vec(self, fn, a, n_out):
        a = _f32(a).reshape(-1); o = np.empty(n_out, np.float32)
        getattr(self.lib, fn)(_p(a), _p(o))
        return o

    def rodrigues(self, w): return self._vec("ref_so3_exp", w, 9).reshape(3, 3)
    def so3_log(self, R): return self._vec("ref_so3_log", R, 3)
    def se3_exp(self, xi): return self._vec("ref_se3_exp", xi, 16).reshape(4, 4)
    def se3_log(self, T): return self._vec("ref_se3_log", T, 6)
    def inverse_pose(self, T): return self._vec("ref_inverse_pose", T, 16).reshape(4, 4)

    def se3_concatenate(self, a, b):
        a = _f32(a); b = _f32(b); o = np.empty(6, np.float32)
        self.lib.ref_se3_concatenate(_p(a), _p(b), _p(o))
        return o

    def norm(self, v):
        v = _f32(v).reshape(-1)
        return self.lib.ref_norm(_p(v), v.size)

    # ---- gaussian
    def rng_reset(self):
        self.lib.ref_rng_reset()

    def gaussian_fuse(self, depth, sigma, d, s):
        a = np.array([depth], np.float32); b = np.array([sigma], np.float32)
        ok = self.lib.ref_gaussian_fuse(_p(a), _p(b), d, s)
        return float(a[0]), float(b[0]), bool(ok)

    def gaussian_update(self, depth, sigma, d, s):
        a = np.array([depth], np.float32); b = np.array([sigma], np.float32)
        ok = self.lib.ref_gaussian_update(_p(a), _p(b), d, s)
        return float(a[0]), float(b[0]), bool(ok)

    # ---- core
    def warp_point_xi(self, xi, K, x, y, d):
        xi = _f32(xi); K = _f32(K).reshape(-1); o = np.empty(2, np.float32)
        self.lib.ref_warp_point(_p(xi), _p(K), x, y, d, _p(o))
        return o

    def get_subpixel(self, img, px, py, dense=False):
        img = _f32(img); h, w = img.shape
        f = self.lib.ref_get_subpixel_dense if dense else self.lib.ref_get_subpixel
        return f(_p(img), w, h, px, py)

    def get_subpixel_many(self, img, pts, dense=False):
        img = _f32(img); h, w = img.shape; pts = _f32(pts).reshape(-1, 2); out = np.empty(len(pts), np.float32)
        self.lib.ref_get_subpixel_many(_p(img), w, h, _p(pts), len(pts), 1 if dense else 0, _p(out))
        return out

    def cull_image(self, img, times):
        img = _f32(img); h, w = img.shape
        out = np.empty((h >> times, w >> times), np.float32)
        self.lib.ref_cull_image(_p(img), C.c_int(w), C.c_int(h), C.c_int(times), _p(out))
        return out

    def cull_intrinsic(self, K, times):
        K = _f32(K).reshape(-1); o = np.empty(9, np.float32)
        self.lib.ref_cull_intrinsic(_p(K), C.c_int(times), _p(o))
        return o.reshape(3, 3)

    def gradiate(self, gray, is_x):
        gray = _f32(gray); h, w = gray.shape; out = np.empty_like(gray)
        self.lib.ref_gradiate(_p(gray), C.c_int(w), C.c_int(h), C.c_int(1 if is_x else 0), _p(out))
        return out

    def warp_image(self, xi, gray, depth, K):
        xi = _f32(xi); gray = _f32(gray); depth = _f32(depth); K = _f32(K).reshape(-1)
        h, w = gray.shape; out = np.empty_like(gray)
        self.lib.ref_warp_image(_p(xi), _p(gray), _p(depth), _p(K), C.c_int(w), C.c_int(h), _p(out))
        return out

    # ---- track
    def optimize(self, obj_gray, ref_gray, ref_depth, ref_sigma, K, xi, level_index):
        arrs = [_f32(a) for a in (obj_gray, ref_gray, ref_depth, ref_sigma)]
        h, w = arrs[0].shape
        K = _f32(K).reshape(-1); xi = _f32(xi)
        upd = np.empty(6, np.float32); res = np.empty(1, np.float32); valid = C.c_int(0)
        self.lib.ref_optimize(*[_p(a) for a in arrs], _p(K), _p(xi), w, h, level_index, _p(upd), _p(res), C.byref(valid))
        return upd, float(res[0]), valid.value

    def frame_create(self, gray, K, levels, culls, depth=None, sigma=None):
        gray = _f32(gray); h, w = gray.shape; K = _f32(K).reshape(-1)
        d = _f32(depth) if depth is not None else None
        s = _f32(sigma) if sigma is not None else None
        return RefFrame(self, self.lib.ref_frame_create(_p(gray), _p(d), _p(s), w, h, _p(K), levels, culls))

    def track(self, obj, ref, log_cap=256):
        xi = np.empty(6, np.float32)
        log = (RefIterLog * log_cap)()
        n = self.lib.ref_track(obj.h, ref.h, _p(xi), log, log_cap)
        return xi, _logs(log, min(n, log_cap))

    # ---- map
    def map_update_pixel(self, obj_gray, ref_gray, gx, gy, rel_xi, K, px, py, depth, sigma):
        arrs = [_f32(a) for a in (obj_gray, ref_gray, gx, gy)]
        h, w = arrs[0].shape
        rel_xi = _f32(rel_xi); K = _f32(K).reshape(-1)
        nd = np.empty(1, np.float32); ns = np.empty(1, np.float32)
        self.lib.ref_map_update_pixel(*[_p(a) for a in arrs], w, h, _p(rel_xi), _p(K), int(px), int(py), depth, sigma, _p(nd), _p(ns))
        return float(nd[0]), float(ns[0])

    def propagate(self, depth, sigma, age, xi, K):
        depth = _f32(depth); sigma = _f32(sigma); age = _f32(age); xi = _f32(xi); K = _f32(K).reshape(-1)
        h, w = depth.shape
        od = np.empty_like(depth); os_ = np.empty_like(depth); oa = np.empty_like(depth)
        self.lib.ref_propagate(_p(depth), _p(sigma), _p(age), _p(xi), _p(K), C.c_int(w), C.c_int(h), _p(od), _p(os_), _p(oa))
        return od, os_, oa

    def regularize(self, depth, sigma):
        depth = _f32(depth); sigma = _f32(sigma); h, w = depth.shape; out = np.empty_like(depth)
        self.lib.ref_regularize(_p(depth), _p(sigma), C.c_int(w), C.c_int(h), _p(out))
        return out

    def mapper_update(self, hist_frames, obj_frame):
        """Mapper::update on RefFrame objects; hist_frames[age]; returns 'valid update'."""
        arr = (C.c_void_p * len(hist_frames))(*[f.h for f in hist_frames])
        return self.lib.ref_mapper_update(arr, len(hist_frames), obj_frame.h)

    # ---- shim self-checks (restated OpenCV pieces)
    def shim_randn(self, seed, n, mean, stddev):
        out = np.empty(n, np.float32)
        self.lib.ref_shim_randn(C.c_ulonglong(seed), C.c_int(n), C.c_double(mean), C.c_double(stddev), _p(out))
        return out

    def shim_gemm(self, A, B, alpha=1.0, Cm=None, beta=0.0):
        A = _f32(A); B = _f32(B); m, k = A.shape; n = B.shape[1]
        Cm = _f32(Cm) if Cm is not None else None
        out = np.empty((m, n), np.float32)
        self.lib.ref_shim_gemm(_p(A), m, k, _p(B), n, C.c_double(alpha), _p(Cm), C.c_double(beta), _p(out))
        return out

    def shim_solve_svd(self, A, b):
        A = _f32(A); b = _f32(b).reshape(-1); m, n = A.shape; x = np.empty(n, np.float32)
        self.lib.ref_shim_solve_svd(_p(A), _p(b), m, n, _p(x))
        return x

    def shim_addex(self, A, B, alpha, beta):
        A = _f32(A).reshape(-1); B = _f32(B).reshape(-1) if B is not None else None
        out = np.empty_like(A)
        self.lib.ref_shim_addex(_p(A), _p(B), A.size, C.c_double(alpha), C.c_double(beta), _p(out))
        return out

    def vo_create(self, K):
        K = _f32(K).reshape(-1)
        return RefVO(self, self.lib.ref_vo_create(_p(K)))


class RefFrame:
    def __init__(self, ref, handle):
        self.ref = ref; self.h = handle

    def __del__(self):
        if self.h:
            self.ref.lib.ref_frame_destroy(self.h)
            self.h = None

    def level_size(self, level):
        w = C.c_int(); h = C.c_int()
        if self.ref.lib.ref_frame_level_size(self.h, level, C.byref(w), C.byref(h)):
            raise IndexError(level)
        return w.value, h.value

    def image(self, level, which):
        idx = {"gray": 0, "depth": 1, "sigma": 2, "gradx": 3, "grady": 4, "age": 5}[which]
        w, h = self.level_size(level)
        out = np.empty((h, w), np.float32)
        if self.ref.lib.ref_frame_image(self.h, level, idx, _p(out)):
            raise IndexError((level, which))
        return out

    def K(self, level):
        o = np.empty(9, np.float32)
        self.ref.lib.ref_frame_K(self.h, level, _p(o))
        return o.reshape(3, 3)

    def set_depth_sigma_age(self, depth=None, sigma=None, age=None):
        d = _f32(depth) if depth is not None else None
        s = _f32(sigma) if sigma is not None else None
        a = _f32(age) if age is not None else None
        self.ref.lib.ref_frame_set_depth_sigma_age(self.h, _p(d), _p(s), _p(a))

    def xi(self):
        a = np.empty(6, np.float32); b = np.empty(6, np.float32)
        self.ref.lib.ref_frame_get_xi(self.h, _p(a), _p(b))
        return a, b

    def set_xi(self, xi=None, rel_xi=None, ref_frame=None):
        a = _f32(xi) if xi is not None else None
        b = _f32(rel_xi) if rel_xi is not None else None
        self.ref.lib.ref_frame_set_xi(self.h, _p(a), _p(b), ref_frame.h if ref_frame is not None else None)


class RefVO:
    def __init__(self, ref, handle):
        self.ref = ref; self.h = handle

    def __del__(self):
        if self.h:
            self.ref.lib.ref_vo_destroy(self.h)
            self.h = None

    def odometrize(self, gray, init_depth=None, init_sigma=None):
        gray = _f32(gray); h, w = gray.shape
        d = _f32(init_depth) if init_depth is not None else None
        if init_sigma is not None:
            assert np.all(np.asarray(init_sigma) == 0.5), "Scene's constructor fixes sigma to 0.5 (frame.hpp:21)"
        if d is not None:
            assert d.shape == (h >> 2, w >> 2) and float(d.min()) >= 0.5
        T = np.empty(16, np.float32)
        self.ref.lib.ref_vo_odometrize(self.h, _p(gray), w, h, _p(d), _p(T))
        return T.reshape(4, 4)

    def history_size(self): return self.ref.lib.ref_vo_history_size(self.h)
    def last_was_keyframe(self): return bool(self.ref.lib.ref_vo_last_was_keyframe(self.h))
    def last_valid_update(self): return self.ref.lib.ref_vo_last_valid_update(self.h)

    def last_log(self):
        log = (RefIterLog * 256)()
        n = self.ref.lib.ref_vo_last_log(self.h, log, 256)
        return _logs(log, min(n, 256))

    def history_frame(self, age=0):
        h = self.ref.lib.ref_vo_history_frame(self.h, age)
        if not h:
            raise IndexError(age)
        return RefFrame(self.ref, h)

    def ref_frame(self):
        return self.history_frame(0)
