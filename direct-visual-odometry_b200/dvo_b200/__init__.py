"""dvo_b200 — thin ctypes loader for libdvo_cuda.so (the C ABI declared in include/dvo_cuda.h).

This module is plumbing for tests and bench.py only: the product is the CUDA library and the C++
host classes in include/dvo/.  There is no CPU fallback: if the library is missing or no CUDA
device is present, construction fails loudly.
"""
import ctypes as C
import os

import numpy as np

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("DVO_LIB") or os.path.join(PKG_ROOT, "libdvo_cuda.so")  # DVO_LIB: experiment builds
F32P = C.POINTER(C.c_float)
I32P = C.POINTER(C.c_int)
U8P = C.POINTER(C.c_uint8)

IMG = {"gray": 0, "depth": 1, "sigma": 2, "gradx": 3, "grady": 4, "age": 5}
VIS = {"gray": 0, "depth": 1, "depth_sigma": 2, "depth_raw": 3, "sigma": 4, "gradx": 5, "grady": 6, "age": 7}  # dvo_vis_kind


class DvoParams(C.Structure):
    _fields_ = [
        ("n_streams", C.c_int), ("width", C.c_int), ("height", C.c_int), ("levels", C.c_int), ("culls", C.c_int),
        ("K", C.c_float * 9), ("max_frames", C.c_int), ("max_keyframes", C.c_int),
        ("min_residual", C.c_float), ("min_update", C.c_float), ("max_iterations", C.c_int),
        ("roi_level", C.c_int), ("roi_x0", C.c_int), ("roi_x1", C.c_int), ("roi_y0", C.c_int), ("roi_y1", C.c_int),
        ("min_depth", C.c_double), ("sigma_lo", C.c_float), ("sigma_hi", C.c_float),
        ("step_default", C.c_float), ("step_level1", C.c_float), ("step_level2", C.c_float),
        ("map_roi_x0", C.c_int), ("map_roi_x1", C.c_int), ("map_roi_y0", C.c_int), ("map_roi_y1", C.c_int),
        ("min_movement", C.c_float), ("max_forward", C.c_int),
        ("accept_depth_lo", C.c_float), ("accept_depth_hi", C.c_float), ("accept_sigma_hi", C.c_float),
        ("rng_table", C.c_int), ("huber_delta", C.c_float), ("min_gradient", C.c_float),
    ]


class DvoKinectParams(C.Structure):
    _fields_ = [("color_width", C.c_int), ("color_height", C.c_int), ("channels", C.c_int), ("rgb_K", C.c_float * 9), ("invT", C.c_float * 16)]


class DvoIterLog(C.Structure):
    _fields_ = [("level", C.c_int), ("iteration", C.c_int), ("valid_pixels", C.c_int), ("residual", C.c_float),
                ("update", C.c_float * 6), ("xi_after", C.c_float * 6)]


EXPORTS = [
    "dvo_device_count", "dvo_set_device", "dvo_params_default", "dvo_create", "dvo_destroy", "dvo_last_error", "dvo_sync", "dvo_level_size", "dvo_level_K",
    "dvo_host_alloc", "dvo_host_free", "dvo_frame_create", "dvo_frame_create_dev", "dvo_frame_release",
    "dvo_frame_set_depth", "dvo_frame_download", "dvo_visualize", "dvo_visualize_image", "dvo_inverse_pose", "dvo_write_trajectory_csv", "dvo_write_png", "dvo_frame_set_xi", "dvo_frame_get_xi", "dvo_warp_image",
    "dvo_gn_step", "dvo_track", "dvo_map_propagate", "dvo_map_update", "dvo_map_regularize", "dvo_rng_position",
    "dvo_mask_compact", "dvo_mask_compact_noalloc", "dvo_vo_reset", "dvo_vo_reserve", "dvo_vo_force_pose", "dvo_vo_push", "dvo_vo_push_dev", "dvo_vo_state", "dvo_vo_ref_slot",
    "dvo_vo_last_slot", "dvo_vo_last_log", "dvo_vo_counters", "dvo_vo_matched", "dvo_vo_error", "dvo_synth_render_dev", "dvo_dev_alloc", "dvo_dev_free",
    "dvo_copy_to_host", "dvo_copy_to_dev", "dvo_timer_start", "dvo_timer_stop", "dvo_launch_count", "dvo_gn_timing", "dvo_kernel_timing",
    "dvo_map_update_points", "dvo_se3_exp", "dvo_se3_log", "dvo_se3_concatenate", "dvo_vo_submit", "dvo_vo_collect",
    "dvo_frame_create_gray8", "dvo_vo_push_gray8", "dvo_vo_submit_gray8",
    "dvo_set_undistort_map", "dvo_frame_create_bgr8", "dvo_vo_push_bgr8", "dvo_vo_submit_bgr8",
    "dvo_kinect_configure", "dvo_frame_create_kinect",
    "dvo_vo_push_gray8_dev", "dvo_synth_pose", "dvo_synth_scene", "dvo_event_record", "dvo_event_elapsed",
    "dvo_synth_render_gray8_dev",
    "dvo_transform_point", "dvo_project_point", "dvo_back_project_point", "dvo_gaussian_fuse", "dvo_gaussian_update", "dvo_gaussian_reseed",
    "dvo_map_depth_to_gray", "dvo_select_pixels",
]


class DvoKernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("total_ms", C.c_float), ("launches", C.c_int)]


def load_library(path=None):
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise RuntimeError("libdvo_cuda.so not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                           "there is no CPU fallback" % path)
    lib = C.CDLL(path)
    lib.dvo_last_error.restype = C.c_char_p
    lib.dvo_last_error.argtypes = [C.c_void_p]
    lib.dvo_set_device.argtypes = [C.c_int]
    lib.dvo_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(DvoParams)]
    lib.dvo_destroy.argtypes = [C.c_void_p]
    lib.dvo_sync.argtypes = [C.c_void_p]
    lib.dvo_level_size.argtypes = [C.c_void_p, C.c_int, I32P, I32P]
    lib.dvo_level_K.argtypes = [C.c_void_p, C.c_int, F32P]
    lib.dvo_host_alloc.restype = C.c_void_p
    lib.dvo_host_alloc.argtypes = [C.c_size_t]
    lib.dvo_host_free.argtypes = [C.c_void_p]
    lib.dvo_frame_create.argtypes = [C.c_void_p, I32P, C.c_void_p]
    lib.dvo_frame_create_dev.argtypes = [C.c_void_p, I32P, C.c_void_p]
    lib.dvo_frame_release.argtypes = [C.c_void_p, C.c_int]
    lib.dvo_frame_set_depth.argtypes = [C.c_void_p, C.c_int, F32P, F32P, F32P, C.c_int]
    lib.dvo_frame_download.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, F32P]
    lib.dvo_visualize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.dvo_visualize_image.argtypes = [C.c_int, F32P, F32P, C.c_int, C.c_int, C.c_void_p]
    lib.dvo_inverse_pose.argtypes = [F32P, F32P]
    lib.dvo_write_trajectory_csv.argtypes = [C.c_char_p, F32P, C.c_int, C.c_int, C.c_int]
    lib.dvo_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    lib.dvo_frame_set_xi.argtypes = [C.c_void_p, C.c_int, F32P, F32P]
    lib.dvo_frame_get_xi.argtypes = [C.c_void_p, C.c_int, F32P, F32P]
    lib.dvo_warp_image.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P, F32P]
    lib.dvo_gn_step.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, F32P, F32P, F32P, I32P, U8P]
    lib.dvo_track.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P, C.POINTER(DvoIterLog), C.c_int, I32P, C.c_double,
                              C.c_float]
    lib.dvo_map_propagate.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P]
    lib.dvo_map_update.argtypes = [C.c_void_p, C.c_int, C.c_int, I32P, C.c_int, I32P, U8P]
    lib.dvo_map_regularize.argtypes = [C.c_void_p, C.c_int]
    lib.dvo_rng_position.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.dvo_mask_compact.argtypes = [C.c_void_p, U8P, C.c_int, I32P, I32P]
    lib.dvo_mask_compact_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.dvo_vo_reset.argtypes = [C.c_void_p]
    lib.dvo_vo_reserve.argtypes = [C.c_void_p, C.c_int]
    lib.dvo_vo_push.argtypes = [C.c_void_p, C.c_void_p, F32P, F32P, F32P]
    lib.dvo_vo_push_dev.argtypes = [C.c_void_p, C.c_void_p, F32P, F32P, F32P]
    lib.dvo_vo_state.argtypes = [C.c_void_p, F32P, F32P, I32P, I32P, I32P, I32P]
    lib.dvo_vo_ref_slot.argtypes = [C.c_void_p, C.c_int, I32P]
    lib.dvo_vo_last_slot.argtypes = [C.c_void_p, C.c_int, I32P]
    lib.dvo_vo_force_pose.argtypes = [C.c_void_p, F32P]
    lib.dvo_vo_error.argtypes = [C.c_void_p]
    U64P = C.POINTER(C.c_uint64)
    lib.dvo_vo_counters.argtypes = [C.c_void_p, U64P, U64P, U64P, U64P]
    lib.dvo_vo_matched.argtypes = [C.c_void_p, U64P]
    lib.dvo_vo_last_log.argtypes = [C.c_void_p, C.POINTER(DvoIterLog), C.c_int, I32P]
    lib.dvo_synth_render_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_float, C.c_float]
    lib.dvo_dev_alloc.restype = C.c_void_p
    lib.dvo_dev_alloc.argtypes = [C.c_size_t]
    lib.dvo_dev_free.argtypes = [C.c_void_p]
    lib.dvo_copy_to_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.dvo_copy_to_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.dvo_timer_start.argtypes = [C.c_void_p]
    lib.dvo_timer_stop.argtypes = [C.c_void_p, F32P]
    lib.dvo_launch_count.restype = C.c_uint64
    lib.dvo_launch_count.argtypes = [C.c_void_p]
    lib.dvo_gn_timing.argtypes = [C.c_void_p, C.c_int, F32P, I32P]
    lib.dvo_kernel_timing.argtypes = [C.c_void_p, C.c_int, C.POINTER(DvoKernelTime), C.c_int, I32P]
    lib.dvo_map_update_points.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P, C.c_int, I32P, I32P, F32P, F32P, F32P, F32P]
    lib.dvo_se3_exp.argtypes = [C.c_void_p, F32P, F32P, C.c_int]
    lib.dvo_se3_log.argtypes = [C.c_void_p, F32P, F32P, C.c_int]
    lib.dvo_se3_concatenate.argtypes = [C.c_void_p, F32P, F32P, F32P, C.c_int]
    lib.dvo_vo_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dvo_vo_collect.argtypes = [C.c_void_p]
    lib.dvo_frame_create_gray8.argtypes = [C.c_void_p, I32P, C.c_void_p, C.c_int]
    lib.dvo_vo_push_gray8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, F32P, F32P, F32P]
    lib.dvo_vo_submit_gray8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.dvo_set_undistort_map.argtypes = [C.c_void_p, F32P, F32P]
    lib.dvo_kinect_configure.argtypes = [C.c_void_p, C.POINTER(DvoKinectParams), F32P, F32P, F32P, F32P]
    lib.dvo_frame_create_kinect.argtypes = [C.c_void_p, I32P, C.c_void_p, C.c_void_p]
    lib.dvo_frame_create_bgr8.argtypes = [C.c_void_p, I32P, C.c_void_p, C.c_int]
    lib.dvo_vo_push_gray8_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, F32P, F32P, F32P]
    lib.dvo_synth_pose.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_float, C.c_float, F32P]
    lib.dvo_synth_scene.argtypes = [C.c_void_p, C.c_uint64, F32P, C.c_int]
    lib.dvo_synth_render_gray8_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_float, C.c_float]
    lib.dvo_transform_point.argtypes = [F32P, F32P, F32P]
    lib.dvo_project_point.argtypes = [F32P, F32P, F32P]
    lib.dvo_back_project_point.argtypes = [F32P, F32P, C.c_float, F32P]
    lib.dvo_gaussian_fuse.argtypes = [F32P, F32P, C.c_float, C.c_float]
    lib.dvo_gaussian_update.argtypes = [F32P, F32P, C.c_float, C.c_float]
    lib.dvo_gaussian_reseed.restype = None
    lib.dvo_map_depth_to_gray.argtypes = [F32P, C.c_int, C.c_int, F32P, C.c_int, C.c_int, F32P, F32P, F32P, F32P, F32P]
    lib.dvo_select_pixels.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, I32P, I32P]
    lib.dvo_event_record.argtypes = [C.c_void_p, C.c_int]
    lib.dvo_event_elapsed.argtypes = [C.c_void_p, C.c_int, C.c_int, F32P]
    lib.dvo_vo_push_bgr8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, F32P, F32P, F32P]
    lib.dvo_vo_submit_bgr8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return lib


def default_params(lib=None, **kw):
    lib = lib or load_library()
    p = DvoParams()
    lib.dvo_params_default(C.byref(p))
    for k, v in kw.items():
        if k == "K":
            v = np.asarray(v, np.float32).reshape(-1)
            for i in range(9):
                p.K[i] = float(v[i])
        else:
            setattr(p, k, v)
    return p


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(F32P) if a is not None else None


class DvoError(RuntimeError):
    pass


class Dvo:
    """One context = n_streams sequences in lock-step on the current CUDA device."""

    def __init__(self, lib=None, **params):
        self.lib = lib or load_library()
        self.params = default_params(self.lib, **params)
        self.ctx = C.c_void_p()
        rc = self.lib.dvo_create(C.byref(self.ctx), C.byref(self.params))
        if rc:
            msg = self.lib.dvo_last_error(self.ctx).decode() if self.ctx else "dvo_create failed"
            if self.ctx:
                self.lib.dvo_destroy(self.ctx)
                self.ctx = None
            raise DvoError("dvo_create rc=%d: %s" % (rc, msg))
        self.S = self.params.n_streams
        self.L = self.params.levels

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.dvo_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        self.close()

    def _ck(self, rc):
        if rc:
            raise DvoError("rc=%d: %s" % (rc, self.lib.dvo_last_error(self.ctx).decode()))

    # ---- geometry
    def level_size(self, level):
        cache = self.__dict__.setdefault("_level_sizes", {})
        if level not in cache:
            w = C.c_int(); h = C.c_int()
            self._ck(self.lib.dvo_level_size(self.ctx, level, C.byref(w), C.byref(h)))
            cache[level] = (w.value, h.value)
        return cache[level]

    def level_K(self, level):
        K = np.empty(9, np.float32)
        self._ck(self.lib.dvo_level_K(self.ctx, level, _p(K)))
        return K.reshape(3, 3)

    # ---- frames
    def _stack(self, img, h, w):
        a = _f32(img)
        if a.ndim == 2:
            a = a[None]
        assert a.shape == (self.S, h, w), (a.shape, (self.S, h, w))
        return np.ascontiguousarray(a)

    def frame_create(self, gray):
        g = self._stack(gray, self.params.height, self.params.width)
        slot = C.c_int(-1)
        self._ck(self.lib.dvo_frame_create(self.ctx, C.byref(slot), g.ctypes.data_as(C.c_void_p)))
        self._ck(self.lib.dvo_sync(self.ctx))
        return slot.value

    def _stack8(self, img):
        a = np.ascontiguousarray(img, dtype=np.uint8)
        if a.ndim == 2:
            a = a[None]
        assert a.shape == (self.S, self.params.height, self.params.width), a.shape
        return a

    def frame_create_gray8(self, gray8, invalid_code=-1):
        """Frame from 8-bit gray: convertTo(CV_32FC1, 1.0/255.0) + cullImage on the device (loader.cpp:59-61)."""
        g = self._stack8(gray8)
        slot = C.c_int(-1)
        self._ck(self.lib.dvo_frame_create_gray8(self.ctx, C.byref(slot), g.ctypes.data_as(C.c_void_p), int(invalid_code)))
        self._ck(self.lib.dvo_sync(self.ctx))
        return slot.value

    def set_undistort_map(self, map_x=None, map_y=None):
        """The CV_32FC1 maps of cv::initUndistortRectifyMap (Loader::createUndistortMap, loader.cpp:15-30); None = identity."""
        if map_x is None:
            self._ck(self.lib.dvo_set_undistort_map(self.ctx, None, None))
            return
        mx = _f32(map_x); my = _f32(map_y)
        assert mx.shape == my.shape == (self.params.height, self.params.width)
        self._ck(self.lib.dvo_set_undistort_map(self.ctx, _p(mx), _p(my)))

    def _stack_bgr(self, img):
        a = np.ascontiguousarray(img, dtype=np.uint8)
        if a.ndim == 3:
            a = a[None]
        assert a.shape[:3] == (self.S, self.params.height, self.params.width) and a.shape[3] in (3, 4), a.shape
        return a

    def frame_create_bgr8(self, bgr8):
        """Frame from the 8-bit BGR / BGRA image: cvtColor + convertTo + remap(NEAREST) + cullImage on the device (loader.cpp:32-63)."""
        g = self._stack_bgr(bgr8)
        slot = C.c_int(-1)
        self._ck(self.lib.dvo_frame_create_bgr8(self.ctx, C.byref(slot), g.ctypes.data_as(C.c_void_p), g.shape[3]))
        self._ck(self.lib.dvo_sync(self.ctx))
        return slot.value

    def vo_push_bgr8(self, bgr8, init_depth=None, init_sigma=None, want_pose=True):
        g = self._stack_bgr(bgr8)
        w, h = self.level_size(self.L - 1)
        d = self._stack(init_depth, h, w) if init_depth is not None else None
        s = self._stack(init_sigma, h, w) if init_sigma is not None else None
        T = np.empty((self.S, 4, 4), np.float32) if want_pose else None
        self._ck(self.lib.dvo_vo_push_bgr8(self.ctx, g.ctypes.data_as(C.c_void_p), g.shape[3], _p(d), _p(s), _p(T)))
        if not want_pose:
            self._ck(self.lib.dvo_sync(self.ctx))
        return T

    def kinect_configure(self, color_size, channels, rgb_K, invT, rgb_maps, depth_maps):
        """KinectLoader::createUndistortMap's four CV_32FC1 maps + the calibration that Transform::mapDepthtoGray uses."""
        kp = DvoKinectParams()
        kp.color_width, kp.color_height = int(color_size[0]), int(color_size[1]); kp.channels = int(channels)
        kp.rgb_K[:] = [float(v) for v in np.asarray(rgb_K, np.float32).reshape(9)]
        kp.invT[:] = [float(v) for v in np.asarray(invT, np.float32).reshape(16)]
        m = [_f32(a) for a in (rgb_maps[0], rgb_maps[1], depth_maps[0], depth_maps[1])]
        assert m[0].shape == m[1].shape == (kp.color_height, kp.color_width)
        assert m[2].shape == m[3].shape == (self.params.height, self.params.width)
        self._ck(self.lib.dvo_kinect_configure(self.ctx, C.byref(kp), _p(m[0]), _p(m[1]), _p(m[2]), _p(m[3])))
        self._kinect = (kp.color_height, kp.color_width, kp.channels)

    def frame_create_kinect(self, bgr8, depth16):
        """Frame from one raw Kinect pair: getMappedImages + mapDepthtoGray + Frame(gray, depth, sigma, K, levels, culls)."""
        ch_, cw_, cc = self._kinect
        b = np.ascontiguousarray(bgr8, dtype=np.uint8); d = np.ascontiguousarray(depth16, dtype=np.uint16)
        if b.ndim == 3:
            b = b[None]
        if d.ndim == 2:
            d = d[None]
        assert b.shape == (self.S, ch_, cw_, cc) and d.shape == (self.S, self.params.height, self.params.width), (b.shape, d.shape)
        slot = C.c_int(-1)
        self._ck(self.lib.dvo_frame_create_kinect(self.ctx, C.byref(slot), b.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p)))
        self._ck(self.lib.dvo_sync(self.ctx))
        return slot.value

    def frame_release(self, slot):
        self._ck(self.lib.dvo_frame_release(self.ctx, slot))

    def frame_set_depth(self, slot, depth=None, sigma=None, age=None, full_res=False):
        if full_res:
            h, w = self.params.height, self.params.width
        else:
            w, h = self.level_size(self.L - 1)
        arrs = [self._stack(a, h, w) if a is not None else None for a in (depth, sigma, age)]
        self._ck(self.lib.dvo_frame_set_depth(self.ctx, slot, _p(arrs[0]), _p(arrs[1]), _p(arrs[2]), 1 if full_res else 0))

    def frame_download(self, slot, level, kind):
        w, h = self.level_size(level)
        out = np.empty((self.S, h, w), np.float32)
        self._ck(self.lib.dvo_frame_download(self.ctx, slot, level, IMG[kind], _p(out)))
        return out

    def visualize(self, slot, level, kind):
        """Draw::visualize* (src/core/draw.cpp:7-100) of one level: (S, h, w, 3) BGR bytes."""
        w, h = self.level_size(level)
        out = np.empty((self.S, h, w, 3), np.uint8)
        self._ck(self.lib.dvo_visualize(self.ctx, slot, level, VIS[kind], out.ctypes.data_as(C.c_void_p)))
        return out

    def visualize_image(self, kind, src, aux=None):
        """Draw::visualize*(image) on a raw host image: (h, w, 3) BGR bytes."""
        src = _f32(src); aux = _f32(aux) if aux is not None else None
        out = np.empty(src.shape + (3,), np.uint8)
        self._ck(self.lib.dvo_visualize_image(VIS[kind], _p(src), _p(aux), src.shape[1], src.shape[0], out.ctypes.data_as(C.c_void_p)))
        return out

    def inverse_pose(self, T):
        """Convert::inversePose (src/core/convert.cpp:31-39)."""
        T = _f32(np.asarray(T, np.float32).reshape(4, 4)); out = np.empty((4, 4), np.float32)
        self._ck(self.lib.dvo_inverse_pose(_p(T), _p(out)))
        return out

    def write_trajectory_csv(self, path, T, first_frame=0, append=False):
        T = _f32(np.asarray(T, np.float32).reshape(-1, 16))
        self._ck(self.lib.dvo_write_trajectory_csv(os.fsencode(path), _p(T), len(T), int(first_frame), 1 if append else 0))

    def write_png(self, path, bgr):
        bgr = np.ascontiguousarray(bgr, np.uint8)
        assert bgr.ndim == 3 and bgr.shape[2] == 3
        self._ck(self.lib.dvo_write_png(os.fsencode(path), bgr.ctypes.data_as(C.c_void_p), bgr.shape[1], bgr.shape[0]))

    def frame_set_xi(self, slot, xi=None, rel_xi=None):
        a = _f32(np.broadcast_to(np.asarray(xi, np.float32).reshape(-1, 6), (self.S, 6))) if xi is not None else None
        b = _f32(np.broadcast_to(np.asarray(rel_xi, np.float32).reshape(-1, 6), (self.S, 6))) if rel_xi is not None else None
        self._ck(self.lib.dvo_frame_set_xi(self.ctx, slot, _p(a), _p(b)))

    def frame_get_xi(self, slot):
        a = np.empty((self.S, 6), np.float32); b = np.empty((self.S, 6), np.float32)
        self._ck(self.lib.dvo_frame_get_xi(self.ctx, slot, _p(a), _p(b)))
        return a, b

    # ---- tracking
    def _xi(self, xi):
        return _f32(np.broadcast_to(np.asarray(xi, np.float32).reshape(-1, 6), (self.S, 6)))

    def warp_image(self, ref_slot, level, xi):
        w, h = self.level_size(level)
        out = np.empty((self.S, h, w), np.float32)
        x = self._xi(xi)
        self._ck(self.lib.dvo_warp_image(self.ctx, ref_slot, level, _p(x), _p(out)))
        return out

    def gn_step(self, obj_slot, ref_slot, level, xi, want_mask=True):
        w, h = self.level_size(level)
        x = self._xi(xi)
        upd = np.empty((self.S, 6), np.float32); res = np.empty(self.S, np.float32); valid = np.empty(self.S, np.int32)
        mask = np.empty((self.S, h, w), np.uint8) if want_mask else None
        self._ck(self.lib.dvo_gn_step(self.ctx, obj_slot, ref_slot, level, _p(x), _p(upd), _p(res),
                                      valid.ctypes.data_as(I32P), mask.ctypes.data_as(U8P) if want_mask else None))
        return upd, res, valid, mask

    def track(self, obj_slot, ref_slot, stop_update=-1.0, stop_residual=-1.0):
        cap = self.L * self.params.max_iterations
        xi = np.empty((self.S, 6), np.float32); n = np.empty(self.S, np.int32)
        log = (DvoIterLog * (cap * self.S))()
        self._ck(self.lib.dvo_track(self.ctx, obj_slot, ref_slot, _p(xi), log, cap, n.ctypes.data_as(I32P),
                                    C.c_double(stop_update), C.c_float(stop_residual)))
        logs = []
        for s in range(self.S):
            lst = []
            for k in range(int(n[s])):
                L = log[s * cap + k]
                lst.append(dict(level=L.level, iteration=L.iteration, valid_pixels=L.valid_pixels, residual=L.residual,
                                update=np.array(L.update[:], np.float32), xi_after=np.array(L.xi_after[:], np.float32)))
            logs.append(lst)
        return xi, logs

    # ---- mapping
    def map_propagate(self, ref_slot, new_slot, rel_xi):
        x = self._xi(rel_xi)
        self._ck(self.lib.dvo_map_propagate(self.ctx, ref_slot, new_slot, _p(x)))

    def map_update(self, obj_slot, ref_slot, history_slots, want_mask=True):
        w, h = self.level_size(self.L - 1)
        hs = np.asarray(history_slots, np.int32)
        vu = np.empty(self.S, np.int32)
        mask = np.empty((self.S, h, w), np.uint8) if want_mask else None
        self._ck(self.lib.dvo_map_update(self.ctx, obj_slot, ref_slot, hs.ctypes.data_as(I32P), len(hs),
                                         vu.ctypes.data_as(I32P), mask.ctypes.data_as(U8P) if want_mask else None))
        return vu, mask

    def map_regularize(self, slot):
        self._ck(self.lib.dvo_map_regularize(self.ctx, slot))

    def map_update_points(self, obj_slot, born_slot, relative_xi, px, py, depth, sigma):
        """Map::Implement::update for [S][n] points -> (depth, sigma), (-1, -1) on failure."""
        S = self.S
        px = np.ascontiguousarray(np.asarray(px, np.int32).reshape(S, -1)); py = np.ascontiguousarray(np.asarray(py, np.int32).reshape(S, -1))
        d = np.ascontiguousarray(np.asarray(depth, np.float32).reshape(S, -1)); sg = np.ascontiguousarray(np.asarray(sigma, np.float32).reshape(S, -1))
        n = px.shape[1]
        od = np.empty((S, n), np.float32); os_ = np.empty((S, n), np.float32)
        xi = self._xi(relative_xi)
        self._ck(self.lib.dvo_map_update_points(self.ctx, obj_slot, born_slot, _p(xi), n, px.ctypes.data_as(I32P), py.ctypes.data_as(I32P),
                                                _p(d), _p(sg), _p(od), _p(os_)))
        return od, os_

    def se3_exp(self, xi):
        xi = np.ascontiguousarray(np.asarray(xi, np.float32).reshape(-1, 6)); T = np.empty((len(xi), 4, 4), np.float32)
        self._ck(self.lib.dvo_se3_exp(self.ctx, _p(xi), _p(T), len(xi)))
        return T

    def se3_log(self, T):
        T = np.ascontiguousarray(np.asarray(T, np.float32).reshape(-1, 16)); xi = np.empty((len(T), 6), np.float32)
        self._ck(self.lib.dvo_se3_log(self.ctx, _p(T), _p(xi), len(T)))
        return xi

    def se3_concatenate(self, a, b):
        a = np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1, 6)); b = np.ascontiguousarray(np.asarray(b, np.float32).reshape(-1, 6))
        out = np.empty_like(a)
        self._ck(self.lib.dvo_se3_concatenate(self.ctx, _p(a), _p(b), _p(out), len(a)))
        return out

    def rng_position(self):
        pos = (C.c_uint64 * self.S)()
        self._ck(self.lib.dvo_rng_position(self.ctx, pos))
        return np.array(pos[:], np.uint64)

    def mask_compact(self, mask):
        m = np.ascontiguousarray(mask, np.uint8).reshape(-1)
        idx = np.empty(m.size, np.int32); cnt = C.c_int(0)
        self._ck(self.lib.dvo_mask_compact(self.ctx, m.ctypes.data_as(U8P), m.size, idx.ctypes.data_as(I32P), C.byref(cnt)))
        return idx[:cnt.value].copy()

    def mask_compact_noalloc(self, mask_dev, n, idx_dev, count_dev):
        """dvo_mask_compact_noalloc: device pointers in and out, no allocation inside; asynchronous (call sync())."""
        self._ck(self.lib.dvo_mask_compact_noalloc(self.ctx, mask_dev, n, idx_dev, count_dev))

    def select_pixels(self, slot, level, min_gradient):
        """Semi-dense selection by gradient threshold, compacted on the device: list of index arrays, one per stream."""
        w, h = self.level_size(level)
        idx = np.empty((self.S, w * h), np.int32); cnt = np.empty(self.S, np.int32)
        self._ck(self.lib.dvo_select_pixels(self.ctx, slot, level, float(min_gradient), idx.ctypes.data_as(I32P), cnt.ctypes.data_as(I32P)))
        return [idx[s, :cnt[s]].copy() for s in range(self.S)]

    # ---- VO
    def vo_reset(self):
        self._ck(self.lib.dvo_vo_reset(self.ctx))

    def vo_reserve(self, n_keyframes):
        self._ck(self.lib.dvo_vo_reserve(self.ctx, n_keyframes))

    def vo_push(self, gray, init_depth=None, init_sigma=None, want_pose=True):
        g = self._stack(gray, self.params.height, self.params.width)
        w, h = self.level_size(self.L - 1)
        d = self._stack(init_depth, h, w) if init_depth is not None else None
        s = self._stack(init_sigma, h, w) if init_sigma is not None else None
        T = np.empty((self.S, 4, 4), np.float32) if want_pose else None
        self._ck(self.lib.dvo_vo_push(self.ctx, g.ctypes.data_as(C.c_void_p), _p(d), _p(s), _p(T)))
        if not want_pose:
            self._ck(self.lib.dvo_sync(self.ctx))  # g must outlive the async copy
        return T

    def vo_push_gray8(self, gray8, invalid_code=-1, init_depth=None, init_sigma=None, want_pose=True):
        g = self._stack8(gray8)
        w, h = self.level_size(self.L - 1)
        d = self._stack(init_depth, h, w) if init_depth is not None else None
        s = self._stack(init_sigma, h, w) if init_sigma is not None else None
        T = np.empty((self.S, 4, 4), np.float32) if want_pose else None
        self._ck(self.lib.dvo_vo_push_gray8(self.ctx, g.ctypes.data_as(C.c_void_p), int(invalid_code), _p(d), _p(s), _p(T)))
        if not want_pose:
            self._ck(self.lib.dvo_sync(self.ctx))
        return T

    def vo_state(self):
        xi = np.empty((self.S, 6), np.float32); rel = np.empty((self.S, 6), np.float32)
        kf = np.empty(self.S, np.int32); nk = np.empty(self.S, np.int32); vu = np.empty(self.S, np.int32)
        ni = np.empty(self.S, np.int32)
        self._ck(self.lib.dvo_vo_state(self.ctx, _p(xi), _p(rel), kf.ctypes.data_as(I32P), nk.ctypes.data_as(I32P),
                                       vu.ctypes.data_as(I32P), ni.ctypes.data_as(I32P)))
        return dict(xi=xi, rel_xi=rel, was_keyframe=kf, n_keyframes=nk, valid_updates=vu, n_iters=ni)

    def vo_ref_slot(self, stream=0):
        s = C.c_int(-1)
        self._ck(self.lib.dvo_vo_ref_slot(self.ctx, stream, C.byref(s)))
        return s.value

    def vo_last_slot(self, stream=0):
        s = C.c_int(-1)
        self._ck(self.lib.dvo_vo_last_slot(self.ctx, stream, C.byref(s)))
        return s.value

    def vo_force_pose(self, rel_xi):
        x = self._xi(rel_xi) if rel_xi is not None else None
        self._ck(self.lib.dvo_vo_force_pose(self.ctx, _p(x)))

    def vo_last_log(self):
        cap = self.L * self.params.max_iterations
        n = np.empty(self.S, np.int32)
        log = (DvoIterLog * (cap * self.S))()
        self._ck(self.lib.dvo_vo_last_log(self.ctx, log, cap, n.ctypes.data_as(I32P)))
        return [[dict(level=log[s * cap + k].level, iteration=log[s * cap + k].iteration,
                      valid_pixels=log[s * cap + k].valid_pixels, residual=log[s * cap + k].residual,
                      update=np.array(log[s * cap + k].update[:], np.float32),
                      xi_after=np.array(log[s * cap + k].xi_after[:], np.float32)) for k in range(int(n[s]))]
                for s in range(self.S)]

    def vo_counters(self):
        arrs = [np.zeros(self.S, np.uint64) for _ in range(4)]
        self._ck(self.lib.dvo_vo_counters(self.ctx, *[a.ctypes.data_as(C.POINTER(C.c_uint64)) for a in arrs]))
        return dict(gn_iters=arrs[0], gn_pixels=arrs[1], n_propagate=arrs[2], n_update=arrs[3])

    def vo_matched(self):
        """Pixels matched and fused by Mapper::update since vo_reset, per stream (sum of valid_update)."""
        a = np.zeros(self.S, np.uint64)
        self._ck(self.lib.dvo_vo_matched(self.ctx, a.ctypes.data_as(C.POINTER(C.c_uint64))))
        return a

    def vo_error(self):
        return self.lib.dvo_vo_error(self.ctx)

    # ---- synthetic inputs / resident device buffers
    def dev_alloc(self, nbytes):
        p = self.lib.dvo_dev_alloc(nbytes)
        if not p:
            raise DvoError("dvo_dev_alloc(%d) failed" % nbytes)
        return p

    def dev_free(self, p):
        self.lib.dvo_dev_free(p)

    def synth_render_dev(self, gray_dev, depth_dev, frame, seed0=1, trans_step=0.004, rot_step_deg=0.15):
        self._ck(self.lib.dvo_synth_render_dev(self.ctx, gray_dev, depth_dev, frame, seed0, trans_step, rot_step_deg))

    def synth_render_gray8_dev(self, gray8_dev, depth_dev, frame, seed0=1, seed_stride=1, trans_step=0.004, rot_step_deg=0.15):
        self._ck(self.lib.dvo_synth_render_gray8_dev(self.ctx, gray8_dev, depth_dev, frame, seed0, seed_stride, trans_step, rot_step_deg))

    def copy_to_host(self, arr, dev_ptr):
        self._ck(self.lib.dvo_copy_to_host(self.ctx, arr.ctypes.data_as(C.c_void_p), dev_ptr, arr.nbytes))
        return arr

    def copy_to_dev(self, dev_ptr, arr):
        arr = np.ascontiguousarray(arr)
        self._ck(self.lib.dvo_copy_to_dev(self.ctx, dev_ptr, arr.ctypes.data_as(C.c_void_p), arr.nbytes))

    def vo_push_dev(self, gray_dev, init_depth=None, init_sigma=None, want_pose=False):
        w, h = self.level_size(self.L - 1)
        d = self._stack(init_depth, h, w) if init_depth is not None else None
        s = self._stack(init_sigma, h, w) if init_sigma is not None else None
        T = np.empty((self.S, 4, 4), np.float32) if want_pose else None
        self._ck(self.lib.dvo_vo_push_dev(self.ctx, gray_dev, _p(d), _p(s), _p(T)))
        return T

    def vo_push_gray8_dev(self, gray8_dev, invalid_code=-1, init_depth=None, init_sigma=None, want_pose=False):
        w, h = self.level_size(self.L - 1)
        d = self._stack(init_depth, h, w) if init_depth is not None else None
        s = self._stack(init_sigma, h, w) if init_sigma is not None else None
        T = np.empty((self.S, 4, 4), np.float32) if want_pose else None
        self._ck(self.lib.dvo_vo_push_gray8_dev(self.ctx, gray8_dev, int(invalid_code), _p(d), _p(s), _p(T)))
        return T

    def synth_pose(self, frame, seed, trans_step=0.004, rot_step_deg=0.15):
        """Ground-truth camera-to-world pose of `frame` of the synthetic sequence `seed` (4x4)."""
        T = np.empty(16, np.float32)
        self._ck(self.lib.dvo_synth_pose(self.ctx, int(frame), int(seed), trans_step, rot_step_deg, _p(T)))
        return T.reshape(4, 4)

    def synth_scene(self, seed):
        """The renderer's scene of `seed`: dict of fu, fv, ph, amp [5][8] and chk [5]."""
        out = np.empty(165, np.float32)
        self._ck(self.lib.dvo_synth_scene(self.ctx, int(seed), _p(out), out.size))
        a = out[:160].reshape(4, 5, 8)
        return dict(fu=a[0], fv=a[1], ph=a[2], amp=a[3], chk=out[160:165])

    # ---- measurement
    def event_record(self, idx):
        self._ck(self.lib.dvo_event_record(self.ctx, int(idx)))

    def event_elapsed(self, i0, i1):
        ms = C.c_float(0)
        self._ck(self.lib.dvo_event_elapsed(self.ctx, int(i0), int(i1), C.byref(ms)))
        return ms.value

    def sync(self):
        self._ck(self.lib.dvo_sync(self.ctx))

    def timer_start(self):
        self._ck(self.lib.dvo_timer_start(self.ctx))

    def timer_stop(self):
        ms = C.c_float(0)
        self._ck(self.lib.dvo_timer_stop(self.ctx, C.byref(ms)))
        return ms.value

    def launch_count(self):
        return int(self.lib.dvo_launch_count(self.ctx))

    def kernel_timing(self, enable):
        """enable=True starts timing every hot-path launch with CUDA events; enable=False stops and
        returns {kernel: (total_ms, launches)}."""
        out = (DvoKernelTime * 16)(); n = C.c_int(0)
        self._ck(self.lib.dvo_kernel_timing(self.ctx, 1 if enable else 0, out, 16, C.byref(n)))
        return {out[k].name.decode(): (out[k].total_ms, out[k].launches) for k in range(n.value)}

    def gn_timing(self, enable):
        ms = C.c_float(0); n = C.c_int(0)
        self._ck(self.lib.dvo_gn_timing(self.ctx, 1 if enable else 0, C.byref(ms), C.byref(n)))
        return ms.value, n.value
