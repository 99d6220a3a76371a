"""Synthetic rendered sequences: a textured box room ray-cast analytically with
numpy — no graphics API.  Gives float32 gray in [0,1], ground-truth depth and a smooth 6-DoF
trajectory.  Used by tests and bench.py to make inputs; it is not part of the hot path.
"""
import numpy as np


def _rodrigues(w):
    th = float(np.linalg.norm(w))
    if th < 1e-12:
        return np.eye(3)
    k = np.asarray(w, np.float64) / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


class Scene:
    """Axis-aligned room: back wall z=zb, side walls x=+-xw, floor/ceiling y=+-yw."""

    def __init__(self, seed=1234, zb=2.6, xw=1.9, yw=1.4, n_waves=10):
        rng = np.random.default_rng(seed)
        self.zb, self.xw, self.yw = zb, xw, yw
        self.planes = []
        for _ in range(5):
            freq = rng.uniform(1.5, 9.0, size=(n_waves, 2)) * rng.choice([-1, 1], size=(n_waves, 2))
            phase = rng.uniform(0, 2 * np.pi, size=n_waves)
            amp = rng.uniform(0.3, 1.0, size=n_waves)
            amp /= amp.sum()
            self.planes.append((freq, phase, amp, rng.uniform(0.8, 2.0)))

    def _tex(self, k, u, v):
        freq, phase, amp, chk = self.planes[k]
        t = np.zeros_like(u)
        for (fu, fv), ph, a in zip(freq, phase, amp):
            t += a * np.sin(fu * u + fv * v + ph)
        checker = np.sign(np.sin(chk * np.pi * u) * np.sin(chk * np.pi * v))
        return np.clip(0.5 + 0.33 * t + 0.10 * checker, 0.02, 0.98)

    def render(self, K, T_wc, w, h):
        """T_wc: camera-to-world 4x4.  Returns (gray[h,w] f32, depth[h,w] f32 = z in the camera frame)."""
        K = np.asarray(K, np.float64).reshape(3, 3)
        R = np.asarray(T_wc, np.float64)[:3, :3]
        t = np.asarray(T_wc, np.float64)[:3, 3]
        xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
        dc = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1)
        dw = dc @ R.T
        best = np.full((h, w), np.inf)
        gray = np.zeros((h, w))
        cand = [(2, self.zb, 0), (0, self.xw, 1), (0, -self.xw, 2), (1, self.yw, 3), (1, -self.yw, 4)]
        for axis, val, k in cand:
            with np.errstate(divide="ignore", invalid="ignore"):
                lam = (val - t[axis]) / dw[..., axis]
            P = t + lam[..., None] * dw
            ok = (lam > 1e-6) & (lam < best)
            ok &= (np.abs(P[..., 0]) <= self.xw + 1e-9) & (np.abs(P[..., 1]) <= self.yw + 1e-9) & (P[..., 2] <= self.zb + 1e-9)
            if axis == 2:
                u, v = P[..., 0], P[..., 1]
            elif axis == 0:
                u, v = P[..., 2], P[..., 1]
            else:
                u, v = P[..., 0], P[..., 2]
            tex = self._tex(k, u, v)
            gray = np.where(ok, tex, gray)
            best = np.where(ok, lam, best)
        depth = best  # dc has unit z, so lambda == z in the camera frame
        return gray.astype(np.float32), depth.astype(np.float32)


def trajectory(n, seed=1234, trans_step=0.004, rot_step_deg=0.15):
    """Smooth camera-to-world poses: low-frequency sinusoids in all 6 DoF (incl. roll/pitch/yaw)."""
    rng = np.random.default_rng(seed + 77)
    f = rng.uniform(0.02, 0.08, size=6)
    ph = rng.uniform(0, 2 * np.pi, size=6)
    amp_t = trans_step / (2 * np.pi * f[:3])
    amp_r = np.deg2rad(rot_step_deg) / (2 * np.pi * f[3:])
    poses = []
    for i in range(n):
        tr = amp_t * (np.sin(2 * np.pi * f[:3] * i + ph[:3]) - np.sin(ph[:3]))
        ro = amp_r * (np.sin(2 * np.pi * f[3:] * i + ph[3:]) - np.sin(ph[3:]))
        T = np.eye(4)
        T[:3, :3] = _rodrigues(ro)
        T[:3, 3] = tr
        poses.append(T)
    return poses


def sequence(n, w=640, h=480, K=None, seed=1234, **kw):
    """n frames of one synthetic sequence: (gray[n,h,w], depth[n,h,w], poses[n,4,4], K)."""
    if K is None:
        K = np.array([[0.82 * w, 0, 0.5 * w - 0.5], [0, 0.82 * w, 0.5 * h - 0.5], [0, 0, 1]], np.float32)
    sc = Scene(seed)
    poses = trajectory(n, seed, **kw)
    g = np.empty((n, h, w), np.float32)
    d = np.empty((n, h, w), np.float32)
    for i, T in enumerate(poses):
        g[i], d[i] = sc.render(K, T, w, h)
    return g, d, np.stack(poses).astype(np.float32), np.asarray(K, np.float32)


def render_device_scene(scene, K, T_wc, w, h):
    """The DEVICE renderer's image formation (csrc/dvo_synth.cuh: k_synth_render) restated in float64 numpy from the scene
    parameters dvo_synth_scene returns and a camera-to-world pose: (gray[h,w], depth[h,w]).  Used only to check the kernel."""
    K = np.asarray(K, np.float64).reshape(3, 3)
    R = np.asarray(T_wc, np.float64)[:3, :3]
    t = np.asarray(T_wc, np.float64)[:3, 3]
    zb, xw, yw = 2.6, 1.9, 1.4
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    dc = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1)
    dw = dc @ R.T
    best = np.full((h, w), 1e30)
    gray = np.zeros((h, w))
    hit = np.zeros((h, w), bool)
    for p, (axis, val) in enumerate([(2, zb), (0, xw), (0, -xw), (1, yw), (1, -yw)]):
        with np.errstate(divide="ignore", invalid="ignore"):
            lam = (val - t[axis]) / dw[..., axis]
        P = t + lam[..., None] * dw
        ok = (lam > 1e-6) & (lam < best)
        ok &= ~((np.abs(P[..., 0]) > xw + 1e-4) | (np.abs(P[..., 1]) > yw + 1e-4) | (P[..., 2] > zb + 1e-4))
        if axis == 2:
            u, v = P[..., 0], P[..., 1]
        elif axis == 0:
            u, v = P[..., 2], P[..., 1]
        else:
            u, v = P[..., 0], P[..., 2]
        tsum = np.zeros((h, w))
        for k in range(scene["fu"].shape[1]):
            tsum += scene["amp"][p, k] * np.sin(scene["fu"][p, k] * u + scene["fv"][p, k] * v + scene["ph"][p, k])
        ck = np.sin(scene["chk"][p] * np.pi * u) * np.sin(scene["chk"][p] * np.pi * v)
        g = np.clip(0.5 + 0.33 * (tsum / scene["amp"][p].sum()) + np.where(ck > 0, 0.10, -0.10), 0.02, 0.98)
        gray = np.where(ok, g, gray)
        best = np.where(ok, lam, best)
        hit |= ok
    return gray.astype(np.float32), np.where(hit, best, 0.0).astype(np.float32)


def random_depth_prior(shape, seed=4321):
    """Stand-in for Scene's cv::randn(1.5, 0.5) clipped below at 0.5 and sigma = 0.5
    (include/system/frame.hpp:18-21): same distribution, NumPy PCG64 stream."""
    rng = np.random.default_rng(seed)
    d = np.maximum(rng.normal(1.5, 0.5, size=shape), 0.5).astype(np.float32)
    s = np.full(shape, 0.5, np.float32)
    return d, s
