#!/usr/bin/env python
"""bench.py — VO frames/s (track + map) of the B200 hot path, and the reference's CPU path beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json: "VO frames/sec (track+map) at 640x480", config C5 "batched multi-stream"
at main.cpp's geometry, configs[1]): every GPU runs `--streams` (256) independent synthetic 640x480
sequences in lock-step through System::VisualOdometry::odometrize semantics — Frame(gray, K, 3, 2),
Tracker::track, Mapper::estimate (needNewFrame -> propagate | update, regularize).  One STEP = one
new frame of every stream (dvo_vo_push over the whole batch).  Sequences are independent, so N GPUs
run N x 256 streams with no data-path collective ("scaling": "weak").

  value      frames/s with the frames already resident in HBM (dvo_vo_push_dev), CUDA events
  e2e        the same through dvo_vo_push with pinned HOST frames: the H2D copy of the step's input and
             the D2H read of the step's poses are inside the timed region
  roofline   the dominant kernel of the step (largest share of device time in a separate pass with CUDA
             events around every launch): algorithmic bytes / its device time;
             `kernels` lists the same for every hot-path kernel
  cpu_baseline  the reference's own sources (oracle/_ref) on the host cores, one sequence per core,
             with the C restatement (oracle/dvo_oracle.c) beside it
  cpu_baseline  the CPU oracle (oracle/, a C restatement of the reference; or oracle/_ref = the
             reference's own sources when built) on the host cores, one sequence per core

`--impl reference` times only the CPU arm (rank 0), same metric / config.
"""
import argparse
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "direct-visual-odometry_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

W_IN, H_IN, LEVELS, CULLS = 640, 480, 3, 2      # main.cpp: Frame(gray, K, 3, 2) on 640x480 (system.hpp:47)
TRANS_STEP, ROT_STEP_DEG = 0.004, 0.15          # per-frame motion of the synthetic trajectories
GN_BYTES_PER_PX = 16                            # depth, sigma, obj gray, ref gray read once


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = []; mx = []; reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            hi = [x for x in sm if x >= 0.5 * max(sm)]   # samples under load
            out.update(sm_mhz=statistics.median(hi), sm_max_mhz=max(mx), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (or oracle/_ref) on the host cores, one independent sequence per worker process
# ------------------------------------------------------------------------------------------------
def _build_cpu_lib():
    """Compile the CPU port's library ON THIS HOST (-march=native must match the cores it runs on)."""
    from oracle import oracle as O
    d = tempfile.mkdtemp(prefix="dvo_cpu_")
    return O.build(fast=True, out_dir=d)


def _ref_lib():
    """oracle/_ref/libdvo_ref_fast.so: the reference's OWN translation units (unmodified, compiled in
    the build container from /root/reference over the OpenCV-API shim) with its release optimisation
    level (-Ofast, CMakeLists.txt:43; -march=x86-64-v3 instead of native so the prebuilt file runs on
    the GPU box's host).  None when it was never built."""
    from oracle import ref as R
    try:
        return R.build(fast=True)
    except Exception:
        return None


def _cpu_worker(conn, frames, K, d0, s0, kind, lib_path):
    """frames: [n][h][w] float32 (shared, fork).  Protocol: ("run", first, count) -> seconds."""
    os.environ["OMP_NUM_THREADS"] = "1"   # one sequence per core; forEach stays serial inside a worker
    if kind == "reference":
        from oracle.ref import Ref
        vo = Ref(path=lib_path).vo_create(K)           # System::VisualOdometry(K), Frame(gray, K, 3, 2)
        vo.odometrize(frames[0], d0)
    else:
        from oracle.oracle import Oracle
        orc = Oracle(path=lib_path)   # reference release flags: -Ofast -march=native (CMakeLists.txt:43)
        vo = orc.vo_create(K, LEVELS, CULLS)
        vo.odometrize(frames[0], d0, s0)
    conn.send("ready")
    while True:
        msg = conn.recv()
        if msg[0] == "quit":
            break
        _, first, count = msg
        t0 = time.perf_counter()
        for i in range(first, first + count):
            vo.odometrize(frames[i])
        conn.send(time.perf_counter() - t0)


class CpuArm:
    """cores worker processes, each running the CPU VO on its own synthetic sequence."""

    def __init__(self, frames, K, d0, s0, kind="reference"):
        # frames: [cores][n][h][w]
        self.cores = frames.shape[0]
        ctx = mp.get_context("fork")
        lib_path = _ref_lib() if kind == "reference" else None
        if lib_path is None:   # oracle/_ref absent: fall back to the C restatement (oracle/dvo_oracle.c)
            kind, lib_path = "port", _build_cpu_lib()
        self.kind = kind
        self.pipes = []; self.procs = []
        for c in range(self.cores):
            a, b = ctx.Pipe()
            p = ctx.Process(target=_cpu_worker, args=(b, frames[c], K, d0[c], s0[c], kind, lib_path), daemon=True)
            p.start()
            self.pipes.append(a); self.procs.append(p)
        for a in self.pipes:
            assert a.recv() == "ready"

    def step(self, first, count):
        """every worker processes frames [first, first+count) of its sequence; returns wall seconds"""
        t0 = time.perf_counter()
        for a in self.pipes:
            a.send(("run", first, count))
        for a in self.pipes:
            a.recv()
        return time.perf_counter() - t0

    def close(self):
        for a in self.pipes:
            try:
                a.send(("quit",))
            except Exception:
                pass
        for p in self.procs:
            p.join(timeout=5)


# ------------------------------------------------------------------------------------------------
# inputs
# ------------------------------------------------------------------------------------------------
def default_K():
    return np.array([[0.82 * W_IN, 0, 0.5 * W_IN - 0.5], [0, 0.82 * W_IN, 0.5 * H_IN - 0.5], [0, 0, 1]], np.float32)


def depth_prior(n_streams, seed):
    """Scene's ctor prior (frame.hpp:18-21): max(randn(1.5, 0.5), 0.5), sigma 0.5 — supplied as an
    input to both arms, one independent draw per stream."""
    rng = np.random.default_rng(seed)
    h, w = H_IN >> CULLS, W_IN >> CULLS
    d = np.maximum(rng.normal(1.5, 0.5, size=(n_streams, h, w)), 0.5).astype(np.float32)
    return d, np.full_like(d, 0.5)


def render_host_frames(n_streams, n_frames, seed0):
    """[n_streams][n_frames][h][w] float32 frames for the CPU arm: rendered by the device renderer
    when a GPU is present (input generation only), else by the numpy renderer."""
    import dvo_b200
    out = np.empty((n_streams, n_frames, H_IN, W_IN), np.float32)
    lib = dvo_b200.load_library()
    if lib.dvo_device_count() > 0:
        dv = dvo_b200.Dvo(lib=lib, n_streams=n_streams, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS, K=default_K(),
                          max_frames=2, max_keyframes=1)
        buf = dv.dev_alloc(n_streams * H_IN * W_IN * 4)
        tmp = np.empty((n_streams, H_IN, W_IN), np.float32)
        for i in range(n_frames):
            dv.synth_render_dev(buf, None, i, seed0, TRANS_STEP, ROT_STEP_DEG)
            dv.copy_to_host(tmp, buf)
            out[:, i] = tmp
        dv.dev_free(buf)
        dv.close()
        return out, "device renderer"
    from dvo_b200 import synth
    for s in range(n_streams):
        g, _, _, _ = synth.sequence(n_frames, W_IN, H_IN, K=default_K(), seed=seed0 + s, trans_step=TRANS_STEP, rot_step_deg=ROT_STEP_DEG)
        out[s] = g
    return out, "numpy renderer"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_cpu_arm(steps, warmup, frames_per_step, cores=None, kind="reference"):
    """Times the CPU VO: every step, each of `cores` workers pushes frames_per_step frames of its own
    sequence.  Returns (frames_per_s, ms_per_step, info)."""
    cores = cores or host_cores()
    n_frames = 1 + (steps + warmup) * frames_per_step
    frames, how = render_host_frames(cores, n_frames, seed0=900000)
    d0, s0 = depth_prior(cores, seed=4321)
    arm = CpuArm(frames, default_K(), d0, s0, kind)
    pos = 1
    for _ in range(warmup):
        arm.step(pos, frames_per_step); pos += frames_per_step
    t = 0.0
    for _ in range(steps):
        t += arm.step(pos, frames_per_step); pos += frames_per_step
    arm.close()
    total = cores * frames_per_step * steps
    what = ("the reference's own sources (oracle/_ref, -Ofast)" if arm.kind == "reference" else
            "C restatement of the reference (oracle/dvo_oracle.c, -Ofast -march=native)")
    info = dict(kind=arm.kind, cores=cores,
                sample="%d sequences x %d frames of the 640x480 workload (one single-threaded process per core, %s; input: %s), "
                       "%d frames per step" % (cores, steps * frames_per_step, what, how, cores * frames_per_step))
    return total / t, 1e3 * t / steps, info


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the newest committed `ncu --set full` capture
    (profiles/*_traffic.json, written by scripts/ncu_summary.py traffic).  A captured launch is ONE frame of the same workload."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), reverse=True):
        try:
            k = json.load(open(path))["kernels"].get(kernel)
            if k:
                return k["dram_bytes"], os.path.basename(path), k
        except Exception:
            continue
    return None, None, None


def run_other_configs(lib):
    """The single-stream configurations of BASELINE.json (the metric also names 1080p): device-resident, CUDA events on the
    library's stream.  Reported beside the headline line, never instead of it.
      C3  synthetic 1920x1080 sequence, full VO loop: Frame(gray, K, 3, 2) with the reference's ROIs scaled x3, and culls = 0;
      C4  3840x2160 mapping-only: Mapper::update on dense hypotheses (rel_xi = (-0.1, 0, 0, 0, 0, 0)) + regularize."""
    import dvo_b200
    out = []
    peak, _ = peak_hbm()

    def Kf(w, h):
        return np.array([[0.82 * w, 0, 0.5 * w - 0.5], [0, 0.82 * w, 0.5 * h - 0.5], [0, 0, 1]], np.float32)

    # ---- C3
    W, H, n_res, n_frames = 1920, 1080, 40, 500
    for culls, sc in ((2, 3), (0, 12)):
        roi = dict(roi_x0=20 * sc, roi_x1=140 * sc, roi_y0=20 * sc, roi_y1=100 * sc,
                   map_roi_x0=16 * sc, map_roi_x1=144 * sc, map_roi_y0=12 * sc, map_roi_y1=108 * sc)
        dv = dvo_b200.Dvo(lib=lib, n_streams=1, width=W, height=H, levels=3, culls=culls, K=Kf(W, H), max_frames=2,
                          max_keyframes=n_frames + 8, rng_table=1 << 24, **roi)
        dv.vo_reserve(n_frames + 4)
        fb = W * H * 4
        dev = dv.dev_alloc(fb * n_res); dd = dv.dev_alloc(fb)
        for i in range(n_res):
            dv.synth_render_dev(dev + i * fb, dd if i == 0 else None, i, 1234, 0.002, 0.07)
        d0 = np.empty((1, H, W), np.float32)
        dv.copy_to_host(d0, dd)
        d0 = np.ascontiguousarray(d0[0, ::1 << culls, ::1 << culls])

        def fptr(i):
            per = 2 * (n_res - 1)
            j = i % per
            return dev + (j if j < n_res else per - j) * fb

        dv.vo_push_dev(dev, d0, np.full_like(d0, 0.1))
        for f in range(1, 4):
            dv.vo_push_dev(fptr(f))
        c0 = dv.vo_counters()
        dv.timer_start()
        for f in range(4, 4 + n_frames):
            dv.vo_push_dev(fptr(f))
        ms = dv.timer_stop()
        c1 = dv.vo_counters()
        err = dv.vo_error()
        it = int((c1["gn_iters"] - c0["gn_iters"]).sum())
        out.append(dict(workload="C3: synthetic 1920x1080, %d frames, 1 stream, Frame(gray,K,3,%d) -> %dx%d finest level, full VO loop"
                                 % (n_frames, culls, W >> culls, H >> culls),
                        value=n_frames / (ms * 1e-3), unit="frames/s", ms_per_frame=ms / n_frames, gn_iters_per_s=it / (ms * 1e-3),
                        gn_iters_per_frame=it / float(n_frames), frames_update=int((c1["n_update"] - c0["n_update"]).sum()),
                        frames_propagate=int((c1["n_propagate"] - c0["n_propagate"]).sum()), device_error=int(err)))
        dv.dev_free(dev); dv.dev_free(dd)
        dv.close()
    # ---- C4
    W, H = 3840, 2160
    roi = dict(map_roi_x0=0, map_roi_x1=W - 1, map_roi_y0=0, map_roi_y1=H - 1)
    dv = dvo_b200.Dvo(lib=lib, n_streams=1, width=W, height=H, levels=1, culls=0, K=Kf(W, H), max_frames=3, rng_table=1 << 24, **roi)
    fb = W * H * 4
    gd = dv.dev_alloc(fb)
    g = np.empty((2, 1, H, W), np.float32)
    for i in range(2):   # two views of the C3 scene (frames 0 and 12 of its trajectory)
        dv.synth_render_dev(gd, None, 12 * i, 1234, 0.002, 0.07)
        dv.copy_to_host(g[i], gd)
    dv.dev_free(gd)
    rng = np.random.default_rng(4321)
    depth = np.maximum(rng.normal(1.5, 0.5, (H, W)), 0.5).astype(np.float32)
    sigma = np.full((H, W), 0.5, np.float32); age = np.zeros((H, W), np.float32)
    ref = dv.frame_create(g[0][0]); obj = dv.frame_create(g[1][0])
    rel = np.array([-0.1, 0, 0, 0, 0, 0], np.float32); zero = np.zeros(6, np.float32)
    dv.frame_set_xi(ref, zero, zero); dv.frame_set_xi(obj, rel, rel)
    reps = 5
    tot = {}
    for r in range(reps + 1):
        dv.frame_set_depth(ref, depth, sigma, age)
        if r:
            dv.kernel_timing(True)
        vu, _ = dv.map_update(obj, ref, [ref], want_mask=False)
        dv.map_regularize(ref)
        if r:
            for k, (ms, n) in dv.kernel_timing(False).items():
                a = tot.setdefault(k, [0.0, 0]); a[0] += ms; a[1] += n
    dv.close()
    n = W * H
    algb = {"k_map_update": 40 * n, "k_map_update_apply": n, "k_regularize": 12 * n, "k_regularize_commit": 8 * n}
    ks = []
    for k, (ms, cnt) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        if cnt and k in algb:
            us = 1e3 * ms / cnt
            ks.append(dict(kernel=k, avg_launch_us=us, algorithmic_bytes_per_launch=algb[k], achieved_gbs=algb[k] / us / 1e3,
                           frac=algb[k] / us / 1e3 / peak))
    upd_ms = sum(v[0] for v in tot.values()) / reps
    out.append(dict(workload="C4: mapping-only 3840x2160, dense hypotheses (depth = max(N(1.5, 0.5), 0.5), sigma = 0.5), ROI = frame, "
                             "Mapper::update + regularize, %d repetitions" % reps,
                    value=1e3 / upd_ms, unit="frames/s", ms_per_frame=upd_ms, pixels_per_s=n / (upd_ms * 1e-3), valid_updates=int(vu[0]),
                    kernels=ks))
    return out


def run_gpu_arm(args, rank, world, local_rank, dist):
    import dvo_b200
    lib = dvo_b200.load_library()
    if lib.dvo_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device — libdvo_cuda has no CPU fallback")
    lib.dvo_set_device(local_rank)
    S, K_, W_ = args.streams, args.steps, args.warmup
    E = min(args.e2e_steps, K_)           # end-to-end steps (host frames)
    R = args.roofline_steps               # extra pass with events around every GN launch
    n_total = 1 + W_ + K_ + R             # frames pushed from HBM; frame 0 = first keyframe
    # Resident frames: the camera path is traversed forwards then backwards (still one continuous
    # motion), so a bounded set of distinct frames serves any number of steps.
    n_dev_frames = min(n_total, args.resident_frames)
    dv = dvo_b200.Dvo(lib=lib, n_streams=S, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS, K=default_K(), max_frames=2,
                      max_keyframes=n_total + W_ + E + 4)
    dv.vo_reserve(n_total + W_ + E + 2)
    frame_bytes = S * H_IN * W_IN * 4
    e2e_bgr = (args.e2e_input == "bgr8")
    e2e_u8 = (args.e2e_input == "gray8") or e2e_bgr
    esz = 3 if e2e_bgr else 1 if e2e_u8 else 4   # bytes per pixel of the end-to-end pass's host frames
    host_frame_bytes = S * H_IN * W_IN * esz
    h2d_bytes = S * (H_IN >> CULLS) * W_IN * esz if (CULLS > 0 and H_IN % (1 << CULLS) == 0) else host_frame_bytes  # stage_gray_to()
    if e2e_bgr:
        h2d_bytes = host_frame_bytes   # the undistortion gather can touch any source pixel: whole BGR frames cross PCIe
    # global sequence ids are sharded id mod world (dvo_b200/shard.py); a rank's S local streams are ids
    # rank, rank + world, ... and the renderer seeds stream k of this context with seed0 + k: the bench uses the
    # contiguous relabelling 1000 + rank * S + k (any bijection of independent sequences gives the same workload)
    seed0 = 1000 + rank * S
    # --- inputs: rendered on device, resident in HBM before the timed region starts
    dev = dv.dev_alloc(frame_bytes * n_dev_frames)
    for i in range(n_dev_frames):
        dv.synth_render_dev(dev + i * frame_bytes, None, i, seed0, TRANS_STEP, ROT_STEP_DEG)
    # host frames for the end-to-end pass: the frames that follow, in pinned memory
    import ctypes as C
    host = []
    # the e2e pass continues the motion from wherever the ping-pong traversal stops
    period = 2 * (n_dev_frames - 1)
    jl = (n_total - 1) % period if period else 0
    last = jl if jl < n_dev_frames else period - jl
    e2e_first = last + 1
    tmpd = dv.dev_alloc(frame_bytes)
    ftmp = np.empty((S, H_IN, W_IN), np.float32) if e2e_u8 else None
    for i in range(W_ + E):
        hp = lib.dvo_host_alloc(host_frame_bytes)
        if not hp:
            raise SystemExit("bench.py: pinned host allocation failed")
        dv.synth_render_dev(tmpd, None, e2e_first + i, seed0, TRANS_STEP, ROT_STEP_DEG)
        if e2e_u8:   # 8-bit camera frames: the rendered intensity quantised to 0..255 (what cv::imread + cvtColor deliver)
            arr = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(S, H_IN, W_IN, 3) if e2e_bgr else (S, H_IN, W_IN))
            dv.copy_to_host(ftmp, tmpd)
            np.multiply(ftmp, 255.0, out=ftmp); np.rint(ftmp, out=ftmp); np.clip(ftmp, 0, 255, out=ftmp)
            arr[...] = ftmp[..., None] if e2e_bgr else ftmp   # B = G = R: cvtColor returns the same code
        else:
            arr = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_float)), shape=(S, H_IN, W_IN))
            dv.copy_to_host(arr, tmpd)
        host.append((hp, arr))
    del ftmp
    dv.dev_free(tmpd)
    d0, s0 = depth_prior(S, seed=4321 + rank)
    dv.sync()

    def barrier():
        dv.sync()
        if dist is not None:
            dist.barrier()

    from dvo_b200 import shard

    def maxrank(x):  # slowest rank's time (device-timed), dvo_b200/shard.py
        return shard.max_over_ranks(x, dist, device="cuda")

    # --- first keyframe + warm-up
    def fptr(i):  # ping-pong over the resident frames
        period = 2 * (n_dev_frames - 1)
        j = i % period if period else 0
        return dev + (j if j < n_dev_frames else period - j) * frame_bytes

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    dv.vo_push_dev(dev, d0, s0)
    f = 1
    for _ in range(W_):
        dv.vo_push_dev(fptr(f)); f += 1
    c0 = dv.vo_counters()
    m0 = dv.vo_matched()
    # --- timed: K steps, inputs resident in HBM, CUDA events on the library's stream
    barrier()
    l0 = dv.launch_count()
    dv.timer_start()
    for _ in range(K_):
        dv.vo_push_dev(fptr(f)); f += 1
    ms = dv.timer_stop()
    l1 = dv.launch_count()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = maxrank(ms)
    c1 = dv.vo_counters()
    matched = int((dv.vo_matched() - m0).sum())
    if dv.vo_error() != 0:
        raise SystemExit("bench.py: device-side VO error")
    gn_iters = int((c1["gn_iters"] - c0["gn_iters"]).sum())
    n_prop = int((c1["n_propagate"] - c0["n_propagate"]).sum()); n_upd = int((c1["n_update"] - c0["n_update"]).sum())
    # --- roofline pass: CUDA events around every hot-path launch (adds event records to the stream, so it
    # is a separate pass); the dominant kernel = the largest share of the step's device time
    dv.kernel_timing(True)
    for _ in range(R):
        dv.vo_push_dev(fptr(f)); f += 1
    ktimes = dv.kernel_timing(False)
    c2 = dv.vo_counters()
    gn_px = int((c2["gn_pixels"] - c1["gn_pixels"]).sum())
    rp = int((c2["n_propagate"] - c1["n_propagate"]).sum()); ru = int((c2["n_update"] - c1["n_update"]).sum())
    n_top = (W_IN >> CULLS) * (H_IN >> CULLS)
    P = dv.params
    roi_n = (max(0, min(P.map_roi_x1, (W_IN >> CULLS) - 1) - max(P.map_roi_x0, 0) + 1) *
             max(0, min(P.map_roi_y1, (H_IN >> CULLS) - 1) - max(P.map_roi_y0, 0) + 1))
    # algorithmic bytes moved over the R timed steps (SURVEY §8-d / DESIGN.md §4), per kernel
    alg = {"k_cull_gray": (8 * n_top * S * R, "frame ingest: sampled gray in + finest level out, 8 B/px"),
           "k_track": (GN_BYTES_PER_PX * gn_px, "16 B x pixels swept by the Gauss-Newton iterations"),
           "k_propagate_scatter": (8 * n_top * rp, "depth+sigma in, 8 B/px of a propagating stream"),
           "k_propagate_gather": (16 * n_top * rp, "winner in + depth, sigma, age out, 16 B/px (propagate = 24 B/px in all)"),
           "k_map_update": (40 * roi_n * ru, "40 B x ROI pixels of an updating stream"),
           "k_map_update_apply": (1 * roi_n * ru, "flag byte per ROI pixel"),
           "k_regularize": (12 * n_top * (rp + ru), "depth, sigma in + depth out, 12 B/px"),
           "k_regularize_commit": (8 * n_top * (rp + ru), "depth copy-back, 8 B/px")}
    # --- end-to-end: host frames through dvo_vo_push, H2D of the input and D2H of the poses inside
    # (dvo_vo_submit / dvo_vo_collect: frame n+1's H2D copy overlaps frame n's kernels; every step's
    # poses are read back into pinned host memory and consumed by the host before the step counts)
    tps = [lib.dvo_host_alloc(S * 64) for _ in range(2)]
    Ts = [np.ctypeslib.as_array(C.cast(tp, C.POINTER(C.c_float)), shape=(S, 4, 4)) for tp in tps]
    pose_sum = 0.0
    def submit(i):
        if e2e_bgr:
            dv._ck(lib.dvo_vo_submit_bgr8(dv.ctx, host[i][0], 3, tps[i & 1]))
        elif e2e_u8:
            dv._ck(lib.dvo_vo_submit_gray8(dv.ctx, host[i][0], -1, tps[i & 1]))
        else:
            dv._ck(lib.dvo_vo_submit(dv.ctx, host[i][0], tps[i & 1]))

    dv.kernel_timing(True)   # the warm-up steps double as a device-time probe of THIS part of the sequence
    for i in range(W_):
        submit(i)
        dv._ck(lib.dvo_vo_collect(dv.ctx))
    e2e_kernel_ms = sum(v[0] for v in dv.kernel_timing(False).values()) / max(W_, 1)
    barrier()
    t0 = time.perf_counter()
    pending = []
    for i in range(W_, W_ + E):
        submit(i)
        pending.append(i & 1)
        if len(pending) == 2:
            dv._ck(lib.dvo_vo_collect(dv.ctx))
            pose_sum += float(Ts[pending.pop(0)][:, :3, 3].sum())
    while pending:
        dv._ck(lib.dvo_vo_collect(dv.ctx))
        pose_sum += float(Ts[pending.pop(0)][:, :3, 3].sum())
    dv.sync()
    e2e_s = maxrank(time.perf_counter() - t0)
    barrier()
    if not np.isfinite(pose_sum):
        raise SystemExit("bench.py: non-finite pose in the end-to-end pass")
    for tp in tps:
        lib.dvo_host_free(tp)
    for hp, _ in host:
        lib.dvo_host_free(hp)
    dv.dev_free(dev)
    dv.close()
    peak, peak_src = peak_hbm()
    tot_ms = sum(v[0] for v in ktimes.values()) or 1e-9
    table = []
    for name, (kms, kn) in sorted(ktimes.items(), key=lambda kv: -kv[1][0]):
        b, what = alg.get(name, (0, ""))
        gbs = b / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
        table.append(dict(kernel=name, launches=kn, avg_launch_us=1e3 * kms / kn if kn else None, share=kms / tot_ms,
                          algorithmic_bytes_per_launch=b / kn if kn else None, achieved_gbs=gbs, frac=gbs / peak, bytes_model=what))
    top = table[0]
    traffic, traffic_src, cap = ncu_traffic(top["kernel"])
    # the levels below HBM, from the same ncu capture (one launch early in a sequence, cold cache): what the kernel is bound by when
    # it is not DRAM - L2 traffic, and ncu's percent-of-peak for L2, L1/TEX and the SM
    hierarchy = None
    if cap and "lts_bytes" in cap:
        hierarchy = dict(source=traffic_src, capture_launch_us=cap["duration_us"], l2_bytes=cap["lts_bytes"],
                         l2_gbs=cap["lts_bytes"] / (cap["duration_us"] * 1e-6) / 1e9, l2_pct_of_peak=cap.get("lts_pct_of_peak"),
                         l2_hit_pct=cap.get("l2_hit_pct"), l1tex_pct_of_peak=cap.get("l1tex_pct_of_peak"), l1_hit_pct=cap.get("l1_hit_pct"),
                         sm_pct_of_peak=cap.get("sm_pct_of_peak"), dram_pct_of_measured_peak=100.0 * cap["dram_bytes"] / (cap["duration_us"] * 1e-6) / 1e9 / peak)
    res = dict(
        value=world * S * K_ / (ms * 1e-3), ms_per_step=ms / K_,
        e2e=dict(value=world * S * E / e2e_s, unit="frames/s", h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=S * 64, steps=E,
                 host_input=args.e2e_input, ms_per_step=1e3 * e2e_s / E, kernel_ms_per_step_same_frames=e2e_kernel_ms,
                 note="host frames are full 640x480 %s in pinned memory (%d B/step)%s; Convert::cullImage(gray, 2) samples every 4th row, "
                      "so only those rows cross PCIe (cudaMemcpy2DAsync); 2 frames in flight (%s/dvo_vo_collect); these are frames %d.. of the "
                      "sequences, where the epipolar searches are longer than over the first %d frames `value` averages "
                      "(kernel_ms_per_step_same_frames = device time of the same part of the sequence)"
                      % ("8-bit BGR (cvtColor + remap on the device too, whole frames cross PCIe)" if e2e_bgr else "8-bit gray" if e2e_u8 else "float32",
                         host_frame_bytes,
                         ", converted on the device like Loader's convertTo(CV_32FC1, 1.0/255.0) (loader.cpp:61)" if e2e_u8 else "",
                         "dvo_vo_submit_bgr8" if e2e_bgr else "dvo_vo_submit_gray8" if e2e_u8 else "dvo_vo_submit", e2e_first + W_, n_total)),
        gpu_launches=int(l1 - l0), clocks=clocks,
        roofline=dict(kernel=top["kernel"], bound="hbm", achieved=top["achieved_gbs"], peak=peak, unit="GB/s", frac=top["frac"],
                      traffic=traffic, traffic_source=traffic_src, peak_source=peak_src, launches=top["launches"], avg_launch_us=top["avg_launch_us"],
                      algorithmic_bytes_per_launch=top["algorithmic_bytes_per_launch"], share_of_step=top["share"],
                      bytes_model=top["bytes_model"], hierarchy=hierarchy),
        kernels=table,
        gn_iters_per_s=world * gn_iters / (ms * 1e-3), gn_iters_per_frame=gn_iters / float(S * K_),
        matched_pixels_per_s=world * matched / (ms * 1e-3),  # pixels the epipolar search matched and Mapper::update fused (rank 0's count x ranks)
        frames_propagate=n_prop, frames_update=n_upd)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=256, help="independent sequences per GPU")
    ap.add_argument("--e2e-steps", type=int, default=12)
    ap.add_argument("--e2e-input", default="gray8", choices=["gray8", "float32", "bgr8"],
                    help="host frames of the end-to-end pass: 8-bit gray (camera / cv::imread form, converted on the device) or float32")
    ap.add_argument("--roofline-steps", type=int, default=4)
    ap.add_argument("--resident-frames", type=int, default=48, help="distinct frames per stream kept in HBM")
    ap.add_argument("--cpu-frames-per-step", type=int, default=1)
    ap.add_argument("--cpu-kind", default="reference", choices=["reference", "port"],
                    help="CPU arm: the reference's own sources (oracle/_ref) or the C restatement (oracle/dvo_oracle.c)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the single-stream 1080p / 4K configurations (N = 1 only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)
    # stdout carries exactly ONE line, the JSON: whatever native libraries print on file descriptor 1 meanwhile (NCCL's "NCCL version ..."
    # banner ignores NCCL_DEBUG_FILE) is sent to stderr, and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    config = dict(workload="C5/configs[1]: %d independent synthetic 640x480 sequences per GPU, full VO loop (track + propagate|update + "
                           "regularize), Frame(gray,K,levels=3,culls=2), random depth prior (frame.hpp:18-21)" % args.streams,
                  streams_per_gpu=args.streams, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS,
                  l2="each step reads a new %.0f MB input batch (> 126 MB L2); no explicit flush" % (args.streams * W_IN * H_IN * 4 / 1e6))
    base = dict(metric="VO frames/sec (track+map) at 640x480", unit="frames/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic", config=config)

    if args.impl == "reference":
        if rank != 0:
            return
        fps, ms_step, info = run_cpu_arm(args.steps, args.warmup, args.cpu_frames_per_step, kind=args.cpu_kind)
        info["value"] = fps
        line = dict(base, impl="reference", value=fps, ms_per_step=ms_step, cpu_baseline=info, gpu_launches=0,
                    e2e=dict(value=fps, unit="frames/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        emit(line)
        return

    dist = None
    if world > 1:
        # stdout carries exactly ONE line (the JSON): NCCL's banner / debug lines ("NCCL version ...") go to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch
        import torch.distributed as td
        torch.cuda.set_device(local_rank)
        td.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = td
    res = run_gpu_arm(args, rank, world, local_rank, dist)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fps, ms_step, info = run_cpu_arm(steps=12, warmup=1, frames_per_step=args.cpu_frames_per_step, kind=args.cpu_kind)
        info["value"] = fps; info["unit"] = "frames/s"
        cpu = info
        if info["kind"] == "reference":   # the restatement beside it: what the same algorithm costs without per-pixel se3::exp
            fps2, _, info2 = run_cpu_arm(steps=12, warmup=1, frames_per_step=4, kind="port")
            cpu["port_value"] = fps2; cpu["port_sample"] = info2["sample"]
    other = None
    if rank == 0 and world == 1 and not args.no_other_configs:
        import dvo_b200
        other = run_other_configs(dvo_b200.load_library())
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        line = dict(base, impl="b200", **res)
        line["other_configs"] = other
        line["cpu_baseline"] = cpu
        emit(line)


if __name__ == "__main__":
    main()
