#!/usr/bin/env python
"""bench.py — VO frames/s (track + map) of the B200 hot path, and the reference's CPU path beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json: "VO frames/sec (track+map) at 640x480", config C5 "batched multi-stream" at main.cpp's geometry,
configs[1]): every GPU runs `--streams` (256) independent synthetic 640x480 sequences in lock-step through
System::VisualOdometry::odometrize semantics — Frame(gray, K, 3, 2), Tracker::track, Mapper::estimate (needNewFrame ->
propagate | update, regularize).  One STEP = `--frames-per-step` (32) consecutive frames of every stream, i.e. 8192 frames at
256 streams; K = 20 steps are 640 frames per stream, long enough for the keyframe history (never pruned, frame.hpp:179-185) and
the epipolar searches to reach their steady state.  Sequences are independent, so N GPUs run N x 256 streams with no data-path
collective ("scaling": "weak"); `strong_256` is BASELINE configs[4] as written (256 sequences in total, sharded id mod N).

  value         frames/s over the K timed steps with the 8-bit frames already resident in HBM (dvo_vo_push_gray8_dev), CUDA events
  value_steady  the same over the timed frames with index >= 250 only; value_early: frames 1 .. 100 (round 1's window)
  e2e           THE SAME frames of THE SAME sequences (dvo_vo_reset, replay) through dvo_vo_push_gray8 / dvo_vo_submit_gray8 with
                pinned HOST frames: the H2D copy of every frame and the D2H read of its poses are inside the timed region
  roofline      the dominant kernel of the step (largest share of device time in a separate pass with CUDA events around every
                launch, at the END of the sequences = steady state): algorithmic bytes / its device
                time; roofline.issue = the instruction-issue figures of the same kernel from the committed ncu capture of THIS
                source tree (profiles/traffic_<source hash>.json); `kernels` lists every hot-path kernel
  cpu_baseline  the reference's own sources (oracle/_ref) on the host cores, one sequence per core, bounded sample; the C
                restatement (oracle/dvo_oracle.c) and the single-thread figure beside it

`--impl reference` times only the CPU arm (rank 0), same metric / config; it never loads the CUDA library (inputs are rendered
by a child process).
"""
import argparse
import ctypes as C
import hashlib
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import tempfile
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
SEED_BASE = 1000
STEADY_FROM = 250                               # frames from here on count as steady state


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
# inputs
# ------------------------------------------------------------------------------------------------
def default_K():
    return np.array([[0.82 * W_IN, 0, 0.5 * W_IN - 0.5], [0, 0.82 * W_IN, 0.5 * H_IN - 0.5], [0, 0, 1]], np.float32)


def depth_prior(n_streams, seed):
    """Scene's ctor prior (frame.hpp:18-21): max(randn(1.5, 0.5), 0.5), sigma 0.5 — supplied as an input to both arms
    (SURVEY H6), one independent draw per stream."""
    rng = np.random.default_rng(seed)
    h, w = H_IN >> CULLS, W_IN >> CULLS
    d = np.maximum(rng.normal(1.5, 0.5, size=(n_streams, h, w)), 0.5).astype(np.float32)
    return d, np.full_like(d, 0.5)


def ring_index(i, n_ring):
    """Frame i of a sequence -> index into the ring of n_ring distinct rendered frames: the camera path is traversed forwards then
    backwards (still one continuous motion), so a bounded set of frames serves any number of steps."""
    period = 2 * (n_ring - 1)
    j = i % period if period else 0
    return j if j < n_ring else period - j


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def render_u8_main(path, n_streams, n_frames, seed0):
    """Child-process entry (`bench.py --render-u8 ...`): [n_streams][n_frames][h][w] uint8 frames of sequences seed0 .. from the
    device renderer when a GPU is present, else from the numpy renderer; written to `path` (.npy)."""
    out = np.lib.format.open_memmap(path, mode="w+", dtype=np.uint8, shape=(n_streams, n_frames, H_IN, W_IN))
    how = None
    try:
        import dvo_b200
        lib = dvo_b200.load_library()
        if lib.dvo_device_count() > 0:
            dv = dvo_b200.Dvo(lib=lib, n_streams=n_streams, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS, K=default_K(),
                              max_frames=2, max_keyframes=1)
            buf = dv.dev_alloc(n_streams * H_IN * W_IN)
            tmp = np.empty((n_streams, H_IN, W_IN), np.uint8)
            for i in range(n_frames):
                dv.synth_render_gray8_dev(buf, None, i, seed0, 1, TRANS_STEP, ROT_STEP_DEG)
                dv.copy_to_host(tmp, buf)
                out[:, i] = tmp
            dv.dev_free(buf)
            dv.close()
            how = "device renderer"
    except Exception as e:   # no library / no device: numpy
        log("render: device renderer unavailable (%s), using numpy" % e)
    if how is None:
        from dvo_b200 import synth
        for s in range(n_streams):
            g, _, _, _ = synth.sequence(n_frames, W_IN, H_IN, K=default_K(), seed=seed0 + s, trans_step=TRANS_STEP, rot_step_deg=ROT_STEP_DEG)
            out[s] = np.clip(np.rint(g * 255.0), 0, 255).astype(np.uint8)
        how = "numpy renderer"
    out.flush()
    print(how)


def render_host_frames_u8(n_streams, n_frames, seed0):
    """8-bit frames for the CPU arm, rendered in a CHILD process so that the arm's own process never maps libdvo_cuda.so."""
    d = tempfile.mkdtemp(prefix="dvo_in_")
    path = os.path.join(d, "frames.npy")
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--render-u8", path, str(n_streams), str(n_frames), str(seed0)],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise SystemExit("bench.py: input rendering failed:\n" + r.stderr[-2000:])
    how = r.stdout.strip().splitlines()[-1]
    frames = np.load(path, mmap_mode="r")
    return frames, how, path


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (or oracle/_ref) on the host cores, one independent sequence per worker process
# ------------------------------------------------------------------------------------------------
def _build_cpu_lib():
    """Compile the CPU port's library ON THIS HOST (-march=native must match the cores it runs on)."""
    from oracle import oracle as O
    d = tempfile.mkdtemp(prefix="dvo_cpu_")
    return O.build(fast=True, out_dir=d)


def _ref_lib():
    """oracle/_ref/libdvo_ref_fast.so: the reference's OWN translation units (unmodified, compiled in the build container from
    /root/reference over the OpenCV-API shim) with its release optimisation level (-Ofast, CMakeLists.txt:43; -march=x86-64-v3
    instead of native so the prebuilt file runs on the GPU box's host).  None when it was never built."""
    from oracle import ref as R
    try:
        return R.build(fast=True)
    except Exception:
        return None


def _to_float(u8):
    return (u8.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)   # Loader's convertTo(CV_32FC1, 1.0/255.0), loader.cpp:61


def _cpu_worker(conn, frames, K, d0, s0, kind, lib_path):
    """frames: [n][h][w] uint8 (shared, fork).  Protocol: ("run", first, count) -> seconds."""
    os.environ["OMP_NUM_THREADS"] = "1"   # one sequence per core; forEach stays serial inside a worker
    if kind == "reference":
        from oracle.ref import Ref
        vo = Ref(path=lib_path).vo_create(K)           # System::VisualOdometry(K), Frame(gray, K, 3, 2)
        vo.odometrize(_to_float(frames[0]), d0)
    else:
        from oracle.oracle import Oracle
        orc = Oracle(path=lib_path)   # reference release flags: -Ofast -march=native (CMakeLists.txt:43)
        vo = orc.vo_create(K, LEVELS, CULLS)
        vo.odometrize(_to_float(frames[0]), d0, s0)
    conn.send("ready")
    while True:
        msg = conn.recv()
        if msg[0] == "quit":
            break
        _, first, count = msg
        t0 = time.perf_counter()
        for i in range(first, first + count):
            vo.odometrize(_to_float(frames[i]))
        conn.send(time.perf_counter() - t0)
    conn.close()   # and return: the worker exits through the interpreter's normal shutdown


class CpuArm:
    """cores worker processes, each running the CPU VO on its own synthetic sequence."""

    def __init__(self, frames, K, d0, s0, kind="reference"):
        # frames: [cores][n][h][w] uint8
        self.cores = frames.shape[0]
        ctx = mp.get_context("fork")
        lib_path = _ref_lib() if kind == "reference" else None
        if lib_path is None:   # oracle/_ref absent: fall back to the C restatement (oracle/dvo_oracle.c)
            kind, lib_path = "port", _build_cpu_lib()
        self.kind = kind
        self.pipes = []; self.procs = []
        for c in range(self.cores):
            a, b = ctx.Pipe()
            p = ctx.Process(target=_cpu_worker, args=(b, frames[c], K, d0[c], s0[c], kind, lib_path))
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
            p.join(timeout=30)
            if p.is_alive():
                p.terminate()


def run_cpu_arm(steps, warmup, frames_per_step, cores=None, kind="reference"):
    """Times the CPU VO: every step, each of `cores` workers pushes frames_per_step frames of its own sequence.  Returns
    (frames_per_s, ms_per_step, info)."""
    cores = cores or host_cores()
    n_frames = 1 + (steps + warmup) * frames_per_step
    frames, how, path = render_host_frames_u8(cores, n_frames, seed0=900000)
    d0, s0 = depth_prior(cores, seed=4321)
    arm = CpuArm(frames, default_K(), d0, s0, kind)
    pos = 1
    for _ in range(warmup):
        arm.step(pos, frames_per_step); pos += frames_per_step
    t = 0.0
    for _ in range(steps):
        t += arm.step(pos, frames_per_step); pos += frames_per_step
    arm.close()
    try:
        del frames
        os.unlink(path); os.rmdir(os.path.dirname(path))
    except Exception:
        pass
    total = cores * frames_per_step * steps
    what = ("the reference's own sources (oracle/_ref, -Ofast) over oracle/shim's OpenCV stand-in" if arm.kind == "reference" else
            "C restatement of the reference (oracle/dvo_oracle.c, -Ofast -march=native)")
    info = dict(kind=arm.kind, cores=cores,
                sample="%d sequences x %d frames (frames %d..%d) of the 640x480 workload, one single-threaded process per core, %s; "
                       "input: 8-bit frames from the %s; %d frames per step" % (cores, steps * frames_per_step, 1 + warmup * frames_per_step,
                                                                                n_frames - 1, what, how, cores * frames_per_step))
    return total / t, 1e3 * t / steps, info


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def source_hash():
    """Identifies the CUDA source tree a profile belongs to (the .git directory does not travel to the GPU box)."""
    h = hashlib.sha256()
    for name in ("csrc/dvo_cuda.cu", "csrc/dvo_kernels.cuh", "csrc/dvo_device.cuh", "csrc/dvo_internal.cuh", "csrc/dvo_synth.cuh", "Makefile"):
        h.update(open(os.path.join(PKG, name), "rb").read())
    return h.hexdigest()[:12]


def ncu_capture(kernel):
    """One launch of `kernel` from the committed `ncu --set full` capture OF THIS SOURCE TREE: profiles/traffic_<source hash>.json,
    written by scripts/ncu_summary.py traffic.  (None, None, why) when there is none - said loudly, never replaced by an older build's."""
    name = "traffic_%s.json" % source_hash()
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        log("bench.py: NO ncu capture for this source tree (%s): roofline.traffic / roofline.issue are null. "
            "Capture one: see profiles/README.md" % name)
        return None, name, "no ncu capture of source tree %s is committed" % source_hash()
    k = json.load(open(path))["kernels"].get(kernel)
    return k, name, None if k else "kernel not in the capture"


def make_ring(dv, lib, S, n_ring, seed0, seed_stride):
    """n_ring distinct 8-bit frames of every stream: in pinned host memory (what the end-to-end pass feeds) and copied to HBM
    (what the device-resident pass feeds) - the SAME pixels for both."""
    fb = S * H_IN * W_IN
    dev = dv.dev_alloc(fb * n_ring)
    host = []
    for i in range(n_ring):
        dv.synth_render_gray8_dev(dev + i * fb, None, i, seed0, seed_stride, TRANS_STEP, ROT_STEP_DEG)
        hp = lib.dvo_host_alloc(fb)
        if not hp:
            raise SystemExit("bench.py: pinned host allocation failed")
        arr = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(S, H_IN, W_IN))
        dv.copy_to_host(arr, dev + i * fb)
        host.append(hp)
    return dev, host, fb


class Passes:
    """The two timed passes over one set of sequences: device-resident and end to end."""

    def __init__(self, lib, S, n_frames, n_ring, seed0, seed_stride, prior_seed, extra_frames=0):
        import dvo_b200
        self.lib, self.S, self.n_frames = lib, S, n_frames
        self.n_ring = min(n_ring, n_frames + extra_frames)
        self.dv = dvo_b200.Dvo(lib=lib, n_streams=S, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS, K=default_K(), max_frames=2,
                               max_keyframes=n_frames + extra_frames + 4, rng_table=1 << 24)   # (the table would grow on demand: not inside a timed region)
        # keyframe pool (FrameHistory is never pruned, frame.hpp:179-185): entry k is needed once some stream owns k keyframes - about
        # 0.36 per frame on these trajectories; reserved up front so that no allocation falls into a timed region (what is missing
        # would be allocated on first need)
        self.dv.vo_reserve(int(0.5 * (n_frames + extra_frames)) + 16)
        self.dev, self.host, self.fb = make_ring(self.dv, lib, S, self.n_ring, seed0, seed_stride)
        self.d0, self.s0 = depth_prior(S, prior_seed)
        self.dv.sync()

    def dev_frame(self, i):
        return self.dev + ring_index(i, self.n_ring) * self.fb

    def host_frame(self, i):
        return self.host[ring_index(i, self.n_ring)]

    def close(self):
        for hp in self.host:
            self.lib.dvo_host_free(hp)
        self.dv.dev_free(self.dev)
        self.dv.close()


def run_other_configs(lib):
    """The single-stream configurations of BASELINE.json (the metric also names 1080p; north_star's ">= 1000 frames/s" is a
    single-stream statement): device-resident, CUDA events on the library's stream.  Reported beside the headline line, never
    instead of it.
      S1  ONE 640x480 stream, full VO loop: the latency-bound configuration (frames/s = 1 / frame latency)
      C1  test/track.cpp on the bundled KINECT_50MM pair (tests/golden/bundled_golden.npz): Gauss-Newton iterations/s, us per
          iteration per level
      C2  main.cpp on the bundled data/logicool0 frames (tests/golden/bundled_long.npz, 150 frames), free-running, one stream
      C3  synthetic 1920x1080 sequence, full VO loop: Frame(gray, K, 3, 2) with the reference's ROIs scaled x3, and culls = 0,
          with the trajectory error against the renderer's ground truth
      C4  3840x2160 mapping-only: Mapper::update on dense hypotheses (rel_xi = (-0.1, 0, 0, 0, 0, 0)) + regularize."""
    import dvo_b200
    out = []
    peak, _ = peak_hbm()

    def Kf(w, h):
        return np.array([[0.82 * w, 0, 0.5 * w - 0.5], [0, 0.82 * w, 0.5 * h - 0.5], [0, 0, 1]], np.float32)

    # ---- S1: single stream 640x480
    n_frames, n_ring = 600, 48
    P = Passes(lib, 1, n_frames + 1, n_ring, 1234, 1, 4321)
    dv = P.dv
    dv.vo_push_gray8_dev(P.dev_frame(0), -1, P.d0, P.s0)
    for f in range(1, 9):
        dv.vo_push_gray8_dev(P.dev_frame(f), -1)
    c0 = dv.vo_counters()
    dv.timer_start()
    for f in range(9, 9 + n_frames - 8):
        dv.vo_push_gray8_dev(P.dev_frame(f), -1)
    ms = dv.timer_stop()
    c1 = dv.vo_counters()
    nf = n_frames - 8
    it = int((c1["gn_iters"] - c0["gn_iters"]).sum())
    out.append(dict(workload="S1: ONE synthetic 640x480 stream, %d frames, Frame(gray,K,3,2), full VO loop, device-resident 8-bit frames" % nf,
                    value=nf / (ms * 1e-3), unit="frames/s", ms_per_frame=ms / nf, gn_iters_per_s=it / (ms * 1e-3), gn_iters_per_frame=it / float(nf),
                    device_error=int(dv.vo_error())))
    P.close()
    # ---- C1 / C2: the bundled fixtures
    try:
        B = dict(np.load(os.path.join(ROOT, "tests", "golden", "bundled_golden.npz")))
        up = lambda a, k: np.ascontiguousarray(np.repeat(np.repeat(a, k, axis=0), k, axis=1))
        sg = np.where(B["c1_sigma10"] == 1, np.float32(0.1), np.float32(1.0)).astype(np.float32)
        dv = dvo_b200.Dvo(lib=lib, width=512, height=424, levels=4, culls=1, K=B["c1_K"])
        pre = dv.frame_create(up(B["c1_gray"][0], 2)); dv.frame_set_depth(pre, up(B["c1_depth"][0], 2), up(sg[0], 2), None, full_res=True)
        cur = dv.frame_create(up(B["c1_gray"][1], 2)); dv.frame_set_depth(cur, up(B["c1_depth"][1], 2), up(sg[1], 2), None, full_res=True)
        per_level = []
        for lv in range(4):   # one Track::optimize per launch (dvo_gn_step), event-timed over repetitions
            reps = 50
            x = np.zeros(6, np.float32)
            dv.gn_step(cur, pre, lv, x, want_mask=False)
            t0 = time.perf_counter()
            dv.kernel_timing(True)
            for _ in range(reps):
                dv.gn_step(cur, pre, lv, x, want_mask=False)
            kt = dv.kernel_timing(False)
            wall = (time.perf_counter() - t0) / reps
            w_, h_ = dv.level_size(lv)
            per_level.append(dict(level=lv, size="%dx%d" % (w_, h_), kernel_us_per_iteration=1e3 * kt["k_track"][0] / kt["k_track"][1],
                                  call_us_per_iteration=1e6 * wall))
        reps = 20
        dv.track(cur, pre)
        dv.kernel_timing(True)
        n_it = 0
        for _ in range(reps):
            _, logs = dv.track(cur, pre)
            n_it += len(logs[0])
        kt = dv.kernel_timing(False)
        out.append(dict(workload="C1: test/track.cpp on the bundled KINECT_50MM pair (entries 9 -> 10, 512x424, Frame(.,4,1): 32x26 .. 256x212), "
                                 "Tracker::track as one persistent launch; per level: one Track::optimize per launch",
                        value=n_it / (kt["k_track"][0] * 1e-3), unit="GN iterations/s", iterations_per_track=n_it / reps,
                        track_us=1e3 * kt["k_track"][0] / reps, per_level=per_level))
        dv.close()
        BL = dict(np.load(os.path.join(ROOT, "tests", "golden", "bundled_long.npz")))
        c = BL["l0_codes"]
        inv = np.broadcast_to(BL["l0_invalid"][None], c.shape)
        free = np.flatnonzero(np.bincount(c[~inv].ravel(), minlength=256) == 0)   # a code no valid pixel uses marks the undistortion border
        if len(free) == 0:
            raise FileNotFoundError("no free 8-bit code for the INVALID border")
        icode = int(free[-1])
        g8 = np.where(inv, icode, c).astype(np.uint8)
        n = len(g8)
        dv = dvo_b200.Dvo(lib=lib, width=640, height=480, levels=3, culls=2, K=BL["l0_K"], max_frames=2, max_keyframes=n + 2)
        dv.vo_reserve(n)
        full = np.ascontiguousarray(np.repeat(np.repeat(g8, 4, axis=1), 4, axis=2))   # the pixels Frame(., 3, 2) samples, at their input positions
        fb = 640 * 480
        dev = dv.dev_alloc(fb * n)
        dv.copy_to_dev(dev, full)
        d0 = BL["l0_init_depth"]
        dv.vo_push_gray8_dev(dev, icode, d0, np.full_like(d0, 0.5))
        c0 = dv.vo_counters()
        dv.timer_start()
        for i in range(1, n):
            dv.vo_push_gray8_dev(dev + i * fb, icode)
        ms = dv.timer_stop()
        c1 = dv.vo_counters()
        it = int((c1["gn_iters"] - c0["gn_iters"]).sum())
        st = dv.vo_state()
        out.append(dict(workload="C2: main.cpp's loop on the bundled data/logicool0 frames 0..%d (640x480 -> 160x120), one stream, free-running, "
                                 "device-resident 8-bit frames" % (n - 1),
                        value=(n - 1) / (ms * 1e-3), unit="frames/s", ms_per_frame=ms / (n - 1), gn_iters_per_s=it / (ms * 1e-3),
                        gn_iters_per_frame=it / float(n - 1), keyframes=int(st["n_keyframes"][0]), reference_keyframes=int(BL["l0_n_keyframes"][-1]),
                        device_error=int(dv.vo_error())))
        dv.dev_free(dev)
        dv.close()
    except FileNotFoundError as e:
        out.append(dict(workload="C1/C2: bundled fixtures", error=str(e)))
    # ---- C3
    W, H, n_res, n_frames = 1920, 1080, 40, 500
    for culls, sc in ((2, 3), (0, 12)):
        roi = dict(roi_x0=20 * sc, roi_x1=140 * sc, roi_y0=20 * sc, roi_y1=100 * sc,
                   map_roi_x0=16 * sc, map_roi_x1=144 * sc, map_roi_y0=12 * sc, map_roi_y1=108 * sc)
        dv = dvo_b200.Dvo(lib=lib, n_streams=1, width=W, height=H, levels=3, culls=culls, K=Kf(W, H), max_frames=2,
                          max_keyframes=n_frames + 8, rng_table=(1 << 24) if culls else (1 << 28), **roi)   # grows on demand otherwise: not while timing
        dv.vo_reserve(n_frames + 4)
        fb = W * H * 4
        dev = dv.dev_alloc(fb * n_res); dd = dv.dev_alloc(fb)
        for i in range(n_res):
            dv.synth_render_dev(dev + i * fb, dd if i == 0 else None, i, 1234, 0.002, 0.07)
        d0 = np.empty((1, H, W), np.float32)
        dv.copy_to_host(d0, dd)
        d0 = np.ascontiguousarray(d0[0, ::1 << culls, ::1 << culls])

        def fidx(i):
            return ring_index(i, n_res)

        dv.vo_push_dev(dev, d0, np.full_like(d0, 0.1))
        for f in range(1, 4):
            dv.vo_push_dev(dev + fidx(f) * fb)
        c0 = dv.vo_counters()
        dv.timer_start()
        for f in range(4, 4 + n_frames):
            dv.vo_push_dev(dev + fidx(f) * fb)
        ms = dv.timer_stop()
        c1 = dv.vo_counters()
        err = dv.vo_error()
        it = int((c1["gn_iters"] - c0["gn_iters"]).sum())
        # trajectory error against the renderer's ground truth: the first n_res frames are one forward pass of the camera path
        dv.vo_reset()
        errs = []; path = 0.0; prev = np.zeros(3)
        for f in range(n_res):
            T = dv.vo_push_dev(dev + f * fb, d0 if f == 0 else None, np.full_like(d0, 0.1) if f == 0 else None, want_pose=True)[0]
            gt = dv.synth_pose(f, 1234, 0.002, 0.07)
            path += float(np.linalg.norm(gt[:3, 3] - prev)); prev = gt[:3, 3].copy()
            errs.append(float(np.linalg.norm(T[:3, 3].astype(np.float64) - np.linalg.inv(gt.astype(np.float64))[:3, 3])))
        out.append(dict(workload="C3: synthetic 1920x1080, %d frames, 1 stream, Frame(gray,K,3,%d) -> %dx%d finest level, full VO loop"
                                 % (n_frames, culls, W >> culls, H >> culls),
                        value=n_frames / (ms * 1e-3), unit="frames/s", ms_per_frame=ms / n_frames, gn_iters_per_s=it / (ms * 1e-3),
                        gn_iters_per_frame=it / float(n_frames), frames_update=int((c1["n_update"] - c0["n_update"]).sum()),
                        frames_propagate=int((c1["n_propagate"] - c0["n_propagate"]).sum()), device_error=int(err),
                        vs_ground_truth=dict(frames=n_res, path_length_m=path, max_translation_error_m=max(errs), final_translation_error_m=errs[-1],
                                             note="ground-truth depth prior (sigma 0.1) on frame 0; se3::exp(frame->m_xi) is the world-to-camera "
                                                  "pose (main.cpp:50 inverts it): compared with T_wc(i)^-1 T_wc(0) of the renderer (dvo_synth_pose)")))
        dv.dev_free(dev); dv.dev_free(dd)
        dv.close()
    # ---- C4
    W, H = 3840, 2160
    roi = dict(map_roi_x0=0, map_roi_x1=W - 1, map_roi_y0=0, map_roi_y1=H - 1)
    dv = dvo_b200.Dvo(lib=lib, n_streams=1, width=W, height=H, levels=1, culls=0, K=Kf(W, H), max_frames=3, rng_table=1 << 26, **roi)
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


def bind_to_gpu_cpus(local_rank):
    """Pin this rank's host threads (and therefore its pinned allocations, first-touch) to the CPUs NVML reports as local to its
    GPU.  Returns a short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
            return "cpus %d-%d (%d, NVML affinity of GPU %d)" % (allowed[0], allowed[-1], len(allowed), local_rank)
    except Exception as e:
        return "unbound (%s)" % type(e).__name__
    return "unbound"


def timed_passes(P, W_, K_, F, barrier, maxrank, rank, local_rank, world, roofline_frames, do_roofline=True, do_e2e=True, sample_clocks=True):
    """Device-resident pass + (optionally) per-kernel pass + end-to-end replay over Passes P.  Returns a dict of raw figures."""
    dv, lib, S = P.dv, P.lib, P.S
    n_warm, n_timed = W_ * F, K_ * F
    last = n_warm + n_timed                      # frames 1 .. last are pushed after the first keyframe (frame 0)
    r = {}
    # ---- device-resident: frame 0 (first keyframe, with the prior), W warm-up steps, K timed steps, an event per step boundary
    dv.vo_push_gray8_dev(P.dev_frame(0), -1, P.d0, P.s0)
    dv.event_record(0)
    f = 1
    for k in range(W_):
        for _ in range(F):
            dv.vo_push_gray8_dev(P.dev_frame(f), -1); f += 1
        dv.event_record(1 + k)
    c0 = dv.vo_counters(); m0 = dv.vo_matched()
    barrier()
    # clocks / throttle reasons are sampled during the headline window; the short strong-scaling window (6-50 ms) is timed without the
    # nvidia-smi poller beside it and without the 0.3 s of idle GPU that starting it costs
    sampler = ClockSampler(local_rank)
    if rank == 0 and sample_clocks:
        sampler.start()
        time.sleep(0.3)
    l0 = dv.launch_count()
    dv.event_record(100)
    for k in range(K_):
        for _ in range(F):
            dv.vo_push_gray8_dev(P.dev_frame(f), -1); f += 1
        dv.event_record(101 + k)
    l1 = dv.launch_count()
    step_ms = [dv.event_elapsed(100 + k, 101 + k) for k in range(K_)]
    warm_ms = [dv.event_elapsed(k, k + 1) for k in range(W_)]
    barrier()
    r["clocks"] = sampler.stop() if rank == 0 and sample_clocks else None
    ms = maxrank(sum(step_ms))
    c1 = dv.vo_counters()
    r["matched"] = int((dv.vo_matched() - m0).sum())
    if dv.vo_error() != 0:
        raise SystemExit("bench.py: device-side VO error")
    r.update(ms=ms, step_ms=step_ms, warm_ms=warm_ms, launches=int(l1 - l0),
             gn_iters=int((c1["gn_iters"] - c0["gn_iters"]).sum()), n_prop=int((c1["n_propagate"] - c0["n_propagate"]).sum()),
             n_upd=int((c1["n_update"] - c0["n_update"]).sum()))
    # frames >= STEADY_FROM of the timed window, and frames 1..100 (the early part of the sequences) from the warm-up events
    steady = [(k, t) for k, t in enumerate(step_ms) if 1 + n_warm + k * F >= STEADY_FROM]
    r["steady_ms"] = maxrank(sum(t for _, t in steady)) if steady else None
    r["steady_frames"] = len(steady) * F
    early_steps = [k for k in range(W_) if 1 + (k + 1) * F - 1 <= 100]
    r["early_ms"] = maxrank(sum(warm_ms[k] for k in early_steps)) if early_steps else None
    r["early_frames"] = len(early_steps) * F
    # ---- per-kernel pass: CUDA events around every hot-path launch (adds event records to the stream and keeps the launches apart,
    # so Mapper::update is NOT overlapped with tracking here); the dominant kernel = the largest share of the step's device time
    if do_roofline:
        dv.kernel_timing(True)
        for _ in range(roofline_frames):
            dv.vo_push_gray8_dev(P.dev_frame(f), -1); f += 1
        r["ktimes"] = dv.kernel_timing(False)
        c2 = dv.vo_counters()
        r["roof_gn_px"] = int((c2["gn_pixels"] - c1["gn_pixels"]).sum())
        r["roof_prop"] = int((c2["n_propagate"] - c1["n_propagate"]).sum()); r["roof_upd"] = int((c2["n_update"] - c1["n_update"]).sum())
        r["roof_first_frame"] = f - roofline_frames
    # ---- end to end: dvo_vo_reset, then the SAME frames of the SAME sequences from pinned host memory.  Frame 0 goes through
    # dvo_vo_push_gray8 (it carries the prior), every later frame through dvo_vo_submit_gray8 / dvo_vo_collect: frame n+1's H2D copy
    # overlaps frame n's kernels; every frame's poses are read back into pinned host memory and consumed before it counts.
    if do_e2e:
        dv.vo_reset()
        tps = [lib.dvo_host_alloc(S * 64) for _ in range(2)]
        Ts = [np.ctypeslib.as_array(C.cast(tp, C.POINTER(C.c_float)), shape=(S, 4, 4)) for tp in tps]
        w, h = dv.level_size(LEVELS - 1)
        T0 = np.empty((S, 4, 4), np.float32)
        dv._ck(lib.dvo_vo_push_gray8(dv.ctx, P.host_frame(0), -1, P.d0.ctypes.data_as(C.POINTER(C.c_float)),
                                     P.s0.ctypes.data_as(C.POINTER(C.c_float)), T0.ctypes.data_as(C.POINTER(C.c_float))))
        pose_sum = 0.0

        def run(first, count):
            nonlocal pose_sum
            pending = []
            for i in range(first, first + count):
                dv._ck(lib.dvo_vo_submit_gray8(dv.ctx, P.host_frame(i), -1, tps[i & 1]))
                pending.append(i & 1)
                if len(pending) == 2:
                    dv._ck(lib.dvo_vo_collect(dv.ctx))
                    pose_sum += float(Ts[pending.pop(0)][:, :3, 3].sum())
            while pending:
                dv._ck(lib.dvo_vo_collect(dv.ctx))
                pose_sum += float(Ts[pending.pop(0)][:, :3, 3].sum())

        run(1, n_warm)
        dv.sync()
        barrier()
        t0 = time.perf_counter()
        run(1 + n_warm, n_timed)
        dv.sync()
        e2e_local = time.perf_counter() - t0
        r["e2e_s"] = maxrank(e2e_local)
        r["e2e_local_s"] = e2e_local
        barrier()
        if not np.isfinite(pose_sum):
            raise SystemExit("bench.py: non-finite pose in the end-to-end pass")
        if dv.vo_error() != 0:
            raise SystemExit("bench.py: device-side VO error in the end-to-end pass")
        for tp in tps:
            lib.dvo_host_free(tp)
    return r


def run_gpu_arm(args, rank, world, local_rank, dist):
    import dvo_b200
    from dvo_b200 import shard
    lib = dvo_b200.load_library()
    if lib.dvo_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device — libdvo_cuda has no CPU fallback")
    binding = bind_to_gpu_cpus(local_rank) if not args.no_bind else "unbound (--no-bind)"
    lib.dvo_set_device(local_rank)
    S, K_, W_, F = args.streams, args.steps, args.warmup, args.frames_per_step
    R = args.roofline_frames
    n_frames = 1 + (W_ + K_) * F

    def barrier():
        if dist is not None:
            dist.barrier()

    def maxrank(x):  # slowest rank's time, dvo_b200/shard.py
        return shard.max_over_ranks(x, dist, device="cuda")

    # global sequence ids are sharded id mod world (dvo_b200/shard.py).  The weak-scaling workload gives every rank S
    # sequences of its own: ids rank, rank + world, ... of world x S sequences; the renderer seeds sequence id with SEED_BASE + id.
    P = Passes(lib, S, n_frames, args.resident_frames, SEED_BASE + rank, world, 4321 + rank, extra_frames=R)
    r = timed_passes(P, W_, K_, F, lambda: (P.dv.sync(), barrier()), maxrank, rank, local_rank, world, R)
    n_top = (W_IN >> CULLS) * (H_IN >> CULLS)
    Pm = P.dv.params
    roi_n = (max(0, min(Pm.map_roi_x1, (W_IN >> CULLS) - 1) - max(Pm.map_roi_x0, 0) + 1) *
             max(0, min(Pm.map_roi_y1, (H_IN >> CULLS) - 1) - max(Pm.map_roi_y0, 0) + 1))
    P.close()
    h2d_bytes = S * (H_IN >> CULLS) * W_IN * F   # stage_gray_to(): only the rows Convert::cullImage samples cross PCIe
    # ---- BASELINE configs[4] as written: 256 sequences IN TOTAL, sharded across the N GPUs (strong scaling), 100 frames
    strong = None
    if not args.no_strong:
        ids = shard.shard_ids(args.strong_total, world, rank)
        Ss = len(ids)
        Fs, Ws, Ks = 10, 2, 8                                    # 1 + 100 frames: 20 warm-up + 80 timed
        Ps = Passes(lib, Ss, 1 + (Ws + Ks) * Fs, args.resident_frames, SEED_BASE + 500000 + rank, world, 8642 + rank)
        rs = timed_passes(Ps, Ws, Ks, Fs, lambda: (Ps.dv.sync(), barrier()), maxrank, rank, local_rank, world, 0, do_roofline=False, sample_clocks=False)
        Ps.close()
        strong = dict(total_sequences=args.strong_total, streams_per_gpu=Ss, frames_timed=Ks * Fs, frames_warmup=1 + Ws * Fs,
                      value=args.strong_total * Ks * Fs / (rs["ms"] * 1e-3), e2e=args.strong_total * Ks * Fs / rs["e2e_s"], unit="frames/s",
                      ms_per_frame_step=rs["ms"] / (Ks * Fs),
                      note="BASELINE configs[4]: %d independent sequences x 100 frames sharded id mod N; at small streams_per_gpu the tracking "
                           "kernel is bound by the latency of one stream's serial Gauss-Newton chain (cooperative split over CTAs when "
                           "streams < SM slots), not by throughput" % args.strong_total)
    peak, peak_src = peak_hbm()
    # ---- per-kernel table (steady-state frames)
    ktimes = r["ktimes"]
    gn_px, rp, ru = r["roof_gn_px"], r["roof_prop"], r["roof_upd"]
    alg = {"k_cull_gray": (5 * n_top * S * R, "frame ingest: sampled 8-bit gray in + finest level out, 5 B/px"),
           "k_track": (GN_BYTES_PER_PX * gn_px + 24 * n_top * rp, "16 B x pixels swept by the Gauss-Newton iterations + Implement::propagate "
                                                                  "of the streams that take a keyframe (24 B/px, in its epilogue)"),
           "k_propagate_scatter": (8 * n_top * rp, "depth+sigma in, 8 B/px of a propagating stream"),
           "k_propagate_gather": (16 * n_top * rp, "winner in + depth, sigma, age out, 16 B/px (propagate = 24 B/px in all)"),
           "k_map_update": (40 * roi_n * ru, "40 B x ROI pixels of an updating stream"),
           "k_map_update_apply": (1 * roi_n * ru, "flag byte per ROI pixel"),
           "k_regularize": (20 * n_top * (rp + ru), "depth, sigma in + depth out + Frame::updateDepth copy-back, 20 B/px"),
           "k_regularize_commit": (8 * n_top * (rp + ru), "depth copy-back, 8 B/px")}
    tot_ms = sum(v[0] for v in ktimes.values()) or 1e-9
    table = []
    for name, (kms, kn) in sorted(ktimes.items(), key=lambda kv: -kv[1][0]):
        b, what = alg.get(name, (0, ""))
        gbs = b / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
        table.append(dict(kernel=name, launches=kn, avg_launch_us=1e3 * kms / kn if kn else None, share=kms / tot_ms,
                          algorithmic_bytes_per_launch=b / kn if kn else None, achieved_gbs=gbs, frac=gbs / peak, bytes_model=what))
    top = table[0]
    cap, cap_name, cap_why = ncu_capture(top["kernel"])
    hierarchy = issue = None
    traffic = None
    if cap:
        traffic = cap["dram_bytes"]
        hierarchy = dict(source=cap_name, capture_launch_us=cap["duration_us"], l2_bytes=cap.get("lts_bytes"),
                         l2_gbs=cap["lts_bytes"] / (cap["duration_us"] * 1e-6) / 1e9 if cap.get("lts_bytes") else None,
                         l2_pct_of_peak=cap.get("lts_pct_of_peak"), l2_hit_pct=cap.get("l2_hit_pct"), l1tex_pct_of_peak=cap.get("l1tex_pct_of_peak"),
                         l1_hit_pct=cap.get("l1_hit_pct"), sm_pct_of_peak=cap.get("sm_pct_of_peak"),
                         dram_pct_of_measured_peak=100.0 * cap["dram_bytes"] / (cap["duration_us"] * 1e-6) / 1e9 / peak)
        if cap.get("warp_inst"):
            units = cap.get("units")   # ROI pixels of the updating streams in the captured launch, when the capture recorded them
            issue = dict(source=cap_name, ipc=cap.get("ipc_active"), ipc_peak=4.0, issue_slot_pct=cap.get("issue_active_pct"),
                         warps_active_pct=cap.get("warps_active_pct"), warp_inst_per_launch=cap["warp_inst"],
                         inst_per_unit=(32.0 * cap["warp_inst"] / units) if units else None,
                         unit="thread instructions per ROI pixel of an updating stream" if units else None,
                         note="instruction-issue-bound kernel: IPC per SM against 4 schedulers x 1 instruction/clock; stall reasons and "
                              "per-pipe utilisation of the same capture: profiles/issue_%s.txt" % source_hash())
    # the ceiling of the memory system for the dominant kernel's ACCESS PATTERN (random 128-byte line fills), measured with
    # tools/gather_peak.cu and committed; beside `frac` (algorithmic bytes), never instead of it
    pattern = None
    try:
        gp = json.load(open(os.path.join(ROOT, "profiles", "gather_peak.json")))
        if traffic and cap and top["kernel"] == "k_map_update":
            dram_gbs = traffic / (cap["duration_us"] * 1e-6) / 1e9
            pattern = dict(source="profiles/gather_peak.json", dram_gbs_of_capture=dram_gbs,
                           ceiling_line_walk_gbs=gp["line_walk_33_dram_gbs"], ceiling_random_taps_gbs=gp["random_taps_dram_gbs"],
                           frac_of_line_walk_ceiling=dram_gbs / gp["line_walk_33_dram_gbs"],
                           frac_of_random_ceiling=dram_gbs / gp["random_taps_dram_gbs"],
                           frac_of_hbm_peak_by_traffic=dram_gbs / peak, note=gp["note"])
    except (OSError, KeyError, ValueError):
        pattern = None
    total_frames = world * S * K_ * F
    res = dict(
        value=total_frames / (r["ms"] * 1e-3), ms_per_step=r["ms"] / K_, ms_per_frame_step=r["ms"] / (K_ * F),
        value_steady=(world * S * r["steady_frames"] / (r["steady_ms"] * 1e-3)) if r["steady_ms"] else None,
        value_early=(world * S * r["early_frames"] / (r["early_ms"] * 1e-3)) if r["early_ms"] else None,
        windows=dict(value="frames %d..%d of every sequence (%.2f s of device time)" % (1 + W_ * F, (W_ + K_) * F, r["ms"] * 1e-3),
                     value_steady="frames >= %d of the timed window (%d frames per sequence)" % (STEADY_FROM, r["steady_frames"]),
                     value_early="frames 1..%d (warm-up steps, not part of `value`)" % r["early_frames"],
                     ms_per_frame_by_step=[t / F for t in r["step_ms"]]),
        e2e=dict(value=total_frames / r["e2e_s"], unit="frames/s", h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=S * 64 * F, steps=K_,
                 host_input="gray8", ms_per_step=1e3 * r["e2e_s"] / K_, ms_per_frame_step=1e3 * r["e2e_s"] / (K_ * F),
                 h2d_gbs_this_rank=h2d_bytes * K_ / r["e2e_local_s"] / 1e9, host_binding=binding, frames_in_flight=2,
                 note="the SAME %d frames of the SAME sequences as `value` (dvo_vo_reset + replay), from full 640x480 8-bit gray frames in "
                      "pinned host memory (%d B/frame-step), converted on the device like Loader's convertTo(CV_32FC1, 1.0/255.0) (loader.cpp:61); "
                      "Convert::cullImage(gray, 2) samples every 4th row, so only those rows cross PCIe (cudaMemcpy2DAsync); frame n+1's copy "
                      "overlaps frame n's kernels (dvo_vo_submit_gray8 / dvo_vo_collect), every frame's poses are read back"
                      % (K_ * F, S * H_IN * W_IN)),
        gpu_launches=r["launches"], launches_per_frame=r["launches"] / float(K_ * F), clocks=r["clocks"],
        roofline=dict(kernel=top["kernel"], bound="hbm", achieved=top["achieved_gbs"], peak=peak, unit="GB/s", frac=top["frac"],
                      traffic=traffic, traffic_source=cap_name if cap else None, traffic_note=cap_why, peak_source=peak_src,
                      launches=top["launches"], avg_launch_us=top["avg_launch_us"],
                      algorithmic_bytes_per_launch=top["algorithmic_bytes_per_launch"], share_of_step=top["share"],
                      bytes_model=top["bytes_model"], hierarchy=hierarchy, issue=issue, pattern_ceiling=pattern,
                      measured_on="frames %d..%d of the sequences (steady state), launches kept apart by the event records"
                                  % (r["roof_first_frame"], r["roof_first_frame"] + R - 1)),
        kernels=table,
        gn_iters_per_s=world * r["gn_iters"] / (r["ms"] * 1e-3), gn_iters_per_frame=r["gn_iters"] / float(S * K_ * F),
        matched_pixels_per_s=world * r["matched"] / (r["ms"] * 1e-3),  # pixels the epipolar search matched and Mapper::update fused (rank 0's count x ranks)
        frames_propagate=r["n_prop"], frames_update=r["n_upd"], strong_256=strong, source_hash=source_hash())
    return res


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--render-u8":
        render_u8_main(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=256, help="independent sequences per GPU")
    ap.add_argument("--frames-per-step", type=int, default=32, help="consecutive frames of every stream in one step")
    ap.add_argument("--roofline-frames", type=int, default=6)
    ap.add_argument("--resident-frames", type=int, default=48, help="distinct frames per stream kept in HBM / pinned host memory")
    ap.add_argument("--strong-total", type=int, default=256, help="sequences of the strong-scaling pass (BASELINE configs[4])")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-bind", action="store_true", help="do not pin the rank to its GPU's local CPUs")
    ap.add_argument("--cpu-frames-per-step", type=int, default=1)
    ap.add_argument("--cpu-kind", default="reference", choices=["reference", "port"],
                    help="CPU arm: the reference's own sources (oracle/_ref) or the C restatement (oracle/dvo_oracle.c)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the single-stream / 1080p / 4K / bundled configurations (N = 1 only)")
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
    F = args.frames_per_step
    config = dict(workload="C5/configs[1]: %d independent synthetic 640x480 sequences per GPU, full VO loop (track + propagate|update + "
                           "regularize), Frame(gray,K,levels=3,culls=2), random depth prior (frame.hpp:18-21); one step = %d consecutive "
                           "frames of every sequence" % (args.streams, F),
                  streams_per_gpu=args.streams, frames_per_step=F, width=W_IN, height=H_IN, levels=LEVELS, culls=CULLS,
                  l2="no explicit flush: a frame-step reads a new 78.6 MB batch of 8-bit frames out of a 3.8 GB ring and gathers from "
                     ">= 0.5 GB of keyframe images per step (growing with the history) - far beyond the 126 MB L2")
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
        if info["kind"] == "reference":
            # "1 thread and all host cores" - the single-thread figure of the reference ...
            fps1, _, info1 = run_cpu_arm(steps=6, warmup=1, frames_per_step=1, cores=1, kind=args.cpu_kind)
            cpu["one_thread_value"] = fps1; cpu["one_thread_sample"] = info1["sample"]
            # ... and the restatement beside it: what the same algorithm costs without per-pixel se3::exp
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
