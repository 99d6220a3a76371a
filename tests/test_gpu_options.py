"""GPU tests of what round 2 added around the hot path:
  * the default-off extras north_star names and the reference lacks: variance-normalised Huber weight and semi-dense
    selection by gradient threshold - the SAME definition on the oracle (orc_params) and the GPU (dvo_params), so parity holds with
    them switched on; dvo_select_pixels (selection + compaction on the device);
  * robustness of the C ABI: a recycled frame slot starts clean, the first VO frame needs its prior, the mt19937 replay
    table grows on demand, a history overflow is reported by the call that reads the poses."""
import numpy as np
import pytest

from conftest import make_dvo
from helpers import oracle_frame_pair, oracle_gn_inputs, se3_norm
from test_gpu_vo import _run_pair

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("opts", [dict(huber_delta=1.0), dict(min_gradient=0.08), dict(huber_delta=0.5, min_gradient=0.04)])
def test_extras_match_oracle_single_step(orc, seq160, opts):
    g, d, poses, K = seq160
    sigma = np.full_like(d[0], 0.2)
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, **opts)
    rs = dv.frame_create(g[0]); dv.frame_set_depth(rs, d[0], sigma, None, full_res=True)
    os_ = dv.frame_create(g[3])
    ref, obj = oracle_frame_pair(orc, g[0], g[3], K, 3, 2, d[0], sigma)
    P = orc.params(**opts)
    Pn = orc.params()
    changed = False
    for level in range(3):
        for xi in (np.zeros(6, np.float32), np.array([0.01, -0.004, 0.006, 0.002, -0.003, 0.001], np.float32)):
            o_upd, o_res, o_valid, o_mask = orc.optimize(*oracle_gn_inputs(orc, ref, obj, level, xi), xi, level, params=P)
            n_upd, _, n_valid, _ = orc.optimize(*oracle_gn_inputs(orc, ref, obj, level, xi), xi, level, params=Pn)
            upd, res, valid, mask = dv.gn_step(os_, rs, level, xi)
            assert int(valid[0]) == o_valid and np.array_equal(mask[0], o_mask)
            assert se3_norm(upd[0], o_upd) <= 1e-5
            assert abs(float(res[0]) - o_res) <= 1e-5 * max(1.0, abs(o_res))
            changed |= (n_valid != o_valid) or se3_norm(n_upd, o_upd) > 1e-4
    assert changed, "the options must change the step (otherwise this test proves nothing)"
    dv.close()


def test_extras_match_oracle_vo_run(orc):
    from dvo_b200 import synth
    g, d, poses, K = synth.sequence(8, 640, 480, seed=21, trans_step=0.01, rot_step_deg=0.25)
    d0 = orc.cull_image(d[0], 2)
    s0 = np.full_like(d0, 0.3)
    n_kf, n_upd, _ = _run_pair(orc, g, K, d0, s0, huber_delta=1.2, min_gradient=0.03)
    assert n_kf + n_upd == 8


def test_select_pixels_matches_numpy(orc, seq160):
    g, d, poses, K = seq160
    from helpers import punch_invalid
    gray = np.stack([punch_invalid(g[0], seed=5), g[1]])
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, n_streams=2)
    slot = dv.frame_create(gray)
    for level in (0, 2):
        for thr in (0.0, 0.05, 0.15):
            got = dv.select_pixels(slot, level, thr)
            for s in range(2):
                fr = orc.frame_create(gray[s], K, 3, 2)
                I = fr.image(level, "gray"); gx = fr.image(level, "gradx"); gy = fr.image(level, "grady")
                ok = (I > -2.0) & (gx > -2.0) & (gy > -2.0)
                with np.errstate(all="ignore"):
                    m2 = (gx * gx + gy * gy).astype(np.float32)
                ok &= ~(m2 < np.float32(thr) * np.float32(thr))
                assert np.array_equal(got[s], np.flatnonzero(ok.ravel()).astype(np.int32)), (level, thr, s)
    # and the plain compaction utility on a large mask (linear-time scan over 40 000 CTAs)
    rng = np.random.default_rng(1)
    m = (rng.random(10_000_000) < 0.3).astype(np.uint8)
    assert np.array_equal(dv.mask_compact(m), np.flatnonzero(m).astype(np.int32))
    dv.close()


def test_recycled_slot_starts_clean(seq160):
    """a released slot handed out again must not carry its previous occupant's depth / sigma / age / pose."""
    g, d, poses, K = seq160
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2)
    a = dv.frame_create(g[0])
    top = 2
    w, h = dv.level_size(top)
    dv.frame_set_depth(a, np.full((h, w), 1.5, np.float32), np.full((h, w), 0.3, np.float32), np.full((h, w), 7, np.float32))
    dv.frame_set_xi(a, np.full(6, 0.1, np.float32), np.full(6, 0.2, np.float32))
    dv.frame_release(a)
    b = dv.frame_create(g[1])
    assert b == a
    for kind in ("depth", "sigma", "age"):
        assert not dv.frame_download(b, top, kind).any(), kind
    xi, rel = dv.frame_get_xi(b)
    assert not xi.any() and not rel.any()
    dv.close()


def test_first_vo_frame_needs_its_prior(seq160):
    import dvo_b200
    g, d, poses, K = seq160
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2)
    with pytest.raises(dvo_b200.DvoError) as e:
        dv.vo_push(g[0])
    assert "rc=-6" in str(e.value)   # DVO_ERR_STATE
    # ... and the sequence can still be started properly afterwards
    d0 = np.full((120, 160), 1.5, np.float32)
    T = dv.vo_push(g[0], d0, np.full_like(d0, 0.5))
    assert np.array_equal(T[0], np.eye(4, dtype=np.float32))
    dv.close()


def test_rng_table_grows_on_demand(orc):
    """Gaussian::update's resets replay std::mt19937 from a table that starts at 16 draws here: it must grow, not fail."""
    from dvo_b200 import synth
    g, d, poses, K = synth.sequence(10, 640, 480, seed=5, trans_step=0.01, rot_step_deg=0.2)
    d0 = orc.cull_image(d[0], 2)
    s0 = np.full_like(d0, 0.4)
    n_kf, n_upd, _ = _run_pair(orc, g, K, d0, s0, accept_sigma_hi=5.0, min_movement=0.05, rng_table=16)
    assert n_upd >= 4


def test_rng_bound_of_a_caller_that_runs_ahead(orc):
    """A caller that never reads poses enqueues all frames at once: the replay-table bound then grows by roi.n per frame in flight.
    Before growing the table the library drains the stream and reads the actual position - same poses and same replay position as
    a frame-by-frame run with a table that never needs to grow, and the table stays small."""
    from dvo_b200 import synth
    n = 40
    g, d, poses, K = synth.sequence(n, 640, 480, seed=5, trans_step=0.01, rot_step_deg=0.2)
    d0 = orc.cull_image(d[0], 2)
    s0 = np.full_like(d0, 0.4)
    out = []
    for table, ahead in ((1 << 22, False), (1 << 16, True)):   # roi.n = 128 x 96 = 12288 draws per frame in flight: 2^16 = 5 frames
        dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2, max_keyframes=n + 2, accept_sigma_hi=5.0,
                      min_movement=0.05, rng_table=table)
        dev = dv.dev_alloc(g[0].nbytes * n)
        for i in range(n):
            dv.copy_to_dev(dev + i * g[0].nbytes, g[i])
        T = None
        for i in range(n):
            T = dv.vo_push_dev(dev + i * g[0].nbytes, d0 if i == 0 else None, s0 if i == 0 else None, want_pose=(not ahead) or i == n - 1)
        out.append((T.copy(), dv.rng_position().copy(), dv.vo_counters()["n_update"].copy()))
        assert dv.vo_error() == 0
        dv.dev_free(dev)
        dv.close()
    assert out[0][2].sum() >= 20 and out[0][1].max() > 0        # updates happened and consumed draws
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    print("RNG-RUN-AHEAD: %d frames enqueued without a read-back, %d update frames, furthest replay position %d (table 65536, pessimistic "
          "bound %d)" % (n, int(out[1][2].sum()), int(out[1][1].max()), 12288 * n))


def test_history_overflow_is_reported(orc):
    import dvo_b200
    from dvo_b200 import synth
    g, d, poses, K = synth.sequence(12, 640, 480, seed=77, trans_step=0.012, rot_step_deg=0.3)
    d0 = orc.cull_image(d[0], 2)
    dv = make_dvo(width=640, height=480, levels=3, culls=2, K=K, max_frames=2, max_keyframes=2)
    failed_at = None
    for i in range(len(g)):
        try:
            dv.vo_push(g[i], d0 if i == 0 else None, np.full_like(d0, 0.3) if i == 0 else None)
        except dvo_b200.DvoError as e:
            assert "rc=-4" in str(e)   # DVO_ERR_HISTORY
            failed_at = i
            break
    assert failed_at is not None and failed_at >= 2
    assert dv.vo_error() == -4
    dv.close()


def _vo_run(n_streams, n_frames, env):
    """A batched VO run on rendered sequences; returns (poses [n_frames][S][4][4], final keyframe depth / sigma / age)."""
    import os
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        W, H = 640, 480
        K = np.array([[0.82 * W, 0, 0.5 * W - 0.5], [0, 0.82 * W, 0.5 * H - 0.5], [0, 0, 1]], np.float32)
        dv = make_dvo(n_streams=n_streams, width=W, height=H, levels=3, culls=2, K=K, max_frames=2, max_keyframes=n_frames + 2)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    g = dv.dev_alloc(n_streams * H * W * 4)
    d0 = np.maximum(np.random.default_rng(7).normal(1.5, 0.5, (n_streams, 120, 160)), 0.5).astype(np.float32)
    Ts = []
    for i in range(n_frames):
        dv.synth_render_dev(g, None, i, 500, 0.012, 0.3)
        Ts.append(dv.vo_push_dev(g, d0 if i == 0 else None, np.full_like(d0, 0.5) if i == 0 else None, want_pose=True).copy())
    assert dv.vo_error() == 0
    maps = []
    for s in range(n_streams):
        slot = dv.vo_ref_slot(s)
        maps.append([dv.frame_download(slot, 2, k)[s].copy() for k in ("depth", "sigma", "age")])
    dv.dev_free(g)
    dv.close()
    return np.stack(Ts), maps


def test_vo_is_schedule_independent_and_repeatable():
    """compute-sanitizer is closed on this pool (profiles/r2_sanitize_racecheck.txt), so the hand-rolled synchronisation is checked by
    what a race would break: the SAME 48 sequences must give bit-identical poses and maps (a) run twice, (b) with Mapper::update
    overlapped with tracking (ticket + ready list, programmatic dependent launch) or launched after it, (c) with
    Implement::propagate in the tracking epilogue or as the two wide kernels - and (d) the first 4 of them run alone, where
    k_track splits every stream over several CTAs (inter-CTA barrier), must match within the fp64 summation-order difference."""
    S, n = 48, 9
    # 48 streams are fewer than the GPU's tracking slots, so by default every stream is split over CTAs and neither the overlap nor
    # the fused propagate is in play: DVO_TRK_B=1 selects the one-CTA-per-stream form that 256 streams take.
    one = {"DVO_TRK_B": "1"}
    T0, M0 = _vo_run(S, n, one)
    for env in ({}, {"DVO_OVERLAP": "0"}, {"DVO_FUSE_PROP": "0"}, {"DVO_OVERLAP": "0", "DVO_FUSE_PROP": "0"}):
        T1, M1 = _vo_run(S, n, dict(one, **env))
        assert T1.tobytes() == T0.tobytes(), env
        for s in range(S):
            for a, b in zip(M0[s], M1[s]):
                assert a.tobytes() == b.tobytes(), (env, s)
    # (e) 160 streams take that form by themselves (more streams than slots would split), under a different schedule of the ticket
    # / ready list: the first 48 are the same sequences and must come out bit-identical
    Tb, Mb = _vo_run(160, n, {})
    assert Tb[:, :S].tobytes() == T0.tobytes()
    for s in range(S):
        for a, b in zip(M0[s], Mb[s]):
            assert a.tobytes() == b.tobytes(), ("160 streams", s)
    # (d) split forms: 48 streams x 6 CTAs and 4 streams x 10 CTAs (k_track_staged, inter-CTA barrier)
    T6, M6 = _vo_run(S, n, {})
    T4, M4 = _vo_run(4, n, {})
    assert np.abs(T6 - T0).max() <= 1e-4 and np.abs(T4 - T0[:, :4]).max() <= 1e-4
    kf6 = sum(M6[s][2].tobytes() == M0[s][2].tobytes() for s in range(S))
    kf4 = sum(M4[s][2].tobytes() == M0[s][2].tobytes() for s in range(4))
    print("SCHEDULES: 48 streams x %d frames, one CTA per stream: bit-identical across repeat / overlap / fused-propagate settings and as "
          "the first 48 of 160 streams; split over CTAs: max pose difference %.3g (48 streams x 6 CTAs, age maps identical in %d of 48), "
          "%.3g (4 streams x 10 CTAs, %d of 4)" % (n, float(np.abs(T6 - T0).max()), kf6, float(np.abs(T4 - T0[:, :4]).max()), kf4))

