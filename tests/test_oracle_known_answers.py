"""Known answers for the THIRD-PARTY arithmetic the reference's hot path calls but does not contain:
OpenCV (un-vendored, "OpenCV >=3", README.md:8) and libstdc++ <random>.  The C oracle
(oracle/dvo_oracle.c) and the OpenCV-API shim (oracle/shim) restate these pieces; here they are
checked against opencv-python (4.13 in this image) wherever Python exposes the same primitive, and
against a committed libstdc++ mt19937 stream (tests/golden/mt19937_libstdcxx.json, generated with
g++ 13.3 by scripts in the file's header).

cv2.solve is the one place that cannot be bit-exact: the opencv-python wheel is built with LAPACK
(cv2.getBuildInformation(): "Lapack: YES"), so DECOMP_SVD runs sgesdd instead of OpenCV's own
JacobiSVDImpl_<float>; the restated Jacobi path is held to 5e-5 relative against it and bit-exact
between the two restatements.
"""
import json
import os

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def ref():
    ref_mod = pytest.importorskip("oracle.ref")
    try:
        return ref_mod.Ref()
    except (FileNotFoundError, OSError) as e:  # pragma: no cover
        pytest.skip("oracle/_ref not available: %s" % e)


def test_rodrigues_bit_exact(orc, ref):
    rng = np.random.default_rng(0)
    for _ in range(3000):
        w = (rng.normal(size=3) * rng.choice([0.0, 1e-17, 1e-9, 1e-4, 0.01, 0.3, 2.0, 3.2])).astype(np.float32)
        want = cv2.Rodrigues(w.reshape(3, 1))[0].astype(np.float32)
        assert np.array_equal(want, orc.rodrigues(w))
        assert np.array_equal(want, ref.rodrigues(w))


def test_norm_bit_exact(orc, ref):
    rng = np.random.default_rng(1)
    for n in (3, 6):
        for _ in range(1000):
            v = (rng.normal(size=n) * rng.choice([1e-6, 1e-2, 1.0, 50.0])).astype(np.float32)
            want = cv2.norm(v.reshape(n, 1))
            assert want == orc.norm(v) == ref.norm(v)


@pytest.mark.parametrize("m,k,n", [(3, 3, 1), (3, 3, 3), (4, 4, 4), (2, 2, 2)])
def test_small_gemm_bit_exact(ref, m, k, n):
    """cv::gemm's unrolled small-matrix path: fp32 dot products left to right (no FMA), then
    (float)(t*alpha + c*beta) in double — what every `Mat1f * Mat1f` (+ / - Mat1f) of the reference
    lowers to (transform.cpp:10, se3.cpp:85-88,114-118,130, implement.cpp:62-63)."""
    rng = np.random.default_rng(m * 100 + k * 10 + n)
    for _ in range(1500):
        A = (rng.normal(size=(m, k)) * rng.choice([1e-3, 1.0, 300.0])).astype(np.float32)
        B = (rng.normal(size=(k, n)) * rng.choice([1e-3, 1.0, 300.0])).astype(np.float32)
        Cm = rng.normal(size=(m, n)).astype(np.float32)
        alpha, beta = float(rng.normal()), float(rng.normal())
        assert np.array_equal(cv2.gemm(A, B, 1.0, None, 0.0), ref.shim_gemm(A, B))
        assert np.array_equal(cv2.gemm(A, B, alpha, None, 0.0), ref.shim_gemm(A, B, alpha))
        assert np.array_equal(cv2.gemm(A, B, 1.0, Cm, 1.0), ref.shim_gemm(A, B, 1.0, Cm, 1.0))
        assert np.array_equal(cv2.gemm(A, B, -1.0, Cm, beta), ref.shim_gemm(A, B, -1.0, Cm, beta))


def test_matexpr_lowering_primitives(ref):
    """MatOp_AddEx::assign lowers a*alpha + b*beta to convertTo / add / subtract / scaleAdd /
    addWeighted; the shim's per-element formulas against the cv2 primitives."""
    rng = np.random.default_rng(4)
    a = rng.normal(size=(64, 1)).astype(np.float32); b = rng.normal(size=(64, 1)).astype(np.float32)
    for _ in range(200):
        al, be = float(rng.normal()), float(rng.normal())
        assert np.array_equal(cv2.scaleAdd(b, be, a).ravel(), ref.shim_addex(a, b, 1.0, be))
        assert np.array_equal(cv2.scaleAdd(a, al, b).ravel(), ref.shim_addex(a, b, al, 1.0))
        assert np.array_equal(cv2.addWeighted(a, al, b, be, 0.0).ravel(), ref.shim_addex(a, b, al, be))
    assert np.array_equal(cv2.add(a, b).ravel(), ref.shim_addex(a, b, 1.0, 1.0))
    assert np.array_equal(cv2.subtract(a, b).ravel(), ref.shim_addex(a, b, 1.0, -1.0))
    assert np.array_equal((a * np.float32(0.37)).ravel(), ref.shim_addex(a, None, float(np.float32(0.37)), 0.0))


def test_solve_svd(orc, ref):
    """cv::solve(A, b, x, DECOMP_SVD), optimize.cpp:97 (tall N x 6 systems, incl. zero rows and a
    rank-deficient case: min-norm solution)."""
    rng = np.random.default_rng(2)
    worst = 0.0
    for t in range(120):
        m = int(rng.integers(6, 4000))
        A = (rng.normal(size=(m, 6)) * rng.uniform(0.1, 50, size=(1, 6))).astype(np.float32)
        A[rng.random(m) < 0.3] = 0.0
        if t % 10 == 0:
            A[:, 4] = A[:, 1]  # rank 5
        b = rng.normal(size=(m, 1)).astype(np.float32)
        want = cv2.solve(A, b, flags=cv2.DECOMP_SVD)[1].ravel()
        got = orc.solve_svd(A, b)
        assert np.array_equal(got, ref.shim_solve_svd(A, b))
        if t % 10 == 0:
            # a numerically null singular value sits right at the 2*eps*sum(w) cut-off: LAPACK and the
            # Jacobi path may fall on either side of it, so only the two restatements are compared
            continue
        worst = max(worst, float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-12)))
    assert worst < 5e-5, worst


@pytest.mark.parametrize("seed", [1, 12345, 987654321, 2147483647])
def test_randn_bit_exact(ref, seed):
    """cv::randn (Scene's depth prior, frame.hpp:19): MWC + Ziggurat + scale, several blocks long."""
    cv2.setRNGSeed(seed)
    want = np.empty((1, 5000), np.float32)
    cv2.randn(want, 1.5, 0.5)
    got = ref.shim_randn(seed, 5000, 1.5, 0.5)
    assert np.array_equal(want.ravel(), got)


def test_mt19937_stream_and_distribution(orc):
    """std::mt19937 (default seed) + uniform_real_distribution<float>(2.0, 0.5), gaussian.cpp:8-9,22."""
    gold = json.load(open(os.path.join(HERE, "golden", "mt19937_libstdcxx.json")))
    raw = np.array(gold["raw"], np.uint32)
    assert np.array_equal(orc.rng_u32(len(raw)), raw)
    key = "dist" if "dist" in gold else [k for k in gold if k != "raw"][0]
    dist = np.array(gold[key], np.float32)
    assert np.array_equal(orc.dist_from_u32(raw[:len(dist)]), dist)


def test_ssd_first_term_identity(tmp_path):
    """search_ssd (csrc/dvo_kernels.cuh) evaluates implement.cpp:134's first term, (float)(0.0 + (1.0 * 1 / 3) * (double)(d * d)), as the
    correctly rounded float quotient (d * d) / 3 through two FMAs.  tests/cpp/ssd_third.c compares the two forms bit for bit; here over
    every 61st non-negative finite float bit pattern (stride 1 = all of them: 0 mismatches, ~20 s)."""
    import subprocess
    exe = str(tmp_path / "ssd_third")
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "ssd_third.c")
    subprocess.check_call(["gcc", "-O2", "-mfma", "-ffp-contract=off", "-o", exe, src, "-lm"])
    out = subprocess.run([exe, "61"], capture_output=True, text=True, timeout=120)
    n, bad = map(int, out.stdout.split())
    assert out.returncode == 0 and bad == 0 and n > 30_000_000, out.stdout
