"""CPU-only: the C-ABI library loads and exports every symbol include/dvo_cuda.h declares
(no compute calls — there is no GPU here), and fails loudly without a device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "dvo_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dvo_[A-Za-z0-9_]+)\s*\(", txt)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for needed in ("dvo_create", "dvo_gn_step", "dvo_track", "dvo_warp_image", "dvo_map_update", "dvo_map_propagate",
                   "dvo_map_regularize", "dvo_vo_push", "dvo_mask_compact"):
        assert needed in syms


def test_library_exports_every_declared_symbol(dvolib):
    for s in declared_symbols():
        assert hasattr(dvolib, s), "libdvo_cuda.so does not export %s" % s


def test_python_binding_lists_same_symbols():
    import dvo_b200
    assert sorted(dvo_b200.EXPORTS) == declared_symbols()


def test_params_default_match_reference_literals(dvolib):
    import dvo_b200
    p = dvo_b200.default_params(dvolib)
    assert (p.width, p.height, p.levels, p.culls) == (640, 480, 3, 2)           # system.hpp:47
    assert p.max_iterations == 15 and abs(p.min_update - 5e-4) < 1e-9           # tracker.cpp:16-19
    assert abs(p.min_residual - 5e-3) < 1e-9
    assert (p.roi_level, p.roi_x0, p.roi_x1, p.roi_y0, p.roi_y1) == (2, 20, 140, 20, 100)  # optimize.cpp:33-36
    assert p.min_depth == 0.20                                                  # optimize.cpp:39
    assert (p.map_roi_x0, p.map_roi_x1, p.map_roi_y0, p.map_roi_y1) == (16, 144, 12, 108)  # mapper.cpp:90
    assert abs(p.min_movement - 0.02) < 1e-9 and p.max_forward == 6             # mapper.cpp:12-13


def test_no_cpu_fallback(dvolib):
    """Without a CUDA device dvo_create must fail (DVO_ERR_CUDA), never compute on the host."""
    import dvo_b200
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = os.path.exists("/dev/nvidia0")
    if has_gpu:
        pytest.skip("GPU present")
    with pytest.raises(dvo_b200.DvoError):
        dvo_b200.Dvo(lib=dvolib)


def test_product_does_not_reference_oracle():
    """The product path must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "direct-visual-odometry_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h", ".py")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "dvo_oracle" not in txt and "oracle/" not in txt and "from oracle" not in txt, os.path.join(dp, f)
    for dp, _, fs in os.walk(os.path.join(ROOT, "include")):
        for f in fs:
            txt = open(os.path.join(dp, f), errors="ignore").read()
            assert "dvo_oracle" not in txt, f
