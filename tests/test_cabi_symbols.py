"""CPU: the C-ABI library builds, loads, and exports every symbol that include/pinq.h declares (no compute)."""
import ctypes
import os
import re

from helpers import ROOT


def declared_functions():
    src = open(os.path.join(ROOT, "include", "pinq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pinq_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    names = declared_functions()
    for must in ("pinq_table_clear", "pinq_index_build", "pinq_neighbor_search", "pinq_query_fused_fwd",
                 "pinq_query_fused_bwd", "pinq_gn_accumulate", "pinq_mesher_grid_query", "pinq_adam_step"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from pin_slam_b200 import _pinq
    lib = ctypes.CDLL(_pinq.lib_path())
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing
    # the ctypes prototypes cover the header one to one
    assert sorted(_pinq.SIGNATURES) == declared_functions()
    assert _pinq.load().pinq_version() >= 100


def test_no_cpu_fallback():
    import pytest
    import torch
    from pin_slam_b200 import _pinq
    with pytest.raises(_pinq.PinqError):
        _pinq.ptr(torch.zeros(3))
    from pin_slam_b200.model.neural_points import NeuralPoints
    from pin_slam_b200.utils.config import kitti_config
    with pytest.raises(_pinq.PinqError):
        NeuralPoints(kitti_config("cpu"))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "pin_slam_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(d, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, os.path.join(d, f)
