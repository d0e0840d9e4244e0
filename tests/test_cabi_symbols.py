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


def test_ctypes_structs_match_the_header_layout(tmp_path):
    """sizeof / field offsets of every struct in include/pinq.h as the C compiler lays them out == the ctypes mirrors."""
    import subprocess
    from pin_slam_b200 import _pinq as P
    pairs = {"pinq_map_view": P.MapView, "pinq_decoder_view": P.DecoderView, "pinq_query_args": P.QueryArgs,
             "pinq_bwd_args": P.BwdArgs, "pinq_tracking_args": P.TrackingArgs, "pinq_tracking_loop": P.TrackingLoop, "pinq_adam_tensor": P.AdamTensor,
             "pinq_mapping_args": P.MappingArgs, "pinq_peer_args": P.PeerArgs, "pinq_pool_view": P.PoolView}
    header = open(os.path.join(ROOT, "include", "pinq.h")).read()
    declared = set(re.findall(r"typedef struct (pinq_[a-z_]+)", header))
    assert declared == set(pairs), declared ^ set(pairs)
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "pinq.h"', 'int main(void) {']
    for cname, cls in pairs.items():
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in cls._fields_:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, "map" if fname == "map" else fname))
    lines.append("return 0; }")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, cls in pairs.items():
        assert int(out[cname]) == ctypes.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(out["%s.%s" % (cname, fname)]) == getattr(cls, fname).offset, (cname, fname)
