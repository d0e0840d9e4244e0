"""CPU, build container only (needs /root/reference): the reference driver imports with the mirrors bound, and every
mirrored method has the reference's exact signature (names and defaults)."""
import importlib
import inspect
import sys

import pytest

from oracle import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present (GPU box)")

METHODS = {
    ("model.neural_points", "NeuralPoints"): ["__init__", "update", "reset_local_map", "assign_local_to_global", "prune_map",
                                              "adjust_map", "recreate_hash", "set_search_neighborhood",
                                              "radius_neighborhood_search", "query_certainty", "query_feature", "clear_temp",
                                              "is_empty", "count", "local_count", "print_memory"],
    ("model.decoder", "Decoder"): ["__init__", "forward", "sdf", "occupancy"],
    ("utils.tracker", "Tracker"): ["__init__", "tracking", "query_source_points", "registration_step"],
    ("utils.mapper", "Mapper"): ["__init__", "dynamic_filter", "get_batch", "mapping", "sdf", "get_numerical_gradient",
                                 "free_pool", "bundle_adjustment"],
    ("utils.mesher", "Mesher"): ["__init__", "query_points", "get_query_from_bbx"],
}


def _sig(fn):
    return [(p.name, p.default) for p in inspect.signature(fn).parameters.values()]


def test_signatures_match_reference():
    ref_shim.install()
    import pin_slam_b200.dropin as dropin
    for k in dropin.HOT_PATH_MODULES:
        sys.modules.pop(k, None)
    for (mod, cls), names in METHODS.items():
        ref_cls = getattr(importlib.import_module(mod), cls)
        mine_cls = getattr(importlib.import_module(dropin.HOT_PATH_MODULES[mod]), cls)
        for n in names:
            a, b = _sig(getattr(ref_cls, n)), _sig(getattr(mine_cls, n))
            if (cls, n) == ("Mapper", "process_frame"):
                continue
            assert a == b[:len(a)], (cls, n, a, b)   # mirrors may append optional keyword arguments only
    for k in dropin.HOT_PATH_MODULES:
        sys.modules.pop(k, None)


def test_reference_driver_imports_with_the_mirrors():
    ref_shim.install()
    import pin_slam_b200.dropin as dropin
    dropin.install()
    sys.modules.pop("pin_slam", None)
    pin_slam = importlib.import_module("pin_slam")
    import pin_slam_b200.model.neural_points as mine
    assert pin_slam.NeuralPoints is mine.NeuralPoints
    for name in ("Tracker", "Mapper", "Mesher", "Decoder"):
        assert getattr(pin_slam, name).__module__.startswith("pin_slam_b200.")
    for k in list(dropin.HOT_PATH_MODULES) + ["pin_slam"]:
        sys.modules.pop(k, None)
