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


def test_install_rebinds_the_per_frame_helpers_of_utils_tools():
    """dropin.install() rebinds voxel_down_sample_torch / voxel_down_sample_min_value_torch / deskewing / transform_torch
    inside the reference's OWN utils.tools: same signatures as the reference functions, the reference function kept for
    inputs the kernels do not take (here: CPU tensors give the reference's results bit for bit), and the reference's dataset
    module picks the rebound functions up through its `from utils.tools import ...`."""
    import torch
    ref_shim.install()
    import pin_slam_b200.dropin as dropin
    replaced = dropin.install()
    tools = importlib.import_module("utils.tools")
    for n in dropin.TOOLS_FUNCTIONS:
        assert "utils.tools." + n in replaced
        fn = getattr(tools, n)
        ref_fn = fn.reference_function
        import pin_slam_b200.utils.tools as mine
        a, b = _sig(ref_fn), _sig(getattr(mine, n))
        assert [x[0] for x in a] == [x[0] for x in b], (n, a, b)
    g = torch.Generator().manual_seed(3)
    pts = torch.rand((500, 3), generator=g) * 5.0
    assert torch.equal(tools.voxel_down_sample_torch(pts, 0.5), tools.voxel_down_sample_torch.reference_function(pts, 0.5))
    T = torch.eye(4, dtype=torch.float64)
    T[:3, 3] = torch.tensor([1.0, 2.0, 3.0], dtype=torch.float64)
    assert torch.equal(tools.transform_torch(pts, T), tools.transform_torch.reference_function(pts, T))
    sys.modules.pop("dataset.slam_dataset", None)
    ds = importlib.import_module("dataset.slam_dataset")
    assert ds.deskewing is tools.deskewing and ds.voxel_down_sample_torch is tools.voxel_down_sample_torch
    dropin.install()  # idempotent: the adapters are not wrapped twice
    assert not hasattr(tools.deskewing.reference_function, "reference_function")
    for k in list(dropin.HOT_PATH_MODULES) + ["pin_slam", "dataset.slam_dataset"]:
        sys.modules.pop(k, None)
