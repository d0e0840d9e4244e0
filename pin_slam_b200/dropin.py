"""Drop-in installer: makes the reference's own `pin_slam.py` import the B200 path.

    import pin_slam_b200.dropin as dropin; dropin.install()      # before `import pin_slam`

aliases the five hot-path modules the reference driver imports (pin_slam.py:13-25) --
`model.neural_points`, `model.decoder`, `utils.tracker`, `utils.mapper`, `utils.mesher` -- to the mirrors in
this package.  Everything else (`utils.config`, `utils.tools`, `dataset.*`, `utils.pgo`, ...) keeps coming
from the reference tree on PYTHONPATH.
"""
import importlib
import sys

HOT_PATH_MODULES = {
    "model.neural_points": "pin_slam_b200.model.neural_points",
    "model.decoder": "pin_slam_b200.model.decoder",
    "utils.tracker": "pin_slam_b200.utils.tracker",
    "utils.mapper": "pin_slam_b200.utils.mapper",
    "utils.mesher": "pin_slam_b200.utils.mesher",
}


def install():
    for ref_name, mine in HOT_PATH_MODULES.items():
        sys.modules[ref_name] = importlib.import_module(mine)
    return sorted(HOT_PATH_MODULES)
