"""TEST INFRASTRUCTURE ONLY -- import shim for the *unmodified* reference.

Makes ``/root/reference`` (kelly7707/PIN_SLAM, pure Python/PyTorch) importable in
the build container so that its own classes can be executed on CPU to

  * validate the restatement in ``oracle/pin_oracle.py`` and
  * generate the golden vectors committed under ``tests/golden/``
    (``oracle/make_golden.py``).

``/root/reference`` does not exist on the GPU box; there the shim falls back to
``oracle/_ref/`` -- the byte-for-byte staged copy written by ``oracle/make_ref.py``
(git-ignored, travels with the tree) -- which is what ``bench.py --impl reference``
and the bench's CUDA-eager comparator time.  Nothing under ``pin_slam_b200/``
imports this module; ``available()`` lets tests skip when no tree is present.

What is stubbed: the reference does top-of-file imports
of packages that are not installed here (open3d, matplotlib, roma, natsort,
gtsam, skimage, pyquaternion, pypose, evo, laspy).  None of them is touched by the
hot-path arithmetic (model/neural_points.py, model/decoder.py, utils/tracker.py,
utils/mapper.py); the stubs are permissive placeholder modules.  ``get_time``
(utils/tools.py:269-274) calls ``torch.cuda.synchronize()`` unconditionally and
is replaced by ``time.time`` when no GPU is visible.
"""
from __future__ import annotations

import importlib
import os
import sys
import time
import types

def _find_root():
    """$PINQ_REFERENCE_ROOT, else the read-only reference of the build container, else the byte-for-byte staged copy
    oracle/_ref/ (oracle/make_ref.py; git-ignored, travels to the GPU box like libpinq.so)."""
    env = os.environ.get("PINQ_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/model"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _find_root()

_STUBBED = [
    "open3d", "matplotlib", "matplotlib.cm", "matplotlib.pyplot", "roma", "natsort",
    "gtsam", "skimage", "skimage.measure", "pyquaternion", "pypose", "laspy",
    "evo", "evo.core", "evo.core.trajectory", "evo.tools", "evo.tools.plot", "evo.tools.settings", "wandb",
]


class _Anything:
    """Object whose every attribute / call / index yields another placeholder."""

    def __init__(self, name="stub"):
        self.__dict__["_name"] = name

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Anything(self._name + "." + item)

    def __setattr__(self, key, value):
        self.__dict__[key] = value

    def __call__(self, *a, **k):
        return _Anything(self._name + "()")

    def __getitem__(self, item):
        return _Anything(self._name + "[]")

    def __iter__(self):
        return iter(())

    def __bool__(self):
        return False

    def __len__(self):
        return 0

    def _binary(self, other):
        return _Anything(self._name + "<op>")

    __add__ = __radd__ = __sub__ = __rsub__ = __mul__ = __rmul__ = __truediv__ = __rtruediv__ = _binary

    def __neg__(self):
        return _Anything("-" + self._name)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class _StubModule(types.ModuleType):
    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Anything(self.__name__ + "." + item)


# ---- real (not permissive) stand-ins for two third-party helpers the per-frame driver actually computes with ----------
def _rotmat_slerp(R0, R1, steps):
    """roma.rotmat_slerp (roma 1.4/1.5, `roma/utils.py`; un-vendored dependency of utils/tools.py:560): spherical linear
    interpolation R(t) = R0 exp(t log(R0^T R1)) for every t in `steps` (extrapolates for t outside [0,1]; shortest arc).
    roma goes through unit quaternions and a rotation vector; this is the same map written with Rodrigues' formula."""
    import torch
    Rrel = R0.transpose(-1, -2) @ R1
    cos = ((Rrel[0, 0] + Rrel[1, 1] + Rrel[2, 2]) - 1.0) * 0.5
    cos = torch.clamp(cos, -1.0, 1.0)
    w = torch.stack([Rrel[2, 1] - Rrel[1, 2], Rrel[0, 2] - Rrel[2, 0], Rrel[1, 0] - Rrel[0, 1]])
    n = torch.linalg.norm(w)
    theta = torch.atan2(0.5 * n, cos)  # (roma: 2 atan2(|q_xyz|, q_w) of the relative quaternion -- as well conditioned)
    axis = w / n if float(n) > 1e-12 else torch.tensor([1.0, 0.0, 0.0], dtype=R1.dtype, device=R1.device)
    if float(n) <= 1e-12:
        theta = torch.zeros((), dtype=R1.dtype, device=R1.device)
    ang = steps.reshape(-1) * theta                                   # [N]
    Kx = torch.zeros(3, 3, dtype=R1.dtype, device=R1.device)
    Kx[0, 1], Kx[0, 2], Kx[1, 0], Kx[1, 2], Kx[2, 0], Kx[2, 1] = -axis[2], axis[1], axis[2], -axis[0], -axis[1], axis[0]
    eye = torch.eye(3, dtype=R1.dtype, device=R1.device)
    R = eye[None] + torch.sin(ang)[:, None, None] * Kx[None] + (1.0 - torch.cos(ang))[:, None, None] * (Kx @ Kx)[None]
    return (R0[None] @ R).reshape(steps.shape + (3, 3))


def _natsorted(seq, key=None, reverse=False, alg=None):
    import re

    def nat(s):
        s = key(s) if key is not None else s
        return [int(t) if t.isdigit() else t.lower() for t in re.split(r"(\d+)", str(s))]
    return sorted(seq, key=nat, reverse=reverse)


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "model"))


_installed = False


def install():
    """Put the reference on sys.path behind permissive stubs. Idempotent."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    for name in _STUBBED:
        try:
            importlib.import_module(name)
            continue
        except Exception:
            pass
        mod = _StubModule(name)
        mod.__path__ = []  # behave like a package
        sys.modules[name] = mod
        if "." in name:
            parent, child = name.rsplit(".", 1)
            setattr(sys.modules[parent], child, mod)
        if name == "roma":
            mod.rotmat_slerp = _rotmat_slerp
        if name == "natsort":
            mod.natsorted = _natsorted
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _installed = True


def _patch_get_time():
    import torch

    if torch.cuda.is_available():
        return
    for modname in ("utils.tools", "model.neural_points", "utils.tracker", "utils.mapper",
                    "utils.data_sampler", "utils.loop_detector", "dataset.slam_dataset", "pin_slam"):
        mod = sys.modules.get(modname)
        if mod is not None and hasattr(mod, "get_time"):
            mod.get_time = time.time


def load():
    """Returns a namespace with the reference's hot-path classes/functions."""
    install()
    ns = types.SimpleNamespace()
    ns.Config = importlib.import_module("utils.config").Config
    ns.tools = importlib.import_module("utils.tools")
    ns.neural_points_mod = importlib.import_module("model.neural_points")
    ns.NeuralPoints = ns.neural_points_mod.NeuralPoints
    ns.Decoder = importlib.import_module("model.decoder").Decoder
    ns.loss = importlib.import_module("utils.loss")
    ns.tracker_mod = importlib.import_module("utils.tracker")
    ns.Tracker = ns.tracker_mod.Tracker
    try:
        ns.mapper_mod = importlib.import_module("utils.mapper")
        ns.Mapper = ns.mapper_mod.Mapper
    except Exception as e:  # dataset imports may need more stubs; not needed for kernels
        ns.mapper_mod = None
        ns.Mapper = None
        ns.mapper_import_error = e
    _patch_get_time()
    return ns


def kitti_config(device="cpu"):
    """The reference's own Config, resolved from config/lidar_slam/run_kitti.yaml."""
    ns = load()
    cfg = ns.Config()
    cfg.load(os.path.join(REFERENCE_ROOT, "config/lidar_slam/run_kitti.yaml"))
    cfg.device = device
    cfg.o3d_vis_on = False
    return cfg


def write_kitti_yaml(dst, paths, output_root, drop_pgo=True, extra=None):
    """config/lidar_slam/run_kitti.yaml of the reference with ONLY these edits: the three input
    paths and the output root, `o3d_vis_on: False` (no graphics stack on this box), and the `pgo:` section removed
    (gtsam is not installed; no loop can fire within 100 frames / 90 m: needs > 4 x 82 m, pin_slam.py:172).  `extra`:
    {section: {key: value}} overrides used by tests (stated by the caller)."""
    import yaml
    with open(os.path.join(REFERENCE_ROOT, "config/lidar_slam/run_kitti.yaml")) as f:
        cfg = yaml.safe_load(f)
    cfg["setting"].update({"pc_path": paths["pc_path"], "pose_path": paths["pose_path"], "calib_path": paths["calib_path"],
                           "output_root": output_root})
    cfg["eval"]["o3d_vis_on"] = False
    if drop_pgo:
        cfg.pop("pgo", None)
    for sec, kv in (extra or {}).items():
        cfg.setdefault(sec, {}).update(kv)
    with open(dst, "w") as f:
        yaml.safe_dump(cfg, f, sort_keys=False)
    return dst
