#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- stages the UNMODIFIED reference (kelly7707/PIN_SLAM, pure Python) under oracle/_ref/.

oracle/_ref/ is git-ignored (never part of the history) but NOT gpurun-ignored, so it travels to the GPU box the same way
the built libpinq.so does.  `bench.py --impl reference` and the bench's CUDA-eager comparator import the reference from
there through oracle/ref_shim.py; nothing under pin_slam_b200/ ever does.  The files are copied byte for byte (a sha256
manifest is written next to them so a test can prove it); no reference source enters the repository.

  python oracle/make_ref.py            # copies /root/reference/{model,utils,dataset,eval,config,*.py} -> oracle/_ref/
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("PINQ_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
DIRS = ["model", "utils", "dataset", "eval", "config"]
FILES = ["pin_slam.py", "vis_pin_map.py", "LICENSE"]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(force: bool = False) -> str | None:
    """Returns the staged root, or None when the reference tree is not present (GPU box: use what travelled)."""
    if not os.path.isdir(os.path.join(SRC, "model")):
        return DST if os.path.isdir(os.path.join(DST, "model")) else None
    manifest = {}
    for d in DIRS:
        for root, _, names in os.walk(os.path.join(SRC, d)):
            for n in names:
                if n.endswith((".py", ".yaml", ".yml", ".txt")):
                    p = os.path.join(root, n)
                    manifest[os.path.relpath(p, SRC)] = _sha(p)
    for f in FILES:
        p = os.path.join(SRC, f)
        if os.path.exists(p):
            manifest[f] = _sha(p)
    mpath = os.path.join(DST, "MANIFEST.json")
    if not force and os.path.exists(mpath):
        with open(mpath) as f:
            if json.load(f) == manifest and all(os.path.exists(os.path.join(DST, r)) for r in manifest):
                return DST
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    for rel in manifest:
        out = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), out)
    with open(mpath, "w") as f:
        json.dump(manifest, f, indent=0, sort_keys=True)
    return DST


def verify() -> bool:
    """Every staged file still has the sha256 recorded when it was copied from the reference."""
    mpath = os.path.join(DST, "MANIFEST.json")
    if not os.path.exists(mpath):
        return False
    with open(mpath) as f:
        manifest = json.load(f)
    return all(os.path.exists(os.path.join(DST, r)) and _sha(os.path.join(DST, r)) == h for r, h in manifest.items())


if __name__ == "__main__":
    root = stage(force="--force" in sys.argv)
    print(root, "verified" if root and verify() else "NOT staged")
