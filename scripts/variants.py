#!/usr/bin/env python3
"""Kernel experiments: builds libpinq variants with extra -D flags (here, on CPU) and times them on the GPU box.

  python scripts/variants.py build  name1=-DPINQ_X=1 name2="-DPINQ_Y=2 -DPINQ_Z=0"     # -> build/variants/libpinq_<name>.so
  python scripts/variants.py run [--ncu name]                                         # on the GPU box: bench kernels per variant
The product library (pin_slam_b200/libpinq.so) is always included as variant `base`."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VDIR = os.path.join(ROOT, "build", "variants")


def build(specs):
    from pin_slam_b200 import build as B
    os.makedirs(VDIR, exist_ok=True)
    for old in os.listdir(VDIR):
        os.remove(os.path.join(VDIR, old))
    procs = []
    for spec in specs:
        name, flags = spec.split("=", 1)
        out = os.path.join(VDIR, "libpinq_%s.so" % name)
        cmd = [B.nvcc_path()] + B.NVCC_FLAGS + ["-I" + os.path.join(B.ROOT, "include"), "-I" + B.CSRC] + flags.split()
        cmd += [os.path.join(B.CSRC, s) for s in B.SOURCES] + ["-o", out]
        procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for name, p in procs:
        out, _ = p.communicate()
        print(name, "ok" if p.returncode == 0 else "FAILED\n" + out[-2000:])


def run(extra):
    libs = [("base", None)] + sorted((f[8:-3], os.path.join(VDIR, f)) for f in os.listdir(VDIR) if f.endswith(".so")) \
        if os.path.isdir(VDIR) else [("base", None)]
    res = {}
    for name, path in libs:
        env = dict(os.environ)
        if path:
            env["PINQ_LIB"] = path
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "10", "--warmup", "3", "--no-e2e", "--no-cpu-baseline",
               "--slam-frames", "0", "--no-extra-legs"] + extra
        p = subprocess.run(cmd, env=env, capture_output=True, text=True)
        try:
            line = json.loads(p.stdout.strip().splitlines()[-1])
            k = line["roofline"]["kernels"]
            res[name] = {"ms_per_step": line["ms_per_step"], **{kk: round(v["ms"], 4) for kk, v in k.items()}}
        except Exception:
            res[name] = {"error": (p.stderr or p.stdout)[-600:]}
        print(name, json.dumps(res[name]), flush=True)
    return res


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        out = run(sys.argv[2:])
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "variants.json"), "w") as f:
            json.dump(out, f, indent=1)
