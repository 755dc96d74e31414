#!/usr/bin/env python
"""Build (here) and run (on a B200) stand-alone variants of one emitted kernel under different
(NT, STATE_SMEM, STAGED) configurations, to choose emit/header.py's choose_config policy from measurements.

  python tools/tune_config.py build atlas forward_dynamics_gradient 3 1  128,0,3 128,0,0 256,0,3 32,1,3
  python tools/tune_config.py run [batch [prefix]]   # on the GPU box: runs every binary under _tune/ (tag prefix)
"""
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gridcodegenerator_b200")
TUNE = os.path.join(PKG, "_tune")


def build(robot, algo, incols, grav, cfgs):
    src = open(os.environ.get("TUNE_SRC") or os.path.join(PKG, "_build", robot, "grid.cuh")).read()
    m = re.search(r"struct %s_algo \{.*?OUT = (\d+);.*?SLOTS = (\d+);" % algo, src, re.S)
    out, slots = int(m.group(1)), int(m.group(2))
    jobs = []
    for cfg in cfgs:
        nt, st, sg = (int(x) for x in cfg.split(","))
        tag = f"{robot}_{algo}_{nt}_{st}_{sg}" + os.environ.get("TUNE_SUFFIX", "")
        d = os.path.join(TUNE, tag)
        os.makedirs(d, exist_ok=True)
        pat = re.compile(r"(struct %s_cfg \{\s*static constexpr int NT = )[^;]*(;\s*static constexpr bool STATE_SMEM = )"
                         r"[^;]*(;\s*static constexpr int STAGED = )[^;]*;" % algo)
        assert pat.search(src), algo
        txt = pat.sub(lambda mm: f"{mm.group(1)}{nt}{mm.group(2)}{'true' if st else 'false'}{mm.group(3)}{sg};", src)
        need = nt * (slots + out)
        txt = re.sub(r"GRID_SCRATCH_ELEMS_PER_BLOCK_F32 = \d+;", f"GRID_SCRATCH_ELEMS_PER_BLOCK_F32 = {need};", txt)
        open(os.path.join(d, "grid.cuh"), "w").write(txt)
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo", *os.environ.get("TUNE_FLAGS", "").split(),
               f"-DALGO={algo}", f"-DINCOLS={incols}", f"-DGRAV={grav}", f'-DTAG="{tag}"', "-I", d,
               "-I", os.path.join(PKG, "csrc"), os.path.join(ROOT, "tools", "tune_kernel.cu"), "-o", os.path.join(d, "tune")]
        jobs.append((tag, cmd))
    with ThreadPoolExecutor(8) as ex:
        for tag, r in zip([j[0] for j in jobs], ex.map(lambda j: subprocess.run(j[1], capture_output=True, text=True), jobs)):
            print(tag, "ok" if r.returncode == 0 else r.stderr[-600:], flush=True)


def run(batch, prefix=""):
    for tag in sorted(t for t in os.listdir(TUNE) if t.startswith(prefix)):
        exe = os.path.join(TUNE, tag, "tune")
        if os.path.exists(exe):
            r = subprocess.run(["timeout", "120", exe] + ([batch] if batch else []), capture_output=True, text=True)
            print(r.stdout.strip() or (tag + " FAILED " + r.stderr[-300:]), flush=True)


def emit(robot, outdir):
    sys.path.insert(0, ROOT)
    from gridcodegenerator_b200 import GRiDCodeGenerator, load_robot
    os.makedirs(outdir, exist_ok=True)
    GRiDCodeGenerator(load_robot(robot)).gen_all_code(output_dir=outdir)


if __name__ == "__main__":
    if sys.argv[1] == "emit":       # emit <robot> <outdir>  (emitter env knobs apply), for TUNE_SRC=<outdir>/grid.cuh
        emit(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "build":
        build(sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]), sys.argv[6:])
    else:
        run(sys.argv[2] if len(sys.argv) > 2 else None, sys.argv[3] if len(sys.argv) > 3 else "")
