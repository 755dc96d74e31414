#!/usr/bin/env python
"""Emit one robot's grid.cuh under the current environment (GRID_* emitter switches), compile ONE translation unit
of the C-ABI library to a cubin and print the static SASS model (tools/sass_model.py) of one kernel.

  GRID_EMIT_PAIRS=1 python tools/kernel_sass.py iiwa14 forward_dynamics_gradient [--part 0x04] [--dtype 1] [--keep DIR]

CPU-only: this is how emitter variants are compared before a GPU run.
"""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    robot, kern = sys.argv[1], sys.argv[2]
    part = int(sys.argv[sys.argv.index("--part") + 1], 0) if "--part" in sys.argv else 0x04
    dtype = int(sys.argv[sys.argv.index("--dtype") + 1]) if "--dtype" in sys.argv else 1
    keep = sys.argv[sys.argv.index("--keep") + 1] if "--keep" in sys.argv else None
    buckets = int(sys.argv[sys.argv.index("--buckets") + 1]) if "--buckets" in sys.argv else 0
    from gridcodegenerator_b200 import GRiDCodeGenerator, load_robot
    from gridcodegenerator_b200 import runtime as rt
    import sass_model
    out_dir = keep or tempfile.mkdtemp(prefix="grid_sass_")
    os.makedirs(out_dir, exist_ok=True)
    t0 = time.time()
    gen = GRiDCodeGenerator(load_robot(robot))
    gen.gen_all_code(output_dir=out_dir)
    t1 = time.time()
    cubin = os.path.join(out_dir, f"capi_{part:03x}_{dtype}.cubin")
    cmd = [rt._nvcc()] + [f for f in rt.NVCC_FLAGS if f not in ("-shared", "-Xcompiler", "-fPIC")] + [
        "-cubin", "-Xptxas", "-v", f"-I{out_dir}", f"-I{rt.INCLUDE}", f'-DGRID_ROBOT_NAME="{robot}"',
        f"-DGRID_PARTS={part}", f"-DGRID_DTYPES={dtype}", os.path.join(rt.CSRC, "grid_capi.cu"), "-o", cubin]
    if part in (0x20, 0x40) and os.environ.get("GRID_SO_PTXAS_O1", "1") == "1":
        cmd[1:1] = ["-Xptxas", "-O1"]           # as runtime._compile_part does for the second-order units
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        print(res.stdout[-3000:])
        return 1
    t2 = time.time()
    for line in res.stdout.splitlines():
        if "registers" in line or "spill" in line:
            print(line.strip())
    print(f"emit {t1 - t0:.1f} s, nvcc {t2 - t1:.1f} s -> {cubin}")
    sel = sass_model.dump(cubin, kern + "_kernel")
    for name, ins in sel.items():
        print("==", name)
        sass_model.analyse(ins, buckets)
    return 0


if __name__ == "__main__":
    sys.exit(main())
