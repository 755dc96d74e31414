#!/usr/bin/env python
"""A/B timing of variant builds of one robot's library (tools/build_variants.sh puts them under
gridcodegenerator_b200/_variants/<tag>/).  Times grad FD at the headline size with CUDA events.
Each variant runs in its own process (the build dir is chosen at import time via GRID_BUILD_DIR)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VAR = os.path.join(ROOT, "gridcodegenerator_b200", "_variants")

CHILD = r'''
import sys, json, numpy as np, torch
sys.path.insert(0, %r)
from gridcodegenerator_b200.runtime import GRiDLibrary
lib = GRiDLibrary("iiwa14", build=False)
n, B = lib.n, 65536
rng = np.random.default_rng(0)
xs = [torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda() for _ in range(9)]
outs = [torch.empty((B, 2 * n * n), dtype=torch.float32, device="cuda") for _ in range(9)]
for i in range(20): lib.forward_dynamics_gradient(xs[i %% 9], out=outs[i %% 9])
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph(); side = torch.cuda.Stream()
with torch.cuda.stream(side):
    lib.forward_dynamics_gradient(xs[0], out=outs[0]); torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=side):
        for i in range(200): lib.forward_dynamics_gradient(xs[i %% 9], out=outs[i %% 9])
g.replay(); torch.cuda.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 200)
so = None
if lib.has_second_order:
    xb = xs[0][:32768].contiguous()
    o2 = torch.empty((32768, 4 * n ** 3), dtype=torch.float32, device="cuda")
    so = {}
    for name in ("idsva_so", "fdsva_so"):
        f = getattr(lib, name)
        f(xb, out=o2); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): f(xb, out=o2)
        e1.record(); torch.cuda.synchronize()
        so[name + "_ms_32768"] = e0.elapsed_time(e1) / 5
print(json.dumps({"second_order": so, "us_per_launch": best * 1e3, "G_knots_per_s": B / best / 1e6, "info": lib.kernel_info("forward_dynamics_gradient"),
                  "checksum": float(outs[0].double().abs().sum())}))
''' % ROOT

for tag in sorted(os.listdir(VAR)):
    env = dict(os.environ, GRID_BUILD_DIR=os.path.join(VAR, tag))
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print(tag, (r.stdout.strip() or r.stderr.strip()[-400:]), flush=True)
