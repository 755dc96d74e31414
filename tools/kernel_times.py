#!/usr/bin/env python
"""Launch every iiwa14 fp32 kernel a few times at 65536 knots (for an ncu --metrics pass; nothing is timed here).
The library is taken from GRID_BUILD_DIR when set (variant builds), else from the in-tree build."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gridcodegenerator_b200.runtime import GRiDLibrary
lib = GRiDLibrary("iiwa14", build=False)
n, B = lib.n, 65536
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
only = [a for a in sys.argv[1:] if not a.startswith("--")]
if "--time" in sys.argv: only = ["none"]
def run(name, fn):
    if only and name not in only: return
    for _ in range(4): fn()
    torch.cuda.synchronize()
q, q_qd = x[:, :n].contiguous(), x[:, :2 * n].contiguous()
qdd = lib.forward_dynamics(x); Minv = lib.direct_minv(q)
run("id", lambda: lib.inverse_dynamics(q_qd, qdd))
run("minv", lambda: lib.direct_minv(q))
run("fd", lambda: lib.forward_dynamics(x))
run("id_du", lambda: lib.inverse_dynamics_gradient(q_qd, qdd))
run("fd_du", lambda: lib.forward_dynamics_gradient(x))
run("fd_du_qm", lambda: lib.forward_dynamics_gradient_qdd_minv(q_qd, qdd, Minv))
run("fd_lin", lambda: lib.forward_dynamics_linearization(x))
run("aba", lambda: lib.aba(x))
run("crba", lambda: lib.crba(q))
if hasattr(lib, "has_ee_hessian"):
    run("ee", lambda: lib.end_effector_pose_gradient_hessian(q))

# ---- timing mode (python tools/kernel_times.py --time): CUDA-graph replay of 20 launches per kernel, CUDA events
if "--time" in sys.argv:
    import json
    outs = {}
    def timed(name, fn):
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            fn()
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=s):
                for _ in range(20): fn()
        for _ in range(3): g.replay()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 20)
        print(json.dumps({"kernel": name, "us": best * 1e3}), flush=True)
    o_c = torch.empty((B, n), dtype=torch.float32, device="cuda")
    o_m = torch.empty((B, n * n), dtype=torch.float32, device="cuda")
    o_g = torch.empty((B, 2 * n * n), dtype=torch.float32, device="cuda")
    timed("id", lambda: lib.inverse_dynamics(q_qd, qdd, out=o_c))
    timed("minv", lambda: lib.direct_minv(q, out=o_m))
    timed("fd", lambda: lib.forward_dynamics(x, out=o_c))
    timed("id_du", lambda: lib.inverse_dynamics_gradient(q_qd, qdd, out=o_g))
    timed("fd_du", lambda: lib.forward_dynamics_gradient(x, out=o_g))
    timed("fd_du_qm", lambda: lib.forward_dynamics_gradient_qdd_minv(q_qd, qdd, Minv, out=o_g))
    o_l = torch.empty((B, 3 * n * n + n), dtype=torch.float32, device="cuda")
    timed("fd_lin", lambda: lib.forward_dynamics_linearization(x, out=o_l))
    timed("aba", lambda: lib.aba(x, out=o_c))
    timed("crba", lambda: lib.crba(q, out=o_m))
