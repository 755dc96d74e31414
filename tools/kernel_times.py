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
only = sys.argv[1:] 
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
run("aba", lambda: lib.aba(x))
run("crba", lambda: lib.crba(q))
if hasattr(lib, "has_ee_hessian"):
    run("ee", lambda: lib.end_effector_pose_gradient_hessian(q))
