#!/usr/bin/env python
"""Small invocation of every iiwa14 / branched6 first-order kernel through all tile-loop paths (TMA-staged full
tiles, ragged tail, strided and 4-byte-aligned buffers, host pipeline) for a compute-sanitizer memcheck pass."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gridcodegenerator_b200.runtime import GRiDLibrary
for robot in ("iiwa14", "branched6", "hyq"):
    lib = GRiDLibrary(robot, build=False)
    n, B = lib.n, 1000                      # 31 full tiles + a ragged one
    x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
    q, q_qd = x[:, :n].contiguous(), x[:, :2 * n].contiguous()
    qdd = lib.forward_dynamics(x); Minv = lib.direct_minv(q)
    lib.inverse_dynamics(q_qd, qdd); lib.inverse_dynamics(x)             # dense + strided
    lib.inverse_dynamics_gradient(q_qd, qdd); lib.forward_dynamics_gradient(x)
    lib.forward_dynamics_gradient_qdd_minv(q_qd, qdd, Minv); lib.forward_dynamics_linearization(x)
    lib.aba(x); lib.crba(q); lib.end_effector_pose(q); lib.end_effector_pose_gradient(q)
    flat = torch.zeros(B * 3 * n + 1, dtype=torch.float32, device="cuda"); flat[1:] = x.reshape(-1)
    lib.forward_dynamics_gradient(flat[1:].view(B, 3 * n))               # 4-byte aligned: plain loads
    lib.forward_dynamics_gradient_host(x.cpu().pin_memory())
    xd = x.double()
    lib.forward_dynamics_gradient(xd); lib.forward_dynamics_linearization(xd)
    torch.cuda.synchronize()
    print(robot, "ok", flush=True)
