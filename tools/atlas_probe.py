#!/usr/bin/env python
"""Launch the Atlas grad FD / grad ID kernels a few times at 8192 knots (for an ncu metrics pass)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gridcodegenerator_b200.runtime import GRiDLibrary
robot = sys.argv[1] if len(sys.argv) > 1 else "atlas"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
lib = GRiDLibrary(robot, build=False)
n = lib.n
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
for _ in range(3):
    lib.forward_dynamics_gradient(x)
    lib.inverse_dynamics_gradient(x[:, :2 * n].contiguous())
torch.cuda.synchronize()
print("ok")
