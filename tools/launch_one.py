#!/usr/bin/env python
"""Launch ONE algorithm of ONE robot a few times (for ncu captures; nothing is timed here).

  python tools/launch_one.py <robot> <algorithm> <batch> [--split 0|1] [--reps 4] [--f64]

algorithm: any GRiDLibrary method taking a (B, k n) tensor (forward_dynamics_gradient, inverse_dynamics_gradient,
forward_dynamics_linearization, idsva_so, fdsva_so, crba, direct_minv, aba, forward_dynamics, ...).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

COLS = {"inverse_dynamics_gradient": 2, "inverse_dynamics": 2, "direct_minv": 1, "crba": 1, "end_effector_pose": 1,
        "end_effector_pose_gradient": 1, "joint_transforms": 1}


def main():
    robot, algo, B = sys.argv[1], sys.argv[2], int(sys.argv[3])
    if "--split" in sys.argv:
        os.environ["GRID_SPLIT_MODE"] = sys.argv[sys.argv.index("--split") + 1]
    reps = int(sys.argv[sys.argv.index("--reps") + 1]) if "--reps" in sys.argv else 4
    import numpy as np
    import torch
    from gridcodegenerator_b200.runtime import get_library
    torch.cuda.set_device(0)
    lib = get_library(robot)
    n = lib.n
    dt = np.float64 if "--f64" in sys.argv else np.float32
    x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, COLS.get(algo, 3) * n)).astype(dt)).cuda()
    for _ in range(reps):
        out = getattr(lib, algo)(x)
    torch.cuda.synchronize()
    o = out[0] if isinstance(out, tuple) else out
    print(f"{robot} {algo} B={B}: out {tuple(o.shape)} finite={bool(torch.isfinite(o).all())}")


if __name__ == "__main__":
    main()
