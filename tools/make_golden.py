#!/usr/bin/env python
"""Generate tests/golden/<robot>.npz by running the REFERENCE's own CPU code.

Runs only in the build container (needs /root/reference); the .npz files are
committed so the GPU box, which has no /root/reference, can still pin the
oracle and the CUDA path to the reference.

For each shipped robot and a few seeded knot points it records the outputs of
the reference's `_test.py` methods (bound on `GRiDCodeGenerator`,
reference GRiDCodeGenerator.py:50-51):
  test_rnea (_test.py:109-115), test_minv (:213-226), test_rnea_grad (:490-494),
  test_fd_grad (:496-520).
Inputs: numpy default_rng(seed), U(-1, 1), fp64; GRAVITY = -9.81 (== kernel gravity 9.81).
"""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root")
sys.path.insert(0, ROOT)

from reference import GRiDCodeGenerator as RefGen  # noqa: E402

from gridcodegenerator_b200.robot import available_robots, load_robot  # noqa: E402

KNOTS = {"atlas": 4}
DEFAULT_KNOTS = 8


def main():
    outdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    for name in available_robots():
        robot = load_robot(name)
        g = RefGen(robot)
        n = robot.get_num_pos()
        K = KNOTS.get(name, DEFAULT_KNOTS)
        rng = np.random.default_rng(1234)
        x = rng.uniform(-1, 1, (K, 4 * n))  # q | qd | u | qdd
        rec = dict(x=x, c=[], c0=[], v=[], a=[], f=[], Minv=[], dc_du=[], dc_du0=[], df_du=[])
        for k in range(K):
            q, qd, u, qdd = (x[k, i * n:(i + 1) * n].copy() for i in range(4))
            with contextlib.redirect_stdout(io.StringIO()):
                c, v, a, f = g.test_rnea(q, qd, qdd)
                c0 = g.test_rnea(q, qd, None)[0]
                Minv = g.test_minv(q, True)
                dc_du = g.test_rnea_grad(q, qd, qdd)
                dc_du0 = g.test_rnea_grad(q, qd, None)
                df_du = g.test_fd_grad(q, qd, u)
            rec["c"].append(np.array(c, dtype=np.float64))
            rec["c0"].append(np.array(c0, dtype=np.float64))
            rec["v"].append(np.array(v, dtype=np.float64))
            rec["a"].append(np.array(a, dtype=np.float64))
            rec["f"].append(np.array(f, dtype=np.float64))
            rec["Minv"].append(np.array(Minv, dtype=np.float64))
            rec["dc_du"].append(np.array(dc_du, dtype=np.float64))
            rec["dc_du0"].append(np.array(dc_du0, dtype=np.float64))
            rec["df_du"].append(np.array(df_du, dtype=np.float64))
        out = {k: np.stack(v) if isinstance(v, list) else v for k, v in rec.items()}
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **out)
        print(name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
