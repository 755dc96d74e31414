#!/usr/bin/env python
"""A/B of emitter variants on the larger robots: for the in-tree build and every variant build under
gridcodegenerator_b200/_variants/<tag>/ that contains the robot, time grad FD and grad ID (CUDA-graph replay, CUDA
events) at the BASELINE per-GPU batch and check 64 knots against the CPU oracle.
usage: python tools/ab_robot.py atlas 8192 [hyq 16384 ...]"""
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
from gridcodegenerator_b200.robot import load_robot
from oracle import Oracle
robot, B = sys.argv[1], int(sys.argv[2])
lib = GRiDLibrary(robot, build=False)
n = lib.n
rng = np.random.default_rng(0)
x = torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
orc = Oracle(load_robot(robot))
x64 = x[:64].cpu().numpy().astype(np.float64)
def err(got, ref):
    got = got[:64].cpu().numpy().astype(np.float64)
    return float((np.abs(got - ref).max(axis=1) / np.maximum(1.0, np.abs(ref).max(axis=1))).max())
res = {"robot": robot, "batch": B}
res["fd_du_err"] = err(lib.forward_dynamics_gradient(x), orc.batch_fd_grad(x64))
res["id_du_err"] = err(lib.inverse_dynamics_gradient(x[:, :2 * n].contiguous(), x[:, 2 * n:].contiguous()), orc.batch_rnea_grad(x64, use_qdd=True))
lin = lib.forward_dynamics_linearization(x)
res["fd_lin_err"] = max(err(lin[:, :2 * n * n], orc.batch_fd_grad(x64)), err(lin[:, 2 * n * n:3 * n * n], orc.batch_minv(x64)))
def timed(fn, steps=10):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(steps): fn()
    g.replay(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / steps)
    return best * 1e3
o = torch.empty((B, 2 * n * n), dtype=torch.float32, device="cuda")
q_qd, qdd = x[:, :2 * n].contiguous(), x[:, 2 * n:].contiguous()
res["fd_du_us"] = timed(lambda: lib.forward_dynamics_gradient(x, out=o))
res["id_du_us"] = timed(lambda: lib.inverse_dynamics_gradient(q_qd, qdd, out=o))
res["info"] = lib.kernel_info("forward_dynamics_gradient")
print(json.dumps(res))
''' % ROOT


def main():
    args = sys.argv[1:]
    pairs = list(zip(args[0::2], args[1::2])) or [("atlas", "8192")]
    tags = [None] + (sorted(os.listdir(VAR)) if os.path.isdir(VAR) else [])
    for robot, B in pairs:
        for tag in tags:
            env = dict(os.environ)
            if tag is not None:
                if not os.path.exists(os.path.join(VAR, tag, robot, f"libgrid_{robot}.so")):
                    continue
                env["GRID_BUILD_DIR"] = os.path.join(VAR, tag)
            r = subprocess.run([sys.executable, "-c", CHILD, robot, B], env=env, capture_output=True, text=True)
            print(tag or "in-tree", (r.stdout.strip() or r.stderr.strip()[-600:]), flush=True)


if __name__ == "__main__":
    main()
