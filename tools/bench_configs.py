#!/usr/bin/env python
"""Secondary measurements for the BASELINE.json configs other than the headline (which bench.py owns):
device-resident inputs, CUDA events around a CUDA-graph replay of the launches, rotating buffer sets larger than L2.
One JSON line per (robot, algorithm, batch).  Run on a B200:  python tools/bench_configs.py [--quick]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # (config of BASELINE.json, robot, method, input columns (in units of n), batch)
    (1, "iiwa14", "inverse_dynamics_qdd", 3, 128),
    (2, "iiwa14", "forward_dynamics_gradient", 3, 1024),
    (2, "iiwa14", "inverse_dynamics_gradient", 2, 65536),
    (2, "iiwa14", "direct_minv", 1, 65536),
    (2, "iiwa14", "forward_dynamics", 3, 65536),
    (2, "iiwa14", "forward_dynamics_gradient", 3, 65536),
    (2, "iiwa14", "forward_dynamics_linearization", 3, 65536),   # grad FD + Minv + qdd, one launch
    (3, "hyq", "forward_dynamics_gradient", 3, 16384),
    (3, "hyq", "aba", 3, 16384),
    (4, "atlas", "forward_dynamics_gradient", 3, 8192),     # 65536 over 8 GPUs = 8192 per GPU
    (4, "atlas", "crba", 1, 8192),
    (4, "atlas", "direct_minv", 1, 8192),
    (5, "iiwa14", "idsva_so", 3, 32768),                     # 262144 over 8 GPUs = 32768 per GPU
    (5, "iiwa14", "fdsva_so", 3, 32768),
    (5, "iiwa14", "end_effector_pose_gradient_hessian", 1, 32768),
    # floating-base HyQ (six-joint chain, 18 DoF) at config 3's batch, and a small MPC-horizon batch
    (0, "hyq_floating", "forward_dynamics_gradient", 3, 16384),
    (0, "hyq_floating", "forward_dynamics_gradient", 3, 1024),
    (0, "hyq_floating", "inverse_dynamics_qdd", 3, 16384),
    (0, "hyq_floating", "forward_dynamics", 3, 16384),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--steps", type=int, default=50)
    args = ap.parse_args()
    import numpy as np
    import torch
    from gridcodegenerator_b200.runtime import get_library
    torch.cuda.set_device(0)
    L2 = 126 << 20
    for cfg, robot, method, cols, B in CASES:
        lib = get_library(robot)
        n = lib.n
        rng = np.random.default_rng(0)
        probe = torch.from_numpy(rng.uniform(-1, 1, (32, 3 * n)).astype(np.float32)).cuda()

        def call(x, out=None):
            if method == "inverse_dynamics_qdd":
                return lib.inverse_dynamics(x[:, :2 * n], x[:, 2 * n:].contiguous())
            if method == "end_effector_pose_gradient_hessian":
                return lib.end_effector_pose_gradient_hessian(x)[1]
            return getattr(lib, method)(x)
        out_elems = call(probe[:, :cols * n].contiguous() if method != "inverse_dynamics_qdd" else probe).shape[1]
        bytes_per_knot = 4 * (cols * n + out_elems)
        nsets = max(2, min(64, -(-2 * L2 // (B * bytes_per_knot))))
        xs = [torch.from_numpy(rng.uniform(-1, 1, (B, cols * n)).astype(np.float32)).cuda() for _ in range(nsets)]
        has_out = method != "end_effector_pose_gradient_hessian"      # that one returns three tensors
        outs = [torch.empty((B, out_elems), dtype=torch.float32, device="cuda") for _ in range(nsets)] if has_out else None

        def step(i):
            x = xs[i % nsets]
            if not has_out:
                return call(x)
            if method == "inverse_dynamics_qdd":
                return lib.inverse_dynamics(x[:, :2 * n], x[:, 2 * n:].contiguous(), out=outs[i % nsets])
            return getattr(lib, method)(x, out=outs[i % nsets])
        steps = 10 if args.quick else args.steps
        for i in range(max(3, nsets)):
            step(i)
        torch.cuda.synchronize()
        # the timed steps are captured in one CUDA graph (as bench.py does), so the figure is the kernels' time on
        # the stream, not the Python / ctypes launch path
        side = torch.cuda.Stream()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            step(0)
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=side):
                for i in range(steps):
                    step(i)
        g.replay()
        torch.cuda.synchronize()
        ms = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            ms = min(ms, e0.elapsed_time(e1) / steps)
        del g
        print(json.dumps({"config": cfg, "robot": robot, "algorithm": method, "batch": B, "dtype": "f32",
                          "ms_per_launch": ms, "knots_per_s": B / (ms * 1e-3),
                          "algorithmic_bytes_per_knot": bytes_per_knot,
                          "achieved_GBps": B * bytes_per_knot / (ms * 1e-3) / 1e9,
                          "note": "CUDA-graph replay of %d launches, best of 3; inputs/outputs rotate over %d buffer sets" % (steps, nsets)}),
              flush=True)


if __name__ == "__main__":
    main()
