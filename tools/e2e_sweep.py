#!/usr/bin/env python
"""End-to-end (host buffers in, host buffers out) throughput of grid_forward_dynamics_gradient_host_f32 for one
transport setting, read from the environment by the library at first use (GRID_HOST_ZEROCOPY, GRID_HOST_CHUNK):
run once per setting, e.g.  GRID_HOST_ZEROCOPY=1 python tools/e2e_sweep.py.  One JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    from gridcodegenerator_b200.runtime import get_library
    torch.cuda.set_device(0)
    robot = os.environ.get("E2E_ROBOT", "iiwa14")
    lib = get_library(robot)
    n, B = lib.n, int(os.environ.get("E2E_BATCH", "65536" if robot == "iiwa14" else "16384"))
    rng = np.random.default_rng(0)
    h_in = torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).pin_memory()
    h_out = torch.empty((B, 2 * n * n), dtype=torch.float32).pin_memory()
    ref = lib.forward_dynamics_gradient(h_in.cuda()).cpu()
    for _ in range(5):
        lib.forward_dynamics_gradient_host(h_in, h_out)
    ok = bool(torch.equal(h_out, ref))
    steps = 50
    best = 1e9
    for rep in range(3):
        t0 = time.perf_counter()
        for _ in range(steps):
            lib.forward_dynamics_gradient_host(h_in, h_out)
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) / steps)
    print(json.dumps({"robot": robot, "zero_copy": os.environ.get("GRID_HOST_ZEROCOPY", "0"), "chunk": "doubling from 2048" if os.environ.get("GRID_HOST_CHUNK_DOUBLING") == "1" else os.environ.get("GRID_HOST_CHUNK", "8192"),
                      "batch": B, "ms_per_step": best * 1e3, "knots_per_s": B / best,
                      "pcie_GBps": B * (3 * n + 2 * n * n) * 4 / best / 1e9, "matches_device_path": ok}), flush=True)


if __name__ == "__main__":
    main()
