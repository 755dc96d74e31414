#!/usr/bin/env python
"""Time the REFERENCE's own CPU path (GRiDCodeGenerator.test_fd_grad, reference _test.py:496-520: NumPy in Python) on
the shipped iiwa14 model, in the build container (the only place /root/reference exists).  BASELINE.md section 4.1.

  python tools/time_reference_numpy.py [knots_per_process]

(a) one process over the first K knots of the benchmark batch (default_rng(0), U(-1, 1), [q|qd|u]) -> knots/s/core;
(b) multiprocessing.Pool(os.cpu_count()) over K x cores knots -> knots/s on all cores of THIS container.
Not a same-box number for the GPU runs (the reference cannot travel to the GPU box); the same-box CPU arm is the C
restatement (oracle/rbd_oracle.c), which tests/test_oracle_golden.py pins to this code's outputs.
"""
import contextlib
import io
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root")
sys.path.insert(0, ROOT)

_G = None


def _gen():
    global _G
    if _G is None:
        from reference import GRiDCodeGenerator as RefGen
        from gridcodegenerator_b200.robot import load_robot
        _G = RefGen(load_robot("iiwa14"))
    return _G


def work(x):
    g = _gen()
    n = x.shape[1] // 3
    acc = 0.0
    with contextlib.redirect_stdout(io.StringIO()):
        for row in x:
            acc += float(np.abs(g.test_fd_grad(row[:n].copy(), row[n:2 * n].copy(), row[2 * n:].copy())).sum())
    return acc


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    cores = os.cpu_count()
    x = np.random.default_rng(0).uniform(-1, 1, (K * cores, 21)).astype(np.float32).astype(np.float64)
    work(x[:4])
    t0 = time.perf_counter()
    work(x[:K])
    t1 = time.perf_counter() - t0
    with mp.Pool(cores) as pool:
        pool.map(work, [x[:2]] * cores)                       # import + model construction in every worker
        t0 = time.perf_counter()
        pool.map(work, [x[i * K:(i + 1) * K] for i in range(cores)])
        tp = time.perf_counter() - t0
    print(json.dumps({"what": "reference _test.py test_fd_grad (NumPy), iiwa14, fp64", "knots_per_process": K,
                      "single_core_knots_per_s": K / t1, "ms_per_knot": 1e3 * t1 / K, "cores": cores,
                      "all_cores_knots_per_s": K * cores / tp, "where": "build container (not the GPU box)"}))


if __name__ == "__main__":
    main()
