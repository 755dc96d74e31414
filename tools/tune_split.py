#!/usr/bin/env python
"""Time the two kernel forms (one lane per knot vs split: phase A + columns over the CTA's warps) of the gradient
kernels over batch sizes, to place the launcher's crossover (<algo>_split_max_knots) and to record the latency floor.

  python tools/tune_split.py [robot[:algo] ...]        (default: iiwa14 hyq atlas, forward_dynamics_gradient)

CUDA-graph replay of 20 launches, CUDA events, best of 5; inputs rotate over two buffer sets.  One JSON line per case.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from gridcodegenerator_b200.runtime import get_library  # noqa: E402

SIZES = {"iiwa14": [32, 128, 1024, 4096, 8192, 16384, 32768, 65536],
         "hyq": [128, 1024, 4096, 16384, 65536],
         "atlas": [128, 1024, 8192, 65536]}
OUT = {"forward_dynamics_gradient": lambda n: 2 * n * n, "inverse_dynamics_gradient": lambda n: 2 * n * n,
       "forward_dynamics_linearization": lambda n: 3 * n * n + n}


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best * 1e3


def main():
    torch.cuda.set_device(0)
    cases = sys.argv[1:] or ["iiwa14", "hyq", "atlas"]
    for case in cases:
        robot, _, algo = case.partition(":")
        algo = algo or "forward_dynamics_gradient"
        lib = get_library(robot)
        n = lib.n
        for B in SIZES.get(robot, [1024, 16384]):
            rng = np.random.default_rng(0)
            xs = [torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda() for _ in range(2)]
            if algo == "inverse_dynamics_gradient":
                xs = [x[:, :2 * n].contiguous() for x in xs]
            outs = [torch.empty((B, OUT[algo](n)), dtype=torch.float32, device="cuda") for _ in range(2)]
            res = {}
            chk = {}
            for mode in ("0", "1"):
                os.environ["GRID_SPLIT_MODE"] = mode
                i = [0]

                def fn():
                    getattr(lib, algo)(xs[i[0] % 2], out=outs[i[0] % 2])
                    i[0] += 1
                res[mode] = timed(fn)
                chk[mode] = getattr(lib, algo)(xs[0]).double().abs().sum().item()
            os.environ.pop("GRID_SPLIT_MODE", None)
            print(json.dumps({"robot": robot, "algo": algo, "B": B, "us_lane_per_knot": round(res["0"], 2),
                              "us_split": round(res["1"], 2), "Mknots_s_lane": round(B / res["0"], 1),
                              "Mknots_s_split": round(B / res["1"], 1),
                              "checksum_rel_diff": abs(chk["0"] - chk["1"]) / max(abs(chk["0"]), 1e-30)}), flush=True)


if __name__ == "__main__":
    main()
