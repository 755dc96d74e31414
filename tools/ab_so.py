#!/usr/bin/env python
"""Time the second-order kernels (iiwa14, 32768 knots, fp32) of the library under GRID_BUILD_DIR; one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from gridcodegenerator_b200.runtime import GRiDLibrary  # noqa: E402

lib = GRiDLibrary("iiwa14", build=False)
n, B = lib.n, 32768
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
out = torch.empty((B, 4 * n ** 3), dtype=torch.float32, device="cuda")
res = {"threads": {a: lib.kernel_info(a)["threads"] for a in ("idsva_so", "fdsva_so")}}
for name in ("idsva_so", "fdsva_so"):
    f = getattr(lib, name)
    f(x, out=out)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            f(x, out=out)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 3)
    res[name] = {"ms": best, "Mknots_s": B / best / 1e3, "checksum": float(out.double().abs().sum())}
print(json.dumps(res))
