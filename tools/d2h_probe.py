#!/usr/bin/env python
"""How fast can 25.7 MB (one step's grad-FD results) leave the GPU?  Pinned device->host copies in 1..16 pieces on one
stream, on three rotating streams, and with the 5.5 MB host->device input copy running concurrently."""
import json
import time

import torch

torch.cuda.set_device(0)
OUT, IN = 65536 * 392, 65536 * 84
g = torch.empty(OUT, dtype=torch.uint8, device="cuda")
h = torch.empty(OUT, dtype=torch.uint8).pin_memory()
gi = torch.empty(IN, dtype=torch.uint8, device="cuda")
hi = torch.empty(IN, dtype=torch.uint8).pin_memory()
streams = [torch.cuda.Stream() for _ in range(3)]
s_in = torch.cuda.Stream()


def step(pieces, rotate, with_h2d):
    n = OUT // pieces
    if with_h2d:
        with torch.cuda.stream(s_in):
            gi.copy_(hi, non_blocking=True)
    for p in range(pieces):
        with torch.cuda.stream(streams[p % 3 if rotate else 0]):
            h[p * n:(p + 1) * n].copy_(g[p * n:(p + 1) * n], non_blocking=True)
    torch.cuda.synchronize()


for pieces in (1, 2, 4, 8, 16):
    for rotate in (False, True):
        for with_h2d in (False, True):
            step(pieces, rotate, with_h2d)
            best = 1e9
            for _ in range(5):
                t0 = time.perf_counter()
                for _ in range(10):
                    step(pieces, rotate, with_h2d)
                best = min(best, (time.perf_counter() - t0) / 10)
            print(json.dumps({"pieces": pieces, "three_streams": rotate, "h2d_concurrent": with_h2d, "ms": round(best * 1e3, 4),
                              "d2h_GBps": round(OUT / best / 1e9, 1), "knots_per_s_bound_M": round(65536 / best / 1e6, 1)}), flush=True)
