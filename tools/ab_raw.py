#!/usr/bin/env python
"""A/B a variant libgrid_<robot>.so against the in-tree library through the raw C ABI (ctypes only: works for variant
binaries that predate newer entry points).  Times grid_forward_dynamics_gradient_f32 (or AB_FN=<entry point with the
same signature>, e.g. grid_idsva_so_f32) in both kernel forms.

  python tools/ab_raw.py <robot> <variant.so> [batch ...]
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from gridcodegenerator_b200.runtime import library_path  # noqa: E402


FN = os.environ.get("AB_FN", "grid_forward_dynamics_gradient_f32")   # any entry point with the grad-FD signature


def bind(path):
    L = ctypes.CDLL(path)
    getattr(L, FN).argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                                     ctypes.c_float, ctypes.c_void_p]
    L.grid_last_error.restype = ctypes.c_char_p
    return L


def timed(fn, reps=10):
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
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best * 1e3


def main():
    robot, var = sys.argv[1], sys.argv[2]
    sizes = [int(a) for a in sys.argv[3:]] or [1024, 8192, 65536]
    torch.cuda.set_device(0)
    libs = {"in_tree": bind(library_path(robot)), "variant": bind(var)}
    n = libs["in_tree"].grid_num_joints()
    for B in sizes:
        x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
        row = {"robot": robot, "B": B}
        outs = {}
        for tag, L in libs.items():
            for mode in ("0", "1"):
                os.environ["GRID_SPLIT_MODE"] = mode
                out = torch.empty((B, 4 * n ** 3 if "_so_" in FN else 2 * n * n), dtype=torch.float32, device="cuda")

                def fn():
                    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
                    rc = getattr(L, FN)(x.data_ptr(), 3 * n, out.data_ptr(), B, 9.81, st)
                    assert rc == 0, L.grid_last_error()
                row[f"{tag}_{'split' if mode == '1' else 'lane'}_us"] = round(timed(fn), 2)
                outs[(tag, mode)] = out.clone()
        ref = outs[("in_tree", "0")].double()
        den = ref.abs().amax(dim=1).clamp(min=1.0)
        for k, o in outs.items():
            row[f"maxrel_{k[0]}_{'split' if k[1] == '1' else 'lane'}"] = float(((o.double() - ref).abs().amax(dim=1) / den).max())
        print(json.dumps(row), flush=True)
    os.environ.pop("GRID_SPLIT_MODE", None)


if __name__ == "__main__":
    main()
