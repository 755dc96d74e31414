"""The allocation-free entry point grid_forward_dynamics_gradient_noalloc_f32 (include/grid_b200.h; no counterpart in the
reference, added for callers that cannot call cudaMalloc).

Its global scratch is a file-scope static __device__ array (csrc/grid_noalloc.cu) instead of the cudaMalloc'ed scratch of
a robotModel handle.  Called here through the C ABI (ctypes) on the robots whose grad-FD kernel really uses the scratch:
HyQ (transposed-store output rows) and Atlas (per-thread state), twice in a row with different inputs -- the second call
finds the first call's contents in the shared array -- and compared with the existing entry point
grid_forward_dynamics_gradient_f32 (one-lane-per-knot form) within the fp32 tolerance of the other GPU tests.

This file is collected last on purpose (a failure here must not hide other tests) and passes
both alone in a fresh process and after the rest of the suite.
"""
import ctypes
import os

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL32 = 1e-4


def _lib(name):
    from gridcodegenerator_b200.runtime import get_library, library_path
    if not os.path.exists(library_path(name)):
        pytest.fail(f"libgrid_{name}.so was not prebuilt by build(); refusing to compile inside a GPU test")
    return get_library(name)


def _bind(lib):
    fn = lib.lib.grid_forward_dynamics_gradient_noalloc_f32
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p]
    lib.lib.grid_noalloc_max_knots.restype = ctypes.c_int
    lib.lib.grid_noalloc_scratch_bytes.restype = ctypes.c_long
    return fn


@pytest.mark.parametrize("name,B", [("hyq", 5000), ("atlas", 300), ("iiwa14", 8192)])
def test_no_alloc_entry_point_matches_existing_one(name, B):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    lib = _lib(name)
    fn = _bind(lib)
    n = lib.n
    max_knots = lib.lib.grid_noalloc_max_knots()
    assert max_knots == 8192 and B <= max_knots
    assert 0 < lib.lib.grid_noalloc_scratch_bytes() <= 64 << 20
    if name != "iiwa14":      # the robots this test is about keep state / output rows in the scratch
        info = lib.kernel_info("forward_dynamics_gradient")
        assert info["state_in_smem"] == 0 or info["staged_io"] == 3, "this robot's grad FD does not use the scratch"
    rng = np.random.default_rng(41)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    old = os.environ.get("GRID_SPLIT_MODE")
    os.environ["GRID_SPLIT_MODE"] = "0"           # reference = the same one-lane-per-knot form, through the existing entry point
    try:
        for call in range(2):                     # two calls in a row share the static scratch
            x = torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
            want = lib.forward_dynamics_gradient(x)
            got = torch.full((B, 2 * n * n), float("nan"), dtype=torch.float32, device="cuda")
            rc = fn(x.data_ptr(), 3 * n, got.data_ptr(), B, 9.81, stream)
            assert rc == 0, lib.lib.grid_last_error().decode()
            torch.cuda.synchronize()
            assert bool(torch.isfinite(got).all()), (name, call)
            err = rel_err(got.cpu().numpy(), want.cpu().numpy())
            assert err < TOL32, (name, call, err)
    finally:
        if old is None:
            os.environ.pop("GRID_SPLIT_MODE", None)
        else:
            os.environ["GRID_SPLIT_MODE"] = old


def test_no_alloc_entry_point_rejects_oversize_and_bad_arguments():
    import torch
    torch.cuda.set_device(0)
    lib = _lib("hyq")
    fn = _bind(lib)
    n = lib.n
    B = lib.lib.grid_noalloc_max_knots() + 1
    x = torch.zeros((B, 3 * n), dtype=torch.float32, device="cuda")
    out = torch.zeros((B, 2 * n * n), dtype=torch.float32, device="cuda")
    rc = fn(x.data_ptr(), 3 * n, out.data_ptr(), B, 9.81, None)
    assert rc != 0
    msg = lib.lib.grid_last_error().decode()
    assert "GRID_NOALLOC_MAX_KNOTS" in msg and str(B) in msg
    assert fn(None, 3 * n, out.data_ptr(), 8, 9.81, None) != 0
    assert fn(x.data_ptr(), n, out.data_ptr(), 8, 9.81, None) != 0            # stride too small
    assert fn(x.data_ptr(), 3 * n, out.data_ptr(), 0, 9.81, None) == 0        # empty batch is a no-op
    torch.cuda.synchronize()
