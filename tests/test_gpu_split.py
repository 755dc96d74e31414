"""The SPLIT (warp-specialised) form of the gradient kernels against the oracle and against the one-lane-one-knot
form (run with -m gpu on a B200).

Split form: one CTA owns one 32-knot tile; phase A (RNEA, CRBA, L^T D L, RNEA with qdd) runs on one warp and
publishes the per-knot state in shared memory, then the derivative columns run concurrently on the CTA's warps
(grid_b200_runtime.cuh: tile_loop_split; emit/algos.py: split_*).  The launchers pick it for small batches and for
robots whose one-lane-one-knot body is instruction-fetch bound; GRID_SPLIT_MODE forces either form (read per launch).
Atlas (fp32) holds TWO tiles per CTA pass once a batch has more tiles than CTAs (tile_loop_split_multi): 4500 knots =
141 tiles run one tile per pass, 4775 = 150 tiles (the last one ragged) and 4832 = 151 tiles (the last pass holds a
single tile) run two.
"""
import os

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL32 = 1e-4
TOL64 = 1e-10


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return torch


def _lib(name):
    from gridcodegenerator_b200.runtime import get_library, library_path
    if not os.path.exists(library_path(name)):
        pytest.fail(f"libgrid_{name}.so was not prebuilt by build(); refusing to compile inside a GPU test")
    return get_library(name)


def _oracle(name):
    from gridcodegenerator_b200.robot import load_robot
    from oracle import Oracle
    return Oracle(load_robot(name))


class _mode:
    def __init__(self, m):
        self.m = m

    def __enter__(self):
        self.old = os.environ.get("GRID_SPLIT_MODE")
        os.environ["GRID_SPLIT_MODE"] = self.m

    def __exit__(self, *a):
        if self.old is None:
            os.environ.pop("GRID_SPLIT_MODE", None)
        else:
            os.environ["GRID_SPLIT_MODE"] = self.old


@pytest.mark.parametrize("name,B", [("iiwa14", 1), ("iiwa14", 31), ("iiwa14", 1024), ("iiwa14", 5000), ("hyq", 515),
                                    ("hyq", 16384), ("atlas", 70), ("atlas", 4500), ("atlas", 4775), ("atlas", 4832),
                                    ("branched6", 97), ("slider3", 33), ("hyq_floating", 5000)])
def test_split_form_fp32(name, B, torch_cuda):
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    rng = np.random.default_rng(31)
    x32 = rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)
    X = torch.from_numpy(x32).cuda()
    q_qd, qdd = X[:, :2 * n].contiguous(), X[:, 2 * n:].clone().reshape(B, n)
    idx = np.unique(np.concatenate([np.arange(min(B, 256)), rng.choice(B, min(B, 256), replace=False)]))
    sel = torch.from_numpy(idx).cuda()
    x64 = x32[idx].astype(np.float64)
    f = lambda t: t[sel].cpu().numpy().astype(np.float64)
    G = 64
    res = {}
    for mode in ("1", "0"):
        with _mode(mode):
            m = 2 * n * n
            buf = torch.full((G + B * m + G,), 12345.0, dtype=torch.float32, device="cuda")
            fdg = lib.forward_dynamics_gradient(X, out=buf[G:G + B * m].view(B, m))
            torch.cuda.synchronize()
            assert bool((buf[:G] == 12345.0).all()) and bool((buf[G + B * m:] == 12345.0).all()), (mode, "guard band")
            res[mode] = (fdg.clone(), lib.inverse_dynamics_gradient(q_qd, qdd), lib.inverse_dynamics_gradient(q_qd),
                         lib.forward_dynamics_linearization(X))
    for mode, (fdg, idg, idg0, lin) in res.items():
        assert rel_err(f(fdg), orc.batch_fd_grad(x64)) < TOL32, (mode, "grad FD")
        assert rel_err(f(idg), orc.batch_rnea_grad(x64, use_qdd=True)) < TOL32, (mode, "grad ID")
        assert rel_err(f(idg0), orc.batch_rnea_grad(x64, use_qdd=False)) < TOL32, (mode, "grad ID, qdd = 0")
        ref = np.concatenate([orc.batch_fd_grad(x64), orc.batch_minv(x64), orc.batch_fd(x64)], axis=1)
        assert rel_err(f(lin), ref) < TOL32, (mode, "linearization")
    # the two forms run the same arithmetic in differently scheduled code: equal up to fp32 rounding
    for a, b in zip(res["1"], res["0"]):
        assert rel_err(a.cpu().numpy(), b.cpu().numpy()) < 2e-5
    # strided caller buffer (stride 3n + 5: plain loads) through the split form
    with _mode("1"):
        wide = torch.zeros((B, 3 * n + 5), dtype=torch.float32, device="cuda")
        wide[:, :3 * n] = X
        assert torch.equal(lib.forward_dynamics_gradient(wide[:, :3 * n]), res["1"][0])


@pytest.mark.parametrize("name,B", [("iiwa14", 257), ("hyq", 96), ("hyq", 4810), ("atlas", 40)])   # hyq 4810: 151 tiles, two per pass
def test_split_form_fp64(name, B, torch_cuda):
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    x = np.random.default_rng(32).uniform(-1, 1, (B, 3 * n))
    X = torch.from_numpy(x).cuda()
    with _mode("1"):
        assert rel_err(lib.forward_dynamics_gradient(X).cpu().numpy(), orc.batch_fd_grad(x)) < TOL64
        assert rel_err(lib.inverse_dynamics_gradient(X[:, :2 * n].contiguous(), X[:, 2 * n:].contiguous()).cpu().numpy(),
                       orc.batch_rnea_grad(x, use_qdd=True)) < TOL64
        lin = lib.forward_dynamics_linearization(X).cpu().numpy()
        ref = np.concatenate([orc.batch_fd_grad(x), orc.batch_minv(x), orc.batch_fd(x)], axis=1)
        assert rel_err(lin, ref) < TOL64


def test_launcher_picks_split_for_small_batches(torch_cuda):
    """Default policy: iiwa14 grad FD takes the split form up to <algo>_split_max_knots() knots (bit-identical to the
    forced split form there) and the one-lane-one-knot form above."""
    torch = torch_cuda
    lib = _lib("iiwa14")
    n = lib.n
    rng = np.random.default_rng(33)
    for B, want in ((1024, "1"), (65536, "0")):
        X = torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
        os.environ.pop("GRID_SPLIT_MODE", None)
        auto = lib.forward_dynamics_gradient(X)
        with _mode(want):
            assert torch.equal(lib.forward_dynamics_gradient(X), auto), B
