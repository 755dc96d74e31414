"""GPU parity tests (run with -m gpu on a B200): the sm_100a kernels, called through the C ABI, against
(a) the committed golden vectors of the reference's _test.py and (b) the pinned C oracle on seeded batches.

Tolerances (BASELINE.json north_star): 1e-4 relative in fp32, 1e-10 in fp64, metric
max_k ||out_k - ref_k||_inf / max(1, ||ref_k||_inf)  (SURVEY.md section 8(d)).
"""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL32 = 1e-4
TOL64 = 1e-10
ROBOTS = ["iiwa14", "hyq", "atlas", "branched6", "slider3", "hyq_floating"]   # hyq_floating: SURVEY 8(f4)


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return torch


def _lib(name):
    import os
    from gridcodegenerator_b200.runtime import get_library, library_path
    if not os.path.exists(library_path(name)):
        # libraries are prebuilt by __graft_entry__.build() (nvcc for the big robots takes minutes)
        pytest.fail(f"libgrid_{name}.so was not prebuilt by build(); refusing to compile inside a GPU test")
    return get_library(name)


def _oracle(name):
    from gridcodegenerator_b200.robot import load_robot
    from oracle import Oracle
    return Oracle(load_robot(name))


def _cm(M):
    return np.asarray(M).transpose(0, 2, 1).reshape(len(M), -1)


@pytest.mark.parametrize("name", ROBOTS)
@pytest.mark.parametrize("dtype,tol", [("float32", TOL32), ("float64", TOL64)])
def test_golden_vectors(name, dtype, tol, golden, torch_cuda):
    """Every first-order algorithm against the reference's own outputs (tests/golden, tools/make_golden.py).
    The fixtures' inputs are fp32-representable, so the fp32 kernels see exactly the reference's inputs and the
    north-star tolerances apply as stated: 1e-4 (fp32), 1e-10 (fp64)."""
    torch = torch_cuda
    g = golden(name)
    lib = _lib(name)
    n = lib.n
    td = getattr(torch, dtype)
    x = g["x"]
    assert np.array_equal(x.astype(np.float32).astype(np.float64), x), "fixture inputs must be fp32-representable"
    X = torch.tensor(x, dtype=td, device="cuda")
    q_qd_u = X[:, :3 * n].contiguous()
    q_qd = X[:, :2 * n].contiguous()
    qdd = X[:, 3 * n:4 * n].contiguous()
    q = X[:, :n].contiguous()
    f = lambda t: t.cpu().numpy().astype(np.float64)
    assert rel_err(f(lib.inverse_dynamics(q_qd)), g["c0"]) < tol
    assert rel_err(f(lib.inverse_dynamics(q_qd, qdd)), g["c"]) < tol
    assert rel_err(f(lib.inverse_dynamics(q_qd_u)), g["c0"]) < tol            # strided q_qd_u buffer
    # RNEA's per-joint v, a, f (the reference's s_vaf; golden arrays are (K, 6, n) -> [v | a | f], 6-vector of joint j at 6 j)
    vaf = np.concatenate([g[k_].transpose(0, 2, 1).reshape(len(x), -1) for k_ in ("v", "a", "f")], axis=1)
    assert rel_err(f(lib.inverse_dynamics_vaf(q_qd, qdd)), vaf) < tol
    assert rel_err(f(lib.direct_minv(q)), _cm(g["Minv"])) < tol
    assert rel_err(f(lib.inverse_dynamics_gradient(q_qd)), _cm(g["dc_du0"])) < tol
    assert rel_err(f(lib.inverse_dynamics_gradient(q_qd, qdd)), _cm(g["dc_du"])) < tol
    K = len(x)
    qdd_fd = np.stack([g["Minv"][k] @ (x[k, 2 * n:3 * n] - g["c0"][k]) for k in range(K)])
    assert rel_err(f(lib.forward_dynamics(q_qd_u)), qdd_fd) < tol
    assert rel_err(f(lib.forward_dynamics_gradient(q_qd_u)), _cm(g["df_du"])) < tol
    Minv_t = torch.tensor(_cm(g["Minv"]), dtype=td, device="cuda")
    qdd_t = torch.tensor(qdd_fd, dtype=td, device="cuda")
    out = lib.forward_dynamics_gradient_qdd_minv(q_qd, qdd_t, Minv_t)
    assert rel_err(f(out), _cm(g["df_du"])) < tol
    # end-effector pose and gradient against the sympy-derived fixtures (exact derivatives)
    ne = lib.num_ees
    assert rel_err(f(lib.end_effector_pose(q)), g["ee_pose"]) < tol
    grad = f(lib.end_effector_pose_gradient(q)).reshape(K, ne, n, 6).transpose(0, 1, 3, 2).reshape(K, 6 * ne, n)
    gtol = tol if dtype == "float64" else 1e-3      # RPY derivatives grow like 1/cos(pitch)
    assert rel_err(grad, g["ee_grad"]) < gtol
    if lib.has_ee_hessian and "ee_hess" in g:
        _, hess = lib.end_effector_pose_gradient_hessian(q)
        if dtype == "float64":
            assert rel_err(f(hess), g["ee_hess"].reshape(K, -1)) < 1e-9
    # cumulative joint transforms (reference X_single_thread / X_warp) and the fixed kinematic target's EE functions
    want = g["joint_xforms"].transpose(0, 1, 3, 2).reshape(K, 16 * n)
    assert rel_err(f(lib.joint_transforms(q)), want) < (1e-12 if dtype == "float64" else 1e-5)
    if "ft_name" in g:
        assert lib.fixed_target == str(g["ft_name"])
        assert rel_err(f(lib.fixed_target_pose(q)), g["ft_pose"]) < tol
        fg, fh = lib.fixed_target_pose_gradient_hessian(q)
        fgrad = f(fg).reshape(K, n, 6).transpose(0, 2, 1)
        assert rel_err(fgrad, g["ft_grad"]) < gtol
        if dtype == "float64":
            assert rel_err(f(fh), g["ft_hess"].reshape(K, -1)) < 1e-9
    else:
        assert lib.fixed_target == ""
        with pytest.raises(RuntimeError, match="fixed kinematic target"):
            lib.fixed_target_pose(q)


@pytest.mark.parametrize("name", ["iiwa14", "branched6", "hyq"])
def test_second_order_golden(name, golden, torch_cuda):
    """IDSVA_SO / FDSVA_SO against the exact (complex-step) derivatives of the reference's own first-order code
    (tools/make_golden.py): fp64 kernel 1e-9 (the fixtures differentiate test_fd_grad, whose Minv amplifies
    rounding), fp32 kernel 1e-4 relative to the tensor's own scale per knot."""
    torch = torch_cuda
    g = golden(name)
    lib = _lib(name)
    if not lib.has_second_order:
        pytest.skip(f"{name}: library built without second-order kernels")
    n = lib.n
    x = g["x"]
    K2 = g["so_idsva"].shape[0]
    xi = np.concatenate([x[:K2, :2 * n], x[:K2, 3 * n:4 * n]], axis=1)
    for dtype, tol in (("float64", 1e-9), ("float32", TOL32)):
        td = getattr(torch, dtype)
        got = lib.idsva_so(torch.tensor(xi, dtype=td, device="cuda")).cpu().numpy().astype(np.float64)
        assert rel_err(got, g["so_idsva"].reshape(K2, -1)) < tol, (name, dtype, "idsva_so")
        got = lib.fdsva_so(torch.tensor(x[:K2, :3 * n], dtype=td, device="cuda")).cpu().numpy().astype(np.float64)
        assert rel_err(got, g["so_fdsva"].reshape(K2, -1)) < tol, (name, dtype, "fdsva_so")


@pytest.mark.parametrize("name,B", [("iiwa14", 128), ("iiwa14", 1024), ("iiwa14", 1000), ("hyq", 515), ("atlas", 70), ("hyq_floating", 515),
                                    ("branched6", 33), ("slider3", 31)])
def test_fd_grad_vs_oracle_fp32(name, B, torch_cuda):
    """BASELINE configs 1-2 sizes + ragged batches: fp32 kernel vs fp64 oracle on the fp32-rounded inputs."""
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    rng = np.random.default_rng(1)
    x32 = rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)
    x64 = x32.astype(np.float64)
    X = torch.from_numpy(x32).cuda()
    f = lambda t: t.cpu().numpy().astype(np.float64)
    assert rel_err(f(lib.forward_dynamics_gradient(X)), orc.batch_fd_grad(x64)) < TOL32
    assert rel_err(f(lib.inverse_dynamics(X[:, :2 * n].contiguous(), X[:, 2 * n:].contiguous())),
                   orc.batch_rnea(x64, use_qdd=True)) < TOL32
    assert rel_err(f(lib.direct_minv(X[:, :n].contiguous())), orc.batch_minv(x64)) < TOL32
    assert rel_err(f(lib.forward_dynamics(X)), orc.batch_fd(x64)) < TOL32
    assert rel_err(f(lib.inverse_dynamics_gradient(X[:, :2 * n].contiguous(), X[:, 2 * n:].contiguous())),
                   orc.batch_rnea_grad(x64, use_qdd=True)) < TOL32


@pytest.mark.parametrize("name,B", [("iiwa14", 1024), ("iiwa14", 1000), ("hyq", 515), ("atlas", 70), ("branched6", 33), ("hyq_floating", 97),
                                    ("slider3", 31)])
@pytest.mark.parametrize("dtype,tol", [("float32", TOL32), ("float64", TOL64)])
def test_fd_linearization_vs_oracle(name, B, dtype, tol, torch_cuda):
    """grad FD + Minv + qdd from one launch (include/grid_b200.h
    grid_forward_dynamics_linearization): every block against the pinned oracle functions, the gradient block
    equal to the stand-alone grad FD kernel, Minv exactly symmetric, host entry point == device one."""
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    rng = np.random.default_rng(7)
    x = rng.uniform(-1, 1, (B, 3 * n)).astype(dtype)
    x64 = x.astype(np.float64)
    X = torch.from_numpy(x).cuda()
    out = lib.forward_dynamics_linearization(X)
    assert out.shape == (B, 3 * n * n + n)
    df_du, Minv, qdd = lib.forward_dynamics_linearization(X, split=True)
    f = lambda t: t.cpu().numpy().astype(np.float64)
    assert rel_err(f(df_du), orc.batch_fd_grad(x64)) < tol
    assert rel_err(f(Minv), orc.batch_minv(x64)) < tol
    assert rel_err(f(qdd), orc.batch_fd(x64)) < tol
    # same algorithm in another kernel: equal up to rounding (the two bodies are scheduled / contracted separately)
    assert rel_err(f(df_du), f(lib.forward_dynamics_gradient(X))) < tol * 0.1
    Mm = Minv.reshape(B, n, n)
    assert torch.equal(Mm, Mm.transpose(1, 2))
    h = lib.forward_dynamics_linearization_host(torch.from_numpy(x).pin_memory())
    assert torch.equal(h, out.cpu())
    assert lib.forward_dynamics_linearization(X[:0]).shape == (0, 3 * n * n + n)


@pytest.mark.parametrize("name,B", [("iiwa14", 1024), ("hyq", 16384), ("atlas", 200), ("branched6", 97), ("slider3", 64),
                                    ("hyq_floating", 300)])
def test_aba_crba_fp32(name, B, torch_cuda):
    """ABA and CRBA (reference algorithms/_aba.py, _crba.py; BASELINE configs 3 and 4).  The reference has no
    executable CPU code for them ("parity unpinned"): the oracle's textbook versions are checked
    against the pinned functions in tests/test_oracle_golden.py, and here additionally ABA == FD kernel and
    CRBA * Minv == I with the library's own (pinned) Minv."""
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    rng = np.random.default_rng(7)
    x32 = rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)
    X = torch.from_numpy(x32).cuda()
    f = lambda t: t.cpu().numpy().astype(np.float64)
    sub = np.unique(np.concatenate([np.arange(min(B, 256)), rng.choice(B, min(B, 256), replace=False)]))
    x64 = x32[sub].astype(np.float64)
    qdd = lib.aba(X)
    assert rel_err(f(qdd)[sub], orc.batch_aba(x64)) < TOL32
    assert rel_err(f(qdd), f(lib.forward_dynamics(X))) < TOL32
    M = lib.crba(X[:, :n].contiguous())
    assert rel_err(f(M)[sub], orc.batch_crba(x64)) < TOL32
    assert torch.equal(lib.crba(X), M)                                          # strided q_qd_u buffer
    Minv = lib.direct_minv(X[:, :n].contiguous()).view(B, n, n).double()
    eye = torch.eye(n, dtype=torch.float64, device="cuda")
    assert float((M.view(B, n, n).double() @ Minv - eye).abs().max()) < 2e-3


@pytest.mark.parametrize("name,B", [("iiwa14", 300), ("branched6", 70)])
def test_second_order_fp32_fp64(name, B, torch_cuda):
    """IDSVA_SO / FDSVA_SO (reference algorithms/_idsva_so.py, _fdsva_so.py; BASELINE config 5).  Parity is
    UNPINNED by the reference (no executable CPU code upstream); the oracle is fp64 central differences of the
    pinned first-order functions (oracle/rbd_oracle.c: rbd_idsva_so, rbd_fdsva_so), good to ~1e-8, so:
    fp64 kernel vs oracle 1e-6, fp32 kernel vs fp64 kernel 1e-4."""
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    assert lib.has_second_order
    n = lib.n
    x = np.random.default_rng(11).uniform(-1, 1, (B, 3 * n)).astype(np.float32).astype(np.float64)
    X64 = torch.from_numpy(x).cuda()
    X32 = X64.float()
    sub = np.arange(0, B, 7)
    for fn, ofn in (("idsva_so", orc.batch_idsva_so), ("fdsva_so", orc.batch_fdsva_so)):
        o64 = getattr(lib, fn)(X64).cpu().numpy()
        o32 = getattr(lib, fn)(X32).cpu().numpy().astype(np.float64)
        assert o64.shape == (B, 4 * n ** 3)
        assert rel_err(o64[sub], ofn(x[sub])) < 1e-6, fn
        assert rel_err(o32, o64) < TOL32, fn
    # symmetry of the pure second derivatives
    T = lib.fdsva_so(X64).view(B, 4, n, n, n)
    assert float((T[:, 0] - T[:, 0].transpose(2, 3)).abs().max()) == 0.0
    assert float((T[:, 1] - T[:, 1].transpose(2, 3)).abs().max()) == 0.0
    # d2qdd/dtau dq is the derivative of Minv: contracting with tau-independent checks is covered by the oracle


@pytest.mark.parametrize("name,B", [("iiwa14", 1000), ("hyq", 200), ("atlas", 64), ("branched6", 50)])
def test_end_effector_pose_gradient_hessian(name, B, torch_cuda):
    """EE pose / gradient / Hessian (reference algorithms/_eepose_gradient_hessian.py; BASELINE config 5) against
    the oracle's restatement of the chain product + RPY formulas and its central differences (parity unpinned
    upstream).  fp64 kernel vs oracle: pose 1e-10, gradient 1e-7, Hessian 1e-5 (finite-difference accuracy);
    fp32 kernel vs fp64 kernel 1e-4 for the pose / gradient, Hessian on knots away from the RPY singularity."""
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n, ne = lib.n, lib.num_ees
    x = np.random.default_rng(12).uniform(-1, 1, (B, n)).astype(np.float32).astype(np.float64)
    X64 = torch.from_numpy(x).cuda()
    X32 = X64.float()
    sub = np.arange(0, B, 5)
    p64 = lib.end_effector_pose(X64)
    assert lib.has_ee_hessian        # every shipped robot, Atlas included (27 000 outputs per knot, plain stores)
    g64, h64 = lib.end_effector_pose_gradient_hessian(X64)
    assert p64.shape == (B, 6 * ne) and g64.shape == (B, 6 * ne * n) and h64.shape == (B, 6 * ne * n * n)
    assert rel_err(p64.cpu().numpy()[sub], orc.batch_ee_pose(x[sub])) < TOL64
    assert rel_err(g64.cpu().numpy()[sub], orc.batch_ee_pose_gradient(x[sub])) < 1e-7
    # the oracle's Hessian is a second central difference (~1e-5 accurate near the RPY singularity); the exact Hessian
    # is pinned at 1e-9 against the sympy fixtures in test_golden_vectors
    assert rel_err(h64.cpu().numpy()[sub], orc.batch_ee_pose_hessian(x[sub])) < 1e-4
    assert torch.equal(lib.end_effector_pose_gradient(X64), g64)
    H = h64.view(B, ne * 6, n, n)
    assert float((H - H.transpose(2, 3)).abs().max()) == 0.0
    p32 = lib.end_effector_pose(X32)
    g32, h32 = lib.end_effector_pose_gradient_hessian(X32)
    assert rel_err(p32.cpu().numpy(), p64.cpu().numpy()) < TOL32
    # RPY derivatives grow like 1/cos(pitch)^k: compare relative to the Hessian's own scale per knot
    assert rel_err(g32.cpu().numpy(), g64.cpu().numpy()) < 1e-3
    hs = h64.abs().amax(dim=1).clamp(min=1.0)
    assert float(((h32.double() - h64).abs().amax(dim=1) / hs).median()) < TOL32


def test_second_order_absent_for_large_robots(torch_cuda):
    torch = torch_cuda
    assert _lib("hyq").has_second_order          # 12 joints but four 3-joint trees: generated
    lib = _lib("atlas")                           # largest base-rooted tree has 18 joints: not generated
    assert not lib.has_second_order
    x = torch.zeros((4, 3 * lib.n), dtype=torch.float32, device="cuda")
    with pytest.raises(RuntimeError, match="second-order"):
        lib.idsva_so(x)


@pytest.mark.parametrize("name,B", [("iiwa14", 130), ("hyq", 64), ("atlas", 33)])
def test_aba_crba_fp64(name, B, torch_cuda):
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    x = np.random.default_rng(8).uniform(-1, 1, (B, 3 * n))
    X = torch.from_numpy(x).cuda()
    assert rel_err(lib.aba(X).cpu().numpy(), orc.batch_aba(x)) < TOL64
    assert rel_err(lib.crba(X[:, :n].contiguous()).cpu().numpy(), orc.batch_crba(x)) < TOL64


@pytest.mark.parametrize("name,B", [("iiwa14", 257), ("hyq", 96), ("atlas", 40), ("hyq_floating", 64)])
def test_fd_grad_vs_oracle_fp64(name, B, torch_cuda):
    torch = torch_cuda
    lib, orc = _lib(name), _oracle(name)
    n = lib.n
    rng = np.random.default_rng(2)
    x = rng.uniform(-1, 1, (B, 3 * n))
    X = torch.from_numpy(x).cuda()
    assert rel_err(lib.forward_dynamics_gradient(X).cpu().numpy(), orc.batch_fd_grad(x)) < TOL64
    assert rel_err(lib.forward_dynamics(X).cpu().numpy(), orc.batch_fd(x)) < TOL64
    assert rel_err(lib.direct_minv(X[:, :n].contiguous()).cpu().numpy(), orc.batch_minv(x)) < TOL64


def test_edge_cases(torch_cuda):
    """Empty batch, single knot, strided / unaligned caller buffers (non-TMA path) give the same numbers."""
    torch = torch_cuda
    lib = _lib("iiwa14")
    n = lib.n
    empty = torch.empty((0, 3 * n), dtype=torch.float32, device="cuda")
    assert lib.forward_dynamics_gradient(empty).shape == (0, 2 * n * n)
    rng = np.random.default_rng(3)
    x = torch.from_numpy(rng.uniform(-1, 1, (200, 3 * n)).astype(np.float32)).cuda()
    ref = lib.forward_dynamics_gradient(x)
    one = lib.forward_dynamics_gradient(x[:1].contiguous())
    assert torch.equal(one, ref[:1])
    wide = torch.zeros((200, 3 * n + 5), dtype=torch.float32, device="cuda")
    wide[:, :3 * n] = x
    assert torch.equal(lib.forward_dynamics_gradient(wide[:, :3 * n]), ref)       # stride 3n+5 -> plain loads
    flat = torch.zeros(200 * 3 * n + 1, dtype=torch.float32, device="cuda")
    flat[1:] = x.reshape(-1)
    assert torch.equal(lib.forward_dynamics_gradient(flat[1:].view(200, 3 * n)), ref)  # 4-byte aligned only
    with pytest.raises(RuntimeError):
        lib.forward_dynamics_gradient(x.cpu())                                     # no CPU fallback


def test_full_size_properties(torch_cuda):
    """BASELINE metric size (iiwa14, 65536 knots): size-independent checks.
    (1) sub-batch consistency: any slice recomputed alone is bit-identical (knots are independent);
    (2) df_du == -Minv dc_du with the library's own Minv / grad ID at the FD solution;
    (3) a seeded 1024-knot subsample against the oracle; (4) host-buffer entry point == device entry point."""
    torch = torch_cuda
    lib, orc = _lib("iiwa14"), _oracle("iiwa14")
    n, B = lib.n, 65536
    rng = np.random.default_rng(0)
    x32 = rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)
    X = torch.from_numpy(x32).cuda()
    out = lib.forward_dynamics_gradient(X)
    assert torch.isfinite(out).all()
    sl = slice(12345, 12345 + 777)
    import os
    os.environ["GRID_SPLIT_MODE"] = "0"          # same kernel form as the full batch (small batches default to split)
    try:
        assert torch.equal(lib.forward_dynamics_gradient(X[sl].contiguous()), out[sl])
    finally:
        os.environ.pop("GRID_SPLIT_MODE", None)
    qdd = lib.forward_dynamics(X)
    Minv = lib.direct_minv(X[:, :n].contiguous()).view(B, n, n)                  # symmetric
    dc = lib.inverse_dynamics_gradient(X[:, :2 * n].contiguous(), qdd).view(B, 2 * n, n).transpose(1, 2)
    recon = -(Minv.double() @ dc.double())
    got = out.view(B, 2 * n, n).transpose(1, 2).double()
    den = got.abs().amax(dim=(1, 2)).clamp(min=1.0)
    assert float(((recon - got).abs().amax(dim=(1, 2)) / den).max()) < TOL32
    idx = np.concatenate([np.arange(512), rng.choice(B, 512, replace=False)])
    ref = orc.batch_fd_grad(x32[idx].astype(np.float64))
    assert rel_err(out[torch.from_numpy(idx).cuda()].cpu().numpy().astype(np.float64), ref) < TOL32
    h_in = torch.from_numpy(x32).pin_memory()
    h_out = lib.forward_dynamics_gradient_host(h_in)
    assert torch.equal(h_out, out.cpu())


@pytest.mark.parametrize("env", [{"GRID_HOST_ZEROCOPY": "1"}, {"GRID_HOST_CHUNK": "1000"}, {"GRID_PDL": "0"}])
def test_host_transports_match_device_path(env, torch_cuda):
    """The host-buffer entry point's transports (zero-copy over mapped pinned memory, odd chunk sizes) and the
    launch without programmatic dependent launch are read from the environment once per process, so each runs in
    a child process (tools/e2e_sweep.py) and must reproduce the device path bit for bit."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ, E2E_BATCH="5000", **env)
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "e2e_sweep.py")], env=e, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["matches_device_path"] is True and line["batch"] == 5000


@pytest.mark.parametrize("name,B", [("iiwa14", 1), ("iiwa14", 33), ("iiwa14", 1000), ("hyq", 97), ("atlas", 40), ("hyq_floating", 70)])
def test_output_guard_bands(name, B, torch_cuda):
    """Own bounds check (compute-sanitizer is closed on this pool): every output-writing path (TMA bulk stores,
    ragged-tail plain stores, transposed stores, global-scratch kernels) writes exactly its (B, m) rows -- guard
    words placed directly before and after the output rows stay untouched, and the rows equal an unguarded run."""
    torch = torch_cuda
    lib = _lib(name)
    n = lib.n
    rng = np.random.default_rng(11)
    x = torch.from_numpy(rng.uniform(-1, 1, (B, 3 * n)).astype(np.float32)).cuda()
    G = 64     # guard elements (256 B: keeps the rows 16-byte aligned so the TMA path is the one exercised)
    cases = [("forward_dynamics_gradient", x, 2 * n * n), ("forward_dynamics_linearization", x, 3 * n * n + n),
             ("inverse_dynamics_gradient", x[:, :2 * n].contiguous(), 2 * n * n), ("forward_dynamics", x, n),
             ("crba", x[:, :n].contiguous(), n * n), ("aba", x, n)]
    for method, inp, m in cases:
        buf = torch.full((G + B * m + G,), 12345.0, dtype=torch.float32, device="cuda")
        out = buf[G:G + B * m].view(B, m)
        got = getattr(lib, method)(inp, out=out)
        torch.cuda.synchronize()
        assert got.data_ptr() == out.data_ptr()
        assert bool((buf[:G] == 12345.0).all()) and bool((buf[G + B * m:] == 12345.0).all()), method
        assert torch.equal(out, getattr(lib, method)(inp)), method
