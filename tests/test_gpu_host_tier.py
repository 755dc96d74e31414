"""The emitted grid.cuh's HOST TIER, compiled and run as user code (reference banner
GRiDCodeGenerator.py:312-380, host tier algorithms/_forward_dynamics_gradient.py:186-249).

csrc/host_tier_check.cu is a program written the way the reference documents its header:
init_robotModel / init_grid / init_gridData<T,N> -> fill hd_data->h_* -> ALGORITHM<T,FLAGS>(hd_data, d_robotModel,
gravity, N, block_dimms, thread_dimms, streams) (+ _compute_only, _single_timing, a raw X_kernel<<<>>> launch) -> read
hd_data->h_* -> close_grid.  __graft_entry__.build() compiles it against the emitted header; here it runs on the GPU
and every section of its output must equal, bit for bit, what the C ABI returns for the same inputs (both go through
the same <algo>_launch functions and the same kernels).
"""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KNOTS = 1000


@pytest.mark.parametrize("name", ["iiwa14", "hyq"])
def test_host_tier_program_matches_c_abi(name, tmp_path):
    import torch
    from gridcodegenerator_b200.runtime import get_library, host_tier_check_path
    exe = host_tier_check_path(name)
    if not os.path.exists(exe):
        pytest.fail(f"{exe} was not prebuilt by __graft_entry__.build()")
    torch.cuda.set_device(0)
    lib = get_library(name)
    n, ne, N = lib.n, lib.num_ees, KNOTS
    x = np.random.default_rng(21).uniform(-1, 1, (N, 4 * n)).astype(np.float32)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    x.tofile(fin)
    r = subprocess.run([exe, str(fin), str(fout)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Single Call FD_DU" in r.stdout                       # _single_timing printed its line
    got = np.fromfile(fout, dtype=np.float32)

    X = torch.from_numpy(x).cuda()
    q_qd_u = X[:, :3 * n].contiguous()
    q_qd = X[:, :2 * n].contiguous()
    q = X[:, :n].contiguous()
    qdd_in = X[:, 3 * n:].contiguous()
    qdd_fd = lib.forward_dynamics(q_qd_u)
    Minv = lib.direct_minv(q_qd_u)
    fdg = lib.forward_dynamics_gradient(q_qd_u)
    want = [
        lib.inverse_dynamics(q_qd_u),                                         # 0
        lib.inverse_dynamics(q_qd, qdd_in),                                   # 1
        lib.inverse_dynamics_gradient(q_qd_u, qdd_in),                        # 2
        Minv,                                                                 # 3
        qdd_fd,                                                               # 4
        fdg,                                                                  # 5
        lib.forward_dynamics_gradient_qdd_minv(q_qd_u, qdd_fd, Minv),         # 6
        fdg, fdg,                                                             # 7 (_compute_only), 8 (_single_timing)
        lib.forward_dynamics_linearization(q_qd_u),                           # 9
        lib.aba(q_qd_u),                                                      # 10
        lib.crba(q_qd_u),                                                     # 11
        lib.end_effector_pose(q_qd_u),                                        # 12
        lib.end_effector_pose_gradient(q),                                    # 13
        lib.inverse_dynamics(q_qd_u),                                         # 14: raw kernel launch
        lib.joint_transforms(q),                                              # 15
    ]
    off = 0
    for i, w in enumerate(want):
        w = w.cpu().numpy().reshape(-1)
        sec = got[off:off + w.size]
        off += w.size
        assert np.array_equal(sec, w), f"{name}: host-tier section {i} differs from the C ABI " \
                                       f"(max abs diff {np.abs(sec - w).max():.3e})"
    # 16-24: block-per-knot user kernels calling the reference-signature DEVICE tier (X_device<T>(s_out, s_q, ...,
    # d_robotModel, gravity)): the same per-thread bodies compiled into another kernel -> equal up to fp32 rounding
    device_tier = [
        lib.inverse_dynamics(q_qd),                                           # 16
        lib.inverse_dynamics(q_qd, qdd_in),                                   # 17
        Minv,                                                                 # 18
        qdd_fd,                                                               # 19
        lib.inverse_dynamics_gradient(q_qd, qdd_in),                          # 20
        fdg,                                                                  # 21
        lib.forward_dynamics_gradient_qdd_minv(q_qd, qdd_in, Minv),           # 22
        lib.end_effector_pose(q),                                             # 23
        lib.inverse_dynamics_vaf(q_qd, qdd_in),                               # 24: [v | a | f] per joint
    ]
    from conftest import rel_err
    for i, w in enumerate(device_tier):
        w = w.cpu().numpy()
        sec = got[off:off + w.size].reshape(w.shape)
        off += w.size
        assert rel_err(sec, w) < 2e-6, f"{name}: device-tier section {16 + i} differs from the C ABI ({rel_err(sec, w):.3e})"
    assert off == got.size
