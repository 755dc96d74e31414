"""Pin the C oracle (oracle/rbd_oracle.c) to the reference's own outputs.

tests/golden/*.npz were produced by tools/make_golden.py from the reference's
_test.py (test_rnea, test_minv, test_rnea_grad, test_fd_grad).  fp64 vs fp64: 1e-10.
"""
import numpy as np
import pytest

from conftest import rel_err
from gridcodegenerator_b200.robot import available_robots, load_robot
from oracle import Oracle

ROBOTS = available_robots()
TOL = 1e-10


@pytest.mark.parametrize("name", ROBOTS)
def test_oracle_matches_reference_golden(name, golden):
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"]
    K = x.shape[0]
    for k in range(K):
        q, qd, u, qdd = (x[k, i * n:(i + 1) * n] for i in range(4))
        c, v, a, f = orc.rnea(q, qd, qdd)
        assert rel_err([c], [g["c"][k]]) < TOL
        assert rel_err([v], [g["v"][k]]) < TOL
        assert rel_err([a], [g["a"][k]]) < TOL
        assert rel_err([f], [g["f"][k]]) < TOL
        assert rel_err([orc.rnea(q, qd, None)[0]], [g["c0"][k]]) < TOL
        assert rel_err([orc.minv(q)], [g["Minv"][k]]) < TOL
        assert rel_err([orc.rnea_grad(q, qd, qdd)], [g["dc_du"][k]]) < TOL
        assert rel_err([orc.rnea_grad(q, qd, None)], [g["dc_du0"][k]]) < TOL
        assert rel_err([orc.fd_grad(q, qd, u)], [g["df_du"][k]]) < 1e-9


@pytest.mark.parametrize("name", ROBOTS)
def test_oracle_unpinned_algorithms_self_consistent(name, golden):
    """ABA and CRBA have no executable CPU code in the reference: check them against the pinned ones."""
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"]
    for k in range(x.shape[0]):
        q, qd, u = (x[k, i * n:(i + 1) * n] for i in range(3))
        qdd_fd = orc.fd(q, qd, u)
        assert rel_err([orc.aba(q, qd, u)], [qdd_fd]) < 1e-8
        M = orc.crba(q)
        assert np.abs(M @ g["Minv"][k] - np.eye(n)).max() < 1e-7


@pytest.mark.parametrize("name", ["iiwa14", "hyq"])
def test_oracle_batch_layout(name, golden):
    """Batched drivers write the GPU layouts: column-major n x 2n per knot."""
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"][:, :3 * n]
    out = orc.batch_fd_grad(x, nthreads=2)
    ref = np.stack([m.T.reshape(-1) for m in g["df_du"]])  # column-major flatten
    assert rel_err(out, ref) < 1e-9
    out = orc.batch_minv(x, nthreads=1)
    assert rel_err(out, g["Minv"].reshape(len(x), -1)) < TOL


@pytest.mark.parametrize("name", ["pendulum2", "chain3", "branched6", "iiwa14", "hyq"])
def test_oracle_second_order_matches_reference_derivatives(name, golden):
    """IDSVA_SO / FDSVA_SO: the reference has no executable CPU code for them, so the fixtures are the exact
    (complex-step) derivatives of the reference's own test_rnea_grad / test_fd_grad / test_minv / test_rnea
    (tools/make_golden.py).  The oracle's central differences of its pinned first-order functions must agree
    to finite-difference accuracy."""
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"]
    K2 = g["so_idsva"].shape[0]
    xi = np.concatenate([x[:K2, :2 * n], x[:K2, 3 * n:4 * n]], axis=1)          # q | qd | qdd
    assert rel_err(orc.batch_idsva_so(xi), g["so_idsva"].reshape(K2, -1)) < 1e-7
    assert rel_err(orc.batch_fdsva_so(x[:K2, :3 * n]), g["so_fdsva"].reshape(K2, -1)) < 1e-6


@pytest.mark.parametrize("name", ROBOTS)
def test_oracle_end_effector_matches_sympy_golden(name, golden):
    """EE pose / gradient / Hessian of the oracle against the fixtures differentiated by sympy from the robot's
    homogeneous transforms (reference algorithms/_eepose_gradient_hessian.py:150-157 pose formulas)."""
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"]
    K = len(x)
    ne = len(robot.get_leaf_nodes())
    assert rel_err(orc.batch_ee_pose(x[:, :n]), g["ee_pose"]) < 1e-12
    grad = orc.batch_ee_pose_gradient(x[:, :n]).reshape(K, ne, n, 6).transpose(0, 1, 3, 2).reshape(K, 6 * ne, n)
    assert rel_err(grad, g["ee_grad"]) < 1e-7
    if "ee_hess" in g:
        assert rel_err(orc.batch_ee_pose_hessian(x[:, :n]), g["ee_hess"].reshape(K, -1)) < 1e-5


@pytest.mark.parametrize("name", ROBOTS)
def test_oracle_joint_transforms_and_fixed_target_match_sympy_golden(name, golden):
    """Cumulative joint transforms (reference X_single_thread / X_warp, algorithms/_eepose_gradient_hessian.py:1226-1398)
    and the pose of a fixed kinematic target (:31-38) against the sympy fixtures of tools/make_golden.py."""
    g = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    orc = Oracle(robot)
    x = g["x"]
    K = len(x)
    want = g["joint_xforms"].transpose(0, 1, 3, 2).reshape(K, 16 * n)          # column-major 4 x 4 per joint
    assert rel_err(orc.batch_joint_transforms(x[:, :n]), want) < 1e-12
    if "ft_name" in g:
        assert rel_err(orc.batch_fixed_target_pose(x[:, :n], str(g["ft_name"])), g["ft_pose"]) < 1e-12
