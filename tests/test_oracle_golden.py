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
