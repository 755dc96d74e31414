"""The emitted straight-line bodies, compiled for the host in fp64, must reproduce the reference's golden
vectors (tests/golden, from reference _test.py) to 1e-10: this pins the emitter's algebra (factored
transforms, per-column Minv, column-wise grad RNEA) independently of the GPU."""
import numpy as np
import pytest

from conftest import rel_err
from emit_host_harness import build_host_inners, call
from gridcodegenerator_b200.robot import available_robots, load_robot

TOL = 1e-10


@pytest.mark.parametrize("name", available_robots())
def test_emitted_bodies_match_reference_golden(name, golden, tmp_path):
    g_ = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    lib, gen = build_host_inners(robot, str(tmp_path))
    x = g_["x"]
    for k in range(x.shape[0]):
        q, qd, u, qdd = (x[k, i * n:(i + 1) * n] for i in range(4))
        cm = lambda M: np.asarray(M).T.reshape(-1)  # column-major flatten
        assert rel_err([call(lib, gen, "inverse_dynamics", n, q, qd)], [g_["c0"][k]]) < TOL
        assert rel_err([call(lib, gen, "inverse_dynamics_qdd", n, q, qd, qdd)], [g_["c"][k]]) < TOL
        assert rel_err([call(lib, gen, "direct_minv", n * n, q)], [cm(g_["Minv"][k])]) < TOL
        assert rel_err([call(lib, gen, "inverse_dynamics_gradient", 2 * n * n, q, qd)], [cm(g_["dc_du0"][k])]) < TOL
        assert rel_err([call(lib, gen, "inverse_dynamics_gradient_qdd", 2 * n * n, q, qd, qdd)], [cm(g_["dc_du"][k])]) < TOL
        assert rel_err([call(lib, gen, "forward_dynamics_gradient", 2 * n * n, q, qd, u)], [cm(g_["df_du"][k])]) < 1e-9
        Minv = g_["Minv"][k]
        qdd_fd = Minv @ (u - g_["c0"][k])
        assert rel_err([call(lib, gen, "forward_dynamics", n, q, qd, u)], [qdd_fd]) < 1e-9
        # qdd+Minv input mode: feeding the FD solution reproduces grad FD
        out = call(lib, gen, "forward_dynamics_gradient_qdd_minv", 2 * n * n, q, qd, qdd_fd, cm(Minv))
        assert rel_err([out], [cm(g_["df_du"][k])]) < 1e-9
