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
        # RNEA's per-joint v, a, f: the reference's s_vaf layout [v | a | f], 6-vector of joint j at 6 j (golden: (6, n))
        vaf = np.concatenate([np.asarray(g_[key][k]).T.reshape(-1) for key in ("v", "a", "f")])
        assert rel_err([call(lib, gen, "inverse_dynamics_vaf_qdd", 18 * n, q, qd, qdd)], [vaf]) < TOL
        assert rel_err([call(lib, gen, "direct_minv", n * n, q)], [cm(g_["Minv"][k])]) < TOL
        assert rel_err([call(lib, gen, "inverse_dynamics_gradient", 2 * n * n, q, qd)], [cm(g_["dc_du0"][k])]) < TOL
        assert rel_err([call(lib, gen, "inverse_dynamics_gradient_qdd", 2 * n * n, q, qd, qdd)], [cm(g_["dc_du"][k])]) < TOL
        assert rel_err([call(lib, gen, "forward_dynamics_gradient", 2 * n * n, q, qd, u)], [cm(g_["df_du"][k])]) < 1e-9
        Minv = g_["Minv"][k]
        qdd_fd = Minv @ (u - g_["c0"][k])
        assert rel_err([call(lib, gen, "forward_dynamics", n, q, qd, u)], [qdd_fd]) < 1e-9
        # ABA == FD and CRBA == inv(Minv) (no executable reference code for these two: SURVEY.md 8(c))
        assert rel_err([call(lib, gen, "aba", n, q, qd, u)], [qdd_fd]) < 1e-8
        M = call(lib, gen, "crba", n * n, q).reshape(n, n).T
        assert np.abs(M @ Minv - np.eye(n)).max() < 1e-7
        # grad FD + by-products in one body: [df_du | Minv | qdd], Minv exactly symmetric
        lin = call(lib, gen, "forward_dynamics_linearization", 3 * n * n + n, q, qd, u)
        assert rel_err([lin[:2 * n * n]], [cm(g_["df_du"][k])]) < 1e-9
        assert rel_err([lin[2 * n * n:3 * n * n]], [cm(Minv)]) < 1e-9
        assert rel_err([lin[3 * n * n:]], [qdd_fd]) < 1e-9
        Ml = lin[2 * n * n:3 * n * n].reshape(n, n)
        assert np.array_equal(Ml, Ml.T)
        # split (warp-specialised) forms: phase A + one function per derivative column on the published state
        assert rel_err([call(lib, gen, "split_inverse_dynamics_gradient", 2 * n * n, q, qd)], [cm(g_["dc_du0"][k])]) < TOL
        assert rel_err([call(lib, gen, "split_inverse_dynamics_gradient_qdd", 2 * n * n, q, qd, qdd)], [cm(g_["dc_du"][k])]) < TOL
        assert rel_err([call(lib, gen, "split_forward_dynamics_gradient", 2 * n * n, q, qd, u)], [cm(g_["df_du"][k])]) < 1e-9
        lin = call(lib, gen, "split_forward_dynamics_linearization", 3 * n * n + n, q, qd, u)
        assert rel_err([lin[:2 * n * n]], [cm(g_["df_du"][k])]) < 1e-9
        assert rel_err([lin[2 * n * n:3 * n * n]], [cm(Minv)]) < 1e-9
        assert rel_err([lin[3 * n * n:]], [qdd_fd]) < 1e-9
        # qdd+Minv input mode: feeding the FD solution reproduces grad FD
        out = call(lib, gen, "forward_dynamics_gradient_qdd_minv", 2 * n * n, q, qd, qdd_fd, cm(Minv))
        assert rel_err([out], [cm(g_["df_du"][k])]) < 1e-9


@pytest.mark.parametrize("name", ["pendulum2", "chain3"])
def test_second_order_bodies_match_finite_differences(name, tmp_path):
    """IDSVA_SO / FDSVA_SO bodies (emit-time forward-mode differentiation of the straight-line RNEA / CRBA) vs the
    oracle's central differences of the pinned first-order functions (parity unpinned upstream)."""
    from oracle import Oracle
    robot = load_robot(name)
    n = robot.get_num_pos()
    lib, gen = build_host_inners(robot, str(tmp_path), second_order=True)
    orc = Oracle(robot)
    rng = np.random.default_rng(3)
    for _ in range(3):
        x = rng.uniform(-1, 1, 3 * n)
        a = call(lib, gen, "idsva_so", 4 * n ** 3, x[:n], x[n:2 * n], x[2 * n:])
        assert rel_err([a], [orc.batch_idsva_so(x[None])[0]]) < 1e-7
        a = call(lib, gen, "fdsva_so", 4 * n ** 3, x[:n], x[n:2 * n], x[2 * n:])
        assert rel_err([a], [orc.batch_fdsva_so(x[None])[0]]) < 1e-7


@pytest.mark.parametrize("name", ["chain3", "branched6", "iiwa14"])
def test_end_effector_bodies(name, tmp_path):
    """EE pose / gradient / Hessian bodies vs the oracle's chain product + RPY restatement and its central
    differences (reference algorithms/_eepose_gradient_hessian.py:142-159, :511-538, :985-1058)."""
    from oracle import Oracle
    robot = load_robot(name)
    n = robot.get_num_pos()
    lib, gen = build_host_inners(robot, str(tmp_path))
    orc = Oracle(robot)
    ne = orc.nleaf
    rng = np.random.default_rng(4)
    for _ in range(3):
        q = rng.uniform(-1, 1, n)
        assert rel_err([call(lib, gen, "end_effector_pose", 6 * ne, q)], [orc.batch_ee_pose(q[None])[0]]) < TOL
        assert rel_err([call(lib, gen, "end_effector_pose_gradient", 6 * ne * n, q)],
                       [orc.batch_ee_pose_gradient(q[None])[0]]) < 1e-8
        assert rel_err([call(lib, gen, "end_effector_pose_gradient_hessian", 6 * ne * n * n, q)],
                       [orc.batch_ee_pose_hessian(q[None])[0]]) < 1e-5


@pytest.mark.parametrize("env", [{"GRID_COLUMN_LOOP": "1"}, {"GRID_EMIT_PAIRS_FROM": "1"}, {"GRID_EMIT_PAIRS": "1"},
                                 {"GRID_FDGRAD_SOLVER": "minv"}])
@pytest.mark.parametrize("name", ["branched6", "iiwa14"])
def test_emitter_variants_match_golden(name, env, golden, tmp_path, monkeypatch):
    """The alternative emitter strategies kept behind environment switches (measured on B200, DESIGN.md section 7:
    runtime loop over derivative columns, packed column pairs from a given column on, analytical-Minv grad FD)
    compute the same numbers as the default form."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    g_ = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    lib, gen = build_host_inners(robot, str(tmp_path))
    x = g_["x"]
    cm = lambda M: np.asarray(M).T.reshape(-1)
    for k in range(x.shape[0]):
        q, qd, u, qdd = (x[k, i * n:(i + 1) * n] for i in range(4))
        assert rel_err([call(lib, gen, "inverse_dynamics_gradient_qdd", 2 * n * n, q, qd, qdd)], [cm(g_["dc_du"][k])]) < TOL
        assert rel_err([call(lib, gen, "forward_dynamics_gradient", 2 * n * n, q, qd, u)], [cm(g_["df_du"][k])]) < 1e-9
        lin = call(lib, gen, "forward_dynamics_linearization", 3 * n * n + n, q, qd, u)
        assert rel_err([lin[:2 * n * n]], [cm(g_["df_du"][k])]) < 1e-9
        assert rel_err([lin[2 * n * n:3 * n * n]], [cm(g_["Minv"][k])]) < 1e-9


@pytest.mark.parametrize("name", ["chain3", "branched6", "iiwa14", "hyq"])
def test_joint_transforms_and_fixed_target_bodies(name, golden, tmp_path):
    """joint_transforms / X_single_thread (reference algorithms/_eepose_gradient_hessian.py:1226-1398) and the EE
    functions of a fixed kinematic target (:31-38, :89-93) against the sympy fixtures (exact derivatives)."""
    import ctypes
    g_ = golden(name)
    robot = load_robot(name)
    n = robot.get_num_pos()
    lib, gen = build_host_inners(robot, str(tmp_path))
    x = g_["x"]
    dp = ctypes.POINTER(ctypes.c_double)
    for k in range(x.shape[0]):
        q = np.ascontiguousarray(x[k, :n])
        want = g_["joint_xforms"][k].transpose(0, 2, 1).reshape(-1)              # column-major 4 x 4 per joint
        assert rel_err([call(lib, gen, "joint_transforms", 16 * n, q)], [want]) < 1e-12
        for tid in (n - 1, n // 2, 0, n + 5):
            jx, xh = np.full(16 * n, 7.0), np.full(16 * n, 7.0)
            lib.host_X_single_thread.restype = None
            lib.host_X_single_thread(jx.ctypes.data_as(dp), xh.ctypes.data_as(dp), q.ctypes.data_as(dp), ctypes.c_int(tid))
            upto = min(tid, n - 1) + 1
            assert np.abs(jx[:16 * upto] - want[:16 * upto]).max() < 1e-12
            assert np.all(jx[16 * upto:] == 7.0)                                  # joints beyond tid are not touched
        if "ft_name" in g_:
            ft = str(g_["ft_name"])
            assert gen.fixed_target_name == ft
            assert rel_err([call(lib, gen, f"end_effector_pose_{ft}", 6, q)], [g_["ft_pose"][k]]) < 1e-12
            grad = call(lib, gen, f"end_effector_pose_gradient_{ft}", 6 * n, q).reshape(n, 6).T      # [o][j]
            assert rel_err([grad], [g_["ft_grad"][k]]) < 1e-10
            hess = call(lib, gen, f"end_effector_pose_gradient_hessian_{ft}", 6 * n * n, q).reshape(6, n, n)
            assert rel_err([hess], [g_["ft_hess"][k]]) < 1e-9
