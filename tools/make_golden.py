#!/usr/bin/env python
"""Generate tests/golden/<robot>.npz by running the REFERENCE's own CPU code.

Runs only in the build container (needs /root/reference); the .npz files are
committed so the GPU box, which has no /root/reference, can still pin the
oracle and the CUDA path to the reference.

First order -- for each shipped robot and a few seeded knot points the outputs of the reference's `_test.py`
methods (bound on `GRiDCodeGenerator`, reference GRiDCodeGenerator.py:50-51):
  test_rnea (_test.py:109-115), test_minv (:213-226), test_rnea_grad (:490-494), test_fd_grad (:496-520).
Inputs: numpy default_rng(seed), U(-1, 1), ROUNDED TO FP32 and stored as fp64 (so that the fp32 kernels and the
fp64 reference see bit-identical inputs and the north-star tolerance 1e-4 applies without an input-rounding
allowance); GRAVITY = -9.81 (== kernel gravity 9.81).

Second order (so_* arrays, revolute robots up to 12 joints) -- the reference has no executable CPU code for
IDSVA_SO / FDSVA_SO (algorithms/_idsva_so.py and _fdsva_so.py only print CUDA), so the fixtures are the EXACT
derivatives of the reference's own first-order functions, obtained by complex-step differentiation of the
unmodified `_test.py` code (f(x + ih).imag / h, h = 1e-30: no subtractive cancellation, accurate to rounding):
  so_idsva[0] = d/dq  of test_rnea_grad[:, :n]   (d2 tau / dq dq)
  so_idsva[1] = d/dqd of test_rnea_grad[:, n:]   (d2 tau / dqd dqd)
  so_idsva[2] = d/dq  of test_rnea_grad[:, n:]   (d2 tau / dqd dq)
  so_idsva[3] = d/dq  of M(q), M[:, j] = test_rnea(q, 0, e_j, GRAVITY=0)      (dM / dq)
  so_fdsva[0..2] = the same three derivatives of test_fd_grad;  so_fdsva[3] = d/dq of test_minv  (d2 qdd / dtau dq)
each stored as [t][i][j][k] = d(entry (i, j)) / d(variable k) -- the reference's tensor layout
(algorithms/_idsva_so.py:156-159).  To run `_test.py` on complex inputs its module-level `np` is wrapped so that
np.zeros allocates complex arrays, and the robot is wrapped so that inertias are complex and X(q) comes from the
lambdified sympy matrices (what URDFParser itself does); the arithmetic executed is the reference's.

End-effector pose (ee_* arrays) -- from the robot's sympy homogeneous transforms
(`get_Xmats_hom_ordered_by_id`, the reference's input: helpers/_topology_helpers.py:76-78): chain product down to
each leaf, pose = [x y z roll pitch yaw] with the reference's formulas
(algorithms/_eepose_gradient_hessian.py:150-157), gradient and Hessian by sympy differentiation.  joint_xforms: the
cumulative transforms base <- joint j (reference X_single_thread / X_warp, :1226-1398).  ft_*: pose / gradient / Hessian
of the robot's first fixed joint as a fixed kinematic target (:31-38, :89-93).
"""
import contextlib
import io
import os
import sys

import numpy as np
import sympy as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root")
sys.path.insert(0, ROOT)

import reference._test as ref_test  # noqa: E402
from reference import GRiDCodeGenerator as RefGen  # noqa: E402

from gridcodegenerator_b200.robot import available_robots, load_robot  # noqa: E402
from gridcodegenerator_b200.robot.model import THETA  # noqa: E402

KNOTS = {"atlas": 4}
DEFAULT_KNOTS = 8
SO_KNOTS = 3
SO_MAX_JOINTS = 12
EE_HESSIAN_MAX_JOINTS = 64
H = 1e-30


class _ComplexNumpy:
    """`np` as seen by the reference's _test.py while it is being differentiated: zeros() allocates complex."""

    def __getattr__(self, k):
        return getattr(np, k)

    @staticmethod
    def zeros(shape, *a, **k):
        return np.zeros(shape, dtype=np.complex128)


class _ComplexRobot:
    """Robot wrapper for the complex-step runs: complex inertias (the reference accumulates into them in place,
    _test.py:168) and X(q) from the lambdified sympy matrices."""

    def __init__(self, robot):
        self._r = robot
        self._xf = [sp.lambdify(THETA, M, "numpy") for M in robot.get_Xmats_ordered_by_id()]

    def __getattr__(self, k):
        return getattr(self._r, k)

    def get_Imats_dict_by_id(self):
        return {k: v.astype(np.complex128) for k, v in self._r.get_Imats_dict_by_id().items()}

    def get_Imats_ordered_by_id(self, *a, **k):
        return [np.asarray(m).astype(np.complex128) for m in self._r.get_Imats_ordered_by_id(*a, **k)]

    def get_Xmat_Funcs_ordered_by_id(self, reverse=False):
        fs = [lambda q, f=f: np.array(f(q), dtype=np.complex128) for f in self._xf]
        return list(reversed(fs)) if reverse else fs

    def get_Xmat_Func_by_id(self, jid):
        f = self._xf[jid]
        return lambda q: np.array(f(q), dtype=np.complex128)


@contextlib.contextmanager
def _complex_reference(g):
    real_np, real_robot = ref_test.np, g.robot
    ref_test.np = _ComplexNumpy()
    g.robot = _ComplexRobot(real_robot)
    try:
        yield
    finally:
        ref_test.np = real_np
        g.robot = real_robot


def second_order(g, n, q, qd, qdd, u):
    """(idsva[4][n][n][n], fdsva[4][n][n][n]) at one knot; see the module docstring."""
    ids = np.zeros((4, n, n, n))
    fds = np.zeros((4, n, n, n))
    cz = lambda a: np.asarray(a, dtype=np.complex128)
    with _complex_reference(g), contextlib.redirect_stdout(io.StringIO()):
        # sanity: the complexified run reproduces the real one
        for k in range(n):
            e = np.zeros(n, dtype=np.complex128)
            e[k] = 1j * H
            dq = np.asarray(g.test_rnea_grad(cz(q) + e, cz(qd), cz(qdd))).imag / H        # d/dq_k
            dv = np.asarray(g.test_rnea_grad(cz(q), cz(qd) + e, cz(qdd))).imag / H        # d/dqd_k
            ids[0][:, :, k] = dq[:, :n]
            ids[2][:, :, k] = dq[:, n:]
            ids[1][:, :, k] = dv[:, n:]
            M = np.stack([np.asarray(g.test_rnea(cz(q) + e, cz(np.zeros(n)), cz(np.eye(n)[j]), GRAVITY=0)[0])
                          for j in range(n)], axis=1)
            ids[3][:, :, k] = M.imag / H
            fq = np.asarray(g.test_fd_grad(cz(q) + e, cz(qd), cz(u))).imag / H
            fv = np.asarray(g.test_fd_grad(cz(q), cz(qd) + e, cz(u))).imag / H
            fds[0][:, :, k] = fq[:, :n]
            fds[2][:, :, k] = fq[:, n:]
            fds[1][:, :, k] = fv[:, n:]
            fds[3][:, :, k] = np.asarray(g.test_minv(cz(q) + e, True)).imag / H
    return ids, fds


def _pose_exprs(M):
    roll = sp.atan2(M[2, 1], M[2, 2])
    pitch = -sp.atan2(M[2, 0], sp.sqrt(M[2, 1] ** 2 + M[2, 2] ** 2))
    yaw = sp.atan2(M[1, 0], M[0, 0])
    return [M[0, 3], M[1, 3], M[2, 3], roll, pitch, yaw]


def chain_transforms(robot):
    """sympy base <- joint j transforms (4 x 4) in symbols q0..q{n-1}, following the parent topology."""
    n = robot.get_num_pos()
    qs = sp.symbols(f"q0:{n}")
    Xh = [M.subs(THETA, qs[j]) for j, M in enumerate(robot.get_Xmats_hom_ordered_by_id())]
    T = {}
    for j in range(n):
        p = robot.get_parent_id(j)
        T[j] = Xh[j] if p < 0 else T[p] * Xh[j]
    return qs, T


def pose_functions(qs, poses, order):
    """Lambdified (pose, gradient, Hessian) of a list of pose expressions; derivatives are only formed with respect to
    the joints an expression depends on (the rest are structural zeros), which keeps Atlas tractable."""
    n = len(qs)
    fpose = sp.lambdify(qs, poses, "numpy", cse=True)
    fgrad = fhess = None
    if order >= 1:
        grads = [[(sp.diff(p, qs[j]) if qs[j] in p.free_symbols else sp.Integer(0)) for j in range(n)] for p in poses]
        fgrad = sp.lambdify(qs, grads, "numpy", cse=True)
        if order >= 2:
            hess = [[[(sp.diff(grads[o][i], qs[j]) if qs[j] in grads[o][i].free_symbols else sp.Integer(0))
                      for j in range(n)] for i in range(n)] for o in range(len(poses))]
            fhess = sp.lambdify(qs, hess, "numpy", cse=True)
    return fpose, fgrad, fhess


def ee_functions(robot, order):
    """Lambdified (pose, gradient, Hessian) of every leaf link, from the sympy homogeneous transforms."""
    qs, T = chain_transforms(robot)
    poses = []
    for e in robot.get_leaf_nodes():
        poses += _pose_exprs(T[e])
    return pose_functions(qs, poses, order)


def fixed_target_functions(robot, fj, order):
    """The same for ONE fixed kinematic target (reference algorithms/_eepose_gradient_hessian.py:31-38, :89-93): the
    chain of the fixed joint's parent body times the fixed joint's constant transform."""
    qs, T = chain_transforms(robot)
    p = fj.parent_joint_id
    M = fj.Xhom_sympy() if p < 0 else T[p] * fj.Xhom_sympy()
    return pose_functions(qs, _pose_exprs(M), order)


def main():
    outdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    only = sys.argv[1:]
    for name in available_robots():
        if only and name not in only:
            continue
        robot = load_robot(name)
        g = RefGen(robot)
        n = robot.get_num_pos()
        K = KNOTS.get(name, DEFAULT_KNOTS)
        rng = np.random.default_rng(1234)
        x = rng.uniform(-1, 1, (K, 4 * n)).astype(np.float32).astype(np.float64)  # q | qd | u | qdd, fp32-representable
        rec = dict(x=x, c=[], c0=[], v=[], a=[], f=[], Minv=[], dc_du=[], dc_du0=[], df_du=[])
        for k in range(K):
            q, qd, u, qdd = (x[k, i * n:(i + 1) * n].copy() for i in range(4))
            with contextlib.redirect_stdout(io.StringIO()):
                c, v, a, f = g.test_rnea(q, qd, qdd)
                c0 = g.test_rnea(q, qd, None)[0]
                Minv = g.test_minv(q, True)
                dc_du = g.test_rnea_grad(q, qd, qdd)
                dc_du0 = g.test_rnea_grad(q, qd, None)
                df_du = g.test_fd_grad(q, qd, u)
            rec["c"].append(np.array(c, dtype=np.float64))
            rec["c0"].append(np.array(c0, dtype=np.float64))
            rec["v"].append(np.array(v, dtype=np.float64))
            rec["a"].append(np.array(a, dtype=np.float64))
            rec["f"].append(np.array(f, dtype=np.float64))
            rec["Minv"].append(np.array(Minv, dtype=np.float64))
            rec["dc_du"].append(np.array(dc_du, dtype=np.float64))
            rec["dc_du0"].append(np.array(dc_du0, dtype=np.float64))
            rec["df_du"].append(np.array(df_du, dtype=np.float64))
        out = {k: np.stack(v) if isinstance(v, list) else v for k, v in rec.items()}

        revolute = all(int(np.argmax(np.abs(robot.get_S_by_id(j)))) < 3 for j in range(n))
        if revolute and n <= SO_MAX_JOINTS:
            K2 = min(SO_KNOTS, K)
            ids, fds = [], []
            for k in range(K2):
                q, qd, u, qdd = (x[k, i * n:(i + 1) * n].copy() for i in range(4))
                # complexified run == real run (guards the wrappers)
                with _complex_reference(g), contextlib.redirect_stdout(io.StringIO()):
                    chk = np.asarray(g.test_fd_grad(q.astype(complex), qd.astype(complex), u.astype(complex)))
                assert np.abs(chk.real - out["df_du"][k]).max() < 1e-9 * max(1.0, np.abs(out["df_du"][k]).max())
                a_, b_ = second_order(g, n, q, qd, qdd, u)
                ids.append(a_)
                fds.append(b_)
            out["so_idsva"] = np.stack(ids)      # inputs: x[:K2, :2n] and qdd = x[:K2, 3n:4n]
            out["so_fdsva"] = np.stack(fds)      # inputs: x[:K2, :3n]

        order = 2 if n <= EE_HESSIAN_MAX_JOINTS else 1
        fpose, fgrad, fhess = ee_functions(robot, order)
        ne = len(robot.get_leaf_nodes())
        out["ee_pose"] = np.stack([np.array(fpose(*x[k, :n]), dtype=np.float64) for k in range(K)])          # [k][6*ee + o]
        out["ee_grad"] = np.stack([np.array(fgrad(*x[k, :n]), dtype=np.float64) for k in range(K)])          # [k][6*ee + o][j]
        if fhess is not None:
            out["ee_hess"] = np.stack([np.array(fhess(*x[k, :n]), dtype=np.float64) for k in range(K)])      # [k][6*ee + o][i][j]
        assert out["ee_pose"].shape == (K, 6 * ne)
        # cumulative joint transforms (what the reference's X_single_thread / X_warp accumulate, :1226-1398)
        qs, T = chain_transforms(robot)
        fT = sp.lambdify(qs, [T[j] for j in range(n)], "numpy", cse=True)
        out["joint_xforms"] = np.stack([np.array(fT(*x[k, :n]), dtype=np.float64).reshape(n, 4, 4) for k in range(K)])
        # first fixed joint of the robot as a fixed kinematic target
        if robot.fixed_joints:
            fj = robot.fixed_joints[0]
            fp, fg, fh = fixed_target_functions(robot, fj, 2)
            out["ft_name"] = np.array(fj.get_name())
            out["ft_pose"] = np.stack([np.array(fp(*x[k, :n]), dtype=np.float64) for k in range(K)])
            out["ft_grad"] = np.stack([np.array(fg(*x[k, :n]), dtype=np.float64) for k in range(K)])
            out["ft_hess"] = np.stack([np.array(fh(*x[k, :n]), dtype=np.float64) for k in range(K)])
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **out)
        print(name, {k: v.shape for k, v in out.items()}, flush=True)


if __name__ == "__main__":
    main()
