"""ctypes wrapper around librbd_oracle.so (see rbd_oracle.c for reference citations)."""
import ctypes
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)


def build_oracle(force=False):
    so = os.path.join(_DIR, "librbd_oracle.so")
    src = os.path.join(_DIR, "rbd_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        # -march=native binaries do not travel between hosts: rebuild when the source is newer
        subprocess.check_call(["make", "-C", _DIR, "-B", "librbd_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def _lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_DIR, "librbd_oracle.so")
        try:
            build_oracle()
            L = ctypes.CDLL(so)
            L.rbd_max_threads()
        except Exception:
            build_oracle(force=True)
            L = ctypes.CDLL(so)
        L.rbd_model_create.restype = ctypes.c_void_p
        L.rbd_model_create.argtypes = [ctypes.c_int, _ip, _ip, _dp, _dp, _dp, _dp, _dp, _dp]
        L.rbd_model_free.argtypes = [ctypes.c_void_p]
        L.rbd_max_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _d(a):
    return a.ctypes.data_as(_dp)


def _arr(x):
    return np.ascontiguousarray(x, dtype=np.float64)


def _hom_tables(robot):
    """Xhom_j(q) = H0 + sin(q) Hs + cos(q) Hc + q Hq per joint (4x4), by evaluating the robot's sympy homogeneous
    transforms at three angles (revolute) / two displacements (prismatic).  Pinned by tests/test_oracle_golden.py
    against the sympy-differentiated end-effector fixtures of tools/make_golden.py."""
    out = []
    for j, M in enumerate(robot.get_Xmats_hom_ordered_by_id()):
        syms = list(M.free_symbols)
        ev = lambda val: np.array(M.subs({s_: val for s_ in syms}).evalf(30), dtype=np.float64).reshape(4, 4)
        Z = np.zeros((4, 4))
        if int(np.argmax(np.abs(robot.get_S_by_id(j)))) < 3:
            A, B, C = ev(0), ev(np.pi / 2), ev(np.pi)
            H0, Hc = (A + C) / 2, (A - C) / 2
            Hs = B - H0
            for m in (H0, Hs, Hc):
                m[np.abs(m) < 1e-12] = 0.0
            out.append((H0, Hs, Hc, Z))
        else:
            A, B = ev(0), ev(1)
            D = B - A
            D[np.abs(D) < 1e-12] = 0.0
            out.append((A, Z, Z, D))
    return out


class Oracle:
    def __init__(self, robot):
        L = _lib()
        t = robot.numeric_tables()
        self.n = n = int(t["n"])
        self._keep = [np.ascontiguousarray(t[k], dtype=np.float64) for k in ("X0", "Xs", "Xc", "Xq", "I", "damping")]
        par = np.ascontiguousarray(t["parent"], dtype=np.int32)
        sind = np.ascontiguousarray(t["S_ind"], dtype=np.int32)
        self._h = ctypes.c_void_p(
            L.rbd_model_create(n, par.ctypes.data_as(_ip), sind.ctypes.data_as(_ip), *[_d(a) for a in self._keep])
        )
        if not self._h:
            raise RuntimeError("rbd_model_create failed (n too large?)")
        self.L = L
        # homogeneous transforms + leaves for the end-effector functions: sampled here from the robot's sympy
        # matrices (the reference's input, helpers/_topology_helpers.py:76-78), NOT taken from the product's emitter
        H = _hom_tables(robot)
        hs = [np.ascontiguousarray(np.stack([h[i] for h in H]), dtype=np.float64) for i in range(4)]
        leaves = np.ascontiguousarray(robot.get_leaf_nodes(), dtype=np.int32)
        self.nleaf = len(leaves)
        self._hom = hs                                   # H0, Hs, Hc, Hq stacked per joint (numpy functions below)
        self._parent = [int(p) for p in par]
        self._robot = robot
        L.rbd_model_set_hom.argtypes = [ctypes.c_void_p, _dp, _dp, _dp, _dp, ctypes.c_int, _ip]
        L.rbd_model_set_hom(self._h, *[_d(a) for a in hs], ctypes.c_int(self.nleaf), leaves.ctypes.data_as(_ip))

    def __del__(self):
        try:
            self.L.rbd_model_free(self._h)
        except Exception:
            pass

    @staticmethod
    def max_threads():
        return int(_lib().rbd_max_threads())

    # ---------------------------------------------------------------- single knot
    def rnea(self, q, qd, qdd=None, gravity=9.81):
        n = self.n
        q, qd = _arr(q), _arr(qd)
        c = np.zeros(n); v = np.zeros((n, 6)); a = np.zeros((n, 6)); f = np.zeros((n, 6))
        qddp = _d(_arr(qdd)) if qdd is not None else None
        self.L.rbd_rnea(self._h, _d(q), _d(qd), qddp, ctypes.c_double(gravity), _d(c), _d(v), _d(a), _d(f))
        return c, v.T.copy(), a.T.copy(), f.T.copy()  # (6,n) like the reference

    def minv(self, q):
        n = self.n
        M = np.zeros((n, n))
        self.L.rbd_minv(self._h, _d(_arr(q)), _d(M))
        return M

    def rnea_grad(self, q, qd, qdd=None, gravity=9.81):
        n = self.n
        out = np.zeros((n, 2 * n))
        qddp = _d(_arr(qdd)) if qdd is not None else None
        self.L.rbd_rnea_grad(self._h, _d(_arr(q)), _d(_arr(qd)), qddp, ctypes.c_double(gravity), _d(out))
        return out

    def fd(self, q, qd, u, gravity=9.81):
        out = np.zeros(self.n)
        self.L.rbd_fd(self._h, _d(_arr(q)), _d(_arr(qd)), _d(_arr(u)), ctypes.c_double(gravity), _d(out))
        return out

    def fd_grad(self, q, qd, u, gravity=9.81):
        n = self.n
        out = np.zeros((n, 2 * n))
        self.L.rbd_fd_grad(self._h, _d(_arr(q)), _d(_arr(qd)), _d(_arr(u)), ctypes.c_double(gravity), _d(out))
        return out

    def aba(self, q, qd, tau, gravity=9.81):
        out = np.zeros(self.n)
        self.L.rbd_aba(self._h, _d(_arr(q)), _d(_arr(qd)), _d(_arr(tau)), ctypes.c_double(gravity), _d(out))
        return out

    def crba(self, q):
        n = self.n
        M = np.zeros((n, n))
        self.L.rbd_crba(self._h, _d(_arr(q)), _d(M))
        return M

    # -------------------------------------------------------------------- batched
    # x: (B, stride) fp64, packed [q | qd | u-or-qdd]; outputs use the GPU layouts
    def _batch(self, fn, x, out_per_knot, *extra, nthreads=0):
        x = _arr(x)
        B, stride = x.shape
        out = np.zeros((B, out_per_knot))
        fn(self._h, ctypes.c_int(B), _d(x), ctypes.c_int(stride), *extra, _d(out), ctypes.c_int(nthreads))
        return out

    def batch_fd_grad(self, x, gravity=9.81, nthreads=0):
        return self._batch(self.L.rbd_batch_fd_grad, x, 2 * self.n * self.n, ctypes.c_double(gravity), nthreads=nthreads)

    def batch_rnea_grad(self, x, use_qdd=True, gravity=9.81, nthreads=0):
        return self._batch(self.L.rbd_batch_rnea_grad, x, 2 * self.n * self.n, ctypes.c_int(int(use_qdd)),
                           ctypes.c_double(gravity), nthreads=nthreads)

    def batch_rnea(self, x, use_qdd=True, gravity=9.81, nthreads=0):
        return self._batch(self.L.rbd_batch_rnea, x, self.n, ctypes.c_int(int(use_qdd)), ctypes.c_double(gravity),
                           nthreads=nthreads)

    def batch_minv(self, x, nthreads=0):
        return self._batch(self.L.rbd_batch_minv, x, self.n * self.n, nthreads=nthreads)

    def batch_fd(self, x, gravity=9.81, nthreads=0):
        return self._batch(self.L.rbd_batch_fd, x, self.n, ctypes.c_double(gravity), nthreads=nthreads)

    def batch_aba(self, x, gravity=9.81, nthreads=0):
        return self._batch(self.L.rbd_batch_aba, x, self.n, ctypes.c_double(gravity), nthreads=nthreads)

    def batch_crba(self, x, nthreads=0):
        return self._batch(self.L.rbd_batch_crba, x, self.n * self.n, nthreads=nthreads)

    def batch_idsva_so(self, x, gravity=9.81, nthreads=0):
        """x: [q | qd | qdd]; out: 4 n^3 per knot (layout in rbd_oracle.c)."""
        return self._batch(self.L.rbd_batch_idsva_so, x, 4 * self.n ** 3, ctypes.c_double(gravity), nthreads=nthreads)

    def batch_fdsva_so(self, x, gravity=9.81, nthreads=0):
        """x: [q | qd | u]; out: 4 n^3 per knot (layout in rbd_oracle.c)."""
        return self._batch(self.L.rbd_batch_fdsva_so, x, 4 * self.n ** 3, ctypes.c_double(gravity), nthreads=nthreads)

    def batch_ee_pose(self, x, nthreads=0):
        """x: rows starting with q; out[6*ee + o]."""
        return self._batch(self.L.rbd_batch_ee_pose, x, 6 * self.nleaf, nthreads=nthreads)

    def batch_ee_pose_gradient(self, x, nthreads=0):
        """out[(ee*n + j)*6 + o] (central differences of the pose)."""
        return self._batch(self.L.rbd_batch_ee_pose_gradient, x, 6 * self.nleaf * self.n, nthreads=nthreads)

    def batch_ee_pose_hessian(self, x, nthreads=0):
        """out[(ee*6 + o)*n*n + i*n + j] (second central differences of the pose, ~1e-7 accurate)."""
        return self._batch(self.L.rbd_batch_ee_pose_hessian, x, 6 * self.nleaf * self.n ** 2, nthreads=nthreads)

    # ------------------------------------------------- homogeneous chain (numpy; small and not timed)
    def _chain(self, x):
        """T[j]: (B, 4, 4) transform base <- frame of joint j: product of Xhom_j(q_j) down the parent chain
        (reference algorithms/_eepose_gradient_hessian.py:76-140; X_single_thread :1226-1298 for serial chains)."""
        x = _arr(x)
        q = x[:, :self.n]
        H0, Hs, Hc, Hq = self._hom
        T = []
        for j in range(self.n):
            Xj = (H0[j][None] + np.sin(q[:, j])[:, None, None] * Hs[j][None] + np.cos(q[:, j])[:, None, None] * Hc[j][None]
                  + q[:, j][:, None, None] * Hq[j][None])
            p = self._parent[j]
            T.append(Xj if p < 0 else np.einsum("bik,bkj->bij", T[p], Xj))
        return T

    def batch_joint_transforms(self, x):
        """(B, >= n) rows starting with q -> (B, 16 n): out[16 j + r + 4 c] = (base <- joint j)[r][c]."""
        T = self._chain(x)
        return np.concatenate([t.transpose(0, 2, 1).reshape(len(t), 16) for t in T], axis=1)

    def batch_fixed_target_pose(self, x, fixed_name):
        """Pose [x y z roll pitch yaw] of the fixed joint `fixed_name` (reference :31-38, :89-93, :150-157) -> (B, 6)."""
        fj = self._robot.get_fixed_joint_by_name(fixed_name)
        Xf = np.array(fj.Xhom_sympy().evalf(30), dtype=np.float64).reshape(4, 4)
        T = self._chain(x)
        B = len(_arr(x))
        M = np.broadcast_to(Xf, (B, 4, 4)) if fj.parent_joint_id < 0 else np.einsum("bik,kj->bij", T[fj.parent_joint_id], Xf)
        roll = np.arctan2(M[:, 2, 1], M[:, 2, 2])
        pitch = -np.arctan2(M[:, 2, 0], np.sqrt(M[:, 2, 1] ** 2 + M[:, 2, 2] ** 2))
        yaw = np.arctan2(M[:, 1, 0], M[:, 0, 0])
        return np.stack([M[:, 0, 3], M[:, 1, 3], M[:, 2, 3], roll, pitch, yaw], axis=1)
