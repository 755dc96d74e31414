"""Spatial-algebra emitters: joint transforms in factored form, cross products, inertias.

X_j(q) = XJ_j(q_j) * XT_j with XT_j the constant tree transform and XJ_j the
principal-axis joint transform.  The reference builds the product numerically
per knot point and multiplies dense 6x6 columns (helpers/_topology_helpers.py:
155-331, helpers/_spatial_algebra_helpers.py:124-143); here XT_j is folded into
the code as immediates and XJ_j costs 8 multiply-adds per 6-vector.

Cross-product conventions follow the reference oracle: mxS == column S of the
motion cross matrix (_test.py:522-608), fxv == Fx(a) b (_test.py:649-664).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

from .core import Em, V, ZERO, const

SNAP = 1e-12


def _snap(A):
    A = np.array(A, dtype=np.float64)
    A[np.abs(A) < SNAP] = 0.0
    return A


class JointData:
    """Numeric description of one joint + its body, taken from the robot-model layer."""

    def __init__(self, robot, j: int):
        jt = robot.get_joint_by_id(j)
        self.id = j
        self.parent = robot.get_parent_id(j)
        self.S = int(np.argmax(np.abs(robot.get_S_by_id(j))))
        self.revolute = self.S < 3
        self.axis = self.S % 3
        self.damping = float(robot.get_damping_by_id(j))
        self.I = _snap(robot.get_Imat_by_id(j))
        f = robot.get_Xmat_Func_by_id(j)
        X0 = np.array(f(0.0), dtype=np.float64)
        # XT = XJ(0)^-1 X(0) and XJ(0) = identity for both joint types
        self.XT = _snap(X0)
        # XJ(q) = J0 + sin(q) Js + cos(q) Jc + q Jq, recovered from samples of X(q) XT^-1
        XTi = np.linalg.inv(X0)
        if self.revolute:
            Ja = np.array(f(0.0)) @ XTi
            Jb = np.array(f(np.pi / 2)) @ XTi
            Jc_ = np.array(f(np.pi)) @ XTi
            self.J0 = _snap((Ja + Jc_) / 2)
            self.Jc = _snap((Ja - Jc_) / 2)
            self.Js = _snap(Jb - self.J0)
            self.Jq = np.zeros((6, 6))
        else:
            Ja = np.array(f(0.0)) @ XTi
            Jb = np.array(f(1.0)) @ XTi
            self.J0 = _snap(Ja)
            self.Jq = _snap(Jb - Ja)
            self.Js = np.zeros((6, 6))
            self.Jc = np.zeros((6, 6))
        # self-check of the factorisation at a generic angle
        th = 0.37
        XJ = self.J0 + np.sin(th) * self.Js + np.cos(th) * self.Jc + th * self.Jq
        assert np.abs(XJ @ self.XT - np.array(f(th))).max() < 1e-9, "X(q) is not J(q) XT for joint %d" % j


class Model:
    def __init__(self, robot):
        self.robot = robot
        self.n = robot.get_num_pos()
        self.joints = [JointData(robot, j) for j in range(self.n)]
        self.parent = [jd.parent for jd in self.joints]
        self.children = [[] for _ in range(self.n)]
        for j, p in enumerate(self.parent):
            if p >= 0:
                self.children[p].append(j)
        self.subtree = [robot.get_subtree_by_id(j) for j in range(self.n)]
        self.ancestors = [sorted(robot.get_ancestors_by_id(j)) for j in range(self.n)]
        self.name = robot.name


class Trig:
    """Per-joint runtime scalars sin(q), cos(q), q as V's."""

    def __init__(self, s: Optional[V], c: Optional[V], q: V):
        self.s, self.c, self.q = s, c, q


def _tm_const(A):
    """numeric matrix -> term matrix: TM[i][k] = list of (coef, factor)"""
    return [[([(float(A[i, k]), None)] if A[i, k] != 0.0 else []) for k in range(A.shape[1])] for i in range(A.shape[0])]


def _tm_joint(jd: JointData, tr: Trig):
    TM = []
    for i in range(6):
        row = []
        for k in range(6):
            e = []
            if jd.J0[i, k] != 0.0:
                e.append((float(jd.J0[i, k]), None))
            if jd.Js[i, k] != 0.0:
                e.append((float(jd.Js[i, k]), tr.s))
            if jd.Jc[i, k] != 0.0:
                e.append((float(jd.Jc[i, k]), tr.c))
            if jd.Jq[i, k] != 0.0:
                e.append((float(jd.Jq[i, k]), tr.q))
            row.append(e)
        TM.append(row)
    return TM


def tm_matvec(em: Em, TM, x: Sequence[V], transpose=False, acc: Optional[Sequence[V]] = None) -> List[V]:
    n_out = len(TM[0]) if transpose else len(TM)
    n_in = len(TM) if transpose else len(TM[0])
    out = []
    for i in range(n_out):
        terms = []
        for k in range(n_in):
            ent = TM[k][i] if transpose else TM[i][k]
            for coef, f in ent:
                terms.append((coef, f, x[k]))
        out.append(em.lincomb(terms, None if acc is None else acc[i]))
    return out


def tm_congruence(em: Em, TM, M: Sequence[Sequence[V]]) -> List[List[V]]:
    """A^T M A for symmetric M (only the upper triangle is computed, then mirrored)."""
    n = len(M)
    MA = [[None] * n for _ in range(n)]
    for r in range(n):
        for k in range(n):
            terms = []
            for m in range(n):
                for coef, f in TM[m][k]:
                    terms.append((coef, f, M[r][m]))
            MA[r][k] = em.lincomb(terms)
    out = [[None] * n for _ in range(n)]
    for i in range(n):
        for k in range(i, n):
            terms = []
            for m in range(n):
                for coef, f in TM[m][i]:
                    terms.append((coef, f, MA[m][k]))
            out[i][k] = em.lincomb(terms)
            out[k][i] = out[i][k]
    return out


class Xform:
    """Emit-time handle on X_j(q_j) in factored form."""

    def __init__(self, jd: JointData, tr: Trig):
        self.jd = jd
        self.TJ = _tm_joint(jd, tr)
        self.TT = _tm_const(jd.XT)

    def apply(self, em: Em, m: Sequence[V]) -> List[V]:
        """X m (motion vector parent -> child)."""
        return tm_matvec(em, self.TJ, tm_matvec(em, self.TT, m))

    def applyT(self, em: Em, f: Sequence[V], acc: Optional[Sequence[V]] = None) -> List[V]:
        """X^T f (force vector child -> parent) (+ acc)."""
        return tm_matvec(em, self.TT, tm_matvec(em, self.TJ, f, transpose=True), transpose=True, acc=acc)

    def congruence(self, em: Em, M) -> List[List[V]]:
        """X^T M X for symmetric M."""
        return tm_congruence(em, self.TT, tm_congruence(em, self.TJ, M))


# column S of the motion cross-product matrix crm(v) = [[w x, 0], [v_lin x, w x]]: entry i = sign * v[k]
MXS_TABLE = {
    0: {1: (1, 2), 2: (-1, 1), 4: (1, 5), 5: (-1, 4)},
    1: {0: (-1, 2), 2: (1, 0), 3: (-1, 5), 5: (1, 3)},
    2: {0: (1, 1), 1: (-1, 0), 3: (1, 4), 4: (-1, 3)},
    3: {4: (1, 2), 5: (-1, 1)},
    4: {3: (-1, 2), 5: (1, 0)},
    5: {3: (1, 1), 4: (-1, 0)},
}


def mxS(em: Em, S: int, v: Sequence[V], alpha: Optional[V] = None) -> List[V]:
    """Column S of the motion cross-product matrix of v, times alpha (reference _test.py:522-608)."""
    a = alpha
    table = MXS_TABLE[S]
    out = []
    for i in range(6):
        if i in table:
            sg, k = table[i]
            out.append(em.lincomb([(float(sg), v[k], a)]))
        else:
            out.append(ZERO)
    return out


def fxv(em: Em, a: Sequence[V], b: Sequence[V], acc: Optional[Sequence[V]] = None) -> List[V]:
    """Fx(a) b = a x* b (reference _test.py:649-664)."""
    rows = [
        [(-1, 2, 1), (1, 1, 2), (-1, 5, 4), (1, 4, 5)],
        [(1, 2, 0), (-1, 0, 2), (1, 5, 3), (-1, 3, 5)],
        [(-1, 1, 0), (1, 0, 1), (-1, 4, 3), (1, 3, 4)],
        [(-1, 2, 4), (1, 1, 5)],
        [(1, 2, 3), (-1, 0, 5)],
        [(-1, 1, 3), (1, 0, 4)],
    ]
    return [em.lincomb([(float(sg), a[i], b[k]) for sg, i, k in rows[r]], None if acc is None else acc[r])
            for r in range(6)]


def inertia_mul(em: Em, I, v: Sequence[V], acc: Optional[Sequence[V]] = None) -> List[V]:
    """I v for a constant 6x6 spatial inertia (zeros folded)."""
    return [em.lincomb([(float(I[r, k]), v[k]) for k in range(6)], None if acc is None else acc[r]) for r in range(6)]


def unit(S: int, val: V) -> List[V]:
    return [val if k == S else ZERO for k in range(6)]


def hom_tables(robot):
    """Per joint the homogeneous child->parent transform as Xhom_j(q) = H0 + sin(q) Hs + cos(q) Hc + q Hq
    (4x4 numeric), recovered by sampling the robot's sympy matrices (reference consumer:
    helpers/_topology_helpers.py:333-504, robot.get_Xmats_hom_ordered_by_id)."""
    mats = robot.get_Xmats_hom_ordered_by_id()
    out = []
    for j, M in enumerate(mats):
        syms = list(M.free_symbols)
        ev = lambda val: np.array(M.subs({s_: val for s_ in syms}).evalf(), dtype=np.float64).reshape(4, 4)
        revolute = int(np.argmax(np.abs(robot.get_S_by_id(j)))) < 3
        Z = np.zeros((4, 4))
        if revolute:
            A, B, C = ev(0.0), ev(np.pi / 2), ev(np.pi)
            H0, Hc = _snap((A + C) / 2), _snap((A - C) / 2)
            Hs = _snap(B - H0)
            out.append((H0, Hs, Hc, Z))
            th = 0.37
            assert np.abs(H0 + np.sin(th) * Hs + np.cos(th) * Hc - ev(th)).max() < 1e-9
        else:
            A, B = ev(0.0), ev(1.0)
            out.append((_snap(A), Z, Z, _snap(B - A)))
    return out
