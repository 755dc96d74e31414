"""Robot model layer: the `robot` object the reference expects from the
un-vendored URDFParser package, re-implemented from the API the reference
consumes (first call sites cited per method).

The model is a fixed-base kinematic tree of 1-DoF joints with principal-axis
motion subspaces (one-hot S, reference helpers/_spatial_algebra_helpers.py:13).
Ids are a DFS pre-order so that parent id < child id and every subtree is a
contiguous id range (required by reference _test.py:190-200 and
algorithms/_direct_minv.py:179-183).
"""
from __future__ import annotations

import copy
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import sympy as sp

from . import spatial as spx

THETA = sp.symbols("theta")

_MOVABLE = ("revolute", "continuous", "prismatic")


@dataclass
class LinkDesc:
    name: str
    mass: float = 0.0
    com: Sequence[float] = (0.0, 0.0, 0.0)
    com_rpy: Sequence[float] = (0.0, 0.0, 0.0)
    inertia: Sequence[float] = (0.0, 0.0, 0.0, 0.0, 0.0, 0.0)  # ixx ixy ixz iyy iyz izz

    def spatial_inertia(self) -> np.ndarray:
        ixx, ixy, ixz, iyy, iyz, izz = self.inertia
        Ic = np.array([[ixx, ixy, ixz], [ixy, iyy, iyz], [ixz, iyz, izz]], dtype=np.float64)
        E = spx.rpy_to_E(self.com_rpy)  # link frame -> inertial frame coordinates
        Ic = E.T @ Ic @ E
        return spx.spatial_inertia(float(self.mass), np.asarray(self.com, dtype=np.float64), Ic)


@dataclass
class JointDesc:
    name: str
    jtype: str
    parent: str
    child: str
    xyz: Sequence[float] = (0.0, 0.0, 0.0)
    rpy: Sequence[float] = (0.0, 0.0, 0.0)
    axis: Sequence[float] = (0.0, 0.0, 1.0)
    damping: float = 0.0
    limits: Sequence[float] = (0.0, 0.0, 0.0, 0.0)  # lower upper effort velocity


class Link:
    """Rigid body attached to a movable joint (fixed-joint children merged in)."""

    def __init__(self, name: str, lid: int, Imat: np.ndarray):
        self.name = name
        self.id = lid
        self.Imat = Imat

    def get_name(self):
        return self.name

    def get_id(self):
        return self.id

    def get_spatial_inertia(self):
        return self.Imat


class Joint:
    def __init__(self, desc: JointDesc, jid: int, parent_id: int, E: np.ndarray, r: np.ndarray, axis_index: int):
        self.name = desc.name
        self.jtype = "revolute" if desc.jtype == "continuous" else desc.jtype
        self.parent = desc.parent
        self.child = desc.child
        self.damping = float(desc.damping)
        self.limits = tuple(float(x) for x in desc.limits)
        self.id = jid
        self.parent_id = parent_id
        # tree transform parent-body frame -> this joint's zero-position frame
        self.E = E
        self.r = r
        # 0..2 principal axis; motion subspace index is axis (revolute) or 3+axis (prismatic)
        self.axis_index = axis_index
        self.S_ind = axis_index if self.jtype == "revolute" else 3 + axis_index
        self.S = np.zeros(6)
        self.S[self.S_ind] = 1.0
        self.bfs_level = 0
        self.XT = spx.snap(spx.X_from_E_r(E, r))

    # ---- URDFParser-style accessors -------------------------------------
    def get_name(self):
        return self.name

    def get_id(self):
        return self.id

    def get_parent(self):
        return self.parent

    def get_child(self):
        return self.child

    def get_type(self):
        return self.jtype

    def get_damping(self):
        return self.damping

    def get_joint_limits(self):
        return list(self.limits)

    def get_S(self):
        return self.S.copy()

    # ---- transforms -------------------------------------------------------
    def XJ_numeric(self, q: float) -> np.ndarray:
        if self.jtype == "revolute":
            E = spx.rot_axis(self.axis_index, q)
            return spx.X_from_E_r(E, np.zeros(3))
        d = np.zeros(3)
        d[self.axis_index] = q
        return spx.X_from_E_r(np.eye(3), d)

    def X_numeric(self, q: float) -> np.ndarray:
        return self.XJ_numeric(q) @ self.XT

    def XJ_sympy(self) -> sp.Matrix:
        th = THETA
        if self.jtype == "revolute":
            c, s = sp.cos(th), sp.sin(th)
            a = self.axis_index
            if a == 0:
                E = sp.Matrix([[1, 0, 0], [0, c, s], [0, -s, c]])
            elif a == 1:
                E = sp.Matrix([[c, 0, -s], [0, 1, 0], [s, 0, c]])
            else:
                E = sp.Matrix([[c, s, 0], [-s, c, 0], [0, 0, 1]])
            X = sp.zeros(6, 6)
            X[:3, :3] = E
            X[3:, 3:] = E
            return X
        d = [0, 0, 0]
        d[self.axis_index] = th
        X = sp.eye(6)
        X[3:, :3] = -sp.Matrix([[0, -d[2], d[1]], [d[2], 0, -d[0]], [-d[1], d[0], 0]])
        return X

    def X_sympy(self) -> sp.Matrix:
        XT = sp.Matrix(6, 6, lambda i, j: _sp_num(self.XT[i, j]))
        return (self.XJ_sympy() * XT).applyfunc(sp.expand)

    # homogeneous child -> parent transform (position of child origin in the
    # parent frame in the last column), q-dependent part on the right
    def Xhom_sympy(self) -> sp.Matrix:
        R_T = spx.snap(self.E.T)
        T = sp.eye(4)
        for i in range(3):
            for j in range(3):
                T[i, j] = _sp_num(R_T[i, j])
            T[i, 3] = _sp_num(self.r[i])
        th = THETA
        J = sp.eye(4)
        if self.jtype == "revolute":
            c, s = sp.cos(th), sp.sin(th)
            a = self.axis_index
            if a == 0:
                R = sp.Matrix([[1, 0, 0], [0, c, -s], [0, s, c]])
            elif a == 1:
                R = sp.Matrix([[c, 0, s], [0, 1, 0], [-s, 0, c]])
            else:
                R = sp.Matrix([[c, -s, 0], [s, c, 0], [0, 0, 1]])
            J[:3, :3] = R
        else:
            J[self.axis_index, 3] = th
        return (T * J).applyfunc(sp.expand)


class FixedJoint:
    """A fixed joint kept as a kinematic target (reference
    algorithms/_eepose_gradient_hessian.py:37, :92-93)."""

    def __init__(self, name: str, fid: int, parent_link: str, parent_joint_id: int, E: np.ndarray, r: np.ndarray):
        self.name = name
        self.id = fid
        self.parent = parent_link
        self.parent_joint_id = parent_joint_id
        self.E = E
        self.r = r
        self.jtype = "fixed"

    def get_name(self):
        return self.name

    def get_id(self):
        return self.id

    def get_parent(self):
        return self.parent

    def get_type(self):
        return "fixed"

    def Xhom_sympy(self) -> sp.Matrix:
        R_T = spx.snap(self.E.T)
        T = sp.eye(4)
        for i in range(3):
            for j in range(3):
                T[i, j] = _sp_num(R_T[i, j])
            T[i, 3] = _sp_num(self.r[i])
        return T


def _sp_num(x: float):
    x = float(x)
    if x == 0.0:
        return sp.Integer(0)
    if x == 1.0:
        return sp.Integer(1)
    if x == -1.0:
        return sp.Integer(-1)
    return sp.Float(x, 17)


def _axis_index(axis: Sequence[float], name: str) -> int:
    a = spx.snap(np.asarray(axis, dtype=np.float64))
    for k in range(3):
        e = np.zeros(3)
        e[k] = 1.0
        if np.allclose(a, e):
            return k
    raise ValueError(
        f"joint {name}: axis {tuple(axis)} is not a positive principal axis; the reference "
        "requires one-hot motion subspaces (helpers/_spatial_algebra_helpers.py:13)"
    )


def _axis_frame(axis: Sequence[float], name: str) -> Tuple[int, np.ndarray]:
    """(k, Q): principal index k and the coordinate rotation Q with Q @ axis = e_k.  The reference requires one-hot
    motion subspaces (helpers/_spatial_algebra_helpers.py:13); a joint about any other axis -- negative principal
    axes are common in real URDFs -- gets its joint frame re-oriented by Q (and the child link's frame by Q^T), which
    leaves the kinematics and dynamics in terms of q unchanged."""
    a = np.asarray(axis, dtype=np.float64)
    nrm = float(np.linalg.norm(a))
    if nrm < 1e-12:
        raise ValueError(f"joint {name}: zero axis")
    a = a / nrm
    k = int(np.argmax(np.abs(a)))
    e = np.zeros(3)
    e[k] = 1.0
    if np.allclose(a, e, atol=1e-12):
        return k, np.eye(3)
    if np.allclose(a, -e, atol=1e-12):
        Q = -np.eye(3)
        Q[(k + 1) % 3, (k + 1) % 3] = 1.0          # half turn about the next principal axis
        return k, Q
    # general axis: Rodrigues rotation about a x e_k taking a onto e_k
    v = np.cross(a, e)
    s_, c_ = float(np.linalg.norm(v)), float(a @ e)
    K = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]]) / s_
    return k, np.eye(3) + s_ * K + (1.0 - c_) * (K @ K)


def _compose(E1, r1, E2, r2):
    """Frame A->B (E1, r1) followed by B->C (E2, r2) gives A->C."""
    return spx.snap(E2 @ E1), spx.snap(r1 + E1.T @ r2)


class Robot:
    """Fixed-base kinematic tree with the query API of URDFParser's robot."""

    floating_base = False

    def __init__(self, name: str, links: Sequence[LinkDesc], joints: Sequence[JointDesc]):
        self.name = name
        self._link_desc: Dict[str, LinkDesc] = {l.name: l for l in links}
        self._joint_desc = list(joints)
        self.joints: List[Joint] = []
        self.links: List[Link] = []
        self.fixed_joints: List[FixedJoint] = []
        self._build()

    # ------------------------------------------------------------------ build
    def _build(self):
        child_links = {j.child for j in self._joint_desc}
        roots = [l for l in self._link_desc if l not in child_links]
        if len(roots) != 1:
            raise ValueError(f"robot {self.name}: expected exactly one root link, found {roots}")
        self.root_link = roots[0]
        by_parent: Dict[str, List[JointDesc]] = {}
        for j in self._joint_desc:
            if j.parent not in self._link_desc or j.child not in self._link_desc:
                raise ValueError(f"joint {j.name}: unknown link {j.parent} / {j.child}")
            by_parent.setdefault(j.parent, []).append(j)

        body_I: Dict[int, np.ndarray] = {-1: np.zeros((6, 6))}
        body_name: Dict[int, str] = {-1: self.root_link}

        def visit(link_name: str, body_id: int, E: np.ndarray, r: np.ndarray):
            # (E, r): body frame -> frame of `link_name`
            X = spx.X_from_E_r(E, r)
            body_I[body_id] = body_I[body_id] + spx.transform_inertia_to_parent(
                X, self._link_desc[link_name].spatial_inertia()
            )
            for jd in by_parent.get(link_name, []):
                Ej, rj = _compose(E, r, spx.rpy_to_E(jd.rpy), np.asarray(jd.xyz, dtype=np.float64))
                if jd.jtype in _MOVABLE:
                    jid = len(self.joints)
                    k, Q = _axis_frame(jd.axis, jd.name)
                    joint = Joint(jd, jid, body_id, spx.snap(Q @ Ej), rj, k)
                    self.joints.append(joint)
                    body_I[jid] = np.zeros((6, 6))
                    body_name[jid] = jd.child
                    visit(jd.child, jid, spx.snap(Q.T), np.zeros(3))
                elif jd.jtype == "fixed":
                    self.fixed_joints.append(FixedJoint(jd.name, len(self.fixed_joints), jd.parent, body_id, Ej, rj))
                    visit(jd.child, body_id, Ej, rj)
                else:
                    raise ValueError(f"joint {jd.name}: unsupported type {jd.jtype}")

        visit(self.root_link, -1, np.eye(3), np.zeros(3))
        n = len(self.joints)
        self.n = n
        self.base_Imat = spx.snap(body_I[-1])
        self.links = [Link(body_name[j], j, _sym(spx.snap(body_I[j]))) for j in range(n)]
        self.parent = [j.parent_id for j in self.joints]
        self.children: List[List[int]] = [[] for _ in range(n)]
        for j, p in enumerate(self.parent):
            if p >= 0:
                self.children[p].append(j)
        for j in self.joints:
            j.bfs_level = 0 if j.parent_id < 0 else self.joints[j.parent_id].bfs_level + 1
        self._ancestors = []
        for j in range(n):
            a, p = [], self.parent[j]
            while p >= 0:
                a.append(p)
                p = self.parent[p]
            self._ancestors.append(sorted(a))
        self._subtree = [[j] for j in range(n)]
        for j in range(n - 1, -1, -1):
            for c in self.children[j]:
                self._subtree[j] += self._subtree[c]
            self._subtree[j].sort()
            st = self._subtree[j]
            assert st == list(range(j, j + len(st))), "ids must be DFS pre-order"
        self._Xsym = None
        self._Xfuncs = None
        self._Xhom = None

    # ----------------------------------------------------------------- counts
    def get_num_pos(self):  # reference helpers/_topology_helpers.py:11
        return self.n

    def get_num_vel(self):
        return self.n

    def get_num_joints(self):  # reference helpers/_topology_helpers.py:5
        return self.n

    def get_num_fixed_joints(self):
        return len(self.fixed_joints)

    def get_name(self):
        return self.name

    def get_leaf_nodes(self):  # reference algorithms/_eepose_gradient_hessian.py:35
        return [j for j in range(self.n) if not self.children[j]]

    def get_total_leaf_nodes(self):  # reference GRiDCodeGenerator.py:96
        return len(self.get_leaf_nodes())

    def get_total_ancestor_count(self):  # reference helpers/_topology_helpers.py:527
        return sum(len(a) for a in self._ancestors)

    def get_total_subtree_count(self):
        return sum(len(s) for s in self._subtree)

    def get_max_num_ancestors(self):  # reference algorithms/_crba.py:174
        return max(len(a) for a in self._ancestors)

    def get_max_bfs_level(self):  # reference _test.py:15
        return max(j.bfs_level for j in self.joints)

    def get_max_bfs_width(self):  # reference algorithms/_direct_minv.py:3
        return max(len(self.get_ids_by_bfs_level(l)) for l in range(self.get_max_bfs_level() + 1))

    # --------------------------------------------------------------- topology
    def get_parent_id(self, jid):  # reference helpers/_topology_helpers.py:569
        return self.parent[jid]

    def get_parent_id_array(self):  # reference _test.py:16
        return list(self.parent)

    def get_parent_ids(self, jids):
        return [self.parent[j] for j in jids]

    def get_unique_parent_ids(self, jids):  # reference _test.py:454
        return sorted(set(self.parent[j] for j in jids))

    def has_repeated_parents(self, jids):  # reference algorithms/_inverse_dynamics.py:290
        p = [self.parent[j] for j in jids]
        return len(set(p)) != len(p)

    def get_ids_by_bfs_level(self, level):  # reference _test.py:27
        return [j.id for j in self.joints if j.bfs_level == level]

    def get_bfs_level_by_id(self, jid):  # reference algorithms/_inverse_dynamics_gradient.py:357
        return self.joints[jid].bfs_level

    def get_ancestors_by_id(self, jid):  # fresh mutable list: reference _test.py:355-356 appends to it
        return list(self._ancestors[jid])

    def get_subtree_by_id(self, jid):  # includes jid: reference _test.py:151
        return list(self._subtree[jid])

    def get_is_ancestor_of(self, a, b):  # a is an ancestor of b
        return a in self._ancestors[b]

    def get_is_in_subtree_of(self, a, b):  # a is in the subtree of b
        return a in self._subtree[b]

    def is_serial_chain(self):  # reference helpers/_topology_helpers.py:138
        return all(self.parent[j] == j - 1 for j in range(self.n))

    def get_jid_ancestor_ids(self, include_joint=False):  # reference algorithms/_idsva_so.py:13
        jids, ancs = [], []
        for j in range(self.n):
            cols = list(self._ancestors[j]) + ([j] if include_joint else [])
            for a in cols:
                jids.append(j)
                ancs.append(a)
        return jids, ancs

    def get_jid_ancestor_st_ids(self, include_joint=False):  # reference algorithms/_idsva_so.py:569
        jids, ancs, sts = [], [], []
        for j in range(self.n):
            cols = list(self._ancestors[j]) + ([j] if include_joint else [])
            for a in cols:
                for s in self._subtree[j]:
                    jids.append(j)
                    ancs.append(a)
                    sts.append(s)
        return jids, ancs, sts

    # ------------------------------------------------------------- joint data
    def get_joint_by_id(self, jid):
        return self.joints[jid]

    def get_joint_by_name(self, name):
        for j in self.joints:
            if j.name == name:
                return j
        return None

    def get_joints_ordered_by_id(self, reverse=False):  # reference helpers/_topology_helpers.py:740
        return list(reversed(self.joints)) if reverse else list(self.joints)

    def get_link_by_id(self, lid):
        return self.links[lid]

    def get_link_by_name(self, name):
        for l in self.links:
            if l.name == name:
                return l
        return None

    def get_links_ordered_by_id(self, reverse=False):
        return list(reversed(self.links)) if reverse else list(self.links)

    def get_fixed_joint_by_name(self, name):
        for j in self.fixed_joints:
            if j.name == name:
                return j
        return None

    def get_fixed_joint_by_id(self, fid):
        return self.fixed_joints[fid]

    def get_S_by_id(self, jid):  # one-hot 6-vector
        return self.joints[jid].S.copy()

    def get_S_ind_by_id(self, jid):
        return self.joints[jid].S_ind

    def get_S_inds(self, n=None):  # list of str: reference helpers/_topology_helpers.py:579
        n = self.n if n is None else n
        return [str(self.joints[j].S_ind) for j in range(n)]

    def are_Ss_identical(self, jids):  # reference helpers/_spatial_algebra_helpers.py:9
        return len(set(self.joints[j].S_ind for j in jids)) <= 1

    def get_damping_by_id(self, jid):  # reference _test.py:105
        return self.joints[jid].damping

    # --------------------------------------------------------------- matrices
    def get_Xmats_ordered_by_id(self, reverse=False):  # sympy 6x6 in `theta`
        if self._Xsym is None:
            self._Xsym = [j.X_sympy() for j in self.joints]
        return list(reversed(self._Xsym)) if reverse else list(self._Xsym)

    def get_Xmat_by_id(self, jid):
        return self.get_Xmats_ordered_by_id()[jid]

    def get_Xmat_Funcs_ordered_by_id(self, reverse=False):  # reference _test.py:20-21
        if self._Xfuncs is None:
            self._Xfuncs = [_XFunc(j) for j in self.joints]
        return list(reversed(self._Xfuncs)) if reverse else list(self._Xfuncs)

    def get_Xmat_Func_by_id(self, jid):  # reference _test.py:96
        return self.get_Xmat_Funcs_ordered_by_id()[jid]

    def get_Imats_ordered_by_id(self, reverse=False):  # base inertia at index 0: reference _test.py:17
        out = [self.base_Imat.copy()] + [l.Imat.copy() for l in self.links]
        return list(reversed(out)) if reverse else out

    def get_Imat_by_id(self, lid):  # reference _test.py:274
        return self.links[lid].Imat

    def get_Imats_dict_by_id(self):  # reference _test.py:127 (deep-copied by the caller)
        return {l.id: l.Imat.copy() for l in self.links}

    def _hom(self):
        if self._Xhom is None:
            th = THETA
            X = [j.Xhom_sympy() for j in self.joints]
            dX = [m.diff(th) for m in X]
            d2X = [m.diff(th) for m in dX]
            self._Xhom = (X, dX, d2X)
        return self._Xhom

    def get_Xmats_hom_ordered_by_id(self, reverse=False, include_fixed_joints=False):
        X = list(self._hom()[0])
        if include_fixed_joints:
            X += [f.Xhom_sympy() for f in self.fixed_joints]
        return list(reversed(X)) if reverse else X

    def get_dXmats_hom_ordered_by_id(self, reverse=False):
        X = list(self._hom()[1])
        return list(reversed(X)) if reverse else X

    def get_d2Xmats_hom_ordered_by_id(self, reverse=False):
        X = list(self._hom()[2])
        return list(reversed(X)) if reverse else X

    # ------------------------------------------------ numeric tables for kernels
    def numeric_tables(self) -> dict:
        """Flat numeric description consumed by the C oracle and the emitter.

        X_j(q) = X0 + sin(q) Xs + cos(q) Xc + q Xq  (exact for principal-axis
        revolute and prismatic joints)."""
        n = self.n
        X0 = np.zeros((n, 6, 6))
        Xs = np.zeros((n, 6, 6))
        Xc = np.zeros((n, 6, 6))
        Xq = np.zeros((n, 6, 6))
        for j in self.joints:
            XT = j.XT
            if j.jtype == "revolute":
                # XJ(q) = A0 + s As + c Ac
                J0 =j.XJ_numeric(0.0)            # c=1, s=0
                J90 = j.XJ_numeric(np.pi / 2)     # c=0, s=1
                J180 = j.XJ_numeric(np.pi)        # c=-1, s=0
                A0 = spx.snap((J0 + J180) / 2)
                Ac = spx.snap((J0 - J180) / 2)
                As = spx.snap(J90 - A0)
                X0[j.id] = spx.snap(A0 @ XT)
                Xs[j.id] = spx.snap(As @ XT)
                Xc[j.id] = spx.snap(Ac @ XT)
            else:
                J0 = j.XJ_numeric(0.0)
                J1 = j.XJ_numeric(1.0)
                X0[j.id] = spx.snap(J0 @ XT)
                Xq[j.id] = spx.snap((J1 - J0) @ XT)
        return dict(
            n=n,
            parent=np.array(self.parent, dtype=np.int32),
            S_ind=np.array([j.S_ind for j in self.joints], dtype=np.int32),
            X0=X0, Xs=Xs, Xc=Xc, Xq=Xq,
            I=np.stack([l.Imat for l in self.links]),
            damping=np.array([j.damping for j in self.joints], dtype=np.float64),
        )


class _XFunc:
    """Callable q_j -> 6x6 numpy motion transform (replaces URDFParser's
    lambdified sympy matrix; same values, no sympy in the hot loop)."""

    def __init__(self, joint: Joint):
        self.joint = joint

    def __call__(self, q):
        return self.joint.X_numeric(float(q))


def _sym(I):
    return 0.5 * (I + I.T)
