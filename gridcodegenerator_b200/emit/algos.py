"""Per-knot-point algorithm bodies, emitted as straight-line CUDA for one thread.

Thread mapping (B200 design): one lane owns one knot point, one
warp owns a tile of 32 knot points; quantities that must survive between
passes live in a per-thread state area `S[slot*NT]` (shared memory, lane
interleaved => conflict-free, immediate-offset addressing), everything else in
registers.  The algorithms follow the reference's CPU oracle:

  RNEA            reference _test.py:5-107      (emitted twin: algorithms/_inverse_dynamics.py:33-321)
  direct Minv     reference _test.py:117-226    (algorithms/_direct_minv.py:23-453)
  FD              reference _test.py:498-501    (algorithms/_forward_dynamics.py:21-115)
  grad RNEA       reference _test.py:229-494    (algorithms/_inverse_dynamics_gradient.py:27-775)
  grad FD         reference _test.py:496-520    (algorithms/_forward_dynamics_gradient.py:7-62)
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

from .core import Em, V, ZERO, ONE, const
from .spatial import Model, Trig, Xform, fxv, inertia_mul, mxS, unit


def fence_policy(n: int) -> bool:
    """Fences at joint / column boundaries for robots whose live state would not fit the register file."""
    mode = os.environ.get("GRID_EMIT_FENCES", "auto")
    if mode in ("0", "1"):
        return mode == "1"
    return n > 8


def regs_only_policy(n: int) -> bool:
    """Robots small enough that the whole per-knot state fits registers + a little local memory."""
    mode = os.environ.get("GRID_EMIT_REGS_ONLY", "auto")
    if mode in ("0", "1"):
        return mode == "1"
    return n <= 8


class Ctx:
    def __init__(self, model: Model, em: Optional[Em] = None):
        self.m = model
        self.em = em or Em()
        self.em.fences = fence_policy(model.n)
        self.em.regs_only = regs_only_policy(model.n)
        self.n = model.n
        self.X: List[Xform] = []
        self.q: List[V] = []
        self.qd: List[V] = []
        self.gravity = V(s="gravity")

    # -------------------------------------------------------------- inputs
    def read_vec(self, ptr: str, prefix: str, joints=None) -> List[V]:
        """Named copies of a per-knot input vector (all joints, or only `joints`: None elsewhere)."""
        return [self.em.named(f"{prefix}{j}", f"{ptr}[{j}]") if (joints is None or j in joints) else None
                for j in range(self.n)]

    def prologue(self, q_ptr: str = "q", seeded: bool = False, joints=None):
        """sin/cos of every revolute joint angle (reference helpers/_topology_helpers.py:194-195) and the
        factored transform handles.  With `seeded`, q_j is AD seed j: sin/cos (or q for prismatic joints) carry
        their first and second derivatives with respect to q_j (used when Em.ad_order > 0)."""
        em = self.em
        self.q = self.read_vec(q_ptr, "q_", joints)
        if joints is None or not getattr(self, "X", None):
            self.X = [None] * self.n
            self.trig = [None] * self.n
        for j, jd in enumerate(self.m.joints):
            if joints is not None and j not in joints:
                continue
            if jd.revolute:
                em.line(f"T sn{j}, cs{j}; grid_sincos(q_{j}, &sn{j}, &cs{j});")
                sn, cs = V(s=f"sn{j}"), V(s=f"cs{j}")
                if seeded:
                    nsn, ncs = em.tmp(f"-sn{j}"), em.tmp(f"-cs{j}")
                    sn.d = {j: V(s=cs.s, d={j: V(s=nsn.s)})}
                    cs.d = {j: V(s=nsn.s, d={j: V(s=ncs.s)})}
                tr = Trig(sn, cs, self.q[j])
            else:
                if seeded:
                    self.q[j] = V(s=self.q[j].s, d={j: ONE})
                tr = Trig(None, None, self.q[j])
            self.X[j] = Xform(jd, tr)
            self.trig[j] = (tr.s, tr.c, tr.q)

    def seed_vec(self, vec: Sequence[V], first_seed: int) -> List[V]:
        """Mark vec[j] as AD seed first_seed + j."""
        return [V(s=v.s, d={first_seed + j: ONE}) for j, v in enumerate(vec)]


class Pusher:
    """Accumulate child -> parent contributions in a fixed order.

    A contribution to the joint processed next (parent == j-1) stays in registers; any other is a
    read-modify-write on the parent's state slots.  This replaces the reference's shared-memory atomicAdd
    for repeated parents (algorithms/_inverse_dynamics.py:290-294) with a deterministic order."""

    def __init__(self, em: Em, name, width: int = 6):
        self.em, self.name, self.width = em, name, width
        self.carry: Dict[int, List[V]] = {}
        self.in_state: Dict[int, bool] = {}

    def set_base(self, j: int, vec: Sequence[V]):
        self.em.store_vec(self.name, j, vec)
        self.in_state[j] = True

    def push(self, p: int, j: int, fn):
        """fn(acc) -> list of V computing contribution (+ acc)."""
        em = self.em
        if p == j - 1:
            self.carry[p] = fn(self.carry.get(p))
        else:
            base = em.load_vec(self.name, p, self.width) if self.in_state.get(p) else None
            em.store_vec(self.name, p, fn(base))
            self.in_state[p] = True

    def pull(self, j: int) -> List[V]:
        em = self.em
        base = em.load_vec(self.name, j, self.width) if self.in_state.get(j) else None
        c = self.carry.pop(j, None)
        if base is None and c is None:
            return [ZERO] * self.width
        if base is None:
            return list(c)
        if c is None:
            return base
        return em.vadd(base, c)


# ====================================================================== RNEA
def rnea_forward(cx: Ctx, qd: Sequence[V], qdd: Optional[Sequence[V]], keep: bool, fname: str = "f", joints=None,
                 keep_Iv: bool = True, on_va=None):
    """v, a, f per joint in id order (ids are topologically sorted).  f_j -> state `fname`; with `keep`
    also v_j, I v_j and X a_parent (needed by the gradient).  `joints`: restrict to the joints of some base-rooted
    trees (they do not interact with the rest of a fixed-base robot)."""
    em, m = cx.em, cx.m
    V_: Dict[int, List[V]] = {}
    A_: Dict[int, List[V]] = {}
    grav = [ZERO] * 5 + [cx.gravity]
    nonadj_parent = {p for j, p in enumerate(m.parent) if p >= 0 and p != j - 1}
    for j, jd in enumerate(m.joints):
        if joints is not None and j not in joints:
            continue
        em.comment(f"RNEA forward, joint {j}")
        p, S = jd.parent, jd.S
        if p >= 0:
            if p != j - 1:
                vp, ap = em.load_vec("v", p), em.load_vec("a", p)
            else:
                vp, ap = V_[p], A_[p]
            vJ = cx.X[j].apply(em, vp)
            aJ = cx.X[j].apply(em, ap)
        else:
            vJ = [ZERO] * 6
            aJ = cx.X[j].apply(em, grav)
        v = list(vJ)
        v[S] = em.add(vJ[S], qd[j])
        a = list(aJ)
        if qdd is not None:
            a[S] = em.add(a[S], qdd[j])
        if p >= 0:
            cr = mxS(em, S, v, qd[j])
            a = [em.add(x, y) for x, y in zip(a, cr)]
        Iv = inertia_mul(em, jd.I, v)
        f = fxv(em, v, Iv, acc=inertia_mul(em, jd.I, a))
        em.store_vec(fname, j, f)
        if keep:
            em.store_vec("v", j, v)
            if keep_Iv:
                em.store_vec("Iv", j, Iv)
            em.store_vec("Xa", j, aJ)
            if j in nonadj_parent:
                em.store_vec("a", j, a)
        elif j in nonadj_parent:
            em.store_vec("v", j, v)
            em.store_vec("a", j, a)
        if on_va is not None:
            on_va(j, v, a)
        V_[j], A_[j] = v, a
        V_.pop(j - 2, None)
        A_.pop(j - 2, None)


def rnea_backward(cx: Ctx, qd: Sequence[V], keep_f: bool, fname: str = "f", joints=None, on_f=None) -> List[V]:
    """f_parent += X^T f, c_j = S^T f_j + damping qd_j (reference _test.py:78-107)."""
    em, m = cx.em, cx.m
    ps = Pusher(em, fname)
    for j in range(m.n):
        ps.in_state[j] = True
    c: List[V] = [ZERO] * m.n
    for j in range(m.n - 1, -1, -1):
        if joints is not None and j not in joints:
            continue
        jd = m.joints[j]
        em.comment(f"RNEA backward, joint {j}")
        f = ps.pull(j)
        if keep_f:
            em.store_vec(fname, j, f)
        if on_f is not None:
            on_f(j, f)
        c[j] = em.lincomb([(1.0, f[jd.S]), (jd.damping, qd[j])])
        if jd.parent >= 0:
            ps.push(jd.parent, j, lambda acc, f=f, j=j: cx.X[j].applyT(em, f, acc=acc))
    return c


# ================================================================ direct Minv
def _sym_store(em: Em, name, j, M):
    for r in range(6):
        for k in range(r, 6):
            em.store((name, j, r * 6 + k), M[r][k])


def _sym_load(em: Em, name, j):
    M = [[None] * 6 for _ in range(6)]
    for r in range(6):
        for k in range(r, 6):
            M[r][k] = em.load((name, j, r * 6 + k))
            M[k][r] = M[r][k]
    return M


def articulated_inertias(cx: Ctx, bias=None):
    """Backward recursion IA_p += X^T (IA - U D^-1 U^T) X (reference _test.py:117-184 without the F part,
    algorithms/_direct_minv.py:127-148, :226-337).  Stores U_j (6) and 1/D_j in the state.

    bias(j, IA, U, Dinv, UD, Ia): optional hook called once per joint with the final articulated inertia
    (Ia and UD are None for base-rooted joints); ABA uses it to propagate the bias force alongside."""
    em, m = cx.em, cx.m
    carry: Dict[int, List[List[V]]] = {}
    in_state: Dict[int, bool] = {}
    for j in range(m.n - 1, -1, -1):
        jd = m.joints[j]
        em.comment(f"articulated inertia, joint {j}")
        S, p = jd.S, jd.parent
        IA = [[const(jd.I[r, k]) for k in range(6)] for r in range(6)]
        extra = []
        if in_state.get(j):
            extra.append(_sym_load(em, "IA", j))
        if j in carry:
            extra.append(carry.pop(j))
        for E in extra:
            for r in range(6):
                for k in range(r, 6):
                    IA[r][k] = em.add(IA[r][k], E[r][k])
                    IA[k][r] = IA[r][k]
        U = [IA[r][S] for r in range(6)]
        Dinv = em.recip(U[S])
        em.store_vec("U", j, U)
        em.store(("Dinv", j, 0), Dinv)
        if p < 0:
            if bias is not None:
                bias(j, IA, U, Dinv, None, None)
            continue
        UD = [em.mul(U[r], Dinv) for r in range(6)]
        Ia = [[None] * 6 for _ in range(6)]
        for r in range(6):
            for k in range(r, 6):
                if r == S or k == S:
                    Ia[r][k] = ZERO  # row/column S of IA - U U^T / D vanishes identically
                else:
                    Ia[r][k] = em.lincomb([(1.0, IA[r][k]), (-1.0, UD[r], U[k])])
                Ia[k][r] = Ia[r][k]
        if bias is not None:
            bias(j, IA, U, Dinv, UD, Ia)
        XtIaX = cx.X[j].congruence(em, Ia)
        if p == j - 1:
            carry[p] = XtIaX
        else:
            if in_state.get(p):
                old = _sym_load(em, "IA", p)
                XtIaX = [[em.add(old[r][k], XtIaX[r][k]) if k >= r else None for k in range(6)] for r in range(6)]
                for r in range(6):
                    for k in range(r):
                        XtIaX[r][k] = XtIaX[k][r]
            _sym_store(em, "IA", p, XtIaX)
            in_state[p] = True


def minv_columns(cx: Ctx) -> Dict[tuple, V]:
    """Minv[j][s] for j >= s, one column s at a time: the response qdd to a unit torque at joint s with zero
    velocity and gravity (articulated-body form of the reference's F-matrix passes, _test.py:117-202).
    Entries between different base-rooted trees are structural zeros."""
    em, m = cx.em, cx.m
    n = m.n
    Minv: Dict[tuple, V] = {}

    def UD(j):
        return em.load_vec("U", j), em.load(("Dinv", j, 0))

    for s in range(n):
        em.comment(f"Minv column {s}")
        path = [s] + sorted(m.ancestors[s], reverse=True)  # s, parent, ..., root
        uu: Dict[int, V] = {s: ONE}
        U, Dinv = UD(s)
        pa = [em.mul(U[r], Dinv) for r in range(6)]
        for idx in range(len(path) - 1):
            j, p = path[idx], path[idx + 1]
            pA = cx.X[j].applyT(em, pa)
            U, Dinv = UD(p)
            uu[p] = em.neg(pA[m.joints[p].S])
            w = em.mul(uu[p], Dinv)
            pa = [em.lincomb([(1.0, pA[r]), (1.0, U[r], w)]) for r in range(6)]
        # forward: joints on the path (ancestors first), then every j > s of the same tree
        root = path[-1]
        tree_end = root + len(m.subtree[root])
        todo = sorted(m.ancestors[s]) + list(range(s, tree_end))
        acc: Dict[int, List[V]] = {}
        needed_later = set()
        for j in todo:
            pj = m.parent[j]
            if pj >= 0:
                needed_later.add(pj)
        for j in todo:
            jd = m.joints[j]
            U, Dinv = UD(j)
            if jd.parent >= 0:
                aJ = cx.X[j].apply(em, acc[jd.parent])
                ua = em.dot(U, aJ)
            else:
                aJ = [ZERO] * 6
                ua = ZERO
            qddj = em.lincomb([(1.0, uu.get(j, ZERO), Dinv), (-1.0, ua, Dinv)])
            if j >= s:
                Minv[(j, s)] = qddj
            if j in needed_later:
                a = list(aJ)
                a[jd.S] = em.add(a[jd.S], qddj)
                acc[j] = a
    return Minv


def minv_get(Minv: Dict[tuple, V], r: int, c: int) -> V:
    return Minv.get((max(r, c), min(r, c)), ZERO)


def store_minv(cx: Ctx, Minv):
    for (j, s), v in Minv.items():
        cx.em.store(("Minv", j, s), v)


def load_minv(cx: Ctx, keys) -> Dict[tuple, V]:
    return {k: cx.em.load(("Minv", k[0], k[1])) for k in keys}


# ===================================================================== FD
def fd_finish(cx: Ctx, Minv, u: Sequence[V], c: Sequence[V]) -> List[V]:
    """qdd = Minv (u - c) with symmetric indexing (reference algorithms/_forward_dynamics.py:21-50)."""
    em, n = cx.em, cx.n
    umc = [em.sub(u[j], c[j]) for j in range(n)]
    return [em.lincomb([(1.0, minv_get(Minv, r, k), umc[k]) for k in range(n)]) for r in range(n)]


# ============================================================== grad RNEA
def pairs_policy() -> bool:
    """Carry the dq and dqd derivative columns of a joint as one packed pair (FFMA2 on sm_100a).
    Off by default: measured on B200 (profiles/r01c_ab.log) FFMA2 occupies the FMA pipe for two cycles, so packing
    saves issue slots and code bytes (8160 -> 6808 SASS instructions for iiwa14 grad FD) but not pipe time, and the
    pair (un)packing MOVs make it 6 % slower than the scalar form (32.7 vs 30.7 us at 65536 knots)."""
    return os.environ.get("GRID_EMIT_PAIRS", "0") != "0"


def pairs_from_column(n: int) -> int:
    """First derivative column emitted in packed form (columns before it are scalar); n = none.
    Hybrid use: the scalar form is faster while the instruction stream is served from the instruction cache (first
    64 KB of the kernel), but code beyond that is instruction-fetch bound (~2x the time per instruction,
    profiles/r01b_noinst_by_address.txt) -- there the packed form's one-third fewer instructions win and the FFMA2's
    second pipe cycle is hidden behind the fetch."""
    if pairs_policy():
        return 0
    v = os.environ.get("GRID_EMIT_PAIRS_FROM", "")
    return int(v) if v else n


def rnea_grad_columns(cx: Ctx, qd: Sequence[V], sink, cols: Optional[Sequence[int]] = None):
    """Analytical dc/dq and dc/dqd, one derivative column c at a time (all columns, or only `cols`).  Needs v, Iv, Xa and the accumulated f in
    the state (rnea_forward(keep=True) + rnea_backward(keep_f=True)).  Follows reference _test.py:229-488; the
    column of joint c touches only subtree(c) (forward) and subtree(c) + ancestors(c) (backward), which is the
    reference's column compression (algorithms/_inverse_dynamics_gradient.py:55-98) made static.

    The q column and the qd column of joint c obey the same linear recursions with different seeds, so they are
    carried as ONE pair-valued column (x = d/dq_c, y = d/dqd_c): every multiply-add on it is a packed FFMA2.

    sink(c, dc): list of n V's, pair-valued (x: dc/dq_c, y: dc/dqd_c); structural zeros are constants."""
    em, m = cx.em, cx.m
    n = m.n
    packed = pairs_policy()
    first_packed = pairs_from_column(n)

    def pk(a: Sequence[V], b: Sequence[V]) -> List:
        return [em.mkpair(x, y) for x, y in zip(a, b)] if packed else [(x, y) for x, y in zip(a, b)]

    # unpacked fallback = two scalar columns run through the same code (tuple-valued entries)
    def lift(fn):
        def run(*vecs, **kw):
            if packed:
                return fn(*vecs, **kw)
            outs = []
            for h in (0, 1):
                args = [[e[h] if isinstance(e, tuple) else e for e in vec] if isinstance(vec, list) else vec for vec in vecs]
                kws = {k: ([e[h] if isinstance(e, tuple) else e for e in v] if isinstance(v, list) else v) for k, v in kw.items()}
                outs.append(fn(*args, **kws))
            return [(x, y) for x, y in zip(*outs)]
        return run

    for c in (range(n) if cols is None else cols):
        packed = c >= first_packed
        em.comment(f"grad RNEA column {c}")
        sub = m.subtree[c]
        DV: Dict[int, tuple] = {}
        ps = Pusher(em, ("df", c))
        if not packed:
            ps = _TuplePusher(em, ("df", c))
        last = sub[-1]
        for j in sub:
            jd = m.joints[j]
            S = jd.S
            if j != c:
                em.comment(f"grad RNEA column {c}, forward joint {j}")
            v = em.load_vec("v", j)
            Iv = em.load_vec("Iv", j) if getattr(cx, "has_Iv", True) else inertia_mul(em, jd.I, v)
            if j == c:
                Xa = em.load_vec("Xa", j)
                dvq = mxS(em, S, v) if jd.parent >= 0 else [ZERO] * 6     # MxXv == Mxv since S x S = 0
                dvd = unit(S, ONE)
                daq = [em.add(x, y) for x, y in zip(mxS(em, S, dvq, qd[j]), mxS(em, S, Xa))]
                dad = mxS(em, S, v)                                         # mxS(S) qd = 0
                dv, da = pk(dvq, dvd), pk(daq, dad)
            else:
                pdv, pda = DV[jd.parent]
                dv = lift(lambda x: cx.X[j].apply(em, x))(pdv)
                da = lift(lambda xa, xv: [em.add(x, y) for x, y in zip(cx.X[j].apply(em, xa), mxS(em, S, xv, qd[j]))])(pda, dv)
            if m.children[j]:
                DV[j] = (dv, da)
            # df = dv x* Iv + I da + v x* (I dv)      (reference _test.py:389-424)
            df = lift(lambda xv, xa: fxv(em, v, inertia_mul(em, jd.I, xv),
                                         acc=fxv(em, xv, Iv, acc=inertia_mul(em, jd.I, xa))))(dv, da)
            if j == last and not m.children[j]:
                ps.carry[j] = df   # consumed immediately by the backward sweep
            else:
                ps.set_base(j, df)
        dc: List = [ZERO] * n
        rel = sorted(set(sub) | set(m.ancestors[c]), reverse=True)
        em.comment(f"grad RNEA column {c}, backward sweep")
        for j in rel:
            jd = m.joints[j]
            f = ps.pull(j)
            dc[j] = f[jd.S]
            if j == c and jd.damping != 0.0:
                if packed:
                    dc[j] = em.add(f[jd.S], em.mkpair(ZERO, const(jd.damping)))
                else:
                    dc[j] = (f[jd.S][0], em.lincomb([(1.0, f[jd.S][1])], const(jd.damping)))
            if jd.parent >= 0:
                if j == c:
                    Mxf = mxS(em, jd.S, em.load_vec("f", j))                  # - X^T mxS(f) (reference :434-437)
                    if packed:
                        f = [em.sub(x, em.mkpair(y, ZERO)) for x, y in zip(f, Mxf)]
                    else:
                        f = [(em.sub(x[0], y), x[1]) for x, y in zip(f, Mxf)]
                ps.push(jd.parent, j, lambda acc, f=f, j=j: lift(lambda ff, acc=None: cx.X[j].applyT(em, ff, acc=acc))(f, acc=acc))
        for j in rel:
            em.release_vec(("df", c), j)
            if not packed:
                em.release_vec(("df", c, 1), j)
        sink(c, dc)


def column_loop_policy() -> bool:
    """Emit the derivative-column recursion of grad RNEA as ONE runtime loop over columns (rnea_grad_columns_loop)
    instead of n unrolled copies (rnea_grad_columns)."""
    return os.environ.get("GRID_COLUMN_LOOP", "0") != "0"


class MVec:
    """A vector that is assigned inside uniform branches of the column loop: plain C variables when the per-knot
    state is register-resident (small robots) or `force_regs`, else per-thread state slots."""

    def __init__(self, em: Em, name: str, size: int, force_regs: bool = False):
        self.em, self.name, self.size = em, name, size
        self.regs = em.regs_only or force_regs
        if self.regs:
            em.line("T " + ", ".join(f"{name}_{k}" for k in range(size)) + ";")
        else:
            self.idx = [em.slot("mv", name, k) for k in range(size)]

    def _ref(self, k: int) -> str:
        return f"{self.name}_{k}" if self.regs else f"{self.em.state_name}[{self.idx[k]}*{self.em.stride_name}]"

    def get(self, k: int) -> V:
        return self.em.tmp(self._ref(k))          # a copy: later assignments must not change what was read

    def getv(self, lo: int = 0, hi: Optional[int] = None) -> List[V]:
        return [self.get(k) for k in range(lo, self.size if hi is None else hi)]

    def set(self, k: int, v: V):
        from .core import lit
        self.em.line(f"{self._ref(k)} = {lit(v.c) if v.is_const else v.s};")

    def setv(self, vec: Sequence[V], lo: int = 0):
        for k, v in enumerate(vec):
            self.set(lo + k, v)


def rnea_grad_columns_loop(cx: Ctx, qd: Sequence[V], sink: str):
    """Same mathematics as rnea_grad_columns, emitted as one runtime loop over the derivative column c_.

    Column c_ touches subtree(c_) going forward and subtree(c_) + ancestors(c_) going backward.  The loop body holds
    one block per joint for each sweep, guarded by a condition on c_ that is uniform over the CTA (every lane
    works on the same column), so no lane ever executes a block its column does not need: the executed operation
    count equals the unrolled form's up to the seed-specific constant folding, while the code shrinks from
    sum_c |subtree(c)| joint blocks to n.  For iiwa14 grad FD that is ~110 KB -> ~60 KB of SASS, under the 64 KB
    beyond which ncu shows instruction-fetch stalls doubling the time per instruction (profiles/r01b_noinst_by_address.txt).

    sink: "id"  -- write [dc/dq | dc/dqd] column c_ to o;
          "fd"  -- -M^-1 dc by the sparse L^T D L solve (factors in the state: ("L", k, i), ("Dinv2", i, 0)), fused into
                   the sweeps: L^-T during the backward sweep (same support), D^-1, then L^-1 ascending."""
    em, m, n = cx.em, cx.m, cx.n
    from .core import lit
    bits = lambda js: sum(1 << j for j in js)
    msk = lambda js: f"(({bits(js)}ull >> c_) & 1ull)"
    last = [max(m.subtree[j]) for j in range(n)]
    root = [min(m.ancestors[j]) if m.ancestors[j] else j for j in range(n)]
    nonadj_parent = {p for j, p in enumerate(m.parent) if p >= 0 and p != j - 1}
    chain_child = [(j + 1 < n and m.parent[j + 1] == j) for j in range(n)]
    O = lambda col, row: f"o[({col}*{n} + {row})*OS]"

    base_ops = em.n_ops
    executed = 0

    _acquire_out(em)
    em.line("#pragma unroll 1")
    em.line(f"for (int c_ = 0; c_ < {n}; ++c_) {{")
    F = [MVec(em, f"cF{j}", 12) for j in range(n)]
    DV = {p: MVec(em, f"cDV{p}", 12) for p in sorted(nonadj_parent)}
    DA = {p: MVec(em, f"cDA{p}", 12) for p in sorted(nonadj_parent)}
    cdv, cda = MVec(em, "ccv", 12, force_regs=True), MVec(em, "cca", 12, force_regs=True)
    if sink == "fd":
        XA = {j: MVec(em, f"cXA{j}", 2) for j in range(n) if m.children[j]}
        Y = [MVec(em, f"cY{j}", 2) for j in range(n)]
        XX = {j: MVec(em, f"cXX{j}", 2) for j in range(n) if m.children[j]}
        for j in XA:
            XA[j].setv([ZERO, ZERO])

    def finish_forward(j, jd, v, Iv, dv, da):
        # dv, da: [q half (6), qd half (6)];  df = dv x* Iv + I da + v x* (I dv)   (reference _test.py:389-424)
        df = []
        for h in (0, 1):
            xv, xa = dv[h], da[h]
            df += fxv(em, v, inertia_mul(em, jd.I, xv), acc=fxv(em, xv, Iv, acc=inertia_mul(em, jd.I, xa)))
        F[j].setv(df)
        if chain_child[j]:
            cdv.setv(dv[0] + dv[1])
            cda.setv(da[0] + da[1])
        if j in nonadj_parent:
            DV[j].setv(dv[0] + dv[1])
            DA[j].setv(da[0] + da[1])

    # ---------------------------------------------------------------- forward sweep
    for j in range(n):
        jd = m.joints[j]
        S, p = jd.S, jd.parent
        t0 = em.n_ops
        em.line(f"if (c_ == {j}) {{")
        em.comment(f"grad RNEA column loop: seed at joint {j}")
        v, Iv, Xa = em.load_vec("v", j), em.load_vec("Iv", j), em.load_vec("Xa", j)
        dvq = mxS(em, S, v) if p >= 0 else [ZERO] * 6                  # MxXv == Mxv since S x S = 0
        dvd = unit(S, ONE)
        daq = [em.add(x, y) for x, y in zip(mxS(em, S, dvq, qd[j]), mxS(em, S, Xa))]
        dad = mxS(em, S, v)                                            # mxS(S) qd = 0
        finish_forward(j, jd, v, Iv, [dvq, dvd], [daq, dad])
        em.line("}")
        executed += em.n_ops - t0
        if m.ancestors[j]:
            t0 = em.n_ops
            em.line(f"else if {msk(m.ancestors[j])} {{")
            em.comment(f"grad RNEA column loop: forward joint {j}")
            v, Iv = em.load_vec("v", j), em.load_vec("Iv", j)
            if p == j - 1:
                pdv, pda = cdv.getv(), cda.getv()
            else:
                pdv, pda = DV[p].getv(), DA[p].getv()
            dv, da = [], []
            for h in (0, 1):
                xv = cx.X[j].apply(em, pdv[6 * h:6 * h + 6])
                xa = [em.add(x, y) for x, y in zip(cx.X[j].apply(em, pda[6 * h:6 * h + 6]), mxS(em, S, xv, qd[j]))]
                dv.append(xv)
                da.append(xa)
            finish_forward(j, jd, v, Iv, dv, da)
            em.line("}")
            executed += (em.n_ops - t0) * len(m.ancestors[j])
        if m.children[j]:
            em.line(f"else if (c_ > {j} && c_ <= {last[j]}) {{")     # strict ancestor of the column's joint
            F[j].setv([ZERO] * 12)
            em.line("}")

    # ---------------------------------------------------------------- backward sweep (+ L^-T and D^-1 of the solve)
    for j in range(n - 1, -1, -1):
        jd = m.joints[j]
        S, p = jd.S, jd.parent
        cond = msk(m.ancestors[j] + [j])
        ncols = len(m.ancestors[j]) + 1
        if m.children[j]:
            cond += f" || (c_ > {j} && c_ <= {last[j]})"
            ncols += last[j] - j
        t0 = em.n_ops
        em.line(f"if ({cond}) {{")
        em.comment(f"grad RNEA column loop: backward joint {j}")
        f = F[j].getv()
        fq, fd_ = f[:6], f[6:]
        em.line(f"T cdq_ = {fq[S].s}, cdd_ = {fd_[S].s};")
        special = p >= 0 or jd.damping != 0.0
        if p >= 0:
            em.line("T " + ", ".join(f"cg{k}_ = {fq[k].s}" for k in range(6)) + ";")
        if special:
            t1 = em.n_ops
            em.line(f"if (c_ == {j}) {{")
            if p >= 0:
                Mxf = mxS(em, S, em.load_vec("f", j))                  # - X^T mxS(f) (reference _test.py:434-437)
                for k in range(6):
                    if not Mxf[k].is_zero():
                        em.line(f"cg{k}_ = cg{k}_ - {lit(Mxf[k].c) if Mxf[k].is_const else Mxf[k].s};")
                        em.n_ops += 1
            if jd.damping != 0.0:
                em.line(f"cdd_ = cdd_ + {lit(jd.damping)};")
                em.n_ops += 1
            em.line("}")
            executed -= (em.n_ops - t1) * (ncols - 1)                  # the special case runs for one column only
        dcq, dcd = V(s="cdq_"), V(s="cdd_")
        if p >= 0:
            gq = [V(s=f"cg{k}_") for k in range(6)]
            acc = F[p].getv()
            F[p].setv(cx.X[j].applyT(em, gq, acc=acc[:6]) + cx.X[j].applyT(em, fd_, acc=acc[6:]))
        if sink == "id":
            em.line(f"{O('c_', j)} = cdq_; {O(f'({n} + c_)', j)} = cdd_;")
        else:
            x = [dcq, dcd]
            if m.children[j]:
                xa = XA[j].getv()
                x = [em.add(dcq, xa[0]), em.add(dcd, xa[1])]
            for a in m.ancestors[j]:
                Lja = em.load(("L", j, a))
                cur = XA[a].getv()
                XA[a].setv([em.lincomb([(1.0, cur[h]), (-1.0, Lja, x[h])]) for h in (0, 1)])
            Dj = em.load(("Dinv2", j, 0))
            Y[j].setv([em.lincomb([(-1.0, x[h], Dj)]) for h in (0, 1)])
        em.line("}")
        executed += (em.n_ops - t0) * ncols
        if ncols < n:
            if sink == "id":
                em.line(f"else {{ {O('c_', j)} = T(0); {O(f'({n} + c_)', j)} = T(0); }}")
            else:
                em.line("else {")
                Y[j].setv([ZERO, ZERO])
                em.line("}")

    # ---------------------------------------------------------------- L^-1 ascending and the output column
    if sink == "fd":
        for i in range(n):
            lo, hi = root[i], last[root[i]]
            whole = lo == 0 and hi == n - 1
            t0 = em.n_ops
            if not whole:
                em.line(f"if (c_ >= {lo} && c_ <= {hi}) {{")
            else:
                em.line("{")
            y = Y[i].getv()
            x = []
            for h in (0, 1):
                terms = [(1.0, y[h])]
                for a in m.ancestors[i]:
                    terms.append((-1.0, em.load(("L", i, a)), XX[a].get(h)))
                x.append(em.lincomb(terms))
            if m.children[i]:
                XX[i].setv(x)
            from_v = lambda v_: lit(v_.c) if v_.is_const else v_.s
            em.line(f"{O('c_', i)} = {from_v(x[0])}; {O(f'({n} + c_)', i)} = {from_v(x[1])};")
            em.line("}")
            executed += (em.n_ops - t0) * (hi - lo + 1)
            if not whole:
                em.line(f"else {{ {O('c_', i)} = T(0); {O(f'({n} + c_)', i)} = T(0); }}")
    em.line("}")
    em.n_ops = base_ops + executed          # executed (not static) operation count per knot


class _TuplePusher(Pusher):
    """Pusher over (q, qd) tuples of scalars: the unpacked (GRID_EMIT_PAIRS=0) form of a pair column."""

    def __init__(self, em, name):
        super().__init__(em, name)
        self.h = [Pusher(em, name), Pusher(em, name + (1,))]

    def set_base(self, j, vec):
        for h in (0, 1):
            self.h[h].set_base(j, [e[h] for e in vec])

    def push(self, p, j, fn):
        # fn(acc) works on tuple vectors; run it once with tuple accumulators and scatter the halves
        if p == j - 1:
            self.carry[p] = fn(self.carry.get(p))
            return
        base = None
        if self.h[0].in_state.get(p):
            base = [(x, y) for x, y in zip(self.em.load_vec(self.h[0].name, p), self.em.load_vec(self.h[1].name, p))]
        out = fn(base)
        for h in (0, 1):
            self.em.store_vec(self.h[h].name, p, [e[h] for e in out])
            self.h[h].in_state[p] = True

    def pull(self, j):
        base = None
        if self.h[0].in_state.get(j):
            base = [(x, y) for x, y in zip(self.em.load_vec(self.h[0].name, j), self.em.load_vec(self.h[1].name, j))]
        c = self.carry.pop(j, None)
        if base is None and c is None:
            return [(ZERO, ZERO)] * 6
        if base is None:
            return list(c)
        if c is None:
            return base
        return [(self.em.add(a[0], b[0]), self.em.add(a[1], b[1])) for a, b in zip(base, c)]

    @property
    def carry(self):
        return self.__dict__.setdefault("_carry", {})

    @carry.setter
    def carry(self, v):
        self.__dict__["_carry"] = v


def _halves(em: Em, e):
    """(x, y) scalar views of a pair-valued entry (pair V, tuple of scalars, or a broadcast scalar)."""
    if isinstance(e, tuple):
        return e
    return em.half(e, 0), em.half(e, 1)


# ============================================================ kernel bodies
def _acquire_out(em: Em):
    """Once per body, before its first output write: the tile's output staging buffer may still be read by the
    previous tile's TMA bulk store; the runtime waits for that here (GRID_BEFORE_FIRST_OUTPUT), i.e. after the
    passes that produce no output, instead of before the whole body."""
    if not getattr(em, "_out_acquired", False):
        em._out_acquired = True
        em.line("GRID_BEFORE_FIRST_OUTPUT();")


def _out(em: Em, idx: int, v: V, ptr: str = "o"):
    """Output element idx of this knot.  OS is the element stride of the output row the body was given: 1 for a
    row in shared memory / global memory, 32 for the lane-interleaved global scratch of the transposed-store mode."""
    from .core import lit
    _acquire_out(em)
    em.line(f"{ptr}[{idx}*OS] = {lit(v.c) if v.is_const else v.s};")


def _sink_product(em: Em, n: int, c_: int, dc, minv):
    """Columns c_ and n + c_ of df_du = -Minv dc_du (reference algorithms/_forward_dynamics_gradient.py:48-62)."""
    if all(not isinstance(e, tuple) for e in dc):
        for r in range(n):
            x, y = _halves(em, em.lincomb([(-1.0, minv(r, k), dc[k]) for k in range(n)]))
            _out(em, c_ * n + r, x)
            _out(em, (n + c_) * n + r, y)
        return
    for r in range(n):
        for h in (0, 1):
            _out(em, (h * n + c_) * n + r, em.lincomb([(-1.0, minv(r, k), _halves(em, dc[k])[h]) for k in range(n)]))


class Body:
    """Result of emitting one algorithm: C statements + number of per-thread state slots."""

    def __init__(self, name: str, args: str, cx: Ctx, doc: str):
        self.name, self.args, self.doc = name, args, doc
        self.code = cx.em.code()
        self.nslots = cx.em.nslots
        self.n_ops = cx.em.n_ops


def body_inverse_dynamics_vaf(model: Model, use_qdd: bool) -> Body:
    """RNEA's spatial velocities, accelerations and (accumulated) forces per joint: the reference's s_vaf array,
    [v (6n) | a (6n) | f (6n)], 6-vector of joint j at 6 j (inverse_dynamics_inner_vaf, algorithms/_inverse_dynamics.py
    :33-321; oracle _test.py:5-107, which returns the same three arrays)."""
    cx = Ctx(model)
    cx.prologue()
    em, n = cx.em, model.n
    qd = cx.read_vec("qd", "qd_")
    qdd = cx.read_vec("qdd", "qdd_") if use_qdd else None

    def on_va(j, v, a):
        for r in range(6):
            _out(em, 6 * j + r, v[r])
            _out(em, 6 * n + 6 * j + r, a[r])

    def on_f(j, f):
        for r in range(6):
            _out(em, 12 * n + 6 * j + r, f[r])

    rnea_forward(cx, qd, qdd, keep=False, on_va=on_va)
    rnea_backward(cx, qd, keep_f=False, on_f=on_f)
    return Body("inverse_dynamics_vaf_inner", "", cx, "RNEA v, a, f per joint: [v | a | f], 6 n values each")


def body_inverse_dynamics(model: Model, use_qdd: bool) -> Body:
    cx = Ctx(model)
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    qdd = cx.read_vec("qdd", "qdd_") if use_qdd else None
    rnea_forward(cx, qd, qdd, keep=False)
    c = rnea_backward(cx, qd, keep_f=False)
    for j, v in enumerate(c):
        _out(cx.em, j, v)
    return Body("inverse_dynamics_inner", "", cx, "c = RNEA(q, qd, qdd)")


def body_direct_minv(model: Model) -> Body:
    cx = Ctx(model)
    cx.prologue()
    articulated_inertias(cx)
    Minv = minv_columns(cx)
    n = model.n
    for col in range(n):
        for r in range(n):
            _out(cx.em, col * n + r, minv_get(Minv, r, col))
    return Body("direct_minv_inner", "", cx, "Minv(q), dense symmetric, column-major")


def body_forward_dynamics(model: Model) -> Body:
    cx = Ctx(model)
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    u = cx.read_vec("u", "u_")
    rnea_forward(cx, qd, None, keep=False)
    c = rnea_backward(cx, qd, keep_f=False)
    articulated_inertias(cx)
    Minv = minv_columns(cx)
    qdd = fd_finish(cx, Minv, u, c)
    for j, v in enumerate(qdd):
        _out(cx.em, j, v)
    return Body("forward_dynamics_inner", "", cx, "qdd = Minv (u - RNEA(q, qd, 0))")


def body_inverse_dynamics_gradient(model: Model, use_qdd: bool) -> Body:
    cx = Ctx(model)
    n = model.n
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    qdd = cx.read_vec("qdd", "qdd_") if use_qdd else None
    rnea_forward(cx, qd, qdd, keep=True)
    rnea_backward(cx, qd, keep_f=True)

    def sink(c, dc):
        for r in range(n):
            x, y = _halves(cx.em, dc[r])
            _out(cx.em, c * n + r, x)
            _out(cx.em, (n + c) * n + r, y)

    if column_loop_policy():
        rnea_grad_columns_loop(cx, qd, "id")
    else:
        rnea_grad_columns(cx, qd, sink)
    return Body("inverse_dynamics_gradient_inner", "", cx, "[dc/dq | dc/dqd], column-major n x 2n")


def ltdl_factor(cx: Ctx, Mv: Dict[tuple, V], joints=None):
    """Sparse L^T D L factorisation of the joint-space inertia matrix (Featherstone RBDA table 6.3): M = L^T D L with
    L unit lower triangular and L[k][i] != 0 only for i an ancestor of k, so branch sparsity is kept.  Returns
    (L: dict (k, i) -> V, Dinv: list of V).  Used by grad FD to apply M^-1 without forming it: the reference
    computes Minv explicitly (algorithms/_direct_minv.py) and multiplies (_forward_dynamics_gradient.py:48-62);
    factor + solves give the same result in ~half the operations of the analytical Minv."""
    em, m, n = cx.em, cx.m, cx.n
    H: Dict[tuple, V] = {}
    act = range(n) if joints is None else sorted(joints)
    for i in act:
        H[(i, i)] = Mv[(i, i)]
        for j in m.ancestors[i]:
            H[(i, j)] = Mv[(min(i, j), max(i, j))]          # lower triangle, row i, column j = ancestor
    L: Dict[tuple, V] = {}
    Dinv: List[V] = [ZERO] * n
    for k in reversed(act):
        Dinv[k] = em.recip(H[(k, k)])
        i = m.parent[k]
        while i >= 0:
            a = em.mul(H[(k, i)], Dinv[k])
            j = i
            while j >= 0:
                H[(i, j)] = em.lincomb([(1.0, H[(i, j)]), (-1.0, a, H[(k, j)])])
                j = m.parent[j]
            L[(k, i)] = a
            i = m.parent[i]
    return L, Dinv


def ltdl_solve(cx: Ctx, L, Dinv, b: Sequence[V], sign: float = 1.0, joints=None) -> List[V]:
    """x = sign * M^-1 b with M = L^T D L (b scalar- or pair-valued).  With `joints` (whole base-rooted trees) only
    that diagonal block is solved; the other entries of x are structural zeros."""
    em, m, n = cx.em, cx.m, cx.n
    if joints is not None:
        act = sorted(joints)
        x = [b[i] if i in joints else ZERO for i in range(n)]
        for i in reversed(act):
            j = m.parent[i]
            while j >= 0:
                x[j] = em.lincomb([(1.0, x[j]), (-1.0, L[(i, j)], x[i])])
                j = m.parent[j]
        for i in act:
            x[i] = em.lincomb([(sign, x[i], Dinv[i])])
        for i in act:
            j = m.parent[i]
            terms = [(1.0, x[i])]
            while j >= 0:
                terms.append((-1.0, L[(i, j)], x[j]))
                j = m.parent[j]
            x[i] = em.lincomb(terms)
        return x
    x = list(b)
    for i in range(n - 1, -1, -1):                      # x = L^-T b
        j = m.parent[i]
        while j >= 0:
            x[j] = em.lincomb([(1.0, x[j]), (-1.0, L[(i, j)], x[i])])
            j = m.parent[j]
    x = [em.lincomb([(sign, x[i], Dinv[i])]) for i in range(n)]
    for i in range(n):                                  # x = L^-1 x
        j = m.parent[i]
        terms = [(1.0, x[i])]
        while j >= 0:
            terms.append((-1.0, L[(i, j)], x[j]))
            j = m.parent[j]
        x[i] = em.lincomb(terms)
    return x


def fdgrad_solver_policy() -> str:
    return os.environ.get("GRID_FDGRAD_SOLVER", "ltdl")


def body_forward_dynamics_gradient(model: Model, linearization: bool = False) -> Body:
    """grad FD.  With `linearization` the same pass also writes M^-1 (= dqdd/du) and qdd behind the gradient:
    o = [dqdd/dq | dqdd/dqd | Minv | qdd], 3 n^2 + n elements -- the (A, B, f) triple an MPC / DDP caller needs
    from one launch.  The reference attempts the Minv side output inside its grad FD body
    (algorithms/_forward_dynamics_gradient.py:56-59) without a barrier."""
    cx = Ctx(model)
    em, n = cx.em, model.n
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    u = cx.read_vec("u", "u_")
    em.comment("---- pass 1: bias forces c = RNEA(q, qd, 0)")
    rnea_forward(cx, qd, None, keep=False)
    c = rnea_backward(cx, qd, keep_f=False)
    if fdgrad_solver_policy() == "minv":
        em.comment("---- Minv")
        articulated_inertias(cx)
        Minv = minv_columns(cx)
        qdd = fd_finish(cx, Minv, u, c)
        store_minv(cx, Minv)
        keys = list(Minv.keys())
    else:
        em.comment("---- M = CRBA(q), M = L^T D L")
        Mv = crba_matrix(cx)
        L, Dinv = ltdl_factor(cx, Mv)
        for k_, v_ in L.items():
            em.store(("L",) + k_, v_)
        for i_, v_ in enumerate(Dinv):
            em.store(("Dinv2", i_, 0), v_)
        qdd = ltdl_solve(cx, L, Dinv, [em.sub(u[j], c[j]) for j in range(n)])
    if linearization:
        em.comment("---- side outputs: Minv (dense symmetric, column-major) and qdd")
        if fdgrad_solver_policy() == "minv":
            cols = [[minv_get(Minv, r, k) for r in range(n)] for k in range(n)]
        else:
            # column k of M^-1 = solve against unit vector k; structural zeros of the sparse right-hand side fold away
            cols = [ltdl_solve(cx, L, Dinv, [ONE if r == k else ZERO for r in range(n)]) for k in range(n)]
        for k in range(n):
            for r in range(n):
                # one value per symmetric pair, so that the stored matrix is exactly symmetric
                _out(em, (2 * n + k) * n + r, cols[k][r] if r <= k else cols[r][k])
        for j in range(n):
            _out(em, 3 * n * n + j, qdd[j])
    em.comment("---- pass 2: RNEA(q, qd, qdd) keeping v, Iv, X a_parent, f")
    rnea_forward(cx, qd, qdd, keep=True)
    rnea_backward(cx, qd, keep_f=True)
    em.comment("---- grad RNEA columns and df_du = -Minv dc_du")

    def sink(c_, dc):
        if fdgrad_solver_policy() == "minv":
            Mi = load_minv(cx, keys)
            _sink_product(em, n, c_, dc, lambda r, k: minv_get(Mi, r, k))
            return
        Ll = {k_: em.load(("L",) + k_) for k_ in L}
        Dl = [em.load(("Dinv2", i_, 0)) for i_ in range(n)]
        if all(not isinstance(e, tuple) for e in dc):
            x = ltdl_solve(cx, Ll, Dl, dc, sign=-1.0)
            for r in range(n):
                a, b = _halves(em, x[r])
                _out(em, c_ * n + r, a)
                _out(em, (n + c_) * n + r, b)
        else:
            for h in (0, 1):
                x = ltdl_solve(cx, Ll, Dl, [_halves(em, e)[h] for e in dc], sign=-1.0)
                for r in range(n):
                    _out(em, (h * n + c_) * n + r, x[r])

    if column_loop_policy() and fdgrad_solver_policy() != "minv":
        rnea_grad_columns_loop(cx, qd, "fd")
    else:
        rnea_grad_columns(cx, qd, sink)
    if linearization:
        return Body("forward_dynamics_linearization_inner", "", cx,
                    "[dqdd/dq | dqdd/dqd | Minv | qdd]: grad FD with its Minv and qdd by-products")
    return Body("forward_dynamics_gradient_inner", "", cx, "[dqdd/dq | dqdd/dqd] = -Minv [dc/dq | dc/dqd]")


def body_forward_dynamics_linearization(model: Model) -> Body:
    return body_forward_dynamics_gradient(model, linearization=True)


def body_forward_dynamics_gradient_qdd_minv(model: Model) -> Body:
    """grad FD when the caller already has qdd and Minv (reference USE_QDD_MINV_FLAG mode,
    algorithms/_forward_dynamics_gradient.py:64-105): RNEA(q, qd, qdd) -> grad RNEA -> -Minv dc_du."""
    cx = Ctx(model)
    em, n = cx.em, model.n
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    qdd = cx.read_vec("qdd", "qdd_")
    rnea_forward(cx, qd, qdd, keep=True)
    rnea_backward(cx, qd, keep_f=True)

    def sink(c_, dc):
        Mi = [[em.tmp(f"Minv[{k * n + r}]") for k in range(n)] for r in range(n)]
        _sink_product(em, n, c_, dc, lambda r, k: Mi[r][k])

    rnea_grad_columns(cx, qd, sink)
    return Body("forward_dynamics_gradient_inner", "", cx, "-Minv [dc/dq | dc/dqd] with qdd and Minv supplied")


# ===================================================================== ABA
def body_aba(model: Model) -> Body:
    """qdd by the articulated-body algorithm (reference algorithms/_aba.py:46-407; Featherstone RBDA table 7.1).
    Differences from the reference's emitted code, all in favour of the oracle: the
    velocity-product term uses each joint's own axis (the reference hard-codes mx2, _aba.py:123), joint damping
    is included (so that ABA == Minv (u - c) with the oracle's c), child contributions are accumulated in a
    fixed order instead of shared-memory atomics (_aba.py:290, :307)."""
    cx = Ctx(model)
    em, m, n = cx.em, model, model.n
    cx.prologue()
    qd = cx.read_vec("qd", "qd_")
    tau = cx.read_vec("u", "tau_")
    nonadj_parent = {p for j, p in enumerate(m.parent) if p >= 0 and p != j - 1}
    V_: Dict[int, List[V]] = {}
    pAs = Pusher(em, "pA")
    for j, jd in enumerate(m.joints):
        em.comment(f"ABA forward 1, joint {j}")
        p, S = jd.parent, jd.S
        if p >= 0:
            vp = V_[p] if p == j - 1 else em.load_vec("v", p)
            v = cx.X[j].apply(em, vp)
            v[S] = em.add(v[S], qd[j])
            c = mxS(em, S, v, qd[j])
        else:
            v = unit(S, qd[j])
            c = [ZERO] * 6
        em.store_vec("c", j, c)
        pAs.set_base(j, fxv(em, v, inertia_mul(em, jd.I, v)))
        if j in nonadj_parent:
            em.store_vec("v", j, v)
        V_[j] = v
        V_.pop(j - 2, None)

    def bias(j, IA, U, Dinv, UD, Ia):
        jd = m.joints[j]
        pA = pAs.pull(j)
        uu = em.lincomb([(1.0, tau[j]), (-jd.damping, qd[j]), (-1.0, pA[jd.S])])
        em.store(("uu", j, 0), uu)
        if jd.parent < 0:
            return
        c = em.load_vec("c", j)
        w = em.mul(uu, Dinv)
        pa = [em.lincomb([(1.0, Ia[r][k], c[k]) for k in range(6)] + [(1.0, U[r], w)], pA[r]) for r in range(6)]
        pAs.push(jd.parent, j, lambda acc, pa=pa, j=j: cx.X[j].applyT(em, pa, acc=acc))

    articulated_inertias(cx, bias=bias)
    A_: Dict[int, List[V]] = {}
    grav = [ZERO] * 5 + [cx.gravity]
    for j, jd in enumerate(m.joints):
        em.comment(f"ABA forward 2, joint {j}")
        p, S = jd.parent, jd.S
        if p >= 0:
            apar = A_[p] if p == j - 1 else em.load_vec("a", p)
        else:
            apar = grav
        ap = em.vadd(cx.X[j].apply(em, apar), em.load_vec("c", j))
        U, Dinv, uu = em.load_vec("U", j), em.load(("Dinv", j, 0)), em.load(("uu", j, 0))
        qddj = em.mul(em.lincomb([(-1.0, U[r], ap[r]) for r in range(6)], uu), Dinv)
        _out(em, j, qddj)
        a = list(ap)
        a[S] = em.add(a[S], qddj)
        if j in nonadj_parent:
            em.store_vec("a", j, a)
        A_[j] = a
        A_.pop(j - 2, None)
    return Body("aba_inner", "", cx, "qdd = ABA(q, qd, tau)")


# ===================================================================== CRBA
def crba_matrix(cx: Ctx, joints=None) -> Dict[tuple, V]:
    """Joint-space inertia matrix M(q), full symmetric, column-major (reference algorithms/_crba.py:30-209;
    Featherstone RBDA table 6.2).  The composite inertia of joint j is final once all its descendants (ids > j,
    DFS pre-order) have been folded in, so column j of M is produced right there by walking f_h = IC_j S up the
    ancestor chain; composite inertias live in their own storage (the reference accumulates them into the shared
    I table, _crba.py:81-143, and hard-codes n = 7 offsets)."""
    em, m, n = cx.em, cx.m, cx.n
    carry: Dict[int, List[List[V]]] = {}
    in_state: Dict[int, bool] = {}
    Mv: Dict[tuple, V] = {}
    for j in range(n - 1, -1, -1):
        if joints is not None and j not in joints:
            continue
        jd = m.joints[j]
        em.comment(f"CRBA, joint {j}")
        IC = [[const(jd.I[r, k]) for k in range(6)] for r in range(6)]
        extra = []
        if in_state.get(j):
            extra.append(_sym_load(em, "IC", j))
        if j in carry:
            extra.append(carry.pop(j))
        for E in extra:
            for r in range(6):
                for k in range(r, 6):
                    IC[r][k] = em.add(IC[r][k], E[r][k])
                    IC[k][r] = IC[r][k]
        fh = [IC[r][jd.S] for r in range(6)]
        Mv[(j, j)] = fh[jd.S]
        i = j
        while m.parent[i] >= 0:
            fh = cx.X[i].applyT(em, fh)
            i = m.parent[i]
            Mv[(i, j)] = fh[m.joints[i].S]
        p = jd.parent
        if p < 0:
            continue
        XtIX = cx.X[j].congruence(em, IC)
        if p == j - 1:
            carry[p] = XtIX
        else:
            if in_state.get(p):
                old = _sym_load(em, "IC", p)
                XtIX = [[em.add(old[r][k], XtIX[r][k]) if k >= r else None for k in range(6)] for r in range(6)]
                for r in range(6):
                    for k in range(r):
                        XtIX[r][k] = XtIX[k][r]
            _sym_store(em, "IC", p, XtIX)
            in_state[p] = True
    return Mv


def body_crba(model: Model) -> Body:
    cx = Ctx(model)
    em, n = cx.em, model.n
    cx.prologue()
    Mv = crba_matrix(cx)
    for col in range(n):
        for r in range(n):
            _out(em, col * n + r, Mv.get((min(r, col), max(r, col)), ZERO))
    return Body("crba_inner", "", cx, "M(q), dense symmetric, column-major")


# ============================================================ second order
def _d1(v: V, k: int) -> V:
    return v.d[k] if (v.d and k in v.d) else ZERO


def _d2(v: V, k: int, l: int) -> V:
    """d2 v / d seed_k d seed_l (tangents are stored for l <= k only)."""
    if l > k:
        k, l = l, k
    return _d1(_d1(v, k), l)


def rnea_second_order(cx: Ctx, qd_in: Sequence[V], qdd: Sequence[V]):
    """c = RNEA(q, qd, qdd) differentiated twice at emit time (forward-mode tangents through the same straight-line
    RNEA as the first-order kernels) with seeds q_k -> k and qd_k -> n + k, qdd held fixed.  Returns c (n V's whose
    .d / .d.d hold dc/du and d2c/du2).  Stands in for the reference's hand-derived IDSVA recursion
    (algorithms/_idsva_so.py:167-912): same tensors, derived mechanically from the pinned RNEA."""
    em, n = cx.em, cx.n
    save = em.ad_order
    em.ad_order = 2
    qd = cx.seed_vec(qd_in, n)
    rnea_forward(cx, qd, qdd, keep=False, fname="f2")
    c = rnea_backward(cx, qd, keep_f=False, fname="f2")
    em.ad_order = save
    return c


def so_blocks_policy() -> bool:
    """Second-order tangents in seed-pair blocks (one pass of the RNEA per pair of joints) instead of all seed pairs
    at once.  All at once, every f_j is live with all its first- and second-order tangents at the end of the forward
    pass (iiwa14: 2 226 values per thread -> a 9.5 KB stack, the kernels run at the L2's spill bandwidth, DESIGN.md
    section 6); in blocks the live set of a pass is <= ~450 values for ~1.7x the arithmetic.  GRID_SO_BLOCKS=0 restores
    the one-pass form."""
    return os.environ.get("GRID_SO_BLOCKS", "1") != "0"


def seed_pair_blocks(model: Model):
    """(a, b), b <= a: the joint pairs whose (q, qd) seeds can give a non-zero second derivative of RNEA -- joints of
    the same base-rooted tree (the trees of a fixed-base robot do not interact)."""
    return [(a, b) for a in range(model.n) for b in range(a + 1) if b in _tree_of(model, a)]


def _so_pass_ctx(cx: Ctx, seed_joints, tree, tag: str, qd_in, qdd):
    """Inputs of one seed-pair pass: copies of the runtime scalars the RNEA reads, passed through GRID_LAUNDER so that
    the compiler cannot merge the value-level work of different passes (it would keep those values live across the
    whole kernel), with tangents only on the seeds of `seed_joints`: q_j -> seed j (through sin / cos), qd_j -> n + j."""
    em, n = cx.em, cx.n
    X = [None] * n
    qd = [None] * n
    qdd2 = [None] * n if qdd is not None else None
    for j in sorted(tree):
        jd = cx.m.joints[j]
        sn0, cs0, q0 = cx.trig[j]
        seeded = j in seed_joints
        if jd.revolute:
            sn = em.named(f"sn{j}_{tag}", f"GRID_LAUNDER({sn0.s})")
            cs = em.named(f"cs{j}_{tag}", f"GRID_LAUNDER({cs0.s})")
            if seeded:
                nsn, ncs = em.tmp(f"-{sn.s}"), em.tmp(f"-{cs.s}")
                sn = V(s=sn.s, d={j: V(s=cs.s, d={j: V(s=nsn.s)})})
                cs = V(s=cs.s, d={j: V(s=nsn.s, d={j: V(s=ncs.s)})})
            tr = Trig(sn, cs, q0)
        else:
            qv = em.named(f"q{j}_{tag}", f"GRID_LAUNDER({q0.s})")
            if seeded:
                qv = V(s=qv.s, d={j: ONE})
            tr = Trig(None, None, qv)
        X[j] = Xform(jd, tr)
        v = em.named(f"qd{j}_{tag}", f"GRID_LAUNDER({qd_in[j].s})")
        qd[j] = V(s=v.s, d={n + j: ONE}) if seeded else v
        if qdd is not None:
            qdd2[j] = em.named(f"qdd{j}_{tag}", f"GRID_LAUNDER({qdd[j].s})")
    return X, qd, qdd2


def rnea_second_order_blocks(cx: Ctx, qd_in: Sequence[V], qdd: Sequence[V], sink):
    """The tangents of rnea_second_order, produced block by block: for every joint pair (a, b), b <= a, of a tree one
    RNEA pass over that tree with the four seeds q_a, qd_a, q_b, qd_b and only the second-order pairs
    {q_a, qd_a} x {q_b, qd_b}; sink(a, b, c) receives c (c[i] for the joints of the tree; c[i].d[k].d[l], l <= k)."""
    em, n = cx.em, cx.n
    save = (em.ad_order, em.ad_pairs, cx.X)
    for (a, b) in seed_pair_blocks(cx.m):
        tree = _tree_of(cx.m, a)
        em.comment(f"==== second-order pass: seeds of joints {a} x {b}")
        K, L = (a, n + a), (b, n + b)
        em.ad_order = 2
        em.ad_pairs = {(max(k, l), min(k, l)) for k in K for l in L}
        cx.X, qd, qdd2 = _so_pass_ctx(cx, {a, b}, tree, f"p{a}_{b}", qd_in, qdd)
        rnea_forward(cx, qd, qdd2, keep=False, fname="f2", joints=tree)
        c = rnea_backward(cx, qd, keep_f=False, fname="f2", joints=tree)
        em.ad_order, em.ad_pairs, cx.X = save
        sink(a, b, c)
        for j in tree:
            em.release_vec("f2", j)


def mass_matrix_gradient(cx: Ctx) -> Dict[tuple, V]:
    """M(q) by CRBA with first-order tangents: Mv[(i, j)].d[k] = dM_ij/dq_k, i <= j
    (reference output dM_dq, algorithms/_idsva_so.py:159)."""
    em = cx.em
    save = em.ad_order
    em.ad_order = 1
    Mv = crba_matrix(cx)
    em.ad_order = save
    return Mv


def _second_order_ctx(model: Model) -> Ctx:
    cx = Ctx(model)
    cx.em.regs_only = True      # tangents ride on the SSA values; the compiler places them (registers / local memory)
    cx.em.fences = False
    cx.prologue(seeded=True)
    return cx


def body_idsva_so(model: Model) -> Body:
    """Second-order inverse-dynamics tensors (reference algorithms/_idsva_so.py; kernel :958-1028), 4 n^3 outputs:
       out[0*n^3 + i*n^2 + j*n + k] = d2 tau_i / dq_j dq_k
       out[1*n^3 + i*n^2 + j*n + k] = d2 tau_i / dqd_j dqd_k
       out[2*n^3 + i*n^2 + j*n + k] = d2 tau_i / dqd_j dq_k
       out[3*n^3 + i*n^2 + j*n + k] = d M_ij / dq_k
    (tensor order of the reference: d2tau_dq2, d2tau_dqd2, d2tau_dvdq, dM_dq, _idsva_so.py:156-159)."""
    cx = _second_order_ctx(model)
    em, n = cx.em, model.n
    qd = cx.read_vec("qd", "qd_")
    qdd = cx.read_vec("qdd", "qdd_")
    n3 = n * n * n
    if so_blocks_policy():
        done = set()

        def put(idx, v):
            _out(em, idx, v)
            done.add(idx)

        def sink(a, b, c):
            for i in sorted(_tree_of(model, a)):
                for (j, k) in ((a, b), (b, a)) if a != b else ((a, a),):
                    idx = i * n * n + j * n + k
                    put(idx, _d2(c[i], j, k))
                    put(n3 + idx, _d2(c[i], n + j, n + k))
                    put(2 * n3 + idx, _d2(c[i], n + j, k))

        rnea_second_order_blocks(cx, qd, qdd, sink)
        em.comment("---- dM/dq")
        Mv = mass_matrix_gradient(cx)
        for idx in range(n3):
            for t in range(3):
                if t * n3 + idx not in done:       # joints of different base-rooted trees: structurally zero
                    _out(em, t * n3 + idx, ZERO)
            i, j, k = idx // (n * n), (idx // n) % n, idx % n
            _out(em, 3 * n3 + idx, _d1(Mv.get((min(i, j), max(i, j)), ZERO), k))
        return Body("idsva_so_inner", "", cx, "second-order inverse dynamics tensors (4 n^3)")
    c = rnea_second_order(cx, qd, qdd)
    Mv = mass_matrix_gradient(cx)
    for i in range(n):
        for j in range(n):
            for k in range(n):
                idx = i * n * n + j * n + k
                _out(em, idx, _d2(c[i], j, k))
                _out(em, n3 + idx, _d2(c[i], n + j, n + k))
                _out(em, 2 * n3 + idx, _d2(c[i], n + j, k))
                _out(em, 3 * n3 + idx, _d1(Mv.get((min(i, j), max(i, j)), ZERO), k))
    return Body("idsva_so_inner", "", cx, "second-order inverse dynamics tensors (4 n^3)")


def body_fdsva_so(model: Model) -> Body:
    """Second-order forward-dynamics tensors (reference algorithms/_fdsva_so.py:4-85; kernel :161-244), 4 n^3:
       out[0*n^3 + i*n^2 + j*n + k] = d2 qdd_i / dq_j dq_k     = -Minv (T_qq[:,j,k] + dM_k A_q[:,j] + dM_j A_q[:,k])
       out[1*n^3 + i*n^2 + j*n + k] = d2 qdd_i / dqd_j dqd_k   = -Minv  T_vv[:,j,k]
       out[2*n^3 + i*n^2 + j*n + k] = d2 qdd_i / dqd_j dq_k    = -Minv (T_vq[:,j,k] + dM_k A_v[:,j])
       out[3*n^3 + i*n^2 + j*n + k] = d2 qdd_i / dtau_j dq_k   = -Minv (dM_k Minv[:,j])
    with T_* the IDSVA tensors at qdd = FD(q, qd, tau), dM_k = dM/dq_k, A_q = dqdd/dq, A_v = dqdd/dqd
    (implicit differentiation of ID(q, qd, FD(q, qd, tau)) = tau; the reference's inner_dq / inner_cross /
    inner_tau terms, _fdsva_so.py:52-83)."""
    cx = _second_order_ctx(model)
    em, n = cx.em, model.n
    qd = cx.read_vec("qd", "qd_")
    u = cx.read_vec("u", "u_")
    em.comment("---- qdd = Minv (u - RNEA(q, qd, 0))")
    rnea_forward(cx, qd, None, keep=False)
    c0 = rnea_backward(cx, qd, keep_f=False)
    articulated_inertias(cx)
    Minv = minv_columns(cx)
    qdd = fd_finish(cx, Minv, u, c0)
    Mi = [[minv_get(Minv, r, k) for k in range(n)] for r in range(n)]
    if so_blocks_policy():
        return _fdsva_so_blocks(cx, qd, qdd, Mi)
    em.comment("---- RNEA(q, qd, qdd) with first and second order tangents")
    c = rnea_second_order(cx, qd, qdd)
    em.comment("---- dM/dq")
    Mv = mass_matrix_gradient(cx)
    dM = [[[_d1(Mv.get((min(i, L), max(i, L)), ZERO), k) for k in range(n)] for L in range(n)] for i in range(n)]
    em.comment("---- first-order FD gradient A = -Minv dc/du")
    Aq = [[em.lincomb([(-1.0, Mi[L][m_], _d1(c[m_], j)) for m_ in range(n)]) for j in range(n)] for L in range(n)]
    Av = [[em.lincomb([(-1.0, Mi[L][m_], _d1(c[m_], n + j)) for m_ in range(n)]) for j in range(n)] for L in range(n)]
    n3 = n * n * n
    em.comment("---- d2qdd/dq2")
    for j in range(n):
        for k in range(j + 1):
            inner = [em.lincomb([(1.0, dM[L][m_][k], Aq[m_][j]) for m_ in range(n)] +
                                [(1.0, dM[L][m_][j], Aq[m_][k]) for m_ in range(n)], _d2(c[L], j, k)) for L in range(n)]
            for i in range(n):
                v = em.lincomb([(-1.0, Mi[i][L], inner[L]) for L in range(n)])
                _out(em, i * n * n + j * n + k, v)
                if k != j:
                    _out(em, i * n * n + k * n + j, v)
    em.comment("---- d2qdd/dqd2")
    for j in range(n):
        for k in range(j + 1):
            for i in range(n):
                v = em.lincomb([(-1.0, Mi[i][L], _d2(c[L], n + j, n + k)) for L in range(n)])
                _out(em, n3 + i * n * n + j * n + k, v)
                if k != j:
                    _out(em, n3 + i * n * n + k * n + j, v)
    em.comment("---- d2qdd/dqd dq and d2qdd/dtau dq")
    for j in range(n):
        for k in range(n):
            inner = [em.lincomb([(1.0, dM[L][m_][k], Av[m_][j]) for m_ in range(n)], _d2(c[L], n + j, k)) for L in range(n)]
            inner_t = [em.lincomb([(1.0, dM[L][m_][k], Mi[m_][j]) for m_ in range(n)]) for L in range(n)]
            for i in range(n):
                _out(em, 2 * n3 + i * n * n + j * n + k, em.lincomb([(-1.0, Mi[i][L], inner[L]) for L in range(n)]))
                _out(em, 3 * n3 + i * n * n + j * n + k, em.lincomb([(-1.0, Mi[i][L], inner_t[L]) for L in range(n)]))
    return Body("fdsva_so_inner", "", cx, "second-order forward dynamics tensors (4 n^3)")


def _fdsva_so_blocks(cx: Ctx, qd, qdd, Mi) -> Body:
    """body_fdsva_so with the RNEA tangents produced in seed-pair blocks: every pass (a, b) carries the first-order
    tangents of its four seeds, so the columns A_q[:, a], A_q[:, b], A_v[:, a], A_v[:, b] and the contractions of the
    pairs (a, b) / (b, a) are finished inside the pass; only Minv and dM/dq stay live between passes."""
    em, n, model = cx.em, cx.n, cx.m
    n3 = n * n * n
    em.comment("---- dM/dq")
    Mv = mass_matrix_gradient(cx)
    dM = [[[_d1(Mv.get((min(i, L), max(i, L)), ZERO), k) for k in range(n)] for L in range(n)] for i in range(n)]
    done = set()

    def put(idx, v):
        _out(em, idx, v)
        done.add(idx)

    def sink(a, b, c):
        tree = sorted(_tree_of(model, a))
        ci = lambda i: c[i] if isinstance(c[i], V) else ZERO

        def A(seed):       # -Minv dc/du_seed (rows of the tree; the other rows are structurally zero)
            return {L: em.lincomb([(-1.0, Mi[L][m_], _d1(ci(m_), seed)) for m_ in tree]) for L in tree}

        Aq = {a: A(a)}
        Av = {a: A(n + a)}
        if b != a:
            Aq[b], Av[b] = A(b), A(n + b)
        em.comment(f"---- contractions of the pairs ({a}, {b})")
        # d2qdd/dq_j dq_k (symmetric in j, k)
        inner = {L: em.lincomb([(1.0, dM[L][m_][b], Aq[a][m_]) for m_ in tree] +
                               [(1.0, dM[L][m_][a], Aq[b][m_]) for m_ in tree], _d2(ci(L), a, b)) for L in tree}
        for i in tree:
            v = em.lincomb([(-1.0, Mi[i][L], inner[L]) for L in tree])
            put(i * n * n + a * n + b, v)
            if a != b:
                put(i * n * n + b * n + a, v)
        # d2qdd/dqd_j dqd_k (symmetric)
        for i in tree:
            v = em.lincomb([(-1.0, Mi[i][L], _d2(ci(L), n + a, n + b)) for L in tree])
            put(n3 + i * n * n + a * n + b, v)
            if a != b:
                put(n3 + i * n * n + b * n + a, v)
        # d2qdd/dqd_j dq_k
        for (j, k) in ((a, b), (b, a)) if a != b else ((a, a),):
            inner = {L: em.lincomb([(1.0, dM[L][m_][k], Av[j][m_]) for m_ in tree], _d2(ci(L), n + j, k)) for L in tree}
            for i in tree:
                put(2 * n3 + i * n * n + j * n + k, em.lincomb([(-1.0, Mi[i][L], inner[L]) for L in tree]))

    em.comment("---- RNEA(q, qd, qdd) with first and second order tangents, in seed-pair blocks")
    rnea_second_order_blocks(cx, qd, qdd, sink)
    em.comment("---- d2qdd/dtau dq, structural zeros")
    for j in range(n):
        for k in range(n):
            inner_t = [em.lincomb([(1.0, dM[L][m_][k], Mi[m_][j]) for m_ in range(n)]) for L in range(n)]
            for i in range(n):
                _out(em, 3 * n3 + i * n * n + j * n + k, em.lincomb([(-1.0, Mi[i][L], inner_t[L]) for L in range(n)]))
    for idx in range(3 * n3):
        if idx not in done:
            _out(em, idx, ZERO)
    return Body("fdsva_so_inner", "", cx, "second-order forward dynamics tensors (4 n^3)")


# ================================================================= EE pose
def hom_chain(cx: Ctx, order: int, before_joint=None, after_joint=None) -> Dict[int, List[List[V]]]:
    """Tc[j] = rows 0..2 of the homogeneous transform base <- frame of joint j: the product of the joint transforms
    Xhom(q) down the chain, following the parent topology (reference algorithms/_eepose_gradient_hessian.py:76-140
    chains them leaf to root; :1226-1298 X_single_thread accumulates them root to leaf for serial chains).  With
    order 1 / 2 the entries carry their gradient / Hessian with respect to q (seed k = joint k)."""
    from .spatial import hom_tables
    em, m, n = cx.em, cx.m, cx.n
    save = em.ad_order
    em.ad_order = order
    H = hom_tables(m.robot)
    Tc: Dict[int, List[List[V]]] = {}
    for j in range(n):
        H0, Hs, Hc, Hq = H[j]
        sn, cs, qv = cx.trig[j]

        def ent(r, c):
            e = []
            if H0[r, c] != 0.0:
                e.append((float(H0[r, c]), None))
            if Hs[r, c] != 0.0:
                e.append((float(Hs[r, c]), sn))
            if Hc[r, c] != 0.0:
                e.append((float(Hc[r, c]), cs))
            if Hq[r, c] != 0.0:
                e.append((float(Hq[r, c]), qv))
            return e
        p = m.parent[j]
        if before_joint is not None:
            before_joint(j)
        T = [[None] * 4 for _ in range(3)]
        for r in range(3):
            for c in range(4):
                if p < 0:
                    T[r][c] = em.lincomb([(coef, f) if f is not None else (coef, ONE) for coef, f in ent(r, c)])
                else:
                    terms = []
                    for k in range(3):
                        for coef, f in ent(k, c):
                            terms.append((coef, f, Tc[p][r][k]))
                    T[r][c] = em.lincomb(terms, Tc[p][r][3] if c == 3 else None)
        Tc[j] = T
        if after_joint is not None:
            local = [[em.lincomb([(coef, f) if f is not None else (coef, ONE) for coef, f in ent(r, c)])
                      for c in range(4)] for r in range(3)]
            after_joint(j, local, T)
    em.ad_order = save
    return Tc


def fixed_target_transform(cx: Ctx, Tc, fixed_name: str, order: int):
    """Transform base <- frame of the fixed joint `fixed_name` (a fixed kinematic target, reference
    algorithms/_eepose_gradient_hessian.py:31-38, :89-93): chain of its parent body times its constant transform."""
    em, m = cx.em, cx.m
    fj = m.robot.get_fixed_joint_by_name(fixed_name)
    import numpy as np
    Xf = np.array(fj.Xhom_sympy().evalf(30), dtype=np.float64).reshape(4, 4)
    Xf[np.abs(Xf) < 1e-12] = 0.0
    p = fj.parent_joint_id
    save = em.ad_order
    em.ad_order = order
    T = [[None] * 4 for _ in range(3)]
    for r in range(3):
        for c in range(4):
            if p < 0:
                T[r][c] = const(Xf[r, c])
            else:
                terms = [(float(Xf[k, c]), Tc[p][r][k]) for k in range(3) if Xf[k, c] != 0.0]
                T[r][c] = em.lincomb(terms, Tc[p][r][3] if c == 3 else None)
    em.ad_order = save
    return T


def pose_from_transform(cx: Ctx, T, order: int) -> List[V]:
    """[x y z roll pitch yaw]: roll = atan2(R21, R22), pitch = -atan2(R20, sqrt(R21^2 + R22^2)), yaw = atan2(R10, R00)
    (reference algorithms/_eepose_gradient_hessian.py:150-157)."""
    em = cx.em
    save = em.ad_order
    em.ad_order = order
    roll = em.atan2(T[2][1], T[2][2])
    sq = em.sqrt(em.lincomb([(1.0, T[2][1], T[2][1]), (1.0, T[2][2], T[2][2])]))
    pitch = em.neg(em.atan2(T[2][0], sq))
    yaw = em.atan2(T[1][0], T[0][0])
    em.ad_order = save
    return [T[0][3], T[1][3], T[2][3], roll, pitch, yaw]


def ee_poses(cx: Ctx, order: int, fixed_target: Optional[str] = None):
    """Pose of every leaf link frame in the base frame (reference algorithms/_eepose_gradient_hessian.py:31-160), or of
    ONE fixed kinematic target when `fixed_target` names a fixed joint (:34-37).  With order 1 / 2 the pose entries
    carry their gradient / Hessian with respect to q; the reference derives those by hand from dXhom / d2Xhom
    (:511-538, :985-1058), here they are the emit-time tangents of the same chain product."""
    Tc = hom_chain(cx, order)
    if fixed_target:
        return [pose_from_transform(cx, fixed_target_transform(cx, Tc, fixed_target, order), order)]
    return [pose_from_transform(cx, Tc[e_], order) for e_ in cx.m.robot.get_leaf_nodes()]


def body_end_effector_pose(model: Model, fixed_target: Optional[str] = None) -> Body:
    """out[6*ee + o], o in x y z roll pitch yaw (reference :142-159)."""
    cx = _second_order_ctx(model)
    for e_, pose in enumerate(ee_poses(cx, 0, fixed_target)):
        for o, v in enumerate(pose):
            _out(cx.em, 6 * e_ + o, v)
    return Body("end_effector_pose_inner", "", cx, "end-effector poses")


def body_end_effector_pose_gradient(model: Model, fixed_target: Optional[str] = None) -> Body:
    """out[(ee*n + j)*6 + o] = d pose_o / d q_j (reference s_deePos[6*deeInd + outputInd], :517-535)."""
    cx = _second_order_ctx(model)
    n = model.n
    for e_, pose in enumerate(ee_poses(cx, 1, fixed_target)):
        for j in range(n):
            for o, v in enumerate(pose):
                _out(cx.em, (e_ * n + j) * 6 + o, _d1(v, j))
    return Body("end_effector_pose_gradient_inner", "", cx, "gradient of the end-effector poses")


def body_end_effector_pose_gradient_hessian(model: Model, fixed_target: Optional[str] = None) -> Body:
    """Hessian[(ee*6 + o)*n*n + i*n + j] = d2 pose_o / dq_i dq_j (reference s_d2eePos[outputInd*n*n + i*n + j],
    :1024, :1056).  The gradient half of the reference kernel's two outputs (:1108) comes from the gradient kernel:
    the host tier / C ABI launch both.  An end effector depends only on the joints of its own chain, so for big
    branched robots most of the 6 EE n^2 entries are structural zeros (Atlas: 27 000 outputs, 5 % non-zero): those
    are written by one zero-fill loop instead of one store statement each."""
    cx = _second_order_ctx(model)
    em, n = cx.em, model.n
    poses = ee_poses(cx, 2, fixed_target)
    entries = []
    for e_, pose in enumerate(poses):
        for o, v in enumerate(pose):
            for i in range(n):
                for j in range(n):
                    entries.append(((e_ * 6 + o) * n * n + i * n + j, _d2(v, i, j)))
    zeros = sum(1 for _, v in entries if v.is_zero())
    if len(entries) > 2048 and 2 * zeros > len(entries):
        _acquire_out(em)
        em.line(f"for (int z_ = 0; z_ < {len(entries)}; ++z_) o[z_*OS] = T(0);")
        entries = [(idx, v) for idx, v in entries if not v.is_zero()]
    for idx, v in entries:
        _out(em, idx, v)
    return Body("end_effector_pose_gradient_hessian_inner", "", cx, "Hessian of the end-effector poses")


def code_X_single_thread(model: Model) -> str:
    """Statements of X_single_thread(s_jointXforms, s_XmatsHom, s_q, tid) (reference signature,
    algorithms/_eepose_gradient_hessian.py:1226-1298): fills the local homogeneous transforms Xhom_j(q_j) and the
    cumulative transforms base <- joint j for every joint id <= tid, 16 values each, column-major."""
    cx = Ctx(model)
    em = cx.em
    em.regs_only, em.fences, em.no_sync = True, False, True
    em._out_acquired = True              # a plain device function: no tile staging buffer to acquire
    cx.prologue(q_ptr="s_q")

    def before(j):
        em.line(f"if (tid < {j}) return;")

    def after(j, local, T):
        for c in range(4):
            for r in range(3):
                _out(em, 16 * j + r + 4 * c, local[r][c], ptr="s_XmatsHom")
                _out(em, 16 * j + r + 4 * c, T[r][c], ptr="s_jointXforms")
            _out(em, 16 * j + 3 + 4 * c, ONE if c == 3 else ZERO, ptr="s_XmatsHom")
            _out(em, 16 * j + 3 + 4 * c, ONE if c == 3 else ZERO, ptr="s_jointXforms")

    hom_chain(cx, 0, before_joint=before, after_joint=after)
    return em.code()


def body_joint_transforms(model: Model) -> Body:
    """Cumulative homogeneous transforms of every joint frame: out[16*j + r + 4*c] = (base <- joint j)[r][c],
    column-major 4 x 4 blocks (what the reference's X_single_thread / X_warp leave in s_jointXforms,
    algorithms/_eepose_gradient_hessian.py:1226-1398; those assume a serial chain, here the parent topology is
    followed)."""
    cx = _second_order_ctx(model)
    em = cx.em
    Tc = hom_chain(cx, 0)
    for j in range(model.n):
        for c in range(4):
            for r in range(3):
                _out(em, 16 * j + r + 4 * c, Tc[j][r][c])
            _out(em, 16 * j + 3 + 4 * c, ONE if c == 3 else ZERO)
    return Body("joint_transforms_inner", "", cx, "cumulative joint transforms (4 x 4, column-major, per joint)")


# ============================================================ split (warp-specialised) bodies
class SplitBody:
    """One algorithm emitted as  phase A (first-order passes; one function per PART = group of base-rooted trees of the
    robot, the parts run concurrently on different warps and publish the per-knot state in shared memory)
    +  one function per derivative column (phase B: the columns of a tile run concurrently on the warps of the CTA).

    The per-knot state (v, Iv, X a_parent, f, L, 1/D, sin/cos, qd: 217 values for iiwa14) is what the reference
    keeps in shared memory between its __syncthreads phases (algorithms/_inverse_dynamics_gradient.py:27-140); here
    it is written once by phase A as S[slot*32 + lane] and read by the column functions."""

    def __init__(self, name: str, cx: Ctx, doc: str, parts, columns):
        self.name, self.doc = name, doc
        self.parts = parts                               # [(code, n_ops)] of phase A
        self.code = "".join(c for c, _ in parts)
        self.nslots = cx.em.nslots
        self.phase_a_ops = max(o for _, o in parts)      # the parts run concurrently
        self.n_ops = sum(o for _, o in parts) + sum(e.n_ops for _, e in columns)
        self.columns = [(c, e.code(), e.n_ops) for c, e in columns]


def tree_parts(model: Model, max_parts: int = 4):
    """Joint sets that phase A can process independently: the base-rooted trees of a fixed-base robot (HyQ: its four
    legs; Atlas: torso + arms, left leg, right leg) do not interact in RNEA, CRBA or the factorisation.  Trees are
    grouped into at most max_parts parts of similar size; a single tree gives [None] (= all joints)."""
    if os.environ.get("GRID_SPLIT_TREE_PARTS", "1") == "0":
        return [None]
    roots = [j for j in range(model.n) if model.parent[j] < 0]
    if len(roots) < 2:
        return [None]
    trees = sorted((sorted(model.subtree[r]) for r in roots), key=len, reverse=True)
    parts = [[] for _ in range(min(max_parts, len(trees)))]
    for t in trees:
        min(parts, key=len).extend(t)
    return [frozenset(p) for p in parts if p]


def _publish_state(cx: Ctx, qd: Sequence[V], extra=(), joints=None):
    """End of (a part of) phase A: make every value the column functions read a shared state slot.  Small robots run
    phase A on SSA values (registers), so the values are stored now; big robots already keep them in slots."""
    em, m, n = cx.em, cx.m, cx.n
    em.comment("---- publish the per-knot state for the column functions")
    act = [j for j in range(n) if joints is None or j in joints]
    keys = []
    for j in act:
        for nm in (("v", "Iv", "Xa", "f") if getattr(cx, "has_Iv", True) else ("v", "Xa", "f")):
            keys += [(nm, j, k) for k in range(6)]
    keys += list(extra)
    pending = []
    if em.regs_only:
        for key in keys:
            if key in em.const_slots or key not in em.reg_state:       # already a state slot (hybrid state)
                continue
            pending.append((key, em.reg_state[key]))
        em.regs_only = False
    for key, v in pending:
        em.store(key, v)
    for j in act:
        sn, cs, qv = cx.trig[j]
        if sn is not None:
            em.store(("sn", j, 0), V(s=sn.s))
            em.store(("cs", j, 0), V(s=cs.s))
        else:
            em.store(("qv", j, 0), V(s=qv.s))
        em.store(("qd", j, 0), V(s=qd[j].s))


def _column_ctx(cx: Ctx, c: int) -> Ctx:
    """Context of the function of derivative column c: reloads sin/cos and qd of the joints the column touches."""
    m = cx.m
    emc = cx.em.fork()
    cc = Ctx.__new__(Ctx)
    cc.m, cc.em, cc.n, cc.gravity = m, emc, cx.n, cx.gravity
    cc.has_Iv = getattr(cx, "has_Iv", True)
    touched = sorted(set(m.subtree[c]) | set(m.ancestors[c]))
    cc.X = [None] * m.n
    cc.trig = [None] * m.n
    cc.q = [None] * m.n
    qd: List[Optional[V]] = [None] * m.n
    for j in touched:
        jd = m.joints[j]
        if jd.revolute:
            sn = emc.named(f"sn{j}", emc.load(("sn", j, 0)).s)
            cs = emc.named(f"cs{j}", emc.load(("cs", j, 0)).s)
            tr = Trig(sn, cs, None)
        else:
            tr = Trig(None, None, emc.named(f"q_{j}", emc.load(("qv", j, 0)).s))
        cc.X[j] = Xform(jd, tr)
        if j in m.subtree[c]:
            qd[j] = emc.named(f"qd_{j}", emc.load(("qd", j, 0)).s)
    cc.qd = qd
    return cc


def split_enabled(n: int) -> bool:
    """Split bodies are emitted for grad ID / grad FD / the FD linearisation of every robot unless GRID_SPLIT=0."""
    return os.environ.get("GRID_SPLIT", "1") != "0"


def split_phase_a_regs_policy(n: int) -> bool:
    """Phase A on SSA values (registers) for small robots, on state slots otherwise."""
    mode = os.environ.get("GRID_SPLIT_PHASEA_REGS", "auto")
    if mode in ("0", "1"):
        return mode == "1"
    return regs_only_policy(n)


def split_publish_Iv_policy(n: int) -> bool:
    """Publish I v per joint (6 n state slots) or let the column functions recompute it from v (~5 % more operations).
    Big robots (n > 16) recompute: Atlas' published state shrinks from 931 to 756 slots (97 KB per tile), so a split
    CTA holds TWO tiles at once (header.split_ntile, grid_rt::tile_loop_split_multi) and every fetched column body is
    executed twice.  (The other use of the smaller state, two resident CTAs per SM at 128 registers, was measured and
    rejected: profiles/r02d_atlas_ab.log, the 128-register column functions spill.)  GRID_SPLIT_PUBLISH_IV=0/1 forces."""
    mode = os.environ.get("GRID_SPLIT_PUBLISH_IV", "auto")
    if mode in ("0", "1"):
        return mode == "1"
    return n <= 16


def _split_columns(cx: Ctx, sink_factory):
    cols = []
    for c in range(cx.n):
        cc = _column_ctx(cx, c)
        rnea_grad_columns(cc, cc.qd, sink_factory(cc), cols=[c])
        cols.append((c, cc.em))
    return cols


def _begin_part(cx: Ctx):
    """Start the function of the next phase-A part on the shared emitter: fresh text, NO reuse of state slots freed
    by the previous part (the parts run concurrently), SSA mode as configured."""
    em = cx.em
    em.lines = []
    em.free = []
    em._out_acquired = False
    em.regs_only = split_phase_a_regs_policy(cx.n)
    return em.n_ops


def _tree_of(model: Model, j: int):
    """Joints of the base-rooted tree that contains joint j."""
    r = min(model.ancestors[j]) if model.ancestors[j] else j
    return frozenset(model.subtree[r])


def split_inverse_dynamics_gradient(model: Model, use_qdd: bool) -> SplitBody:
    cx = Ctx(model)
    em, n = cx.em, model.n
    em.no_sync = True
    cx.has_Iv = split_publish_Iv_policy(n)
    parts = []
    for js in tree_parts(model):
        ops0 = _begin_part(cx)
        cx.prologue(joints=js)
        qd = cx.read_vec("qd", "qd_", js)
        qdd = cx.read_vec("qdd", "qdd_", js) if use_qdd else None
        rnea_forward(cx, qd, qdd, keep=True, joints=js, keep_Iv=cx.has_Iv)
        rnea_backward(cx, qd, keep_f=True, joints=js)
        _publish_state(cx, qd, joints=js)
        parts.append((em.code(), em.n_ops - ops0))

    def sink_factory(cc):
        def sink(c, dc):
            for r in range(n):
                x, y = _halves(cc.em, dc[r])
                _out(cc.em, c * n + r, x)
                _out(cc.em, (n + c) * n + r, y)
        return sink

    cols = _split_columns(cx, sink_factory)
    return SplitBody("inverse_dynamics_gradient", cx, "[dc/dq | dc/dqd], column-major n x 2n", parts, cols)


def split_forward_dynamics_gradient(model: Model, linearization: bool = False) -> SplitBody:
    """grad FD (optionally with the Minv and qdd side outputs) in split form: phase A = RNEA(q, qd, 0), CRBA,
    L^T D L, qdd, RNEA(q, qd, qdd), one part per group of base-rooted trees; phase B = one function per column c
    computing dc/dq_c, dc/dqd_c and the two solves -M^-1 dc against the published factors."""
    cx = Ctx(model)
    em, n = cx.em, model.n
    em.no_sync = True
    cx.has_Iv = split_publish_Iv_policy(n)
    parts = []
    L_all: Dict[tuple, V] = {}
    for js in tree_parts(model):
        ops0 = _begin_part(cx)
        act = list(range(n)) if js is None else sorted(js)
        cx.prologue(joints=js)
        qd = cx.read_vec("qd", "qd_", js)
        u = cx.read_vec("u", "u_", js)
        em.comment("---- phase A, pass 1: bias forces c = RNEA(q, qd, 0)")
        rnea_forward(cx, qd, None, keep=False, joints=js)
        c = rnea_backward(cx, qd, keep_f=False, joints=js)
        em.comment("---- M = CRBA(q), M = L^T D L")
        Mv = crba_matrix(cx, joints=js)
        L, Dinv = ltdl_factor(cx, Mv, joints=js)
        for k_, v_ in L.items():
            em.store(("L",) + k_, v_)
        for i_ in act:
            em.store(("Dinv2", i_, 0), Dinv[i_])
        L_all.update(L)
        rhs = [em.sub(u[j], c[j]) if j in act else ZERO for j in range(n)]
        qdd = ltdl_solve(cx, L, Dinv, rhs, joints=frozenset(act))
        if linearization:
            em.comment("---- side outputs: Minv (dense symmetric, column-major) and qdd")
            cols_ = {k: ltdl_solve(cx, L, Dinv, [ONE if r == k else ZERO for r in range(n)], joints=frozenset(act))
                     for k in act}
            for k in act:
                for r in range(n):
                    # one value per symmetric pair, so that the stored matrix is exactly symmetric; entries between
                    # different trees are structural zeros
                    val = (cols_[k][r] if r <= k else cols_[r][k]) if r in cols_ else ZERO
                    _out(em, (2 * n + k) * n + r, val)
            for j in act:
                _out(em, 3 * n * n + j, qdd[j])
        em.comment("---- pass 2: RNEA(q, qd, qdd) keeping v, Iv, X a_parent, f")
        rnea_forward(cx, qd, qdd, keep=True, joints=js, keep_Iv=cx.has_Iv)
        rnea_backward(cx, qd, keep_f=True, joints=js)
        lkeys = [("L",) + k_ for k_ in L] + [("Dinv2", i_, 0) for i_ in act]
        _publish_state(cx, qd, extra=lkeys, joints=js)
        parts.append((em.code(), em.n_ops - ops0))

    def sink_factory(cc):
        def sink(c_, dc):
            e = cc.em
            tree = _tree_of(model, c_)
            Ll = {k_: e.load(("L",) + k_) for k_ in L_all if k_[0] in tree}
            Dl = [e.load(("Dinv2", i_, 0)) if i_ in tree else ZERO for i_ in range(n)]
            if all(not isinstance(d_, tuple) for d_ in dc):
                # packed columns (GRID_EMIT_PAIRS): the dq and dqd halves go through the solve as one pair
                x = ltdl_solve(cc, Ll, Dl, dc, sign=-1.0, joints=tree)
                for r in range(n):
                    a_, b_ = _halves(e, x[r])
                    _out(e, c_ * n + r, a_)
                    _out(e, (n + c_) * n + r, b_)
                return
            for h in (0, 1):
                x = ltdl_solve(cc, Ll, Dl, [_halves(e, d_)[h] for d_ in dc], sign=-1.0, joints=tree)
                for r in range(n):
                    _out(e, (h * n + c_) * n + r, x[r])
        return sink

    cols = _split_columns(cx, sink_factory)
    if linearization:
        return SplitBody("forward_dynamics_linearization", cx,
                         "[dqdd/dq | dqdd/dqd | Minv | qdd]: grad FD with its Minv and qdd by-products", parts, cols)
    return SplitBody("forward_dynamics_gradient", cx, "[dqdd/dq | dqdd/dqd] = -Minv [dc/dq | dc/dqd]", parts, cols)
