"""Expression layer of the kernel emitter.

The emitted kernels are straight-line, robot-specialised code: every quantity
is either a compile-time constant (folded here, in Python) or a named scalar in
the generated CUDA source.  Structural zeros / +-1 of the tree transforms and
inertias therefore never reach the GPU, and the joint loop is fully unrolled
with compile-time topology (the reference gets the same effect by printing
robot-specific code: helpers/_topology_helpers.py:155-331, but keeps dense 6x6
dot products at run time, helpers/_spatial_algebra_helpers.py:124-143).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple, Union

Number = Union[int, float]
# a lockstep barrier (GRID_SYNC) at every SYNC_EVERY-th section marker of the emitted bodies (0 = none)
SYNC_EVERY = int(os.environ.get("GRID_SYNC_EVERY", "8"))
# Hybrid state of the register-resident (small-robot) bodies: state keys whose first element is listed here live in
# per-thread shared-memory slots instead of registers (A/B knob, e.g. "f,L,Dinv2"): fewer registers per thread buys
# more resident warps per SM, and a pass over the body costs the same however many warps share it
HYBRID_KEYS = frozenset(k for k in os.environ.get("GRID_HYBRID_KEYS", "").split(",") if k)


class V:
    """A scalar: compile-time constant (`c`) or a C identifier/expression (`s`).
    With `p` the identifier names a PAIR of scalars (type P = grid_rt::pair_t<T>: float2 for T = float, so that
    linear combinations of pairs with scalar coefficients become packed FFMA2 / FMUL2 / FADD2 on sm_100a).
    A scalar used where a pair is expected means the broadcast (s, s)."""

    __slots__ = ("c", "s", "p", "d")

    def __init__(self, c: Optional[float] = None, s: Optional[str] = None, p: bool = False, d=None):
        self.c = None if c is None else float(c)
        self.s = s
        self.p = p
        # emit-time forward-mode AD (Em.ad_order > 0): d[seed] is the V holding d(self)/d(seed); those V's carry
        # their own d (second order) for seeds <= seed.  None = no dependence on any seed.
        self.d = d

    @property
    def is_const(self):
        return self.c is not None

    def is_zero(self):
        return self.c is not None and self.c == 0.0

    def __repr__(self):
        return f"V({self.c if self.is_const else self.s})"


ZERO = V(0.0)
ONE = V(1.0)


def const(x: Number) -> V:
    return V(c=float(x))


def lit(x: float) -> str:
    """C literal usable for T = float or double."""
    if x == int(x) and abs(x) < 1e6:
        return f"T({int(x)})"
    return f"T({x!r})"


class Em:
    """Code buffer + SSA temp allocator + per-thread state-slot allocator."""

    def __init__(self, state_name: str = "S", stride_name: str = "NT"):
        self.lines: List[str] = []
        self.ctr = 0
        self.state_name = state_name
        self.stride_name = stride_name
        self.slots: Dict[Tuple, int] = {}
        self.free: List[int] = []
        self.nslots = 0
        self.const_slots: Dict[Tuple, float] = {}  # values known at emit time are never stored
        self.n_ops = 0  # multiply/add operations emitted (for the executed-FLOP audit)
        self.fences = False
        # small robots: "state" is plain SSA values (registers; ptxas spills the coldest to local memory = L1)
        self.regs_only = False
        self.reg_state: Dict[Tuple, V] = {}
        # emit-time forward-mode AD: 0 = off, 1 = first-order tangents, 2 = also second order
        self._nsync = 0
        self.no_sync = False        # split (warp-specialised) bodies: warps run different code, no lockstep barriers
        self.private_regs = False   # column functions of the split mode: column-private values (tuple-named keys) are SSA
        self.frozen = False         # ... and the shared state is read-only
        self.ad_order = 0
        self._ad_depth = 0
        self._ad_cap = None
        # seed-pair blocks (second order in passes): when set, only the second-order tangents (k, l), l <= k, listed
        # here are formed; the first-order seeds are whatever the inputs carry
        self.ad_pairs = None

    # ------------------------------------------------------------------ text
    def line(self, txt: str):
        self.lines.append(txt)

    def comment(self, txt: str):
        """Section marker.  With `fences` on, also a compiler-level memory fence: values loaded from the
        state before the marker are not kept live in registers across it (they are re-read from shared
        memory when needed), which bounds register pressure for big robots (Atlas: 38 KB of spills without)."""
        self.lines.append("// " + txt)
        if self.fences:
            self.lines.append("GRID_FENCE();")
        # lockstep point: the warps of a CTA re-converge here so that they stream the (straight-line, >100 KB)
        # code through the instruction caches together; a no-op for single-warp CTAs
        self._nsync += 1
        if SYNC_EVERY > 0 and self._nsync % SYNC_EVERY == 0 and not self.no_sync:
            self.lines.append("GRID_SYNC();")

    def tmp(self, expr: str) -> V:
        name = f"t{self.ctr}"
        self.ctr += 1
        self.lines.append(f"const T {name} = {expr};")
        return V(s=name)

    def tmp2(self, expr: str) -> V:
        name = f"t{self.ctr}"
        self.ctr += 1
        self.lines.append(f"const P {name} = {expr};")
        return V(s=name, p=True)

    def mkpair(self, a: V, b: V) -> V:
        """The pair (a, b) of two scalars (constants allowed)."""
        if a.p or b.p:
            raise ValueError("mkpair takes scalars")
        if a.is_const and b.is_const and a.c == b.c:
            return a                      # broadcast constant (covers the structural zero)
        ea = lit(a.c) if a.is_const else a.s
        eb = lit(b.c) if b.is_const else b.s
        return self.tmp2(f"grid_rt::mkp<T>({ea}, {eb})")

    def half(self, v: V, which: int) -> V:
        """Component 0 / 1 of a pair (a scalar stands for (s, s))."""
        if not v.p:
            return v
        return V(s=f"{v.s}.{'xy'[which]}")

    def named(self, name: str, expr: str) -> V:
        self.lines.append(f"const T {name} = {expr};")
        return V(s=name)

    def fork(self) -> "Em":
        """Emitter for a function that runs AFTER this one on the same per-knot state (split mode: the derivative
        columns, each on its own warp): same slot map and emit-time constants, read-only; values stored under
        column-private keys (tuple-valued names such as ("df", c)) stay SSA values of the new function."""
        e = Em(self.state_name, self.stride_name)
        e.slots, e.const_slots, e.nslots = self.slots, self.const_slots, self.nslots
        e.fences, e.no_sync = self.fences, True
        e.regs_only, e.private_regs, e.frozen = False, True, True
        return e

    def code(self, indent: str = "    ") -> str:
        return "\n".join(indent + l for l in self.lines) + "\n"

    # ------------------------------------------------------------ arithmetic
    def lincomb(self, terms: Sequence[Tuple], acc: Optional[V] = None) -> V:
        """sum_k coef_k * x_k [* y_k ...] (+ acc), with emit-time forward-mode differentiation: when ad_order > 0
        and some factor carries tangents (V.d), the result carries the tangents given by the product rule, computed
        by recursive calls (the tangent of a tangent gives the second order; only seed pairs l <= k are formed)."""
        r = self._lincomb(terms, acc)
        if self.ad_order <= self._ad_depth:
            return r
        seeds = set()
        for t in terms:
            for v in t[1:]:
                if v is not None and v.d:
                    seeds.update(v.d.keys())
        if acc is not None and acc.d:
            seeds.update(acc.d.keys())
        if self._ad_cap is not None:
            seeds = {k for k in seeds if k <= self._ad_cap}
            if self.ad_pairs is not None:
                seeds = {k for k in seeds if (self._ad_cap, k) in self.ad_pairs}
        if not seeds:
            return r
        r = V(c=r.c, s=r.s, p=r.p)          # fresh object: r may alias an input
        r.d = {}
        depth, cap = self._ad_depth, self._ad_cap
        for k in sorted(seeds):
            tt = []
            for t in terms:
                fs = [f for f in t[1:] if f is not None]
                for i, f in enumerate(fs):
                    fk = f.d.get(k) if f.d else None
                    if fk is not None:
                        tt.append((t[0], fk) + tuple(g for m_, g in enumerate(fs) if m_ != i))
            ak = acc.d.get(k) if (acc is not None and acc.d) else None
            self._ad_depth, self._ad_cap = depth + 1, k
            rk = self.lincomb(tt, ak)
            self._ad_depth, self._ad_cap = depth, cap
            if not rk.is_zero():
                r.d[k] = rk
        if not r.d:
            r.d = None
        return r

    def elem(self, expr: str, inputs: Sequence[V], partials: Sequence[V], second=None) -> V:
        """A differentiable elementary function g(u_1..u_m) emitted as `expr`.  partials[a] = dg/du_a and
        second[a][b] = d2g/du_a du_b are V's WITHOUT tangents (plain expressions of the inputs); the tangents of
        the result follow from the chain rule:  r_k = sum_a g_a u_a,k ;
        r_kl = sum_a g_a u_a,kl + sum_ab g_ab u_a,k u_b,l."""
        r = self.tmp(expr)
        if self.ad_order <= self._ad_depth:
            return r
        seeds = set()
        for u in inputs:
            if u.d:
                seeds.update(u.d.keys())
        if self._ad_cap is not None:
            seeds = {k for k in seeds if k <= self._ad_cap}
        if not seeds:
            return r
        save = (self.ad_order, self._ad_depth, self._ad_cap)
        order = self.ad_order - self._ad_depth
        self.ad_order = 0                     # plain arithmetic below: tangents are assembled explicitly
        r.d = {}
        m = len(inputs)
        for k in sorted(seeds):
            uk = [u.d.get(k) if u.d else None for u in inputs]
            rk = self.lincomb([(1.0, partials[a], uk[a]) for a in range(m) if uk[a] is not None])
            rk = V(c=rk.c, s=rk.s)
            if order >= 2 and second is not None:
                rk.d = {}
                for l in sorted(x for x in seeds if x <= k and (self.ad_pairs is None or (k, x) in self.ad_pairs)):
                    ul = [u.d.get(l) if u.d else None for u in inputs]
                    tt = []
                    for a in range(m):
                        ukl = uk[a].d.get(l) if (uk[a] is not None and uk[a].d) else None
                        if ukl is not None:
                            tt.append((1.0, partials[a], ukl))
                        for b in range(m):
                            if uk[a] is not None and ul[b] is not None:
                                tt.append((1.0, second[a][b], uk[a], ul[b]))
                    rkl = self.lincomb(tt)
                    if not rkl.is_zero():
                        rk.d[l] = rkl
                if not rk.d:
                    rk.d = None
            if not rk.is_zero():
                r.d[k] = rk
        self.ad_order, self._ad_depth, self._ad_cap = save
        if not r.d:
            r.d = None
        return r

    def _lincomb(self, terms: Sequence[Tuple], acc: Optional[V] = None) -> V:
        """sum_k coef_k * x_k [* y_k ...] (+ acc).  terms: (coef, x, ...) with coef a Python float and the rest
        V's.  Constants are folded; emits at most one statement."""
        if (acc is not None and acc.p) or any(v is not None and v.p for t in terms for v in t[1:]):
            return self._lincomb2(terms, acc)
        csum = 0.0
        parts: List[Tuple[float, str]] = []
        if acc is not None:
            if acc.is_const:
                csum += acc.c
            else:
                parts.append((1.0, acc.s))
        for t in terms:
            coef = float(t[0])
            vs = t[1:]
            names = []
            for v in vs:
                if v is None:
                    continue
                if v.is_const:
                    coef *= v.c
                else:
                    names.append(v.s)
            if coef == 0.0:
                continue
            if not names:
                csum += coef
            else:
                parts.append((coef, "*".join(names)))
        if self.ad_order and len(parts) > 1:      # product-rule expansions repeat terms: merge them
            merged: Dict[str, float] = {}
            for coef, name in parts:
                key = "*".join(sorted(name.split("*")))
                merged[key] = merged.get(key, 0.0) + coef
            parts = [(c_, k_) for k_, c_ in merged.items() if c_ != 0.0]
        if not parts:
            return const(csum)
        if len(parts) == 1 and csum == 0.0 and parts[0][0] == 1.0 and "*" not in parts[0][1]:
            return V(s=parts[0][1])  # pure alias, no code
        out = []
        # put a positive term first so the expression does not start with a negation when avoidable
        parts.sort(key=lambda p: p[0] < 0)
        for i, (coef, name) in enumerate(parts):
            mag = abs(coef)
            body = name if mag == 1.0 else f"{lit(mag)}*{name}"
            self.n_ops += name.count("*") + (0 if mag == 1.0 else 1) + (1 if i else 0)
            if i == 0:
                out.append(body if coef > 0 else f"-{body}")
            else:
                out.append((" + " if coef > 0 else " - ") + body)
        if csum != 0.0:
            out.append((" + " if csum > 0 else " - ") + lit(abs(csum)))
            self.n_ops += 1
        return self.tmp("".join(out))

    def _lincomb2(self, terms: Sequence[Tuple], acc: Optional[V]) -> V:
        """Pair-valued lincomb: every term has at most one pair factor, the rest are scalar coefficients."""
        csum = 0.0
        expr: Optional[str] = None
        bparts: List[Tuple[float, str]] = []           # scalar (broadcast) terms
        pparts: List[Tuple[float, List[str], str]] = []
        if acc is not None:
            if acc.is_const:
                csum += acc.c
            elif acc.p:
                expr = acc.s
            else:
                bparts.append((1.0, acc.s))
        for t in terms:
            coef = float(t[0])
            names: List[str] = []
            pn: Optional[str] = None
            for v in t[1:]:
                if v is None:
                    continue
                if v.is_const:
                    coef *= v.c
                elif v.p:
                    if pn is not None:
                        raise ValueError("product of two pairs in a linear combination")
                    pn = v.s
                else:
                    names.append(v.s)
            if coef == 0.0:
                continue
            if pn is None:
                if not names:
                    csum += coef
                else:
                    bparts.append((coef, "*".join(names)))
            else:
                pparts.append((coef, names, pn))
        if expr is None and len(pparts) == 1 and not bparts and csum == 0.0 and pparts[0][0] == 1.0 and not pparts[0][1]:
            return V(s=pparts[0][2], p=True)
        pparts.sort(key=lambda q: q[0] < 0)
        for coef, names, pn in pparts:
            mag = abs(coef)
            unity = mag == 1.0 and not names
            sc = "*".join(([lit(mag)] if mag != 1.0 else []) + names)
            self.n_ops += max(len(names) - 1 + (mag != 1.0), 0)      # scalar multiplies forming the coefficient
            if expr is None:
                if unity:
                    expr = pn if coef > 0 else f"grid_rt::pneg({pn})"
                else:
                    expr = f"grid_rt::pmul({pn}, {sc})" if coef > 0 else f"grid_rt::pmul({pn}, -({sc}))"
                    self.n_ops += 2
            else:
                if unity:
                    expr = f"grid_rt::padd({expr}, {pn})" if coef > 0 else f"grid_rt::psub({expr}, {pn})"
                    self.n_ops += 2
                else:
                    expr = f"grid_rt::pfma({pn}, {sc}, {expr})" if coef > 0 else f"grid_rt::pfma({pn}, -({sc}), {expr})"
                    self.n_ops += 4
        if bparts or csum != 0.0:
            out = []
            bparts.sort(key=lambda q: q[0] < 0)
            for i, (coef, name) in enumerate(bparts):
                mag = abs(coef)
                body = name if mag == 1.0 else f"{lit(mag)}*{name}"
                if i == 0:
                    out.append(body if coef > 0 else f"-{body}")
                else:
                    out.append((" + " if coef > 0 else " - ") + body)
            if csum != 0.0:
                out.append(((" + " if csum > 0 else " - ") if out else ("" if csum > 0 else "-")) + lit(abs(csum)))
            b = "".join(out)
            expr = f"grid_rt::pbc<T>({b})" if expr is None else f"grid_rt::padds({expr}, {b})"
            self.n_ops += 2
        if expr is None:
            return const(0.0)
        return self.tmp2(expr)

    def add(self, a: V, b: V) -> V:
        return self.lincomb([(1.0, a), (1.0, b)])

    def sub(self, a: V, b: V) -> V:
        return self.lincomb([(1.0, a), (-1.0, b)])

    def mul(self, a: V, b: V) -> V:
        return self.lincomb([(1.0, a, b)])

    def neg(self, a: V) -> V:
        return self.lincomb([(-1.0, a)])

    def recip(self, a: V) -> V:
        if a.is_const:
            return const(1.0 / a.c)
        self.n_ops += 1
        if self.ad_order > self._ad_depth and a.d:
            r = self.tmp(f"T(1)/{a.s}")
            save = self.ad_order
            self.ad_order = 0
            g1 = self.lincomb([(-1.0, r, r)])
            g2 = self.lincomb([(-2.0, g1, r)])
            self.ad_order = save
            out = self.elem(r.s, [a], [g1], [[g2]])
            return out
        return self.tmp(f"T(1)/{a.s}")

    def _plain(self, fn):
        """Run fn() with differentiation switched off (for the explicit partials handed to elem)."""
        save = self.ad_order
        self.ad_order = 0
        try:
            return fn()
        finally:
            self.ad_order = save

    def sqrt(self, a: V) -> V:
        if a.is_const:
            return const(a.c ** 0.5)
        r = self.tmp(f"sqrt({a.s})")
        if not (self.ad_order > self._ad_depth and a.d):
            return r
        g1 = self._plain(lambda: self.lincomb([(0.5, self.recip(r))]))              # 1 / (2 sqrt a)
        g2 = self._plain(lambda: self.lincomb([(-0.5, g1, self.recip(a))]))         # -1 / (4 a sqrt a)
        return self.elem(r.s, [a], [g1], [[g2]])

    def atan2(self, y: V, x: V) -> V:
        if y.is_const and x.is_const:
            import math
            return const(math.atan2(y.c, x.c))
        from_v = lambda v: lit(v.c) if v.is_const else v.s
        r = self.tmp(f"atan2({from_v(y)}, {from_v(x)})")
        if not (self.ad_order > self._ad_depth and (y.d or x.d)):
            return r

        def partials():
            w = self.recip(self.lincomb([(1.0, x, x), (1.0, y, y)]))
            gy = self.lincomb([(1.0, x, w)])
            gx = self.lincomb([(-1.0, y, w)])
            gyy = self.lincomb([(-2.0, gy, y, w)])          # -2 x y w^2
            gxx = self.lincomb([(-2.0, gx, x, w)])          #  2 x y w^2
            gxy = self.lincomb([(1.0, y, y, w, w), (-1.0, x, x, w, w)])
            return gy, gx, gyy, gxx, gxy
        gy, gx, gyy, gxx, gxy = self._plain(partials)
        return self.elem(r.s, [y, x], [gy, gx], [[gyy, gxy], [gxy, gxx]])

    def vadd(self, a: Sequence[V], b: Sequence[V]) -> List[V]:
        return [self.add(x, y) for x, y in zip(a, b)]

    def dot(self, a: Sequence[V], b: Sequence[V], acc: Optional[V] = None, sign: float = 1.0) -> V:
        return self.lincomb([(sign, x, y) for x, y in zip(a, b)], acc)

    def matvec(self, M: Sequence[Sequence[V]], x: Sequence[V], acc: Optional[Sequence[V]] = None) -> List[V]:
        return [self.dot(M[r], x, None if acc is None else acc[r]) for r in range(len(M))]

    def matTvec(self, M: Sequence[Sequence[V]], x: Sequence[V], acc: Optional[Sequence[V]] = None) -> List[V]:
        rows, cols = len(M), len(M[0])
        return [self.dot([M[r][c] for r in range(rows)], x, None if acc is None else acc[c]) for c in range(cols)]

    # ------------------------------------------------------------ state slots
    def slot(self, *key) -> int:
        if key not in self.slots:
            if self.free:
                self.slots[key] = self.free.pop()
            else:
                self.slots[key] = self.nslots
                self.nslots += 1
        return self.slots[key]

    def release(self, *key):
        """Return a slot to the pool (its value is dead from here on)."""
        if self.private_regs and isinstance(key[0], tuple):
            self.reg_state.pop(key, None)
            return
        if self.frozen:
            return
        for k in (key, key + ("x",), key + ("y",)):
            idx = self.slots.pop(k, None)
            if idx is not None:
                self.free.append(idx)

    def store(self, key: Tuple, v: V):
        """Store v into the state slot `key`; constants are remembered at emit time, not stored."""
        if self.private_regs and isinstance(key[0], tuple):
            self.reg_state[key] = v
            return
        if v.is_const:
            if self.frozen:
                raise ValueError(f"store to shared state {key} from a read-only (column) emitter")
            self.const_slots[key] = v.c
            self.release(*key)
            return
        if self.frozen:
            raise ValueError(f"store to shared state {key} from a read-only (column) emitter")
        self.const_slots.pop(key, None)
        if self.regs_only and key[0] not in HYBRID_KEYS:
            self.reg_state[key] = v
            return
        if v.p:
            idx = self.slots.pop(key, None)
            if idx is not None:
                self.free.append(idx)
            for h in "xy":
                idx = self.slot(*(key + (h,)))
                self.lines.append(f"{self.state_name}[{idx}*{self.stride_name}] = {v.s}.{h};")
            return
        for h in "xy":
            idx = self.slots.pop(key + (h,), None)
            if idx is not None:
                self.free.append(idx)
        idx = self.slot(*key)
        self.lines.append(f"{self.state_name}[{idx}*{self.stride_name}] = {v.s};")

    def load(self, key: Tuple) -> V:
        if self.private_regs and key in self.reg_state:
            return self.reg_state[key]
        if key in self.const_slots:
            return const(self.const_slots[key])
        if self.regs_only and key in self.reg_state:
            return self.reg_state[key]
        if key not in self.slots and (key + ("x",)) in self.slots:
            ix, iy = self.slots[key + ("x",)], self.slots[key + ("y",)]
            return self.tmp2(f"grid_rt::mkp<T>({self.state_name}[{ix}*{self.stride_name}], "
                             f"{self.state_name}[{iy}*{self.stride_name}])")
        idx = self.slots[key]
        return self.tmp(f"{self.state_name}[{idx}*{self.stride_name}]")

    def store_vec(self, name, j, vec: Sequence[V]):
        for k, v in enumerate(vec):
            self.store((name, j, k), v)

    def load_vec(self, name, j, m: int = 6) -> List[V]:
        return [self.load((name, j, k)) for k in range(m)]

    def release_vec(self, name, j, m: int = 6):
        for k in range(m):
            self.release(name, j, k)


def cmat(A) -> List[List[V]]:
    """numpy matrix -> matrix of constant V's."""
    return [[const(x) for x in row] for row in A]
