"""Text emitters for the four tiers of every algorithm in grid.cuh.

Tier names are the reference's (README.md:25-30; banner GRiDCodeGenerator.py:312-380):
  <algo>_inner   per-thread straight-line body (one lane = one knot point)
  <algo>_device  per-warp tile loop (TMA staging + inner) callable from a user kernel
  <algo>_kernel  __global__ with the reference's parameter list
  <algo>         __host__ wrapper on gridData<T> (+ _single_timing, _compute_only)
plus  <algo>_launch  the stream-asynchronous launcher the C ABI binds.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Tuple

from . import algos
from .spatial import Model

SMEM_LIMIT = 232448 - 1024   # 227 KB opt-in maximum per CTA minus the per-CTA reservation
MAX_WARPS = int(os.environ.get("GRID_MAX_WARPS", "8"))           # 65536 registers / (255 x 32)
MAX_WARPS_SCRATCH = int(os.environ.get("GRID_MAX_WARPS_SCRATCH", "4"))   # warps per CTA when the state lives in the global scratch


@dataclass
class AlgoSpec:
    key: str                 # python-side generator name, e.g. "forward_dynamics_gradient"
    fn: str                  # C++ function stem
    code: str                # shared-memory constant prefix (ID, MINV, FD, ID_DU, FD_DU, ...)
    tag: str                 # unique suffix distinguishing overloads ("", "_qdd", ...)
    body: Callable[[Model], algos.Body]
    inner_inputs: List[str]  # pointer names of the inner function, in order
    streams: List[Tuple[str, str]]  # (kernel parameter name, elements per knot as C expr of n)
    split: List[Tuple[int, int]]    # inner input k comes from (stream index, element offset)
    out_name: str
    out_n: Callable[[int], int]
    uses_gravity: bool = True
    extra_kernel_params: str = ""
    fixed_target: str = ""   # name of the fixed joint this (single-target EE) variant is about


def fixed_target_specs(n: int, name: str) -> List[AlgoSpec]:
    """EE pose / gradient / Hessian of ONE fixed kinematic target (reference gen_all_code(fixed_target_name=...),
    GRiDCodeGenerator.py:309-310, :417; function names carry the target's name as a suffix like the reference's,
    algorithms/_eepose_gradient_hessian.py:46, :176, :207, :264)."""
    out = [
        AlgoSpec("end_effector_pose", f"end_effector_pose_{name}", "EE_FIXED", "",
                 lambda m: algos.body_end_effector_pose(m, name),
                 ["q"], [("d_q", n)], [(0, 0)], "d_eePos", lambda n: 6, uses_gravity=False),
        AlgoSpec("end_effector_pose_gradient", f"end_effector_pose_gradient_{name}", "DEE_FIXED", "",
                 lambda m: algos.body_end_effector_pose_gradient(m, name),
                 ["q"], [("d_q", n)], [(0, 0)], "d_deePos", lambda n: 6 * n, uses_gravity=False),
        AlgoSpec("end_effector_pose_gradient_hessian", f"end_effector_pose_gradient_hessian_{name}", "D2EE_FIXED", "",
                 lambda m: algos.body_end_effector_pose_gradient_hessian(m, name),
                 ["q"], [("d_q", n)], [(0, 0)], "d_d2eePos", lambda n: 6 * n * n, uses_gravity=False),
    ]
    for sp in out:
        sp.fixed_target = name
    return out


def specs(n: int, num_ees: int = 1, max_tree: Optional[int] = None, fixed_target: Optional[str] = None) -> List[AlgoSpec]:
    """max_tree: joints of the largest base-rooted tree of the robot (default n: one tree).
    fixed_target: name of a fixed joint for which the additional single-target EE functions are generated."""
    ee = num_ees
    max_tree = n if max_tree is None else max_tree
    return (fixed_target_specs(n, fixed_target) if fixed_target else []) + [
        AlgoSpec("joint_transforms", "joint_transforms", "JX", "", algos.body_joint_transforms,
                 ["q"], [("d_q", n)], [(0, 0)], "d_jointXforms", lambda n: 16 * n, uses_gravity=False),
    ] + [
        AlgoSpec("end_effector_pose", "end_effector_pose", "EE", "", algos.body_end_effector_pose,
                 ["q"], [("d_q", n)], [(0, 0)], "d_eePos", lambda n: 6 * ee, uses_gravity=False),
        AlgoSpec("end_effector_pose_gradient", "end_effector_pose_gradient", "DEE", "",
                 algos.body_end_effector_pose_gradient,
                 ["q"], [("d_q", n)], [(0, 0)], "d_deePos", lambda n: 6 * ee * n, uses_gravity=False),
    ] + ([
        AlgoSpec("end_effector_pose_gradient_hessian", "end_effector_pose_gradient_hessian", "D2EE", "",
                 algos.body_end_effector_pose_gradient_hessian,
                 ["q"], [("d_q", n)], [(0, 0)], "d_d2eePos", lambda n: 6 * ee * n * n, uses_gravity=False),
    ] if ee_hessian_enabled(n) else []) + [
        AlgoSpec("inverse_dynamics", "inverse_dynamics", "ID", "", lambda m: algos.body_inverse_dynamics(m, False),
                 ["q", "qd"], [("d_q_qd", 2 * n)], [(0, 0), (0, n)], "d_c", lambda n: n),
        AlgoSpec("inverse_dynamics", "inverse_dynamics", "ID", "_qdd", lambda m: algos.body_inverse_dynamics(m, True),
                 ["q", "qd", "qdd"], [("d_q_qd", 2 * n), ("d_qdd", n)], [(0, 0), (0, n), (1, 0)], "d_c", lambda n: n),
        # RNEA's per-joint v, a, f (reference inverse_dynamics_inner_vaf / inverse_dynamics_vaf_device)
        AlgoSpec("inverse_dynamics_vaf", "inverse_dynamics_vaf", "ID_VAF", "", lambda m: algos.body_inverse_dynamics_vaf(m, False),
                 ["q", "qd"], [("d_q_qd", 2 * n)], [(0, 0), (0, n)], "d_vaf", lambda n: 18 * n),
        AlgoSpec("inverse_dynamics_vaf", "inverse_dynamics_vaf", "ID_VAF", "_qdd", lambda m: algos.body_inverse_dynamics_vaf(m, True),
                 ["q", "qd", "qdd"], [("d_q_qd", 2 * n), ("d_qdd", n)], [(0, 0), (0, n), (1, 0)], "d_vaf", lambda n: 18 * n),
        AlgoSpec("direct_minv", "direct_minv", "MINV", "", algos.body_direct_minv,
                 ["q"], [("d_q", n)], [(0, 0)], "d_Minv", lambda n: n * n, uses_gravity=False),
        AlgoSpec("forward_dynamics", "forward_dynamics", "FD", "", algos.body_forward_dynamics,
                 ["q", "qd", "u"], [("d_q_qd_u", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_qdd", lambda n: n),
        AlgoSpec("inverse_dynamics_gradient", "inverse_dynamics_gradient", "ID_DU", "",
                 lambda m: algos.body_inverse_dynamics_gradient(m, False),
                 ["q", "qd"], [("d_q_qd", 2 * n)], [(0, 0), (0, n)], "d_dc_du", lambda n: 2 * n * n),
        AlgoSpec("inverse_dynamics_gradient", "inverse_dynamics_gradient", "ID_DU", "_qdd",
                 lambda m: algos.body_inverse_dynamics_gradient(m, True),
                 ["q", "qd", "qdd"], [("d_q_qd", 2 * n), ("d_qdd", n)], [(0, 0), (0, n), (1, 0)], "d_dc_du",
                 lambda n: 2 * n * n),
        AlgoSpec("forward_dynamics_gradient", "forward_dynamics_gradient", "FD_DU", "",
                 algos.body_forward_dynamics_gradient,
                 ["q", "qd", "u"], [("d_q_qd_u", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_df_du", lambda n: 2 * n * n),
        AlgoSpec("forward_dynamics_gradient", "forward_dynamics_gradient", "FD_DU", "_qdd_minv",
                 algos.body_forward_dynamics_gradient_qdd_minv,
                 ["q", "qd", "qdd", "Minv"], [("d_q_qd", 2 * n), ("d_qdd", n), ("d_Minv", n * n)],
                 [(0, 0), (0, n), (1, 0), (2, 0)], "d_df_du", lambda n: 2 * n * n),
        # grad FD + its Minv and qdd by-products in one launch (what an MPC / DDP caller consumes)
        AlgoSpec("forward_dynamics_linearization", "forward_dynamics_linearization", "FD_LIN", "",
                 algos.body_forward_dynamics_linearization,
                 ["q", "qd", "u"], [("d_q_qd_u", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_fd_lin",
                 lambda n: 3 * n * n + n),
        AlgoSpec("aba", "aba", "ABA", "", algos.body_aba,
                 ["q", "qd", "u"], [("d_q_qd_tau", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_qdd", lambda n: n),
        AlgoSpec("crba", "crba", "CRBA", "", algos.body_crba,
                 ["q"], [("d_q_qd", n)], [(0, 0)], "d_M", lambda n: n * n),
    ] + ([
        AlgoSpec("idsva_so", "idsva_so", "IDSVA_SO", "", algos.body_idsva_so,
                 ["q", "qd", "qdd"], [("d_q_qd_u", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_idsva_so",
                 lambda n: 4 * n * n * n),
        AlgoSpec("fdsva_so", "fdsva_so", "FDSVA_SO", "", algos.body_fdsva_so,
                 ["q", "qd", "u"], [("d_q_qd_u", 3 * n)], [(0, 0), (0, n), (0, 2 * n)], "d_df2",
                 lambda n: 4 * n * n * n),
    ] if second_order_enabled(max_tree) else [])


def ee_hessian_enabled(n: int) -> bool:
    """The EE Hessian has 6 EE n^2 outputs per knot (27 000 for Atlas, 95 % of them structural zeros); it is generated
    for robots up to GRID_EE_HESSIAN_MAX_JOINTS joints (default 64), the pose and its gradient for every robot.  Outputs
    that large go through plain per-lane stores (choose_config), not the transposed-store scratch."""
    return n <= int(os.environ.get("GRID_EE_HESSIAN_MAX_JOINTS", "64"))


def second_order_enabled(max_tree: int) -> bool:
    """The second-order kernels are emitted by differentiating the straight-line RNEA / CRBA at emit time.  A
    quantity only carries tangents for the joints it depends on, so the code grows with the cube of the largest
    base-rooted TREE, not of the joint count: HyQ (12 joints, four 3-joint legs) is a third of iiwa14's code.  They
    are generated for robots whose largest tree has at most SECOND_ORDER_MAX_JOINTS joints (the reference likewise
    switches strategy above 8-10 joints: _idsva_so.py:1, _fdsva_so.py:1); results go straight to global memory
    (transposed-store path) whatever n is, as the reference does for large n (_fdsva_so.py:205-210)."""
    return max_tree <= int(os.environ.get("GRID_SECOND_ORDER_MAX_JOINTS", "8"))


def choose_config(slots: int, in_elems: List[int], out_elems: int, sizeof: int):
    """(NT, STATE_SMEM, STAGED) for one algorithm and scalar size.  STAGED: 0 = plain loads/stores, 2 = TMA-staged
    I/O with double-buffered inputs, 1 = TMA-staged with a single input buffer (chosen when that admits more warps),
    3 = plain loads, outputs through a lane-interleaved global scratch and a shared-memory transposition.

    One CTA per SM with as many warps as shared memory (per-thread state + per-warp TMA staging) and the
    register file allow.  The warps of the CTA run in lockstep through the straight-line body (GRID_SYNC points),
    which is what keeps the kernel off the instruction-fetch limit."""
    forced = os.environ.get("GRID_FORCE_LANE_CONFIG", "")      # A/B: "NT,STATE_SMEM,STAGED" for bodies with state slots
    if forced and slots > 1:
        nt_, st_, sg_ = (int(x) for x in forced.split(","))
        return nt_, bool(st_), sg_
    a16 = lambda x: (x + 15) & ~15
    state_w = a16(slots * 32 * sizeof)
    best = None
    if os.environ.get("GRID_FORCE_UNSTAGED") != "1":
        for nbuf in (2, 1):
            stage_w = sum(nbuf * a16(32 * e * sizeof) for e in in_elems) + a16(32 * out_elems * sizeof) + 16
            fit = (SMEM_LIMIT - 16) // (state_w + stage_w)
            if fit >= 1 and (best is None or min(MAX_WARPS, fit) > best[0]):
                best = (min(MAX_WARPS, fit), nbuf)
    fit = (SMEM_LIMIT - 16) // (state_w + 32 * 33 * sizeof)
    if best is not None and (best[0] >= 4 or fit < 2 * best[0] or os.environ.get("GRID_NO_TRANSPOSED") == "1"):
        return 32 * best[0], True, best[1]
    # Outputs beyond 1 MB per warp tile (Atlas EE Hessian: 27 000 values per knot) would need gigabytes of
    # transposed-store scratch per in-flight launch: each lane stores straight into its own output row instead.
    if 32 * out_elems * sizeof > (1 << 20):
        fit_state = max(1, (SMEM_LIMIT - 16) // max(state_w, 1))
        return 32 * min(MAX_WARPS, fit_state), True, 0
    # Staging a large per-knot output (Atlas matrices, second-order tensors) in shared memory would leave 1-3 warps
    # per SM.  STAGED = 3: every lane writes its output row into a lane-interleaved global scratch (element idx at
    # [idx*32 + lane]: each store instruction is one full 128-byte line), then the CTA transposes the scratch into
    # the caller's knot-major rows through 32 x 33 shared-memory tiles with coalesced reads and writes.
    if fit >= MAX_WARPS_SCRATCH:
        # GRID_LARGE_OUT_PLAIN=1 (A/B): lane-strided plain stores instead of the transposed store
        return 32 * min(MAX_WARPS, fit), True, (0 if os.environ.get("GRID_LARGE_OUT_PLAIN") == "1" else 3)
    # State too large for >= 4 warps in shared memory (Atlas gradients): state in the global scratch, 4 warps,
    # plain stores (measured on B200 with tools/tune_config.py: Atlas grad ID 212 us against 271 us transposed and
    # 294 us with 2 warps of shared-memory state; grad FD 387 / 393 / 662 us; batch 8192).
    return 32 * MAX_WARPS_SCRATCH, False, int(os.environ.get("GRID_SCRATCH_STAGED", "0"))


def algo_struct_name(sp: AlgoSpec) -> str:
    return f"{sp.fn}{sp.tag}_algo"


def emit_inner(sp: AlgoSpec, body: algos.Body) -> str:
    ins = "".join(f"const T* __restrict__ {p}, " for p in sp.inner_inputs)
    grav = "const T gravity" if sp.uses_gravity else "const T /*gravity*/"
    doc = (f"/**\n * {body.doc}\n * One thread = one knot point. o: this knot's outputs; inputs: this knot's "
           f"vectors;\n * S: this thread's state base (element i at S[i*NT], {body.nslots} elements).\n"
           f" * Emitted operations (mul+add): {body.n_ops}\n */\n")
    head = (f"template <typename T, int NT, int OS = 1>\n__device__ __forceinline__\n"
            f"void {sp.fn}_inner(T* __restrict__ o, {ins}T* __restrict__ S, {grav}) {{\n")
    return doc + head + "    using grid_rt::grid_sincos;\n    typedef grid_rt::pair_t<T> P;\n" + body.code + "}\n\n"


def emit_algo_struct(sp: AlgoSpec, body: algos.Body, n: int) -> str:
    IN = [e for _, e in sp.streams] + [0, 0, 0]
    args = []
    for (s, off) in sp.split:
        args.append(f"x{s}" + (f" + {off}" if off else ""))
    call = f"{sp.fn}_inner<T, NT, OS>(o, {', '.join(args)}, S, gravity);"
    return (f"template <typename T>\nstruct {algo_struct_name(sp)} {{\n"
            f"    static constexpr int IN0 = {IN[0]}, IN1 = {IN[1]}, IN2 = {IN[2]};\n"
            f"    static constexpr int OUT = {sp.out_n(n)};\n"
            f"    static constexpr int SLOTS = {max(body.nslots, 1)};\n"
            f"    template <int NT, int OS>\n"
            f"    static __device__ __forceinline__ void run(T* __restrict__ o, const T* __restrict__ x0, "
            f"const T* __restrict__ x1, const T* __restrict__ x2, T* __restrict__ S, const T gravity) {{\n"
            f"        {call}\n    }}\n}};\n\n")


def emit_config(sp: AlgoSpec, body: algos.Body, n: int, split_scratch=(False, False)) -> str:
    IN = [e for _, e in sp.streams]
    out = sp.out_n(n)
    c4 = choose_config(max(body.nslots, 1), IN, out, 4)
    c8 = choose_config(max(body.nslots, 1), IN, out, 8)
    b = lambda x: "true" if x else "false"
    name = f"{sp.fn}{sp.tag}_cfg"
    A = algo_struct_name(sp)
    return (f"template <typename T>\nstruct {name} {{\n"
            f"    static constexpr int NT = sizeof(T) == 4 ? {c4[0]} : {c8[0]};\n"
            f"    static constexpr bool STATE_SMEM = sizeof(T) == 4 ? {b(c4[1])} : {b(c8[1])};\n"
            f"    static constexpr int STAGED = sizeof(T) == 4 ? {int(c4[2])} : {int(c8[2])};\n"
            f"    static constexpr size_t SMEM_BYTES = grid_rt::tile_smem_bytes<{A}<T>, T, NT, STATE_SMEM, STAGED>();\n"
            f"    // the kernel uses d_robotModel->d_scratch: at most ONE such launch may be in flight per robotModel\n"
            f"    static constexpr bool NEEDS_SCRATCH = !STATE_SMEM || STAGED == 3 || (sizeof(T) == 4 ? {b(split_scratch[0])} : {b(split_scratch[1])});   // incl. the split form's\n"
            f"}};\n\n")


# ------------------------------------------------------------------ split (warp-specialised) form
SPLIT_KEYS = {"inverse_dynamics_gradient": lambda m, tag: algos.split_inverse_dynamics_gradient(m, tag == "_qdd"),
              "forward_dynamics_gradient": lambda m, tag: algos.split_forward_dynamics_gradient(m, False),
              "forward_dynamics_linearization": lambda m, tag: algos.split_forward_dynamics_gradient(m, True)}
SPLIT_MAX_WARPS = int(os.environ.get("GRID_SPLIT_MAX_WARPS", "16"))


def has_split(sp: AlgoSpec, n: int) -> bool:
    return algos.split_enabled(n) and sp.key in SPLIT_KEYS and sp.tag in ("", "_qdd")


def split_body(sp: AlgoSpec, model: Model) -> "algos.SplitBody":
    return SPLIT_KEYS[sp.key](model, sp.tag)


def split_assign(sb: "algos.SplitBody", max_warps: int = 0):
    """Columns -> warps, longest-processing-time first.  Warp counts are multiples of four (registers are partitioned
    per scheduler, ceil(warps / 4) warps each); the smallest count is taken whose heaviest warp is within 10 % of the
    heaviest single column (more warps cannot shorten phase B) or below 30 % of phase A (phase B no longer matters):
    small CTAs leave room for several CTAs per SM (measured on B200, profiles/r02a_tune_variants.log: iiwa14 with 4
    warps 12.3 us at 8192 knots against 22.7 us with 8).  Returns a list (one entry per warp) of column lists."""
    max_warps = max_warps or SPLIT_MAX_WARPS
    cols = sorted(((ops, c) for c, _, ops in sb.columns), reverse=True)
    forced = os.environ.get("GRID_SPLIT_WARPS", "")
    best = None
    for nw in ([int(forced)] if forced else range(4, max_warps + 1, 4)):
        load = [0] * nw
        groups = [[] for _ in range(nw)]
        for ops, c in cols:
            w = load.index(min(load))
            load[w] += ops
            groups[w].append(c)
        best = groups
        if max(load) <= max(1.10 * cols[0][0], 0.30 * sb.phase_a_ops):
            break
    return [sorted(g) for g in best if g]


def split_maxnreg(nw: int, ctas: int = 1) -> int:
    """Largest register budget per thread that lets `ctas` CTAs of nw warps reside on an SM: each of the SM's four
    schedulers owns 16384 registers and hosts ceil(nw / 4) warps of every CTA."""
    per_sched = ctas * -(-nw // 4)
    return max(32, min(255, (16384 // (per_sched * 32)) // 8 * 8))


SMEM_PER_SM = 232448            # 227 KB; every resident CTA also reserves 1 KB


def split_smem_plan(sb: "algos.SplitBody", sp: AlgoSpec, n: int, sizeof: int):
    """(OUT_STAGED, NBUF, resident CTAs per SM that shared memory allows, bytes) for the split form.
    Output rows are staged in shared memory (one TMA bulk store per tile) when that fits at all.  Inputs are double
    buffered unless dropping the second buffer buys a second resident CTA: the split kernels of big robots are bound
    by per-warp instruction-fetch latency (Atlas: every warp streams its own ~100 KB of straight-line code), so a
    second CTA per SM is worth far more than overlapping a 12 KB load with phase A."""
    a16 = lambda x: (x + 15) & ~15
    state = a16(max(sb.nslots, 1) * 32 * sizeof)
    ins = sum(a16(32 * e * sizeof) for _, e in sp.streams)
    out = a16(32 * sp.out_n(n) * sizeof)
    half = (SMEM_PER_SM - 2 * 1024) // 2
    forced = os.environ.get("GRID_SPLIT_NBUF", "")
    best = None
    for staged in (True, False):
        for nbuf in ([int(forced)] if forced else (2, 1)):
            need = state + nbuf * ins + (out if staged else 0) + 32
            if need > SMEM_LIMIT:
                continue
            ctas = 2 if need <= half else 1
            cand = (staged, nbuf, ctas, need)
            # preference: staged output first; then two resident CTAs; then double buffering
            if best is None or (cand[0], cand[2], cand[1]) > (best[0], best[2], best[1]):
                best = cand
        if best is not None and best[0]:
            break
    return best or (False, 1, 1, state + ins + 32)


def split_ntile(sb: "algos.SplitBody", sp: AlgoSpec, n: int, sizeof: int, mono_slots: Optional[int] = None) -> int:
    """Tiles whose published state a split CTA holds at once (grid_rt::tile_loop_split_multi).  Two, when shared
    memory holds them, for robots whose split form is the default at EVERY batch size -- their one-lane-per-knot form
    would keep its state in the global scratch (mono_slots = its slot count: Atlas, floating-base HyQ), or their
    output rows do not fit shared memory -- since at size that form is bound by instruction supply and every column
    body fetched from L2 is then executed by two warps at once.  One for the others: there the split form is the
    small-batch, latency-first form and keeps its TMA-staged output rows."""
    forced = os.environ.get("GRID_SPLIT_NTILE", "")
    if forced:
        return max(1, int(forced))
    staged = split_smem_plan(sb, sp, n, sizeof)[0]
    always = mono_slots is not None and not choose_config(max(mono_slots, 1), [e for _, e in sp.streams], sp.out_n(n), sizeof)[1]
    if staged and not always:
        return 1
    a16 = lambda x: (x + 15) & ~15
    need = lambda g: a16(g * max(sb.nslots, 1) * 32 * sizeof) + sum(a16(g * 32 * e * sizeof) for _, e in sp.streams) + 32
    # phase B pairs the CTA's warps over two tiles (emit_split rounds the warp count up to even)
    return 2 if need(2) <= SMEM_LIMIT else 1


def split_out_staged(sb: "algos.SplitBody", sp: AlgoSpec, n: int, sizeof: int) -> bool:
    return split_smem_plan(sb, sp, n, sizeof)[0]


def emit_split(sp: AlgoSpec, sb: "algos.SplitBody", n: int, mono_slots: int = 1) -> str:
    """phase A + column functions + algo struct + config + kernel of the split form of one algorithm."""
    stem = f"{sp.fn}{sp.tag}_split"
    ins = "".join(f"const T* __restrict__ {p}, " for p in sp.inner_inputs)
    grav = "const T gravity" if sp.uses_gravity else "const T /*gravity*/"
    pre = "    using grid_rt::grid_sincos;\n    typedef grid_rt::pair_t<T> P;\n"
    out = []
    for p_, (code, ops) in enumerate(sb.parts):
        out.append(f"/**\n * {sb.doc} -- SPLIT form, phase A, part {p_} of {len(sb.parts)} (parts = groups of base-rooted trees;\n"
                   f" * they run concurrently on different warps): the first-order passes of one knot per lane; publishes the\n"
                   f" * per-knot state as S[slot*NT] ({sb.nslots} slots in all).  Emitted operations (mul+add): {ops}\n */\n"
                   f"template <typename T, int NT, int OS = 1>\n__device__ __forceinline__\n"
                   f"void {stem}_phaseA{p_}(T* __restrict__ o, {ins}T* __restrict__ S, {grav}) {{\n" + pre + code + "}\n\n")
    for c, code, ops in sb.columns:
        out.append(f"/** SPLIT form, derivative column {c} (dq_{c} and dqd_{c}) of {sp.fn}: reads the published state, "
                   f"writes its output entries.  Emitted operations: {ops} */\n"
                   f"template <typename T, int NT, int OS = 1>\n__device__ __forceinline__\n"
                   f"void {stem}_col{c}(T* __restrict__ o, T* __restrict__ S) {{\n" + pre + code + "}\n\n")
    groups = split_assign(sb)
    na = len(sb.parts)
    nw = max(len(groups), na)
    if nw % 2 and max(split_ntile(sb, sp, n, 4, mono_slots), split_ntile(sb, sp, n, 8, mono_slots)) > 1:
        nw += 1                  # the two-tile form pairs the warps (tile_loop_split_multi)
    plan4 = split_smem_plan(sb, sp, n, 4)
    plan8 = split_smem_plan(sb, sp, n, 8)
    IN = [e for _, e in sp.streams] + [0, 0, 0]
    args = []
    for (s_, off) in sp.split:
        args.append(f"x{s_}" + (f" + {off}" if off else ""))
    cases = []
    for w, g in enumerate(groups):
        calls = " ".join(f"{stem}_col{c}<T, NT, OS>(o, S);" for c in g)
        cases.append(f"            case {w}: {calls} break;")
    writes = "GRID_BEFORE_FIRST_OUTPUT" in sb.code
    g4, g8 = split_ntile(sb, sp, n, 4, mono_slots), split_ntile(sb, sp, n, 8, mono_slots)
    xpose = os.environ.get("GRID_SPLIT_XPOSE", "0") == "1"
    mcases = []
    for w in range(0, len(groups), 2):
        # warps w and w+1: with two tiles resident both run the columns of groups w and w+1, each on its own tile (the
        # same code at the same time: one instruction fetch); with one tile resident warp w+h runs group w+h
        halves = []
        for h in (0, 1):
            if w + h < len(groups):
                calls = " ".join(f"{stem}_col{c}<T, NT, OS>(o, St);" for c in groups[w + h])
                halves.append(f"                if (both || h == {h}) {{ {calls} }}")
        mcases.append(f"            case {w // 2}:\n" + "\n".join(halves) + "\n                break;")
    out.append(f"template <typename T>\nstruct {stem}_algo {{\n"
               f"    static constexpr int IN0 = {IN[0]}, IN1 = {IN[1]}, IN2 = {IN[2]};\n"
               f"    static constexpr int OUT = {sp.out_n(n)};\n"
               f"    static constexpr int SLOTS = {max(sb.nslots, 1)};\n"
               f"    static constexpr int NW = {nw};          // warps per CTA = column groups of phase B\n"
               f"    static constexpr int NA = {na};           // warps running phase A (one part each)\n"
               f"    static constexpr int NBUF = sizeof(T) == 4 ? {plan4[1]} : {plan8[1]};   // input staging buffers\n"
               f"    static constexpr bool PHASE_A_WRITES = {'true' if writes else 'false'};\n"
               f"    static constexpr int NTILE = sizeof(T) == 4 ? {g4} : {g8};   // tiles resident per CTA pass (tile_loop_split_multi when > 1)\n"
               f"    static constexpr bool XPOSE = {'true' if xpose else 'false'};  // multi-tile form: transposed store through the global scratch\n"
               f"    template <int NT, int OS>\n"
               f"    static __device__ __forceinline__ void phaseA(int warp, T* __restrict__ o, const T* __restrict__ x0, "
               f"const T* __restrict__ x1, const T* __restrict__ x2, T* __restrict__ S, const T gravity) {{\n"
               f"        switch (warp) {{\n" + "\n".join(
                   f"            case {p_}: {stem}_phaseA{p_}<T, NT, OS>(o, {', '.join(args)}, S, gravity); break;"
                   for p_ in range(na)) + "\n            default: break;\n        }\n    }\n"
               f"    template <int NT, int OS>\n"
               f"    static __device__ __forceinline__ void phaseB(int warp, T* __restrict__ o, T* __restrict__ S) {{\n"
               f"        switch (warp) {{\n" + "\n".join(cases) + "\n            default: break;\n        }\n    }\n"
               f"    // phase B over the nt (1 or 2) resident tiles of a pass: tile t's output rows are o0 / o1, its state S + t*SLOTS*NT\n"
               f"    template <int NT, int OS>\n"
               f"    static __device__ __forceinline__ void phaseB_multi(int warp, int nt, T* __restrict__ o0, T* __restrict__ o1, T* __restrict__ S) {{\n"
               f"        const bool both = nt > 1;\n"
               f"        const int h = warp & 1;\n"
               f"        T* __restrict__ o = (both && h) ? o1 : o0;\n"
               f"        T* __restrict__ St = (both && h) ? S + size_t(SLOTS) * NT : S;\n"
               f"        switch (warp >> 1) {{\n" + "\n".join(mcases) + "\n            default: break;\n        }\n    }\n};\n\n")
    st4, st8 = plan4[0], plan8[0]
    b = lambda x: "true" if x else "false"
    A = f"{stem}_algo"
    regs = int(os.environ.get("GRID_SPLIT_MAXNREG", "0"))
    out.append(f"template <typename T>\nstruct {stem}_cfg {{\n"
               f"    static constexpr int NT = {nw * 32};\n"
               f"    static constexpr bool OUT_STAGED = sizeof(T) == 4 ? {b(st4)} : {b(st8)};\n"
               f"    static constexpr size_t SMEM_BYTES = {A}<T>::NTILE > 1 ? grid_rt::split_multi_smem_bytes<{A}<T>, T>()\n"
               f"                                                          : grid_rt::split_smem_bytes<{A}<T>, T, OUT_STAGED>();\n"
               f"    static constexpr bool NEEDS_SCRATCH = {A}<T>::NTILE > 1 && {A}<T>::XPOSE;   // transposed store through d_robotModel->d_scratch\n"
               f"    static constexpr bool FITS = SMEM_BYTES <= {SMEM_LIMIT};   // else the split kernel is never launched\n"
               f"}};\n\n")
    fill = []
    for i, (name, e) in enumerate(sp.streams):
        stride = f"stride_{sp.streams[0][0][2:]}" if i == 0 else str(e)
        fill.append(f"    a.in[{i}] = {name}; a.stride[{i}] = {stride};")
    for i in range(len(sp.streams), 3):
        fill.append(f"    a.in[{i}] = nullptr; a.stride[{i}] = 0;")
    gv = "gravity" if sp.uses_gravity else "T(0)"
    C = f"{stem}_cfg<T>"
    # register budget for as many resident CTAs as the fp32 shared-memory plan allows (at most two)
    # (only when that leaves >= 128 registers per thread: HyQ's 12-warp CTAs keep the whole register file)
    two = split_maxnreg(nw, 2) if (plan4[2] >= 2 and g4 == 1) else 0
    bounds = f"__maxnreg__({regs or (two if two >= 128 else split_maxnreg(nw, 1))})"
    out.append(f"template <typename T>\n__global__ void {bounds}\n"
               f"{sp.fn}_split_kernel({kernel_params(sp)}) {{\n"
               f"    extern __shared__ __align__(16) unsigned char grid_smem[];\n"
               f"    grid_rt::TileArgs<T> a;\n" + "\n".join(fill) + "\n"
               f"    a.out = {sp.out_name}; a.scratch = d_robotModel ? d_robotModel->d_scratch : nullptr;\n"
               f"    a.gravity = {gv}; a.num_knots = NUM_TIMESTEPS;\n"
               f"    if constexpr ({A}<T>::NTILE > 1) {{ if ({C}::FITS) grid_rt::tile_loop_split_multi<{A}<T>, T>(a, grid_smem); }}\n"
               f"    else {{ if ({C}::FITS) grid_rt::tile_loop_split<{A}<T>, T, {C}::OUT_STAGED>(a, grid_smem); }}\n}}\n\n")
    # default batch size up to which the launcher prefers the split form (tools/tune_split.py measures the crossover)
    mono = choose_config(max(mono_slots, 1), [e for _, e in sp.streams], sp.out_n(n), 4)
    max_knots = (1 << 30) if not mono[1] else SPLIT_DEFAULT_MAX_KNOTS
    out.append(f"inline int {stem}_max_knots() {{ return {max_knots}; }}\n\n")
    return "".join(out)


SPLIT_DEFAULT_MAX_KNOTS = 2 * 148 * 32


def kernel_params(sp: AlgoSpec) -> str:
    ps = [f"T *{sp.out_name}"]
    first = True
    for name, _ in sp.streams:
        ps.append(f"const T *{name}")
        if first:
            ps.append(f"const int stride_{name[2:]}")
            first = False
    ps.append("const robotModel<T> *d_robotModel")
    if sp.uses_gravity:
        ps.append("const T gravity")
    ps.append("const int NUM_TIMESTEPS")
    return ", ".join(ps)


def emit_device_and_kernel(sp: AlgoSpec, n: int) -> str:
    A = algo_struct_name(sp)
    C = f"{sp.fn}{sp.tag}_cfg<T>"
    tl = f"grid_rt::tile_loop<{A}<T>, T, {C}::NT, {C}::STATE_SMEM, {C}::STAGED>"
    fill = []
    for i, (name, e) in enumerate(sp.streams):
        stride = f"stride_{sp.streams[0][0][2:]}" if i == 0 else str(e)
        fill.append(f"    a.in[{i}] = {name}; a.stride[{i}] = {stride};")
    for i in range(len(sp.streams), 3):
        fill.append(f"    a.in[{i}] = nullptr; a.stride[{i}] = 0;")
    grav = "gravity" if sp.uses_gravity else "T(0)"
    dev = (f"/** Per-warp tile loop of {sp.fn}: call from every thread of a kernel launched with\n"
           f" *  {sp.code}_DYNAMIC_SHARED_MEM_COUNT*sizeof(T) bytes of dynamic shared memory. */\n"
           f"template <typename T>\n__device__ __forceinline__\n"
           f"void {sp.fn}_device(const grid_rt::TileArgs<T> &a, unsigned char *s_mem) {{\n"
           f"    {tl}(a, s_mem);\n}}\n\n") if sp.tag == "" else ""
    # overloads of _device are distinguished by the algo struct, so tagged variants get a tagged name
    if sp.tag:
        dev = (f"template <typename T>\n__device__ __forceinline__\n"
               f"void {sp.fn}{sp.tag}_device(const grid_rt::TileArgs<T> &a, unsigned char *s_mem) {{\n"
               f"    {tl}(a, s_mem);\n}}\n\n")
    # register budget: __launch_bounds__(NT, 1) lets ptxas round the per-thread limit down coarsely (128 for 448
    # threads); __maxnreg__ pins it to the largest multiple of 8 that still fits one CTA in the 64 K register file
    bounds = (f"__maxnreg__(({C}::NT * 255 <= 65536) ? 255 : (65536 / {C}::NT / 8 * 8))"
              if os.environ.get("GRID_USE_MAXNREG", "1") == "1" else f"__launch_bounds__({C}::NT, 1)")
    ker = (f"template <typename T>\n__global__ void {bounds}\n"
           f"{sp.fn}_kernel({kernel_params(sp)}) {{\n"
           f"    extern __shared__ __align__(16) unsigned char grid_smem[];\n"
           f"    grid_rt::TileArgs<T> a;\n" + "\n".join(fill) + "\n"
           f"    a.out = {sp.out_name}; a.scratch = d_robotModel ? d_robotModel->d_scratch : nullptr;\n"
           f"    a.gravity = {grav}; a.num_knots = NUM_TIMESTEPS;\n"
           f"    {sp.fn}{sp.tag}_device<T>(a, grid_smem);\n}}\n\n")
    return dev + ker


# algorithms whose reference device tier (banner GRiDCodeGenerator.py:318-352) gets a same-signature entry point here
REFERENCE_DEVICE_TIER = ("inverse_dynamics", "inverse_dynamics_vaf", "direct_minv", "forward_dynamics", "inverse_dynamics_gradient",
                         "forward_dynamics_gradient", "end_effector_pose", "end_effector_pose_gradient",
                         "end_effector_pose_gradient_hessian")


def emit_reference_device(sp: AlgoSpec, n: int) -> str:
    """`X_device<T>(T *s_out, const T *s_q, ..., const robotModel<T> *d_robotModel[, const T gravity])` with the
    REFERENCE's parameter list and calling convention (shared-memory pointers, one knot point per thread block, every
    thread of the block calls it; e.g. algorithms/_inverse_dynamics.py:360-386).  User kernels written against the
    reference's device tier compile and give the right results; the work of the knot is done by the block's first warp
    running this back end's per-thread body (all its lanes compute the same knot, so the shared-memory results are
    written consistently), the other warps wait at the barrier.  For throughput use X_kernel / X_launch instead."""
    if sp.key not in REFERENCE_DEVICE_TIER or sp.fixed_target:
        return ""
    ins = "".join(f"const T *s_{p_}, " for p_ in sp.inner_inputs)
    grav = ", const T gravity" if sp.uses_gravity else ""
    args = ", ".join(f"s_{p_}" for p_ in sp.inner_inputs)
    gcall = "gravity" if sp.uses_gravity else "T(0)"
    return (f"/** Reference-signature device tier of {sp.fn}{sp.tag}: shared-memory pointers, one knot point per thread block,\n"
            f" *  called by every thread of the block (it synchronises the block before and after).  d_robotModel is accepted\n"
            f" *  for source compatibility (robot constants are immediates of the code). */\n"
            f"template <typename T>\n__device__\n"
            f"void {sp.fn}_device(T *s_out, {ins}const robotModel<T> *d_robotModel{grav}) {{\n"
            f"    (void)d_robotModel;\n"
            f"    __syncthreads();\n"
            f"    if (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y < 32) {{\n"
            f"        T S[{algo_struct_name(sp)}<T>::SLOTS];\n"
            f"        {sp.fn}_inner<T, 1, 1>(s_out, {args}, S, {gcall});\n"
            f"    }}\n"
            f"    __syncthreads();\n}}\n\n")


def emit_launch(sp: AlgoSpec, n: int) -> str:
    """Stream-asynchronous launcher: persistent grid sized from the occupancy of this kernel."""
    C = f"{sp.fn}{sp.tag}_cfg<T>"
    ps = [f"T *{sp.out_name}"]
    call = [sp.out_name]
    first = True
    for name, _ in sp.streams:
        ps.append(f"const T *{name}")
        call.append(name)
        if first:
            ps.append(f"const int stride_{name[2:]}")
            call.append(f"stride_{name[2:]}")
            first = False
    ps.append("const robotModel<T> *d_robotModel")
    call.append("d_robotModel")
    if sp.uses_gravity:
        ps.append("const T gravity")
        call.append("gravity")
    ps.append("const int num_timesteps")
    call.append("num_timesteps")
    ps.append("cudaStream_t stream")
    kfn = f"{sp.fn}_kernel<T>"
    split = form_fn = ""
    if has_split(sp, n):
        # `form`: -1 = pick by batch size (<algo>_form_for), 0 = one lane per knot, 1 = split.  The host-buffer entry
        # points decide once for the whole batch and pass the decision to every chunk's launch.
        ps.append("int form = -1")
        SC = f"{sp.fn}{sp.tag}_split_cfg<T>"
        form_fn = (f"/** Kernel form the launcher uses for a batch of num_timesteps knot points: 1 = split (phase A + columns\n"
                   f" *  spread over the CTA's warps), 0 = one lane per knot point. */\n"
                   f"template <typename T>\n__host__\nint {sp.fn}{sp.tag}_form_for(const int num_timesteps) {{\n"
                   f"    return ({SC}::FITS && grid_rt::use_split_kernel(num_timesteps, {sp.fn}{sp.tag}_split_max_knots())) ? 1 : 0;\n}}\n\n")
        split = (f"    if (form < 0) form = {sp.fn}{sp.tag}_form_for<T>(num_timesteps);\n"
                 f"    if (form == 1 && {SC}::FITS) {{\n"
                 f"        auto skern = static_cast<void (*)({_sig_types(sp)})>({sp.fn}_split_kernel<T>);\n"
                 f"        int sblocks = grid_rt::persistent_blocks(reinterpret_cast<const void*>(skern), {SC}::NT, "
                 f"{SC}::SMEM_BYTES, num_timesteps);\n"
                 f"        if (sblocks <= 0) {{ cudaError_t e = cudaGetLastError(); return e != cudaSuccess ? e : cudaErrorLaunchOutOfResources; }}\n"
                 f"        if ({SC}::NEEDS_SCRATCH) {{\n"
                 f"            const int cap = grid_rt::scratch_capacity(GRID_SCRATCH_BLOCKS_PER_SM);\n"
                 f"            if (cap <= 0 || d_robotModel == nullptr) return cudaErrorInvalidValue;\n"
                 f"            if (sblocks > cap) sblocks = cap;\n"
                 f"        }}\n"
                 f"        cudaError_t se = grid_rt::launch_tile_kernel(skern, sblocks, {SC}::NT, {SC}::SMEM_BYTES, stream, {', '.join(call)});\n"
                 f"        return se != cudaSuccess ? se : cudaGetLastError();\n"
                 f"    }}\n")
    return (form_fn + f"template <typename T>\n__host__\ncudaError_t {sp.fn}{sp.tag}_launch({', '.join(ps)}) {{\n"
            f"    if (num_timesteps <= 0) return cudaSuccess;\n" + split +
            f"    auto kern = static_cast<void (*)({_sig_types(sp)})>({kfn});\n"
            f"    int blocks = grid_rt::persistent_blocks(reinterpret_cast<const void*>(kern), {C}::NT, "
            f"{C}::SMEM_BYTES, num_timesteps);\n"
            f"    if (blocks <= 0) {{ cudaError_t e = cudaGetLastError(); return e != cudaSuccess ? e : cudaErrorLaunchOutOfResources; }}\n"
            f"    if (!{C}::STATE_SMEM || {C}::STAGED == 3) {{\n"
            f"        const int cap = grid_rt::scratch_capacity(GRID_SCRATCH_BLOCKS_PER_SM);\n"
            f"        if (cap <= 0 || d_robotModel == nullptr) return cudaErrorInvalidValue;\n"
            f"        if (blocks > cap) blocks = cap;\n"
            f"    }}\n"
            f"    cudaError_t le = grid_rt::launch_tile_kernel(kern, blocks, {C}::NT, {C}::SMEM_BYTES, stream, {', '.join(call)});\n"
            f"    return le != cudaSuccess ? le : cudaGetLastError();\n}}\n\n")


def _sig_types(sp: AlgoSpec) -> str:
    ts = ["T*"]
    first = True
    for _ in sp.streams:
        ts.append("const T*")
        if first:
            ts.append("const int")
            first = False
    ts.append("const robotModel<T>*")
    if sp.uses_gravity:
        ts.append("const T")
    ts.append("const int")
    return ", ".join(ts)


# ------------------------------------------------------------------ host tier
@dataclass
class HostSpec:
    fn: str                         # C++ function stem; for fixed-target EE variants it carries the target suffix
    tmpl: str                       # template parameter list after "typename T"
    label: str                      # name printed by _single_timing
    uses_gravity: bool
    pre: List[str]                  # lines computing stride / pointers
    h2d: List[Tuple[str, str, str, str]]   # (condition or "", d_ptr, h_ptr, elements per knot)
    launch: List[str]               # lines performing the launch into `err`
    d2h: Tuple[str, str, str]       # (d_ptr, h_ptr, elements per knot)
    extra_d2h: List[Tuple[str, str, str]] = field(default_factory=list)
    key: str = ""                   # generator key (gen_<key>_host); defaults to fn


def host_specs(n: int, fixed_target: str = "") -> List[HostSpec]:
    qsel = ["const int stride_q = USE_COMPRESSED_MEM ? NUM_JOINTS : 3*NUM_JOINTS;",
            "T *d_in = USE_COMPRESSED_MEM ? hd_data->d_q : hd_data->d_q_qd_u;",
            "T *h_in = USE_COMPRESSED_MEM ? hd_data->h_q : hd_data->h_q_qd_u;"]
    comp = ["const int stride_q_qd = USE_COMPRESSED_MEM ? 2*NUM_JOINTS : 3*NUM_JOINTS;",
            "T *d_in = USE_COMPRESSED_MEM ? hd_data->d_q_qd : hd_data->d_q_qd_u;",
            "T *h_in = USE_COMPRESSED_MEM ? hd_data->h_q_qd : hd_data->h_q_qd_u;"]
    ft = []
    if fixed_target:
        sfx = "_" + fixed_target
        ft = [
            HostSpec("end_effector_pose" + sfx, ", bool USE_COMPRESSED_MEM = false", "EE_POS", False, qsel,
                     [("", "d_in", "h_in", "stride_q")],
                     [f"err = end_effector_pose{sfx}_launch<T>(hd_data->d_eePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                     ("hd_data->d_eePos", "hd_data->h_eePos", "6"), key="end_effector_pose"),
            HostSpec("end_effector_pose_gradient" + sfx, ", bool USE_COMPRESSED_MEM = false", "DEE_POS", False, qsel,
                     [("", "d_in", "h_in", "stride_q")],
                     [f"err = end_effector_pose_gradient{sfx}_launch<T>(hd_data->d_deePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                     ("hd_data->d_deePos", "hd_data->h_deePos", "6*NUM_JOINTS"), key="end_effector_pose_gradient"),
            HostSpec("end_effector_pose_gradient_hessian" + sfx, ", bool USE_COMPRESSED_MEM = false", "D2EE_POS", False, qsel,
                     [("", "d_in", "h_in", "stride_q")],
                     [f"err = end_effector_pose_gradient{sfx}_launch<T>(hd_data->d_deePos, d_in, stride_q, d_robotModel, num_timesteps, 0);",
                      f"if (err == cudaSuccess) err = end_effector_pose_gradient_hessian{sfx}_launch<T>(hd_data->d_d2eePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                     ("hd_data->d_d2eePos", "hd_data->h_d2eePos", "6*NUM_JOINTS*NUM_JOINTS"),
                     extra_d2h=[("hd_data->d_deePos", "hd_data->h_deePos", "6*NUM_JOINTS")],
                     key="end_effector_pose_gradient_hessian"),
        ]
    return ft + [
        HostSpec("joint_transforms", ", bool USE_COMPRESSED_MEM = false", "JOINT_XFORMS", False, qsel,
                 [("", "d_in", "h_in", "stride_q")],
                 ["err = joint_transforms_launch<T>(hd_data->d_jointXforms, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                 ("hd_data->d_jointXforms", "hd_data->h_jointXforms", "16*NUM_JOINTS")),
        HostSpec("inverse_dynamics", ", bool USE_QDD_FLAG = false, bool USE_COMPRESSED_MEM = false", "ID", True, comp,
                 [("", "d_in", "h_in", "stride_q_qd"), ("USE_QDD_FLAG", "hd_data->d_qdd", "hd_data->h_qdd", "NUM_JOINTS")],
                 ["if (USE_QDD_FLAG) err = inverse_dynamics_qdd_launch<T>(hd_data->d_c, d_in, stride_q_qd, hd_data->d_qdd, d_robotModel, gravity, num_timesteps, 0);",
                  "else err = inverse_dynamics_launch<T>(hd_data->d_c, d_in, stride_q_qd, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_c", "hd_data->h_c", "NUM_JOINTS")),
        HostSpec("direct_minv", ", bool USE_COMPRESSED_MEM = false", "Minv", False,
                 ["const int stride_q = USE_COMPRESSED_MEM ? NUM_JOINTS : 3*NUM_JOINTS;",
                  "T *d_in = USE_COMPRESSED_MEM ? hd_data->d_q : hd_data->d_q_qd_u;",
                  "T *h_in = USE_COMPRESSED_MEM ? hd_data->h_q : hd_data->h_q_qd_u;"],
                 [("", "d_in", "h_in", "stride_q")],
                 ["err = direct_minv_launch<T>(hd_data->d_Minv, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                 ("hd_data->d_Minv", "hd_data->h_Minv", "NUM_JOINTS*NUM_JOINTS")),
        HostSpec("forward_dynamics", "", "FD", True, ["const int stride_q_qd_u = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd_u")],
                 ["err = forward_dynamics_launch<T>(hd_data->d_qdd, hd_data->d_q_qd_u, stride_q_qd_u, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_qdd", "hd_data->h_qdd", "NUM_JOINTS")),
        HostSpec("inverse_dynamics_gradient", ", bool USE_QDD_FLAG = false, bool USE_COMPRESSED_MEM = false", "ID_DU",
                 True, comp,
                 [("", "d_in", "h_in", "stride_q_qd"), ("USE_QDD_FLAG", "hd_data->d_qdd", "hd_data->h_qdd", "NUM_JOINTS")],
                 ["if (USE_QDD_FLAG) err = inverse_dynamics_gradient_qdd_launch<T>(hd_data->d_dc_du, d_in, stride_q_qd, hd_data->d_qdd, d_robotModel, gravity, num_timesteps, 0);",
                  "else err = inverse_dynamics_gradient_launch<T>(hd_data->d_dc_du, d_in, stride_q_qd, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_dc_du", "hd_data->h_dc_du", "2*NUM_JOINTS*NUM_JOINTS")),
        HostSpec("forward_dynamics_gradient", ", bool USE_QDD_MINV_FLAG = false", "FD_DU", True,
                 ["const int stride_q_qd = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd"),
                  ("USE_QDD_MINV_FLAG", "hd_data->d_qdd", "hd_data->h_qdd", "NUM_JOINTS"),
                  ("USE_QDD_MINV_FLAG", "hd_data->d_Minv", "hd_data->h_Minv", "NUM_JOINTS*NUM_JOINTS")],
                 ["if (USE_QDD_MINV_FLAG) err = forward_dynamics_gradient_qdd_minv_launch<T>(hd_data->d_df_du, hd_data->d_q_qd_u, stride_q_qd, hd_data->d_qdd, hd_data->d_Minv, d_robotModel, gravity, num_timesteps, 0);",
                  "else err = forward_dynamics_gradient_launch<T>(hd_data->d_df_du, hd_data->d_q_qd_u, stride_q_qd, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_df_du", "hd_data->h_df_du", "2*NUM_JOINTS*NUM_JOINTS")),
        HostSpec("forward_dynamics_linearization", "", "FD_LIN", True, ["const int stride_q_qd_u = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd_u")],
                 ["err = forward_dynamics_linearization_launch<T>(hd_data->d_fd_lin, hd_data->d_q_qd_u, stride_q_qd_u, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_fd_lin", "hd_data->h_fd_lin", "(3*NUM_JOINTS*NUM_JOINTS + NUM_JOINTS)")),
        HostSpec("aba", "", "ABA", True, ["const int stride_q_qd = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd")],
                 ["err = aba_launch<T>(hd_data->d_qdd, hd_data->d_q_qd_u, stride_q_qd, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_qdd", "hd_data->h_qdd", "NUM_JOINTS")),
        HostSpec("crba", ", bool USE_COMPRESSED_MEM = false", "CRBA", True, comp,
                 [("", "d_in", "h_in", "stride_q_qd")],
                 ["err = crba_launch<T>(hd_data->d_M, d_in, stride_q_qd, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_M", "hd_data->h_M", "NUM_JOINTS*NUM_JOINTS")),
        HostSpec("end_effector_pose", ", bool USE_COMPRESSED_MEM = false", "EE_POS", False, qsel,
                 [("", "d_in", "h_in", "stride_q")],
                 ["err = end_effector_pose_launch<T>(hd_data->d_eePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                 ("hd_data->d_eePos", "hd_data->h_eePos", "6*NUM_EES")),
        HostSpec("end_effector_pose_gradient", ", bool USE_COMPRESSED_MEM = false", "DEE_POS", False, qsel,
                 [("", "d_in", "h_in", "stride_q")],
                 ["err = end_effector_pose_gradient_launch<T>(hd_data->d_deePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                 ("hd_data->d_deePos", "hd_data->h_deePos", "6*NUM_EES*NUM_JOINTS")),
        HostSpec("end_effector_pose_gradient_hessian", ", bool USE_COMPRESSED_MEM = false", "D2EE_POS", False, qsel,
                 [("", "d_in", "h_in", "stride_q")],
                 ["err = end_effector_pose_gradient_launch<T>(hd_data->d_deePos, d_in, stride_q, d_robotModel, num_timesteps, 0);",
                  "if (err == cudaSuccess) err = end_effector_pose_gradient_hessian_launch<T>(hd_data->d_d2eePos, d_in, stride_q, d_robotModel, num_timesteps, 0);"],
                 ("hd_data->d_d2eePos", "hd_data->h_d2eePos", "6*NUM_EES*NUM_JOINTS*NUM_JOINTS"),
                 extra_d2h=[("hd_data->d_deePos", "hd_data->h_deePos", "6*NUM_EES*NUM_JOINTS")]),
        # reference names the IDSVA host function idsva_so_host (banner GRiDCodeGenerator.py:368); both names emitted
        HostSpec("idsva_so", "", "IDSVA_SO", True, ["const int stride_q_qd_u = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd_u")],
                 ["err = idsva_so_launch<T>(hd_data->d_idsva_so, hd_data->d_q_qd_u, stride_q_qd_u, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_idsva_so", "hd_data->h_idsva_so", "4*NUM_JOINTS*NUM_JOINTS*NUM_JOINTS")),
        HostSpec("fdsva_so", "", "FDSVA_SO", True, ["const int stride_q_qd_u = 3*NUM_JOINTS;"],
                 [("", "hd_data->d_q_qd_u", "hd_data->h_q_qd_u", "stride_q_qd_u")],
                 ["err = fdsva_so_launch<T>(hd_data->d_df2, hd_data->d_q_qd_u, stride_q_qd_u, d_robotModel, gravity, num_timesteps, 0);"],
                 ("hd_data->d_df2", "hd_data->h_df2", "4*NUM_JOINTS*NUM_JOINTS*NUM_JOINTS")),
    ]


def emit_host(hs: HostSpec, mode: int) -> str:
    """mode 0: H2D + kernel + D2H (blocking, like the reference); 1: _single_timing; 2: _compute_only.
    block_dimms / thread_dimms are accepted for source compatibility; the persistent grid is sized by
    occupancy (reference host tier: algorithms/_forward_dynamics_gradient.py:186-249)."""
    suffix = {0: "", 1: "_single_timing", 2: "_compute_only"}[mode]
    grav = "const T gravity, " if hs.uses_gravity else ""
    streams = "" if mode == 2 else ", cudaStream_t *streams"
    out = [f"template <typename T{hs.tmpl}>", "__host__",
           f"void {hs.fn}{suffix}(gridData<T> *hd_data, const robotModel<T> *d_robotModel, {grav}const int num_timesteps,",
           f"        const dim3 block_dimms, const dim3 thread_dimms{streams}) {{",
           "    (void)block_dimms; (void)thread_dimms;", "    cudaError_t err = cudaSuccess;"]
    out += ["    " + l for l in hs.pre]
    if mode != 2:
        for i, (cond, d, h, cnt) in enumerate(hs.h2d):
            line = f"gpuErrchk(cudaMemcpyAsync({d}, {h}, size_t({cnt})*num_timesteps*sizeof(T), cudaMemcpyHostToDevice, streams[{i}]));"
            out.append("    " + (f"if ({cond}) " if cond else "") + line)
        out.append("    gpuErrchk(cudaDeviceSynchronize());")
    if mode == 1:
        out.append("    struct timespec start, end; clock_gettime(CLOCK_MONOTONIC, &start);")
    out += ["    " + l for l in hs.launch]
    out.append("    gpuErrchk(err);")
    out.append("    gpuErrchk(cudaDeviceSynchronize());")
    if mode == 1:
        out.append("    clock_gettime(CLOCK_MONOTONIC, &end);")
        out.append(f'    printf("Single Call {hs.label} %fus\\n", time_delta_us_timespec(start, end)/static_cast<double>(num_timesteps));')
    if mode != 2:
        for d, h, cnt in [hs.d2h] + list(hs.extra_d2h):
            out.append(f"    gpuErrchk(cudaMemcpy({h}, {d}, size_t({cnt})*num_timesteps*sizeof(T), cudaMemcpyDeviceToHost));")
        out.append("    gpuErrchk(cudaDeviceSynchronize());")
    out.append("}")
    return "\n".join(out) + "\n\n"
