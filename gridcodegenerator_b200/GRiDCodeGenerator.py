"""GRiDCodeGenerator -- same Python surface as the reference class
(reference GRiDCodeGenerator.py:3-63, :309-437; README.md:10-56), new back end.

`GRiDCodeGenerator(robot).gen_all_code()` still writes one robot-specialised
header `<namespace>.cuh` exposing ALGORITHM_inner / _device / _kernel / host
functions, `robotModel<T>`, `gridData<T>`, `init_grid`, ... ; what it emits is
no longer the reference's block-per-knot generic SIMT code but straight-line
sm_100a code (one lane per knot point, one warp per 32-knot tile, TMA-staged
batch I/O) produced by `gridcodegenerator_b200.emit`.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

from .emit import algos as _algos
from .emit import header as _hdr
from .emit.spatial import Model

_CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")


class GRiDCodeGenerator:
    # ------------------------------------------------------------------ ctor
    def __init__(self, robotObj, DEBUG_MODE=False, NEED_PRINT_MAT=False, USE_DYNAMIC_SHARED_MEM=True,
                 FILE_NAMESPACE="grid", fixed_target_name=""):
        self.robot = robotObj
        self.code_str = ""
        self.indent_level = 0
        self.DEBUG_MODE = DEBUG_MODE
        self.gen_print_mat = DEBUG_MODE or NEED_PRINT_MAT
        self.use_dynamic_shared_mem_flag = True   # the tile kernels always use dynamic shared memory
        self.file_namespace = FILE_NAMESPACE
        if getattr(robotObj, "floating_base", False):
            raise NotImplementedError("floating-base robots are out of scope of the B200 hot path (SURVEY.md 2.1 row 18)")
        self.model = Model(robotObj)
        self._bodies: Dict[str, _algos.Body] = {}
        self._sbodies: Dict[str, _algos.SplitBody] = {}
        roots = [j for j in range(self.model.n) if self.model.parent[j] < 0]
        self.max_tree = max(len(self.model.subtree[r]) for r in roots)      # joints of the largest base-rooted tree
        self._set_fixed_target(fixed_target_name or "")

    def _set_fixed_target(self, name: str):
        """Fixed kinematic target (a fixed joint of the robot) for which the additional single-target EE functions
        are generated (reference gen_all_code(fixed_target_name=...), GRiDCodeGenerator.py:309-310, :417)."""
        if name:
            self.robot.get_fixed_joint_by_name(name)          # raises for an unknown name
        self.fixed_target_name = name
        self.include_fixed_kinematic_targets = name != ""
        self._specs = {sp.fn + sp.tag: sp
                       for sp in _hdr.specs(self.model.n, self.robot.get_total_leaf_nodes(), self.max_tree, name)}

    # ------------------------------------------------- emission core (L1 API)
    # reference helpers/_code_generation_helpers.py:1-40 (same names, same behaviour)
    def gen_add_code_line(self, new_code_line, add_indent_after=False):
        self.code_str += "    " * self.indent_level + new_code_line + "\n"
        if add_indent_after:
            self.indent_level += 1

    def gen_add_code_lines(self, new_code_lines, add_indent_after=False):
        for line in new_code_lines:
            self.gen_add_code_line(line)
        if add_indent_after:
            self.indent_level += 1

    def gen_add_end_control_flow(self):
        self.indent_level -= 1
        self.gen_add_code_line("}")

    def gen_add_end_function(self):
        self.indent_level -= 1
        self.gen_add_code_line("}")
        self.gen_add_code_line("")

    def gen_add_func_doc(self, func_desc, notes=[], params=[], return_val=None):
        self.gen_add_code_line("/**")
        self.gen_add_code_line(" * " + func_desc)
        self.gen_add_code_line(" *")
        for note in notes:
            self.gen_add_code_line(" * " + note)
        for p in params:
            self.gen_add_code_line(" * @param " + p)
        if return_val is not None:
            self.gen_add_code_line(" * @return " + return_val)
        self.gen_add_code_line(" */")

    def _add_block(self, text: str):
        # one append per block: code_str += per line is quadratic on a 16 MB header (Atlas: minutes)
        ind = "    " * self.indent_level
        self.code_str += "".join((ind + line if line else "") + "\n" for line in text.split("\n"))

    # --------------------------------------------------------- body cache
    def _body(self, name: str) -> _algos.Body:
        if name not in self._bodies:
            self._bodies[name] = self._specs[name].body(self.model)
        return self._bodies[name]

    def _variants(self, key: str) -> List[str]:
        return [k for k, sp in self._specs.items() if sp.key == key]

    def _temp_mem_size(self, key: str) -> int:
        """Per-knot state elements (the reference's s_temp size plays this role)."""
        return max([self._body(v).nslots for v in self._variants(key)] or [0])

    # ------------------------------------------------------ generic tiers
    def _gen_inner(self, key: str, variants: Optional[List[str]] = None):
        for v in variants or self._variants(key):
            sp = self._specs[v]
            b = self._body(v)
            self._add_block(_hdr.emit_inner(sp, b))
            self._add_block(_hdr.emit_algo_struct(sp, b, self.model.n))
            ss = (False, False)
            if _hdr.has_split(sp, self.model.n):
                if v not in self._sbodies:
                    self._sbodies[v] = _hdr.split_body(sp, self.model)
                xp = os.environ.get("GRID_SPLIT_XPOSE", "0") == "1"
                ss = tuple(xp and _hdr.split_ntile(self._sbodies[v], sp, self.model.n, sz, b.nslots) > 1 for sz in (4, 8))
            self._add_block(_hdr.emit_config(sp, b, self.model.n, ss))
            if sp.fn == "joint_transforms":
                # the reference's helper names for the same quantity (algorithms/_eepose_gradient_hessian.py:1226-1398)
                self._add_block(
                    "/** Local homogeneous transforms Xhom_j(q_j) -> s_XmatsHom and cumulative transforms base <- joint j\n"
                    " *  -> s_jointXforms for every joint id <= tid (16 values each, column-major), in one thread. */\n"
                    "template <typename T>\n__device__\nvoid X_single_thread(T *s_jointXforms, T *s_XmatsHom, const T *s_q, int tid) {\n"
                    "    using grid_rt::grid_sincos;\n    constexpr int OS = 1;\n"
                    + _algos.code_X_single_thread(self.model) + "}\n\n"
                    "/** Same result with the reference's warp-level entry point: in this back end one lane owns one knot point,\n"
                    " *  so lane 0 of the calling warp runs the chain and the warp re-converges. */\n"
                    "template <typename T>\n__device__ inline void X_warp(T* __restrict__ s_jointXforms, T* __restrict__ s_XmatsHom,\n"
                    "                              const T* __restrict__ s_q, int tid) {\n"
                    "    if ((threadIdx.x & 31) == 0) X_single_thread<T>(s_jointXforms, s_XmatsHom, s_q, tid);\n"
                    "    __syncwarp();\n}\n")
            if _hdr.has_split(sp, self.model.n):
                # second form of the same algorithm: phase A on one warp, derivative columns spread over the CTA's warps
                if v not in self._sbodies:
                    self._sbodies[v] = _hdr.split_body(sp, self.model)
                self._add_block(_hdr.emit_split(sp, self._sbodies[v], self.model.n, b.nslots))

    def _gen_device(self, key: str, variants=None):
        # device and kernel tiers are emitted together (the kernel is a 6-line shell around the device tier)
        for v in variants or self._variants(key):
            self._add_block(_hdr.emit_device_and_kernel(self._specs[v], self.model.n))
            self._add_block(_hdr.emit_reference_device(self._specs[v], self.model.n))

    def _gen_kernel(self, key: str, variants=None):
        for v in variants or self._variants(key):
            self._add_block(_hdr.emit_launch(self._specs[v], self.model.n))

    def _gen_host(self, key: str, mode: int = 0):
        for hs in _hdr.host_specs(self.model.n, self.fixed_target_name):
            if (hs.key or hs.fn) == key:
                self._add_block(_hdr.emit_host(hs, mode))

    def _gen_algo(self, key: str):
        """inner -> device/kernel -> launcher -> host x3, like the reference's gen_<algo> drivers
        (e.g. algorithms/_inverse_dynamics.py:514-533)."""
        if not self._variants(key):
            return      # e.g. the second-order kernels of a robot above GRID_SECOND_ORDER_MAX_JOINTS
        self._gen_inner(key)
        self._gen_device(key)
        self._gen_kernel(key)
        for mode in (0, 1, 2):
            self._gen_host(key, mode)
        if key == "idsva_so":   # the reference's banner calls this host function idsva_so_host (GRiDCodeGenerator.py:368)
            self._add_block("template <typename T>\n__host__\nvoid idsva_so_host(gridData<T> *hd_data, const robotModel<T> "
                            "*d_robotModel, const T gravity, const int num_timesteps,\n        const dim3 block_dimms, "
                            "const dim3 thread_dimms, cudaStream_t *streams) {\n    idsva_so<T>(hd_data, d_robotModel, "
                            "gravity, num_timesteps, block_dimms, thread_dimms, streams);\n}\n")


def _bind(key: str):
    """Create the reference's per-algorithm generator family for `key`."""
    def temp_mem_size(self, *a, **k):
        return self._temp_mem_size(key)

    def inner(self, use_thread_group=False, **k):
        self._gen_inner(key)

    def device(self, use_thread_group=False, **k):
        self._gen_device(key)

    def kernel(self, use_thread_group=False, single_call_timing=False, **k):
        self._gen_kernel(key)

    def host(self, mode=0):
        self._gen_host(key, mode)

    def driver(self, use_thread_group=False):
        self._gen_algo(key)

    def function_call(self, use_thread_group=False, updated_var_names=None, variant=None, **k):
        """Emit the call of <algo>_inner as this back end defines it (reference gen_<algo>_inner_function_call, e.g.
        algorithms/_inverse_dynamics.py:323-346, prints the call of ITS inner function with shared-memory pointers):
            <algo>_inner<T, NT>(o, q, qd, ..., S, gravity);
        one thread = one knot point; o / inputs point at this knot's rows, S is the thread's state base (element i at
        S[i*NT]).  updated_var_names renames arguments: keys "o", "S", "gravity" and the input names of the variant
        (q, qd, qdd, u, Minv).  variant selects an overload ("inverse_dynamics_qdd", ...); default: the first."""
        names = self._variants(key)
        if not names:
            raise ValueError(f"{key} is not generated for this robot")
        sp = self._specs[variant or names[0]]
        ren = dict(updated_var_names or {})
        args = [ren.get("o", "o")] + [ren.get(p_, p_) for p_ in sp.inner_inputs] + [ren.get("S", "S")]
        args.append(ren.get("gravity", "gravity") if sp.uses_gravity else "T(0)")
        self.gen_add_code_line(f"{sp.fn}_inner<T, NT>({', '.join(args)});")

    return temp_mem_size, inner, device, kernel, host, driver, function_call


for _key in ("joint_transforms", "inverse_dynamics", "inverse_dynamics_vaf", "direct_minv", "forward_dynamics", "inverse_dynamics_gradient",
             "forward_dynamics_gradient", "forward_dynamics_linearization", "aba", "crba", "idsva_so", "fdsva_so",
             "end_effector_pose", "end_effector_pose_gradient", "end_effector_pose_gradient_hessian"):
    _t, _i, _d, _k, _h, _drv, _fc = _bind(_key)
    setattr(GRiDCodeGenerator, f"gen_{_key}_inner_temp_mem_size", _t)
    setattr(GRiDCodeGenerator, f"gen_{_key}_inner", _i)
    setattr(GRiDCodeGenerator, f"gen_{_key}_device", _d)
    setattr(GRiDCodeGenerator, f"gen_{_key}_kernel", _k)
    setattr(GRiDCodeGenerator, f"gen_{_key}_host", _h)
    setattr(GRiDCodeGenerator, f"gen_{_key}", _drv)
    setattr(GRiDCodeGenerator, f"gen_{_key}_inner_function_call", _fc)
# names that differ in the reference (algorithms/_forward_dynamics_gradient.py:7, :107; _inverse_dynamics.py:323)
GRiDCodeGenerator.gen_forward_dynamics_gradient_inner_python = GRiDCodeGenerator.gen_forward_dynamics_gradient_inner
GRiDCodeGenerator.gen_forward_dynamics_gradient_kernel_max_temp_mem_size = \
    GRiDCodeGenerator.gen_forward_dynamics_gradient_inner_temp_mem_size
GRiDCodeGenerator.gen_inverse_dynamics_gradient_kernel_max_temp_mem_size = \
    GRiDCodeGenerator.gen_inverse_dynamics_gradient_inner_temp_mem_size
GRiDCodeGenerator.gen_inverse_dynamics_device_temp_mem_size = GRiDCodeGenerator.gen_inverse_dynamics_inner_temp_mem_size
GRiDCodeGenerator.gen_crba_device_temp_mem_size = GRiDCodeGenerator.gen_crba_inner_temp_mem_size  # reference _crba.py:214


# the reference's public code-text helpers for USER kernels (README.md:44-56): thread-strided loops, serial sections,
# barriers, per-thread selects, global <-> shared copies, compressed-column bookkeeping
from .emit import user_helpers as _uh  # noqa: E402

for _name in _uh.PUBLIC:
    setattr(GRiDCodeGenerator, _name, getattr(_uh, _name))

# ----------------------------------------------------------------- file scaffolding
from .emit import scaffold as _sc  # noqa: E402

ALGO_ORDER = ("joint_transforms", "end_effector_pose", "end_effector_pose_gradient", "end_effector_pose_gradient_hessian",
              "inverse_dynamics", "inverse_dynamics_vaf", "direct_minv", "forward_dynamics", "inverse_dynamics_gradient",
              "forward_dynamics_gradient", "forward_dynamics_linearization", "aba", "crba", "idsva_so", "fdsva_so")


def gen_add_includes(self, use_thread_group=False):
    """reference GRiDCodeGenerator.py:66-84; the runtime header is inlined so grid.cuh stays standalone."""
    self._add_block(_sc.INCLUDES)
    from .emit.core import SYNC_EVERY
    self.gen_add_code_line(f"#define GRID_LOCKSTEP {int(SYNC_EVERY > 0)}")
    with open(os.path.join(_CSRC, "grid_b200_runtime.cuh")) as f:
        rt = f.read().replace("#pragma once\n", "")
    self._add_block("// ---- inlined gridcodegenerator_b200/csrc/grid_b200_runtime.cuh ----\n" + rt)


def gen_add_gpu_err(self):
    self._add_block(_sc.GPU_ERR)
    if self.gen_print_mat:
        self._add_block(_sc.PRINT_MAT)


def _scratch_constants(self):
    """Elements of global scratch per CTA, max over algorithms: per-thread state that does not fit shared memory
    plus the lane-interleaved output scratch of the transposed-store mode (STAGED = 3)."""
    per4 = per8 = 0
    for name, sp in self._specs.items():
        b = self._body(name)
        IN = [e for _, e in sp.streams]
        out = sp.out_n(self.model.n)
        for size in (4, 8):
            NT, state_smem, staged = _hdr.choose_config(max(b.nslots, 1), IN, out, size)
            need = (0 if state_smem else NT * max(b.nslots, 1)) + (NT * out if staged == 3 else 0)
            if _hdr.has_split(sp, self.model.n):      # transposed store of the multi-tile split form
                if name not in self._sbodies:
                    self._sbodies[name] = _hdr.split_body(sp, self.model)
                g = _hdr.split_ntile(self._sbodies[name], sp, self.model.n, size, b.nslots)
                if g > 1 and os.environ.get("GRID_SPLIT_XPOSE", "0") == "1":
                    need = max(need, g * 32 * out)
            if size == 4:
                per4 = max(per4, need)
            else:
                per8 = max(per8, need)
    return per4, per8


def gen_add_constants_helpers(self):
    """reference GRiDCodeGenerator.py:86-158 (same constant names; *_DYNAMIC_SHARED_MEM_COUNT is in units of T)."""
    n = self.model.n
    r = self.robot
    lines = [f"const int NUM_JOINTS = {n};", f"const int NUM_VEL = {n};",
             f"const int NUM_EES = {r.get_total_leaf_nodes()};",
             "const int KNOTS_PER_WARP = 32;    // one lane owns one knot point",
             "const int SUGGESTED_THREADS = 32; // single-warp CTAs: warps never synchronise with each other"]
    per4, per8 = _scratch_constants(self)
    ft = self.fixed_target_name
    lines += [f"#define GRID_HAS_FIXED_TARGET {int(bool(ft))}", f'#define GRID_FIXED_TARGET_NAME "{ft}"']
    if ft:   # un-suffixed aliases of the single-target EE launchers (the C ABI binds these)
        lines += [f"#define GRID_FIXED_TARGET_POSE_LAUNCH end_effector_pose_{ft}_launch",
                  f"#define GRID_FIXED_TARGET_GRADIENT_LAUNCH end_effector_pose_gradient_{ft}_launch",
                  f"#define GRID_FIXED_TARGET_HESSIAN_LAUNCH end_effector_pose_gradient_hessian_{ft}_launch",
                  f"#define GRID_FIXED_TARGET_POSE_CFG end_effector_pose_{ft}_cfg",
                  f"#define GRID_FIXED_TARGET_GRADIENT_CFG end_effector_pose_gradient_{ft}_cfg",
                  f"#define GRID_FIXED_TARGET_HESSIAN_CFG end_effector_pose_gradient_hessian_{ft}_cfg"]
    lines += [f"#define GRID_HAS_SECOND_ORDER {int(_hdr.second_order_enabled(self.max_tree))}",
              f"#define GRID_HAS_EE_HESSIAN {int(_hdr.ee_hessian_enabled(n))}",
              f"const size_t GRID_SCRATCH_ELEMS_PER_BLOCK_F32 = {per4};",
              f"const size_t GRID_SCRATCH_ELEMS_PER_BLOCK_F64 = {per8};",
              # the scratch serves SMs x this many CTAs; one per SM when a slab would pass 1 GiB (fp32, 148 SMs)
              f"const int GRID_SCRATCH_BLOCKS_PER_SM = {1 if per4 * 296 * 4 > (1 << 30) else 2};"]
    self.gen_add_code_lines(lines)
    self._add_block(_sc.STRUCTS)
    self._add_block(_uh.mx_device_helpers())


def gen_add_shared_mem_constants(self):
    """<CODE>_DYNAMIC_SHARED_MEM_COUNT for T = float (reference GRiDCodeGenerator.py:97-110), emitted after the
    algorithm configs exist."""
    seen = {}
    for name, sp in self._specs.items():
        seen.setdefault(sp.code, []).append(f"{sp.fn}{sp.tag}_cfg<float>::SMEM_BYTES")
    for code, exprs in seen.items():
        mx = exprs[0]
        for e in exprs[1:]:
            mx = f"({mx} > {e} ? {mx} : {e})"
        self.gen_add_code_line(f"const size_t {code}_DYNAMIC_SHARED_MEM_COUNT = ({mx} + sizeof(float) - 1)/sizeof(float);")


def gen_init_close_grid(self):
    self._add_block(_sc.LIFECYCLE)


def gen_all_code(self, use_thread_group=False, include_base_inertia=False, include_homogenous_transforms=False,
                 fixed_target_name=None, output_dir="."):
    """Emit <namespace>.cuh into output_dir (reference GRiDCodeGenerator.py:309-437 writes it to the CWD).
    fixed_target_name: additionally generate the EE functions of that fixed joint (reference :310, :417)."""
    if fixed_target_name is not None and (fixed_target_name or "") != self.fixed_target_name:
        self._set_fixed_target(fixed_target_name or "")
        self._bodies, self._sbodies = {}, {}
    self.code_str = ""
    self.indent_level = 0
    notes = ["Interface (same names as the reference GRiD header):",
             "    __host__   robotModel<T> *d_robotModel = init_robotModel<T>()",
             "    __host__   cudaStream_t *streams = init_grid<T>()",
             "    __host__   gridData<T> *hd_data = init_gridData<T,NUM_TIMESTEPS>();",
             "    __host__   close_grid<T>(cudaStream_t *streams, robotModel<T> *d_robotModel, gridData<T> *hd_data)",
             "  per ALGORITHM in {inverse_dynamics, direct_minv, forward_dynamics, inverse_dynamics_gradient,",
             "                    forward_dynamics_gradient, forward_dynamics_linearization, aba, crba, idsva_so, fdsva_so}:",
             "    __device__ ALGORITHM_inner<T,NT>(T *o, const T *q, ..., T *S, const T gravity)   one thread = one knot point",
             "    __device__ ALGORITHM_device<T>(const grid_rt::TileArgs<T> &a, unsigned char *s_mem) one warp = 32 knot points",
             "    __global__ ALGORITHM_kernel<T>(T *d_out, const T *d_in, const int stride, ..., const robotModel<T> *, const T gravity, const int NUM_TIMESTEPS)",
             "    __host__   ALGORITHM<T,FLAGS...>(gridData<T> *hd_data, const robotModel<T> *d_robotModel, const T gravity, const int num_timesteps, const dim3 block_dimms, const dim3 thread_dimms, cudaStream_t *streams)",
             "    __host__   cudaError_t ALGORITHM_launch<T>(..., cudaStream_t stream)   asynchronous, no copies",
             "", "Suggested Type T is float", "",
             "Kernels are sm_100a only: compile with -gencode arch=compute_100a,code=sm_100a"]
    self.gen_add_func_doc("This instance of grid.cuh is optimized for the urdf: " + self.robot.name, notes)
    self.gen_add_code_line("#pragma once")
    self.gen_add_includes(use_thread_group)
    self.gen_add_gpu_err()
    self.gen_add_code_line("namespace " + self.file_namespace + " {", True)
    self.gen_add_constants_helpers()
    for key in ALGO_ORDER:
        if self._variants(key):
            getattr(self, "gen_" + key)(use_thread_group)
    self.gen_add_shared_mem_constants()
    self.gen_init_close_grid()
    self.gen_add_end_control_flow()
    path = os.path.join(output_dir, self.file_namespace + ".cuh")
    with open(path, "w") as f:
        f.write(self.code_str)
    return path


for _f in (gen_add_includes, gen_add_gpu_err, gen_add_constants_helpers, gen_add_shared_mem_constants,
           gen_init_close_grid, gen_all_code):
    setattr(GRiDCodeGenerator, _f.__name__, _f)
