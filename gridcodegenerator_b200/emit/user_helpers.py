"""The public code-text helpers of the reference's generator class (reference README.md:44-56; behaviour as in
helpers/_code_generation_helpers.py:34-190 and helpers/_topology_helpers.py:515-542).

The algorithm emitters of this back end do not use them (they print straight-line per-thread bodies through
emit/core.py); they exist for USER code that drives the generator's code buffer the way the reference's README shows:
a block-per-knot kernel assembled from thread-strided loops, serial sections, barriers and global <-> shared copies.
Every function appends CUDA text to `gen.code_str` through gen_add_code_line and keeps `gen.indent_level` consistent.
tests/test_user_helpers.py assembles a kernel from them, compiles it with nvcc and (on a GPU) runs it.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple


# ------------------------------------------------------------------ control flow
def _rank_and_size(use_thread_group: bool, block_level: bool) -> Tuple[str, str]:
    """(linear index of this thread or block, number of them) as C expressions."""
    if block_level:
        if use_thread_group:
            raise NotImplementedError("block-level loops over a cooperative thread group are not defined by the reference either")
        return "blockIdx.x + blockIdx.y*gridDim.x", "gridDim.x*gridDim.y"
    if use_thread_group:
        return "tgrp.thread_rank()", "tgrp.size()"
    return "threadIdx.x + threadIdx.y*blockDim.x", "blockDim.x*blockDim.y"


def gen_add_parallel_loop(self, var_name, max_val, use_thread_group=False, block_level=False):
    """Open `for (var = my rank; var < max_val; var += ranks)` over the threads of the block (or of `tgrp`, or over
    the blocks of the grid with block_level).  Close it with gen_add_end_control_flow()."""
    first, step = _rank_and_size(use_thread_group, block_level)
    self.gen_add_code_line(f"for(int {var_name} = {first}; {var_name} < {max_val}; {var_name} += {step}){{", True)


def gen_add_serial_ops(self, use_thread_group=False):
    """Open a section that only one thread of the block (or group) executes."""
    cond = "tgrp.thread_rank() == 0" if use_thread_group else "threadIdx.x == 0 && threadIdx.y == 0"
    self.gen_add_code_line(f"if({cond}){{", True)


def gen_add_sync(self, use_thread_group=False):
    self.gen_add_code_line("tgrp.sync();" if use_thread_group else "__syncthreads();")


def gen_add_debug_print_code_lines(self, print_code_string_arr, use_thread_group=False):
    """The given statements (printf calls) executed by one thread between two barriers."""
    gen_add_sync(self, use_thread_group)
    gen_add_serial_ops(self, use_thread_group)
    for line in print_code_string_arr:
        self.gen_add_code_line(line)
    self.gen_add_end_control_flow()
    gen_add_sync(self, use_thread_group)


def gen_add_debug_print_code_line(self, print_code_string, use_thread_group=False):
    gen_add_debug_print_code_lines(self, [print_code_string], use_thread_group)


# ------------------------------------------------------------------ small expression builders
def gen_static_array_ind_2d(self, col, row, col_stride=6):
    return col_stride * col + row


def gen_static_array_ind_3d(self, ind, col, row, ind_stride=36, col_stride=6):
    return ind_stride * ind + col_stride * col + row


def _joined(var_name: str, option_list: Sequence[str], op: str, glue: str) -> str:
    tests = [f"({var_name} {op} {o})" for o in option_list]
    return tests[0] if len(tests) == 1 else "(" + glue.join(tests) + ")"


def gen_var_in_list(self, var_name, option_list):
    return _joined(var_name, option_list, "==", " || ")


def gen_var_not_in_list(self, var_name, option_list):
    return _joined(var_name, option_list, "!=", " && ")


def _flip(comparator: str) -> str:
    """The comparator of the complementary test: < -> >=, <= -> >, > -> <=, >= -> <, == -> ==."""
    return {"<": ">=", "<=": ">", ">": "<=", ">=": "<", "==": "=="}[comparator]


def gen_add_multi_threaded_select(self, loop_counter, comparator, counts, select_tuples, USE_NON_BRANCH_ALWAYS=False):
    """Per-thread selection of values by ranges of a loop counter.

    counts = [c0, c1, ...] are thresholds tested in order with `comparator` (e.g. ind < c0, else ind < c1, ..., the last
    entry is the catch-all); select_tuples = [(type, name, [value per range]), ...].  `type` None assigns to an existing
    variable; a type containing "|" wraps the name ("pre|post" -> pre name) post).  Several variables are set in an
    if / else-if chain; a single one (or USE_NON_BRANCH_ALWAYS) by a branch-free sum of (predicate * value) terms."""
    decls: List[str] = []
    for dst_type, dst_var, _ in select_tuples:
        if dst_type is None:
            decls.append(dst_var)
        elif "|" in dst_type:
            pre, post = dst_type.split("|")[:2]
            decls.append(f"{pre}{dst_var}){post}")
        else:
            decls.append(f"{dst_type} {dst_var}")
    n = len(counts)
    if len(select_tuples) > 1 and not USE_NON_BRANCH_ALWAYS:
        self.gen_add_code_line("// branch to get pointer locations")
        self.gen_add_code_line("; ".join(decls) + ";")
        for i in range(n):
            if i == 0:
                head = f"     if ({loop_counter} {comparator} {counts[i]}){{ "
            elif i < n - 1:
                head = f"else if ({loop_counter} {comparator} {counts[i]}){{ "
            else:
                head = "else              { "
            body = "".join(f"{name} = {vals[i]}; " for _, name, vals in select_tuples)
            self.gen_add_code_line(head + body + "}")
        return
    self.gen_add_code_line("// non-branching pointer selector")
    inv = _flip(comparator)
    for decl, (_, _, vals) in zip(decls, select_tuples):
        terms = []
        for i in range(n):
            if comparator == "==" or i == 0:
                pred = f"({loop_counter} {comparator} {counts[i]})"
            elif i < n - 1:
                pred = f"({loop_counter} {comparator} {counts[i]} && {loop_counter} {inv} {counts[i - 1]})"
            else:
                pred = f"({loop_counter} {inv} {counts[i - 1]})"
            terms.append(f"{pred} * {vals[i]}")
        self.gen_add_code_line(f"{decl} = " + " + ".join(terms) + ";")


# ------------------------------------------------------------------ global <-> shared copies of a block-per-knot kernel
def _copy_in(self, name: str, amount, use_thread_group: bool, per_knot_stride=None):
    src = f"d_{name}"
    if per_knot_stride is not None:
        self.gen_add_code_line(f"const T *d_{name}_k = &d_{name}[k*{per_knot_stride}];")
        src = f"d_{name}_k"
    gen_add_parallel_loop(self, "ind", amount, use_thread_group)
    self.gen_add_code_line(f"s_{name}[ind] = {src}[ind];")
    self.gen_add_end_control_flow()


def gen_kernel_load_inputs(self, name, stride, amount, use_thread_group=False,
                           name2=None, stride2=1, amount2=1, name3=None, stride3=1, amount3=1):
    """Knot k's slice of up to three global arrays d_<name> (stride elements apart) -> shared arrays s_<name>."""
    self.gen_add_code_line("// load to shared mem")
    for nm, st, am in ((name, stride, amount), (name2, stride2, amount2), (name3, stride3, amount3)):
        if nm is not None:
            _copy_in(self, nm, str(am), use_thread_group, str(st))
    gen_add_sync(self, use_thread_group)


def gen_kernel_load_inputs_single_timing(self, name, amount, use_thread_group=False,
                                         name2=None, amount2=1, name3=None, amount3=1):
    """The same without the per-knot offset (the reference's single-knot timing kernels)."""
    self.gen_add_code_line("// load to shared mem")
    for nm, am in ((name, amount), (name2, amount2), (name3, amount3)):
        if nm is not None:
            _copy_in(self, nm, str(am), use_thread_group)
    gen_add_sync(self, use_thread_group)


def gen_kernel_save_result(self, store_to_name, stride, amount, use_thread_group=False, load_from_name=None):
    """Shared array (s_<store_to_name> unless load_from_name is given) -> knot k's slice of d_<store_to_name>."""
    src = load_from_name if load_from_name is not None else f"s_{store_to_name}"
    self.gen_add_code_line("// save down to global")
    self.gen_add_code_line(f"T *d_{store_to_name}_k = &d_{store_to_name}[k*{stride}];")
    gen_add_parallel_loop(self, "ind", str(amount), use_thread_group)
    self.gen_add_code_line(f"d_{store_to_name}_k[ind] = {src}[ind];")
    self.gen_add_end_control_flow()
    gen_add_sync(self, use_thread_group)


def gen_kernel_save_result_single_timing(self, store_to_name, amount, use_thread_group=False, load_from_name=None):
    src = load_from_name if load_from_name is not None else f"s_{store_to_name}"
    self.gen_add_code_line("// save down to global")
    gen_add_parallel_loop(self, "ind", str(amount), use_thread_group)
    self.gen_add_code_line(f"d_{store_to_name}[ind] = {src}[ind];")
    self.gen_add_end_control_flow()
    gen_add_sync(self, use_thread_group)


# ------------------------------------------------------------------ topology bookkeeping (pure Python)
def gen_topology_sparsity_helpers_python(self, INIT_MODE=False):
    """Column bookkeeping of the reference's compressed gradient storage (joint j keeps its ancestors + itself for
    dv, da and its ancestors + subtree for df).  INIT_MODE: the four per-joint tables as lists of str; otherwise
    (dva_cols_per_partial, dva_cols_per_jid, running_sum_dva_cols_per_jid,
     df_cols_per_partial, df_cols_per_jid, running_sum_df_cols_per_jid, df_col_that_is_jid)."""
    r = self.robot
    nj = r.get_num_joints()
    anc = [len(r.get_ancestors_by_id(j)) for j in range(nj)]
    sub = [len(r.get_subtree_by_id(j)) for j in range(nj)]
    anc_before = [sum(anc[:j]) for j in range(nj + 1)]          # one extra entry: loops test `< jid + 1`
    sub_before = [sum(sub[:j]) for j in range(nj)]
    if INIT_MODE:
        return [str(v) for v in anc], [str(v) for v in sub], [str(v) for v in anc_before], [str(v) for v in sub_before]
    dva_cols_per_partial = r.get_total_ancestor_count() + nj
    df_cols_per_partial = r.get_total_ancestor_count() + r.get_total_subtree_count()
    dva_cols_per_jid = [a + 1 for a in anc]
    df_cols_per_jid = [a + s for a, s in zip(anc, sub)]
    running_dva = [anc_before[j] + j for j in range(nj + 1)]
    running_df = [anc_before[j] + sub_before[j] for j in range(nj)]
    return dva_cols_per_partial, dva_cols_per_jid, running_dva, df_cols_per_partial, df_cols_per_jid, running_df, anc


# ------------------------------------------------------------------ motion cross-product columns for user kernels
def mx_device_helpers() -> str:
    """C++ text of the reference's device helpers mx<S>[_peq][_scaled]<T>(s_vecX, s_vec[, alpha]) and
    mxX[_peq][_scaled]<T>(s_vecX, s_vec[, alpha], S_ind) (reference helpers/_spatial_algebra_helpers.py:197-255):
    s_vecX (=, +=) [alpha *] column S of the motion cross-product matrix of the 6-vector s_vec.  The algorithm bodies
    of this back end fold these products at emit time (spatial.mxS); the functions exist for user kernels and for
    gen_mx_func_call_for_cpp.  Generated from the same table (spatial.MXS_TABLE)."""
    from .spatial import MXS_TABLE
    out = ["// ---- motion cross-product columns (user-kernel helpers; reference names and argument lists)"]
    variants = (("", "=", False), ("_peq", "+=", False), ("_scaled", "=", True), ("_peq_scaled", "+=", True))
    for S in range(6):
        for suffix, op, scaled in variants:
            args = "T *s_vecX, const T *s_vec" + (", const T alpha" if scaled else "")
            body = []
            for i in range(6):
                if i in MXS_TABLE[S]:
                    sg, k = MXS_TABLE[S][i]
                    val = ("-" if sg < 0 else "") + f"s_vec[{k}]" + ("*alpha" if scaled else "")
                elif op == "+=":
                    continue
                else:
                    val = "static_cast<T>(0)"
                body.append(f"s_vecX[{i}] {op} {val};")
            out.append(f"template <typename T>\n__host__ __device__ __forceinline__\nvoid mx{S}{suffix}({args}) {{ " + " ".join(body) + " }")
    for suffix, _, scaled in variants:
        args = "T *s_vecX, const T *s_vec" + (", const T alpha" if scaled else "") + ", const int S_ind"
        call = "s_vecX, s_vec" + (", alpha" if scaled else "")
        cases = " ".join(f"case {S}: mx{S}{suffix}<T>({call}); break;" for S in range(6))
        out.append(f"template <typename T>\n__host__ __device__ __forceinline__\nvoid mxX{suffix}({args}) {{ switch (S_ind) {{ {cases} }} }}")
    return "\n".join(out) + "\n"


def gen_mx_func_call_for_cpp(self, inds=None, PEQ_FLAG=False, SCALE_FLAG=False, updated_var_names=None):
    """Emit the call of the motion cross-product helper for the joints `inds` (default: all): mx<S> when they share
    one motion-subspace index S, mxX with the run-time S_ind otherwise.  updated_var_names renames the arguments
    (keys S_ind_name, s_dst_name, s_src_name, s_scale_name)."""
    names = dict(S_ind_name="S_ind", s_dst_name="s_dst", s_src_name="s_src", s_scale_name="s_scale")
    names.update(updated_var_names or {})
    if inds is None:
        inds = list(range(self.robot.get_num_pos()))
    same = self.robot.are_Ss_identical(inds)
    which = str(int(list(self.robot.get_S_by_id(inds[0])).index(1))) if same else "X"
    fn = "mx" + which + ("_peq" if PEQ_FLAG else "") + ("_scaled" if SCALE_FLAG else "") + "<T>"
    args = [names["s_dst_name"], names["s_src_name"]]
    if SCALE_FLAG:
        args.append(names["s_scale_name"])
    if not same:
        args.append(names["S_ind_name"])
    self.gen_add_code_line(fn + "(" + ", ".join(args) + ");")


PUBLIC = ("gen_add_parallel_loop", "gen_add_serial_ops", "gen_add_sync", "gen_add_debug_print_code_line",
          "gen_add_debug_print_code_lines", "gen_static_array_ind_2d", "gen_static_array_ind_3d", "gen_var_in_list",
          "gen_var_not_in_list", "gen_add_multi_threaded_select", "gen_kernel_load_inputs",
          "gen_kernel_load_inputs_single_timing", "gen_kernel_save_result", "gen_kernel_save_result_single_timing",
          "gen_topology_sparsity_helpers_python", "gen_mx_func_call_for_cpp")
