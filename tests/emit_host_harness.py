"""Test helper: compile the EMITTED per-thread bodies (<algo>_inner) as plain host C++ (T = double,
NT = 1) so the emitter's algebra can be checked against the golden vectors without a GPU.
Not a product path: the product only ever runs these bodies inside the sm_100a tile kernels."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np

from gridcodegenerator_b200.GRiDCodeGenerator import GRiDCodeGenerator
from gridcodegenerator_b200.emit import header as hdr

PRELUDE = """
#include <math.h>
#include <stddef.h>
#define __device__
#define __forceinline__ inline
#define __restrict__
#define GRID_FENCE() do { } while (0)
#define GRID_SYNC() do { } while (0)
#define GRID_LAUNDER(x) (x)
#define GRID_BEFORE_FIRST_OUTPUT() do { } while (0)
namespace grid_rt {
static inline void grid_sincos(double x, double* s, double* c) { *s = sin(x); *c = cos(x); }
struct double_pair { double x, y; };
template <typename T> using pair_t = double_pair;
template <typename T> static inline double_pair mkp(T a, T b) { double_pair r; r.x = a; r.y = b; return r; }
template <typename T> static inline double_pair pbc(T s) { return mkp<T>(s, s); }
static inline double_pair pfma(double_pair a, double s, double_pair c) { return mkp<double>(a.x * s + c.x, a.y * s + c.y); }
static inline double_pair pmul(double_pair a, double s) { return mkp<double>(a.x * s, a.y * s); }
static inline double_pair padd(double_pair a, double_pair b) { return mkp<double>(a.x + b.x, a.y + b.y); }
static inline double_pair psub(double_pair a, double_pair b) { return mkp<double>(a.x - b.x, a.y - b.y); }
static inline double_pair padds(double_pair a, double s) { return mkp<double>(a.x + s, a.y + s); }
static inline double_pair pneg(double_pair a) { return mkp<double>(-a.x, -a.y); }
}
"""


def build_host_inners(robot, workdir=None, second_order=False):
    ft = robot.fixed_joints[0].get_name() if getattr(robot, "fixed_joints", None) else ""
    g = GRiDCodeGenerator(robot, fixed_target_name=ft)     # like runtime.build_library: first fixed joint as the target
    n = g.model.n
    src = [PRELUDE]
    wrappers = []
    for name, sp in g._specs.items():
        if name in ("idsva_so", "fdsva_so") and not second_order:
            continue          # ~n^3 code: compiled for the host only by the test that checks them
        body = g._body(name)
        src.append(hdr.emit_inner(sp, body))
        ins = ", ".join(f"const double* {p}" for p in sp.inner_inputs)
        call = ", ".join(sp.inner_inputs)
        wrappers.append(
            f'extern "C" void host_{name}(double* o, {ins}, double gravity) {{\n'
            f"    static thread_local double S[{max(body.nslots, 1)}];\n"
            f"    {sp.fn}_inner<double, 1>(o, {call}, S, gravity);\n}}\n")
        if hdr.has_split(sp, n):
            # the split form on one host thread: phase A, then every column function, on the same state array
            sb = hdr.split_body(sp, g.model)
            stem = f"{sp.fn}{sp.tag}_split"
            pre = "    using grid_rt::grid_sincos;\n    typedef grid_rt::pair_t<T> P;\n"
            params = "".join(f"const T* {p}, " for p in sp.inner_inputs)
            for p_, (pcode, _) in enumerate(sb.parts):
                src.append(f"template <typename T, int NT, int OS = 1>\nvoid {stem}_phaseA{p_}(T* o, {params}T* S, const T gravity) {{\n"
                           + pre + pcode + "}\n")
            for c, code, _ in sb.columns:
                src.append(f"template <typename T, int NT, int OS = 1>\nvoid {stem}_col{c}(T* o, T* S) {{\n" + pre + code + "}\n")
            cols = "".join(f"    {stem}_col{c}<double, 1>(o, S);\n" for c, _, _ in sb.columns)
            wrappers.append(
                f'extern "C" void host_split_{name}(double* o, {ins}, double gravity) {{\n'
                f"    static thread_local double S[{max(sb.nslots, 1)}];\n"
                + "".join(f"    {stem}_phaseA{p_}<double, 1>(o, {call}, S, gravity);\n" for p_ in range(len(sb.parts)))
                + f"{cols}}}\n")
    # the reference-signature helper X_single_thread (algorithms/_eepose_gradient_hessian.py:1226-1298)
    from gridcodegenerator_b200.emit import algos as _algos
    src.append("template <typename T>\nvoid X_single_thread(T *s_jointXforms, T *s_XmatsHom, const T *s_q, int tid) {\n"
               "    using grid_rt::grid_sincos;\n    constexpr int OS = 1;\n" + _algos.code_X_single_thread(g.model) + "}\n")
    wrappers.append('extern "C" void host_X_single_thread(double* jx, double* xh, const double* q, int tid) {\n'
                    "    X_single_thread<double>(jx, xh, q, tid);\n}\n")
    text = "\n".join(src) + "\n".join(wrappers)
    workdir = workdir or tempfile.mkdtemp(prefix="grid_host_")
    cpp = os.path.join(workdir, f"inner_{robot.name}.cpp")
    so = os.path.join(workdir, f"inner_{robot.name}.so")
    with open(cpp, "w") as f:
        f.write(text)
    subprocess.check_call(["g++", "-O0", "-shared", "-fPIC", "-o", so, cpp])
    lib = ctypes.CDLL(so)
    return lib, g


def call(lib, g, name, out_n, *inputs, gravity=9.81):
    dp = ctypes.POINTER(ctypes.c_double)
    out = np.zeros(out_n)
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in inputs]
    fn = getattr(lib, "host_" + name)
    fn.restype = None
    fn(out.ctypes.data_as(dp), *[a.ctypes.data_as(dp) for a in arrs], ctypes.c_double(gravity))
    return out
