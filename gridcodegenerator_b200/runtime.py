"""Build + load the per-robot C-ABI library and call it on torch device tensors.

Python host side of the drop-in boundary (include/grid_b200.h).  torch is used only for device memory
and streams; all arithmetic happens in the emitted sm_100a kernels.  There is NO CPU or eager fallback:
if the library cannot be built/loaded or no CUDA device is present, calls raise.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import shutil
import subprocess
from typing import Optional

from .GRiDCodeGenerator import GRiDCodeGenerator
from .robot import load_robot

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
BUILD_DIR = os.environ.get("GRID_BUILD_DIR") or os.path.join(_PKG, "_build")   # override: A/B variant builds
CSRC = os.path.join(_PKG, "csrc")
INCLUDE = os.path.join(_ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-shared"] + os.environ.get("GRID_NVCC_EXTRA", "").split()

ALGO_INDEX = {"inverse_dynamics": 0, "inverse_dynamics_qdd": 1, "direct_minv": 2, "forward_dynamics": 3,
              "inverse_dynamics_gradient": 4, "inverse_dynamics_gradient_qdd": 5, "forward_dynamics_gradient": 6,
              "forward_dynamics_gradient_qdd_minv": 7, "aba": 8, "crba": 9, "idsva_so": 10, "fdsva_so": 11, "end_effector_pose": 12,
              "end_effector_pose_gradient": 13, "end_effector_pose_gradient_hessian": 14,
              "forward_dynamics_linearization": 15}


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the GRiD B200 library cannot be built")


def _source_digest(header_text: str) -> str:
    h = hashlib.sha256()
    h.update(header_text.encode())
    for fn in ("grid_capi.cu", "grid_noalloc.cu"):
        with open(os.path.join(CSRC, fn), "rb") as f:
            h.update(f.read())
    with open(os.path.join(INCLUDE, "grid_b200.h"), "rb") as f:
        h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def _package_digest() -> str:
    """Digest of everything a library is generated from EXCEPT the robot description: emitter, robot layer, CUDA
    runtime, C ABI, compile flags.  Cheap (no emission), so GRiDLibrary checks it at load time and refuses a stale
    binary; build_library() additionally keys on the emitted header itself."""
    h = hashlib.sha256()
    files = []
    for sub in ("emit", "robot", "csrc"):
        for dp, _, fns in os.walk(os.path.join(_PKG, sub)):
            files += [os.path.join(dp, f) for f in fns if f.endswith((".py", ".cu", ".cuh", ".urdf"))]
    files += [os.path.join(_PKG, "GRiDCodeGenerator.py"), os.path.join(INCLUDE, "grid_b200.h")]
    for fn in sorted(files):
        h.update(os.path.relpath(fn, _ROOT).encode())
        with open(fn, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def library_path(robot_name: str) -> str:
    return os.path.join(BUILD_DIR, robot_name, f"libgrid_{robot_name}.so")


# translation units of one library: (GRID_PARTS bitmask, GRID_DTYPES) -- see csrc/grid_capi.cu
CAPI_PARTS = [(0x01, 1), (0x01, 2), (0x02, 1), (0x02, 2), (0x04, 1), (0x04, 2), (0x08, 1), (0x08, 2),
              (0x10, 1), (0x10, 2), (0x20, 1), (0x20, 2), (0x40, 1), (0x40, 2), (0x80, 3), (0x100, 1), (0x100, 2)]


def _compile_part(name: str, out_dir: str, part: int, dtypes: int, verbose: bool) -> str:
    obj = os.path.join(out_dir, f"capi_{part:03x}_{dtypes}.o")
    cmd = [_nvcc()] + [f for f in NVCC_FLAGS if f != "-shared"] + [
        "-c", f"-I{out_dir}", f"-I{INCLUDE}", f'-DGRID_ROBOT_NAME="{name}"', f"-DGRID_PARTS={part}",
        f"-DGRID_DTYPES={dtypes}", os.path.join(CSRC, "grid_capi.cu"), "-o", obj]
    if part in (0x20, 0x40) and os.environ.get("GRID_SO_PTXAS_O1", "1") == "1":
        # second-order bodies are ~1 MB of straight-line code: ptxas -O3 needs ~10 min per translation unit on them,
        # -O1 under 2 min
        cmd[1:1] = ["-Xptxas", "-O1"]
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for robot {name} part {part:#x}/{dtypes}:\n{res.stdout[-4000:]}")
    if verbose:
        with open(os.path.join(out_dir, f"ptxas_{part:03x}_{dtypes}.log"), "w") as f:
            f.write(res.stdout)
    return obj


def _compile_noalloc(name: str, out_dir: str, verbose: bool) -> str:
    """csrc/grid_noalloc.cu: the allocation-free grad-FD entry point with its static device scratch (one more
    translation unit of the same library)."""
    obj = os.path.join(out_dir, "noalloc.o")
    cmd = [_nvcc()] + [f for f in NVCC_FLAGS if f != "-shared"] + [
        "-c", f"-I{out_dir}", f"-I{INCLUDE}", os.path.join(CSRC, "grid_noalloc.cu"), "-o", obj]
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for robot {name}, grid_noalloc.cu:\n{res.stdout[-4000:]}")
    return obj


def build_library(robot, force: bool = False, verbose: bool = False, pool=None) -> str:
    """Emit grid.cuh for `robot` (a robot object or the name of a shipped robot) and compile the C-ABI
    library for sm_100a.  Returns the .so path.  Up-to-date builds are reused (digest of all sources).
    The library's translation units (CAPI_PARTS) are compiled in parallel on `pool` (a
    concurrent.futures executor shared between robots) or a private one."""
    from concurrent.futures import ThreadPoolExecutor
    if isinstance(robot, str):
        robot = load_robot(robot)
    name = robot.name
    out_dir = os.path.join(BUILD_DIR, name)
    os.makedirs(out_dir, exist_ok=True)
    # fixed kinematic target of the library's single-target EE entry points: the robot's first fixed joint, if any
    ft = os.environ.get("GRID_FIXED_TARGET")
    if ft is None:
        ft = robot.fixed_joints[0].get_name() if getattr(robot, "fixed_joints", None) else ""
    gen = GRiDCodeGenerator(robot, fixed_target_name=ft)
    gen.gen_all_code(output_dir=out_dir)
    digest = _source_digest(gen.code_str)
    pkg_digest = _package_digest()          # taken BEFORE compiling: edits made during a long build must not be stamped
    so = library_path(name)
    stamp = os.path.join(out_dir, "digest.txt")
    if not force and os.path.exists(so) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        with open(os.path.join(out_dir, "package_digest.txt"), "w") as f:     # same header => binary still current
            f.write(pkg_digest)
        return so
    own = pool is None
    if own:
        pool = ThreadPoolExecutor(max_workers=min(len(CAPI_PARTS), os.cpu_count() or 1))
    try:
        parts = CAPI_PARTS
        only = os.environ.get("GRID_BUILD_ONLY")       # A/B variant builds: "0x01:1,0x04:1" = just these (part, dtype) units
        if only:
            parts = [(int(a, 0), int(b)) for a, b in (x.split(":") for x in only.split(","))]
        futs = [pool.submit(_compile_part, name, out_dir, part, dt, verbose) for part, dt in parts]
        if not only:
            futs.append(pool.submit(_compile_noalloc, name, out_dir, verbose))
        objs = [f.result() for f in futs]
        if only:
            # the core unit's grid_kernel_info references one (hidden) info function per unit: weak stand-ins for
            # the units that were left out
            stub = os.path.join(out_dir, "partial_stub.cpp")
            names = ["core", "ee", "id_du", "aba_crba", "idsva", "fd_du_qm", "fd_lin", "fdsva", "fd_du"]
            with open(stub, "w") as f:
                for nm in names:
                    for dt in ("f32", "f64"):
                        f.write(f'__attribute__((weak, visibility("hidden"))) int grid_capi_info_{nm}_{dt}(int, int) {{ return -1; }}\n')
            subprocess.check_call(["g++", "-fPIC", "-c", stub, "-o", stub[:-4] + ".o"])
            objs.append(stub[:-4] + ".o")
    finally:
        if own:
            pool.shutdown()
    cmd = [_nvcc(), "-shared", "-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", so]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed for robot {name}:\n{res.stdout[-4000:]}")
    for o in objs:
        os.remove(o)
    with open(stamp, "w") as f:
        f.write(digest)
    with open(os.path.join(out_dir, "package_digest.txt"), "w") as f:
        f.write(pkg_digest)
    return so


def host_tier_check_path(robot_name: str) -> str:
    return os.path.join(BUILD_DIR, robot_name, "host_tier_check")


def build_host_tier_check(robot, knots: int = 1000, force: bool = False) -> str:
    """Compile csrc/host_tier_check.cu -- a user program against the emitted grid.cuh's host tier -- for `robot`
    (build_library must have emitted the header).  tests/test_gpu_host_tier.py runs it and compares its results
    with the C ABI.  Same code-generation flags as the library, so the kernels are the same machine code."""
    name = robot if isinstance(robot, str) else robot.name
    out_dir = os.path.join(BUILD_DIR, name)
    exe = host_tier_check_path(name)
    src = os.path.join(CSRC, "host_tier_check.cu")
    hdr = os.path.join(out_dir, "grid.cuh")
    h = hashlib.sha256()
    for fn in (src, hdr):
        with open(fn, "rb") as f:
            h.update(f.read())
    h.update(f"{knots} {' '.join(NVCC_FLAGS)}".encode())
    stamp = os.path.join(out_dir, "host_tier_digest.txt")
    if not force and os.path.exists(exe) and os.path.exists(stamp) and open(stamp).read().strip() == h.hexdigest():
        return exe
    cmd = [_nvcc()] + [f for f in NVCC_FLAGS if f not in ("-shared", "-Xcompiler", "-fPIC")] + [
        f"-I{out_dir}", f"-DHOST_TIER_KNOTS={knots}", src, "-o", exe]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for host_tier_check ({name}):\n{res.stdout[-4000:]}")
    with open(stamp, "w") as f:
        f.write(h.hexdigest())
    return exe


_f32p = ctypes.c_void_p


class GRiDLibrary:
    """ctypes binding of libgrid_<robot>.so.  Tensors in, tensors out; asynchronous on the current torch stream."""

    def __init__(self, robot, build: bool = True):
        name = robot if isinstance(robot, str) else robot.name
        so = library_path(name)
        if not os.path.exists(so):
            if not build:
                raise RuntimeError(f"{so} is missing and build=False: no fallback path exists")
            so = build_library(robot)
        # a binary built from other sources than the ones in the tree must not be used silently
        stamp = os.path.join(os.path.dirname(so), "package_digest.txt")
        if os.environ.get("GRID_SKIP_DIGEST_CHECK") != "1":
            have = open(stamp).read().strip() if os.path.exists(stamp) else None
            if have != _package_digest():
                if not build:
                    raise RuntimeError(f"{so} is stale (built from different emitter / runtime sources) and build=False")
                so = build_library(robot)
        self.path = so
        self.lib = ctypes.CDLL(so)
        L = self.lib
        L.grid_robot_name.restype = ctypes.c_char_p
        L.grid_last_error.restype = ctypes.c_char_p
        L.grid_kernel_info.restype = ctypes.c_long
        L.grid_kernel_info.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int]
        self.n = int(L.grid_num_joints())
        self.name = L.grid_robot_name().decode()
        vp, ci = ctypes.c_void_p, ctypes.c_int
        for suf, g in (("f32", ctypes.c_float), ("f64", ctypes.c_double)):
            getattr(L, f"grid_inverse_dynamics_{suf}").argtypes = [vp, ci, vp, vp, ci, g, vp]
            getattr(L, f"grid_inverse_dynamics_vaf_{suf}").argtypes = [vp, ci, vp, vp, ci, g, vp]
            getattr(L, f"grid_direct_minv_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_forward_dynamics_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_inverse_dynamics_gradient_{suf}").argtypes = [vp, ci, vp, vp, ci, g, vp]
            getattr(L, f"grid_forward_dynamics_gradient_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_forward_dynamics_gradient_qdd_minv_{suf}").argtypes = [vp, ci, vp, vp, vp, ci, g, vp]
            getattr(L, f"grid_forward_dynamics_gradient_host_{suf}").argtypes = [vp, vp, ci, g]
            getattr(L, f"grid_forward_dynamics_linearization_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_forward_dynamics_linearization_host_{suf}").argtypes = [vp, vp, ci, g]
            getattr(L, f"grid_aba_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_crba_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_idsva_so_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_fdsva_so_{suf}").argtypes = [vp, ci, vp, ci, g, vp]
            getattr(L, f"grid_end_effector_pose_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_end_effector_pose_gradient_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_end_effector_pose_gradient_hessian_{suf}").argtypes = [vp, ci, vp, vp, ci, vp]
            getattr(L, f"grid_joint_transforms_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_fixed_target_pose_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_fixed_target_pose_gradient_{suf}").argtypes = [vp, ci, vp, ci, vp]
            getattr(L, f"grid_fixed_target_pose_gradient_hessian_{suf}").argtypes = [vp, ci, vp, vp, ci, vp]
            getattr(L, f"grid_direct_minv_host_{suf}").argtypes = [vp, vp, ci]
            getattr(L, f"grid_crba_host_{suf}").argtypes = [vp, vp, ci]
            getattr(L, f"grid_aba_host_{suf}").argtypes = [vp, vp, ci, g]
            getattr(L, f"grid_inverse_dynamics_gradient_host_{suf}").argtypes = [vp, vp, vp, ci, g]
            getattr(L, f"grid_idsva_so_host_{suf}").argtypes = [vp, vp, ci, g]
            getattr(L, f"grid_fdsva_so_host_{suf}").argtypes = [vp, vp, ci, g]
        L.grid_fixed_target_name.restype = ctypes.c_char_p
        self.fixed_target = L.grid_fixed_target_name().decode()
        self.has_second_order = bool(L.grid_has_second_order())
        self.has_ee_hessian = bool(L.grid_has_ee_hessian())
        self.num_ees = int(L.grid_num_ees())
        L.grid_forward_dynamics_host_f32.argtypes = [vp, vp, ci, ctypes.c_float]
        L.grid_inverse_dynamics_host_f32.argtypes = [vp, vp, ci, ctypes.c_float]

    # ------------------------------------------------------------------ helpers
    def _check(self, rc: int):
        if rc != 0:
            raise RuntimeError(f"GRiD library error {rc}: {self.lib.grid_last_error().decode()}")

    @staticmethod
    def _suf(t) -> str:
        import torch
        if t.dtype == torch.float32:
            return "f32"
        if t.dtype == torch.float64:
            return "f64"
        raise TypeError(f"unsupported dtype {t.dtype}")

    @staticmethod
    def _dev(t, cols_min: int):
        if not t.is_cuda:
            raise RuntimeError("device entry points need CUDA tensors (no CPU fallback exists)")
        if t.dim() != 2 or t.stride(1) != 1 or t.shape[1] < cols_min:
            raise ValueError(f"expected a (B, >={cols_min}) row-major tensor, got {tuple(t.shape)} strides {t.stride()}")
        # a one-row tensor has no meaningful row stride (torch keeps whatever the parent had)
        stride = int(t.stride(0)) if t.shape[0] > 1 else int(t.shape[1])
        return t.data_ptr(), stride, int(t.shape[0])

    @staticmethod
    def _out(out, like, B: int, m: int):
        """Allocate the (B, m) result, or validate a caller-supplied one: the kernels write B*m contiguous elements."""
        import torch
        if out is None:
            return torch.empty((B, m), dtype=like.dtype, device=like.device)
        if not out.is_cuda or out.device != like.device or out.dtype != like.dtype:
            raise ValueError(f"out must be a {like.dtype} tensor on {like.device}")
        if tuple(out.shape) != (B, m) or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous ({B}, {m}) tensor, got {tuple(out.shape)} strides {out.stride()}")
        return out

    @staticmethod
    def _stream():
        import torch
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def kernel_info(self, algo: str, f64: bool = False) -> dict:
        i = ALGO_INDEX[algo]
        keys = ("smem_bytes", "threads", "state_slots", "state_in_smem", "staged_io")
        return {k: int(self.lib.grid_kernel_info(i, int(f64), w)) for w, k in enumerate(keys)}

    def persistent_blocks(self, algo: str, f64: bool = False) -> int:
        return int(self.lib.grid_kernel_info(ALGO_INDEX[algo], int(f64), 5))

    # ------------------------------------------------------- device entry points
    def inverse_dynamics(self, q_qd, qdd=None, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd, 2 * n)
        suf = self._suf(q_qd)
        out = self._out(out, q_qd, B, n)
        pq = None
        if qdd is not None:
            pq, sq, Bq = self._dev(qdd, n)
            if sq != n or Bq != B:
                raise ValueError("qdd must be a dense (B, n) tensor")
        with torch.cuda.device(q_qd.device):
            self._check(getattr(self.lib, f"grid_inverse_dynamics_{suf}")(p, stride, pq, out.data_ptr(), B, gravity, self._stream()))
        return out

    def inverse_dynamics_vaf(self, q_qd, qdd=None, gravity: float = 9.81, out=None):
        """RNEA's per-joint v, a, f: (B, 18 n) = [v | a | f], the 6-vector of joint j at 6 j of each block."""
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd, 2 * n)
        suf = self._suf(q_qd)
        out = self._out(out, q_qd, B, 18 * n)
        pq = None
        if qdd is not None:
            pq, sq, Bq = self._dev(qdd, n)
            if sq != n or Bq != B:
                raise ValueError("qdd must be a dense (B, n) tensor")
        with torch.cuda.device(q_qd.device):
            self._check(getattr(self.lib, f"grid_inverse_dynamics_vaf_{suf}")(p, stride, pq, out.data_ptr(), B, gravity, self._stream()))
        return out

    def direct_minv(self, q, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q, n)
        suf = self._suf(q)
        out = self._out(out, q, B, n * n)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_direct_minv_{suf}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    def forward_dynamics(self, q_qd_u, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd_u, 3 * n)
        suf = self._suf(q_qd_u)
        out = self._out(out, q_qd_u, B, n)
        with torch.cuda.device(q_qd_u.device):
            self._check(getattr(self.lib, f"grid_forward_dynamics_{suf}")(p, stride, out.data_ptr(), B, gravity, self._stream()))
        return out

    def inverse_dynamics_gradient(self, q_qd, qdd=None, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd, 2 * n)
        suf = self._suf(q_qd)
        out = self._out(out, q_qd, B, 2 * n * n)
        pq = None
        if qdd is not None:
            pq, sq, Bq = self._dev(qdd, n)
            if sq != n or Bq != B:
                raise ValueError("qdd must be a dense (B, n) tensor")
        with torch.cuda.device(q_qd.device):
            self._check(getattr(self.lib, f"grid_inverse_dynamics_gradient_{suf}")(p, stride, pq, out.data_ptr(), B, gravity, self._stream()))
        return out

    def forward_dynamics_gradient(self, q_qd_u, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd_u, 3 * n)
        suf = self._suf(q_qd_u)
        out = self._out(out, q_qd_u, B, 2 * n * n)
        with torch.cuda.device(q_qd_u.device):
            self._check(getattr(self.lib, f"grid_forward_dynamics_gradient_{suf}")(p, stride, out.data_ptr(), B, gravity, self._stream()))
        return out

    def forward_dynamics_linearization(self, q_qd_u, gravity: float = 9.81, out=None, split: bool = False):
        """grad FD with its by-products in one launch: (B, 3n^2 + n) rows [dqdd/dq | dqdd/dqd | Minv | qdd]
        (include/grid_b200.h).  With `split`, returns the views (df_du (B, 2n^2), Minv (B, n^2), qdd (B, n))."""
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd_u, 3 * n)
        suf = self._suf(q_qd_u)
        out = self._out(out, q_qd_u, B, 3 * n * n + n)
        with torch.cuda.device(q_qd_u.device):
            self._check(getattr(self.lib, f"grid_forward_dynamics_linearization_{suf}")(p, stride, out.data_ptr(), B, gravity, self._stream()))
        if split:
            return out[:, :2 * n * n], out[:, 2 * n * n:3 * n * n], out[:, 3 * n * n:]
        return out

    def forward_dynamics_gradient_qdd_minv(self, q_qd, qdd, Minv, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd, 2 * n)
        suf = self._suf(q_qd)
        pq, sq, _ = self._dev(qdd, n)
        pm, sm, _ = self._dev(Minv, n * n)
        if sq != n or sm != n * n:
            raise ValueError("qdd and Minv must be dense")
        out = self._out(out, q_qd, B, 2 * n * n)
        with torch.cuda.device(q_qd.device):
            self._check(getattr(self.lib, f"grid_forward_dynamics_gradient_qdd_minv_{suf}")(
                p, stride, pq, pm, out.data_ptr(), B, gravity, self._stream()))
        return out

    def aba(self, q_qd_tau, gravity: float = 9.81, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q_qd_tau, 3 * n)
        suf = self._suf(q_qd_tau)
        out = self._out(out, q_qd_tau, B, n)
        with torch.cuda.device(q_qd_tau.device):
            self._check(getattr(self.lib, f"grid_aba_{suf}")(p, stride, out.data_ptr(), B, gravity, self._stream()))
        return out

    def _second_order(self, fn: str, x, gravity: float, out):
        import torch
        n = self.n
        p, stride, B = self._dev(x, 3 * n)
        suf = self._suf(x)
        out = self._out(out, x, B, 4 * n ** 3)
        with torch.cuda.device(x.device):
            self._check(getattr(self.lib, f"grid_{fn}_{suf}")(p, stride, out.data_ptr(), B, gravity, self._stream()))
        return out

    def idsva_so(self, q_qd_qdd, gravity: float = 9.81, out=None):
        """(B, 3n) [q|qd|qdd] -> (B, 4 n^3): d2tau/dq2, d2tau/dqd2, d2tau/dqd dq, dM/dq (layout: include/grid_b200.h)."""
        return self._second_order("idsva_so", q_qd_qdd, gravity, out)

    def fdsva_so(self, q_qd_u, gravity: float = 9.81, out=None):
        """(B, 3n) [q|qd|tau] -> (B, 4 n^3): d2qdd/dq2, d2qdd/dqd2, d2qdd/dqd dq, d2qdd/dtau dq."""
        return self._second_order("fdsva_so", q_qd_u, gravity, out)

    def end_effector_pose(self, q, out=None):
        """(B, >=n) rows starting with q -> (B, 6*num_ees) [x y z roll pitch yaw] per end effector."""
        import torch
        p, stride, B = self._dev(q, self.n)
        out = self._out(out, q, B, 6 * self.num_ees)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_end_effector_pose_{self._suf(q)}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    def end_effector_pose_gradient(self, q, out=None):
        """-> (B, 6*num_ees*n), index (ee*n + j)*6 + o."""
        import torch
        p, stride, B = self._dev(q, self.n)
        out = self._out(out, q, B, 6 * self.num_ees * self.n)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_end_effector_pose_gradient_{self._suf(q)}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    def end_effector_pose_gradient_hessian(self, q):
        """-> (gradient (B, 6*num_ees*n), Hessian (B, 6*num_ees*n*n) with index (ee*6 + o)*n*n + i*n + j)."""
        import torch
        n = self.n
        p, stride, B = self._dev(q, n)
        g = torch.empty((B, 6 * self.num_ees * n), dtype=q.dtype, device=q.device)
        h = torch.empty((B, 6 * self.num_ees * n * n), dtype=q.dtype, device=q.device)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_end_effector_pose_gradient_hessian_{self._suf(q)}")(
                p, stride, g.data_ptr(), h.data_ptr(), B, self._stream()))
        return g, h

    def joint_transforms(self, q, out=None):
        """-> (B, 16 n): cumulative transform base <- joint j at [16 j + r + 4 c] (4 x 4 column-major per joint)."""
        import torch
        p, stride, B = self._dev(q, self.n)
        out = self._out(out, q, B, 16 * self.n)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_joint_transforms_{self._suf(q)}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    def fixed_target_pose(self, q, out=None):
        """Pose [x y z roll pitch yaw] of the library's fixed kinematic target (self.fixed_target) -> (B, 6)."""
        import torch
        p, stride, B = self._dev(q, self.n)
        out = self._out(out, q, B, 6)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_fixed_target_pose_{self._suf(q)}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    def fixed_target_pose_gradient_hessian(self, q):
        """-> (gradient (B, 6 n) at [j*6 + o], Hessian (B, 6 n^2) at [o*n*n + i*n + j]) of the fixed target's pose."""
        import torch
        n = self.n
        p, stride, B = self._dev(q, n)
        g = torch.empty((B, 6 * n), dtype=q.dtype, device=q.device)
        h = torch.empty((B, 6 * n * n), dtype=q.dtype, device=q.device)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_fixed_target_pose_gradient_hessian_{self._suf(q)}")(
                p, stride, g.data_ptr(), h.data_ptr(), B, self._stream()))
        return g, h

    def crba(self, q, out=None):
        import torch
        n = self.n
        p, stride, B = self._dev(q, n)
        suf = self._suf(q)
        out = self._out(out, q, B, n * n)
        with torch.cuda.device(q.device):
            self._check(getattr(self.lib, f"grid_crba_{suf}")(p, stride, out.data_ptr(), B, self._stream()))
        return out

    # --------------------------------------------------------- host entry points
    def forward_dynamics_gradient_host(self, h_q_qd_u, h_out=None, gravity: float = 9.81):
        """Host tensors in / out (pinned for full PCIe speed); H2D, kernel and D2H run inside the call."""
        import torch
        n = self.n
        if h_q_qd_u.is_cuda or not h_q_qd_u.is_contiguous() or h_q_qd_u.shape[1] != 3 * n:
            raise ValueError("expected a contiguous host (B, 3n) tensor")
        B = h_q_qd_u.shape[0]
        suf = self._suf(h_q_qd_u)
        if h_out is None:
            h_out = torch.empty((B, 2 * n * n), dtype=h_q_qd_u.dtype, pin_memory=True)
        self._check(getattr(self.lib, f"grid_forward_dynamics_gradient_host_{suf}")(h_q_qd_u.data_ptr(), h_out.data_ptr(), B, gravity))
        return h_out

    def forward_dynamics_linearization_host(self, h_q_qd_u, h_out=None, gravity: float = 9.81):
        import torch
        n = self.n
        if h_q_qd_u.is_cuda or not h_q_qd_u.is_contiguous() or h_q_qd_u.shape[1] != 3 * n:
            raise ValueError("expected a contiguous host (B, 3n) tensor")
        B = h_q_qd_u.shape[0]
        suf = self._suf(h_q_qd_u)
        if h_out is None:
            h_out = torch.empty((B, 3 * n * n + n), dtype=h_q_qd_u.dtype, pin_memory=True)
        self._check(getattr(self.lib, f"grid_forward_dynamics_linearization_host_{suf}")(h_q_qd_u.data_ptr(), h_out.data_ptr(), B, gravity))
        return h_out

    def _host(self, name: str, h_in, in_cols: int, out_cols: int, *extra, h_in2=None):
        """Generic host-buffer call: contiguous host (B, in_cols) tensor in, pinned (B, out_cols) tensor out."""
        import torch
        if h_in.is_cuda or not h_in.is_contiguous() or h_in.dim() != 2 or h_in.shape[1] != in_cols:
            raise ValueError(f"expected a contiguous host (B, {in_cols}) tensor")
        B = h_in.shape[0]
        suf = self._suf(h_in)
        h_out = torch.empty((B, out_cols), dtype=h_in.dtype, pin_memory=True)
        args = [h_in.data_ptr()]
        if name == "inverse_dynamics_gradient":
            if h_in2 is not None and (h_in2.is_cuda or not h_in2.is_contiguous() or tuple(h_in2.shape) != (B, self.n)):
                raise ValueError("qdd must be a contiguous host (B, n) tensor")
            args.append(h_in2.data_ptr() if h_in2 is not None else None)
        self._check(getattr(self.lib, f"grid_{name}_host_{suf}")(*args, h_out.data_ptr(), B, *extra))
        return h_out

    def direct_minv_host(self, h_q):
        return self._host("direct_minv", h_q, self.n, self.n ** 2)

    def crba_host(self, h_q):
        return self._host("crba", h_q, self.n, self.n ** 2)

    def aba_host(self, h_q_qd_tau, gravity: float = 9.81):
        return self._host("aba", h_q_qd_tau, 3 * self.n, self.n, gravity)

    def inverse_dynamics_gradient_host(self, h_q_qd, h_qdd=None, gravity: float = 9.81):
        return self._host("inverse_dynamics_gradient", h_q_qd, 2 * self.n, 2 * self.n ** 2, gravity, h_in2=h_qdd)

    def idsva_so_host(self, h_q_qd_qdd, gravity: float = 9.81):
        return self._host("idsva_so", h_q_qd_qdd, 3 * self.n, 4 * self.n ** 3, gravity)

    def fdsva_so_host(self, h_q_qd_u, gravity: float = 9.81):
        return self._host("fdsva_so", h_q_qd_u, 3 * self.n, 4 * self.n ** 3, gravity)

    def forward_dynamics_host(self, h_q_qd_u, h_out=None, gravity: float = 9.81):
        import torch
        B = h_q_qd_u.shape[0]
        if h_out is None:
            h_out = torch.empty((B, self.n), dtype=torch.float32, pin_memory=True)
        self._check(self.lib.grid_forward_dynamics_host_f32(h_q_qd_u.data_ptr(), h_out.data_ptr(), B, gravity))
        return h_out

    def inverse_dynamics_host(self, h_q_qd_qdd, h_out=None, gravity: float = 9.81):
        import torch
        B = h_q_qd_qdd.shape[0]
        if h_out is None:
            h_out = torch.empty((B, self.n), dtype=torch.float32, pin_memory=True)
        self._check(self.lib.grid_inverse_dynamics_host_f32(h_q_qd_qdd.data_ptr(), h_out.data_ptr(), B, gravity))
        return h_out


_LIBS = {}


def get_library(robot_name: str) -> GRiDLibrary:
    if robot_name not in _LIBS:
        _LIBS[robot_name] = GRiDLibrary(robot_name)
    return _LIBS[robot_name]


# ------------------------------------------------------------------ multi-GPU sharding
def shard_bounds(num_knots: int, world_size: int, rank: int, align: int = 32):
    """Contiguous slice [lo, hi) of the batch owned by `rank`.  Knot points are independent, so the batch is cut
    into world_size contiguous slices with no exchange step; slice starts are multiples of `align` (the 32-knot
    warp tile) so every rank's slice keeps the TMA-aligned fast path, the last rank takes the ragged tail."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    tiles = -(-num_knots // align)
    per = -(-tiles // world_size)
    lo = min(num_knots, rank * per * align)
    hi = min(num_knots, (rank + 1) * per * align)
    return lo, hi


def gather_results(local, num_knots: int, group=None):
    """Optional final gather of per-rank result slices (the only collective of the path; off the compute path).
    local: (hi - lo, m) tensor of this rank's slice from shard_bounds().  Returns the (num_knots, m) tensor on
    every rank.  Uses all_gather_into_tensor over equal-sized padded slices (NCCL on GPU, gloo on CPU)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    per = shard_bounds(num_knots, world, 0)[1]
    m = local.shape[1]
    pad = torch.zeros((per, m), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((world * per, m), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return out[:num_knots]
