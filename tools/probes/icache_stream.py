#!/usr/bin/env python
"""How fast does an SM execute straight-line code that does not fit its instruction caches?

Generates PTX kernels whose loop body is K x 8 FFMAs on 8 dependent chains (K x 128 B of SASS), assembles them with
ptxas -O1 (nvcc's front end is quadratic on basic blocks of this size) and runs them through the CUDA driver API
(cuda-python) on 148 CTAs x 8 warps -- one CTA per SM, as the emitted tile kernels run.  Prints, per body size, the
cycles per warp-instruction per SM sub-partition (2 warps each) for one cold pass (reps = 1) and for eight passes, and
the same for G different 128 KB bodies selected by blockIdx (SMs of one GPC then stream different code).

  python tools/probes/icache_stream.py build <dir>      (CPU box: writes <dir>/icache_K.cubin)
  python tools/probes/icache_stream.py run <dir>        (GPU box)
"""
import os
import subprocess
import sys

SIZES = [256, 512, 1024, 2048, 4096, 8192, 16384]       # K: body = K * 128 B
MULTI = 8

HEAD = """.version 8.7
.target sm_100a
.address_size 64
"""


def steps(k, pa="%a", pb="%b", pc="%c", fast=False):
    if fast:        # two distinct source registers per FMA: one dispatch cycle each (no register-port limit)
        return "".join(f"fma.rn.f32 %x{i}, %x{i}, {pa}, %x{i};\n" for i in range(8)) * k
    one = (f"fma.rn.f32 %x0, %x0, {pa}, {pb};\nfma.rn.f32 %x1, %x1, {pa}, {pc};\nfma.rn.f32 %x2, %x2, {pb}, {pa};\n"
           f"fma.rn.f32 %x3, %x3, {pc}, {pa};\nfma.rn.f32 %x4, %x4, {pa}, %x0;\nfma.rn.f32 %x5, %x5, {pb}, %x1;\n"
           f"fma.rn.f32 %x6, %x6, {pc}, %x2;\nfma.rn.f32 %x7, %x7, {pa}, %x3;\n")
    return one * k


def kernel(name, body):
    return f"""
.visible .entry {name}(.param .u64 p_out, .param .u32 p_reps, .param .f32 p_a, .param .f32 p_b, .param .f32 p_c, .param .u32 p_g)
.maxntid 512, 1, 1
{{
    .reg .f32 %x<8>; .reg .f32 %a, %b, %c, %s; .reg .b32 %r<8>; .reg .b64 %rd<4>; .reg .pred %p<10>;
    ld.param.u64 %rd0, [p_out]; ld.param.u32 %r0, [p_reps]; ld.param.f32 %a, [p_a]; ld.param.f32 %b, [p_b]; ld.param.f32 %c, [p_c];
    ld.param.u32 %r5, [p_g];
    mov.u32 %r1, %tid.x; mov.u32 %r2, %ctaid.x; rem.u32 %r6, %r2, %r5;
    cvt.rn.f32.u32 %x0, %r1; mov.f32 %x1, 0f3F800000; mov.f32 %x2, 0f40000000; mov.f32 %x3, 0f40400000;
    mov.f32 %x4, 0f40800000; mov.f32 %x5, 0f40A00000; mov.f32 %x6, 0f40C00000; mov.f32 %x7, 0f40E00000;
    mov.u32 %r3, 0;
LOOP:
{body}
    add.u32 %r3, %r3, 1; setp.lt.u32 %p0, %r3, %r0; @%p0 bra LOOP;
    add.f32 %s, %x0, %x1; add.f32 %s, %s, %x2; add.f32 %s, %s, %x3; add.f32 %s, %s, %x4; add.f32 %s, %s, %x5;
    add.f32 %s, %s, %x6; add.f32 %s, %s, %x7;
    mad.lo.u32 %r4, %r2, 512, %r1; cvta.to.global.u64 %rd1, %rd0; mul.wide.u32 %rd2, %r4, 4; add.u64 %rd3, %rd1, %rd2;
    st.global.f32 [%rd3], %s;
    ret;
}}
"""


def multi_body():
    perms = [("%a", "%b", "%c"), ("%b", "%c", "%a"), ("%c", "%a", "%b"), ("%a", "%c", "%b"),
             ("%b", "%a", "%c"), ("%c", "%b", "%a"), ("%a", "%a", "%b"), ("%b", "%b", "%c")]
    out = ""
    for g, p in enumerate(perms):
        out += f"    setp.ne.u32 %p{g + 1}, %r6, {g}; @%p{g + 1} bra SKIP{g};\n" + steps(1024, *p) + f"    bra DONE;\nSKIP{g}:\n"
    return out + "DONE:\n"


def build(d):
    os.makedirs(d, exist_ok=True)
    for k in SIZES:
        open(f"{d}/icache_{k}.ptx", "w").write(HEAD + kernel("k", steps(k)))
        open(f"{d}/icache_fast_{k}.ptx", "w").write(HEAD + kernel("k", steps(k, fast=True)))
    open(f"{d}/icache_multi.ptx", "w").write(HEAD + kernel("k", multi_body()))
    for f in [f"icache_{k}" for k in SIZES] + [f"icache_fast_{k}" for k in SIZES] + ["icache_multi"]:
        subprocess.check_call(["ptxas", "-arch=sm_100a", "-O1", f"{d}/{f}.ptx", "-o", f"{d}/{f}.cubin"])
        os.remove(f"{d}/{f}.ptx")
        print("built", f, os.path.getsize(f"{d}/{f}.cubin"))


def run(d):
    import ctypes
    import numpy as np
    from cuda import cuda

    def ck(r):
        if r[0] != cuda.CUresult.CUDA_SUCCESS:
            raise RuntimeError(str(r[0]))
        return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)
    ck(cuda.cuInit(0))
    dev = ck(cuda.cuDeviceGet(0))
    ctx = ck(cuda.cuDevicePrimaryCtxRetain(dev))
    ck(cuda.cuCtxSetCurrent(ctx))
    clk = ck(cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_CLOCK_RATE, dev))
    out = ck(cuda.cuMemAlloc(148 * 512 * 4))
    e0, e1 = ck(cuda.cuEventCreate(0)), ck(cuda.cuEventCreate(0))

    def timed(fn_, reps, groups, threads=256):
        args = (np.array([int(out)], dtype=np.uint64), np.array([reps], dtype=np.uint32), np.array([1.0001], dtype=np.float32),
                np.array([0.5], dtype=np.float32), np.array([0.25], dtype=np.float32), np.array([groups], dtype=np.uint32))
        ptrs = np.array([a.ctypes.data for a in args], dtype=np.uint64)
        best = 1e9
        for i in range(4):
            ck(cuda.cuEventRecord(e0, 0))
            ck(cuda.cuLaunchKernel(fn_, 148, 1, 1, threads, 1, 1, 0, 0, ptrs.ctypes.data, 0))
            ck(cuda.cuEventRecord(e1, 0))
            ck(cuda.cuEventSynchronize(e1))
            ms = ck(cuda.cuEventElapsedTime(e0, e1))
            if i:
                best = min(best, ms)
        return best
    print(f"# nominal SM clock {clk} kHz; cycles are per warp-instruction per SM sub-partition (2 warps each)")
    print("# body_KB reps ms cycles/instr code_GB/s_per_SM")
    for tag in ("", "fast_"):
        print(f"# {'two-register FMAs (1 dispatch cycle)' if tag else 'three-register FMAs (2 dispatch cycles)'}")
        for k in SIZES:
            mod = ck(cuda.cuModuleLoad(f"{d}/icache_{tag}{k}.cubin".encode()))
            fn_ = ck(cuda.cuModuleGetFunction(mod, b"k"))
            for reps in (1, 8):
                ms = timed(fn_, reps, 1)
                cyc = ms * 1e-3 * clk * 1e3
                print(f"{k * 128 // 1024:6d} {reps:2d} {ms:9.4f} {cyc / (k * 8 * reps * 2):8.3f} {k * 128.0 * reps / (ms * 1e-3) / 1e9:8.2f}", flush=True)
    print("# warps per SM sharing one 512 KB body (reps 4): warps ms cycles_per_pass code_bytes_per_clk_per_SM_if_fetched_once")
    mod = ck(cuda.cuModuleLoad(f"{d}/icache_fast_4096.cubin".encode()))
    fn_ = ck(cuda.cuModuleGetFunction(mod, b"k"))
    for warps in (1, 2, 4, 8, 16):
        ms = timed(fn_, 4, 1, 32 * warps)
        cyc = ms * 1e-3 * clk * 1e3 / 4
        print(f"warps {warps:2d} {ms:9.4f} {cyc:10.0f} {4096 * 128 / cyc:6.2f}", flush=True)
    mod = ck(cuda.cuModuleLoad(f"{d}/icache_multi.cubin".encode()))
    fn_ = ck(cuda.cuModuleGetFunction(mod, b"k"))
    for groups in (1, 2, 4, 8):
        ms = timed(fn_, 8, groups)
        cyc = ms * 1e-3 * clk * 1e3
        print(f"multi: {groups} different 128 KB bodies over the CTAs (blockIdx % {groups}), reps 8: {ms:9.4f} ms {cyc / (1024 * 8 * 8 * 2):8.3f} cycles/instr")


if __name__ == "__main__":
    {"build": build, "run": run}[sys.argv[1]](sys.argv[2])
