#!/usr/bin/env python
"""Headline benchmark: grad-FD knot-points/sec, KUKA iiwa14, batch 65536 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (forward_dynamics_gradient, reference
algorithms/_forward_dynamics_gradient.py:113-249) over one batch of 65536 synthetic knot points
[q|qd|u] ~ U(-1,1).

`value` (scaling "weak"): every rank owns its own 65536-knot batch, no collective on the data path; whole-job
knots/s with the inputs resident in HBM, CUDA events on the launching stream, max over ranks.  Between timed steps
the kernel is pointed at the next of several input/output buffer sets whose total size exceeds the 126 MB L2.
`strong_scaling` (N > 1 only): ONE global 65536-knot batch cut into contiguous per-rank slices
(runtime.shard_bounds), same timing rules -- this is the split the metric's "1/2/4/8 B200" means.
`e2e`: the same metric through the C-ABI host entry point (grid_forward_dynamics_gradient_host_f32: pinned host
buffers, H2D and D2H inside the timed region).
`roofline`: the binding roof is the FP32 pipe; achieved = EXECUTED flop per knot (FFMA x 2 + FMUL +
FADD per lane, counted from the kernel's SASS and confirmed by ncu: profiles/executed_flops.json) x knots / launch
time; peak = 148 SMs x 128 lanes x 2 x max SM clock.  The HBM view (algorithmic bytes / time over the measured copy
bandwidth) is kept under roofline.hbm.
Before any number is printed a seeded subsample of the timed outputs is compared with the CPU oracle (1e-4).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROBOT = "iiwa14"
BATCH = 65536
GRAVITY = 9.81
# algorithmic work per knot point for iiwa14 grad FD
N_JOINTS = 7
BYTES_PER_KNOT = 4 * (3 * N_JOINTS + 2 * N_JOINTS * N_JOINTS)   # 476 B (fp32 in + out)
MODEL_FLOPS_PER_KNOT = 44512                                    # dense 6x6 operation count of _test.py
L2_BYTES = 126 * 1024 * 1024
METRIC = "grad-FD knot-points/sec (iiwa14, batch 65536)"
WORKLOAD = "iiwa14 grad-FD, batch 65536 per GPU"
PARITY_TOL = 1e-4


def _load_json(name):
    p = os.path.join(ROOT, "profiles", name)
    if os.path.exists(p):
        try:
            return json.load(open(p))
        except Exception:
            return None
    return None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler(threading.Thread):
    """Polls SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.005):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.active = threading.Event()
        self.stop_flag = threading.Event()
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False

    NAMES = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
             0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            if self.active.is_set():
                try:
                    self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    try:
                        r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                    except Exception:
                        r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                    for bit, name in self.NAMES.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(self.period)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def _dist():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def _oracle():
    from gridcodegenerator_b200.robot import load_robot
    from oracle import Oracle
    return Oracle(load_robot(ROBOT)), Oracle.max_threads()


def _workload(seed: int, batch: int = BATCH):
    import numpy as np
    return np.random.default_rng(seed).uniform(-1, 1, (batch, 3 * N_JOINTS)).astype(np.float32)


def cpu_reference_leg(seconds: float):
    """The reference's CPU algorithm for the path (oracle/rbd_oracle.c, a C restatement of reference
    _test.py:496-520 pinned to its golden vectors), all host threads, whole 65536-knot batches until `seconds`
    have passed (a bounded sample of the same workload)."""
    import numpy as np
    orc, cores = _oracle()
    x = _workload(0).astype(np.float64)
    orc.batch_fd_grad(x[:4096], nthreads=cores)  # warm
    t0 = time.perf_counter()
    done = 0
    while True:
        orc.batch_fd_grad(x, nthreads=cores)
        done += BATCH
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    return done / dt, cores, done, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.
    The reference itself is pure Python + NumPy and cannot travel to the GPU box (no /root/reference there;
    measured in the build container at ~80 knots/s per core), so the timed code is its C restatement
    (cpu_baseline.kind = "port", pinned to the reference's outputs by tests/test_oracle_golden.py).  Same config as
    the GPU arm: every step is the WHOLE 65536-knot batch on all host threads (fp64, the reference's precision);
    the step count is capped so the run ends within minutes."""
    rank, world, _ = _dist()
    if rank != 0:
        return 0
    import numpy as np
    orc, cores = _oracle()
    x = _workload(0).astype(np.float64)
    for _ in range(max(1, min(args.warmup, 3))):
        orc.batch_fd_grad(x, nthreads=cores)
    steps = max(1, min(args.steps, 40))
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.batch_fd_grad(x, nthreads=cores)
    dt = time.perf_counter() - t0
    value = BATCH * steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "knots/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": max(1, min(args.warmup, 3)), "ms_per_step": 1e3 * dt / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "robot": ROBOT, "batch_per_gpu": BATCH, "gravity": GRAVITY,
                   "knots_per_step": BATCH},
        "cpu_baseline": {"value": value, "unit": "knots/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} steps x {BATCH} knots (the full batch of the GPU arm), C port of "
                                   f"reference _test.py test_fd_grad (oracle/rbd_oracle.c), {cores} threads"},
        "e2e": {"value": value, "unit": "knots/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def _timed_graph(torch, dist, world, stream, step_fn, steps):
    """Capture `steps` launches into one CUDA graph, replay once untimed, then time one replay with CUDA events on
    `stream` between barriers.  Returns (local ms, max-over-ranks ms)."""
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        step_fn(0)
        torch.cuda.synchronize()
        with torch.cuda.graph(graph, stream=side):
            for i in range(steps):
                step_fn(i)
    graph.replay()                        # untimed replay: graph upload + instruction caches
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    graph.replay()
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return ms, float(t.item())


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    rank, world, local = _dist()
    os.dup2(2, 1)     # stray library output (NCCL banner) goes to stderr; the JSON line is written by _emit()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from gridcodegenerator_b200.runtime import get_library, shard_bounds
    lib = get_library(ROBOT)
    n = lib.n
    assert n == N_JOINTS

    # rotating buffer sets: total footprint > L2 so that no step finds its inputs in cache
    set_bytes = BATCH * BYTES_PER_KNOT
    nsets = max(2, -(-2 * L2_BYTES // set_bytes))
    rng = np.random.default_rng(1000 + rank)
    host_ins, ins, outs = [], [], []
    for s in range(nsets):
        x = rng.uniform(-1, 1, (BATCH, 3 * n)).astype(np.float32)
        host_ins.append(x)
        ins.append(torch.from_numpy(x).cuda())
        outs.append(torch.empty((BATCH, 2 * n * n), dtype=torch.float32, device="cuda"))
    stream = torch.cuda.current_stream()

    def step(i):
        lib.forward_dynamics_gradient(ins[i % nsets], GRAVITY, out=outs[i % nsets])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    warmup = max(args.warmup, 3)
    for i in range(warmup):
        step(i)
    barrier()
    # The K timed steps are captured once into a CUDA graph (K kernel nodes, each on the next buffer set) and
    # replayed: at ~24 us per launch the Python/ctypes call path (~28 us per call) would otherwise be what is timed.
    sampler.active.set()
    ms, ms_max = _timed_graph(torch, dist, world, stream, step, args.steps)
    sampler.active.clear()
    value = world * BATCH * args.steps / (ms_max * 1e-3)
    kernel_ms = ms / args.steps                         # one launch per step, back to back on one stream

    # ---- parity gate: a seeded subsample of what the timed launches wrote, against the oracle
    parity_err = None
    if rank == 0:
        orc, _ = _oracle()
        prng = np.random.default_rng(7)
        worst = 0.0
        for s in range(min(nsets, args.steps)):
            idx = np.unique(np.concatenate([np.arange(64), prng.choice(BATCH, 192, replace=False)]))
            got = outs[s][torch.from_numpy(idx).cuda()].cpu().numpy().astype(np.float64)
            ref = orc.batch_fd_grad(host_ins[s][idx].astype(np.float64))
            e = np.abs(got - ref).max(axis=1) / np.maximum(1.0, np.abs(ref).max(axis=1))
            worst = max(worst, float(e.max()))
        parity_err = worst
        if not (parity_err < PARITY_TOL):
            raise RuntimeError(f"parity gate failed: max rel err {parity_err:.3e} >= {PARITY_TOL}; no number is reported")

    # ---- strong scaling: ONE global batch of 65536 knots cut into contiguous per-rank slices
    strong = None
    if world > 1:
        lo, hi = shard_bounds(BATCH, world, rank)
        s_ins = [t[lo:hi] for t in ins]
        s_outs = [t[lo:hi] for t in outs]

        def sstep(i):
            lib.forward_dynamics_gradient(s_ins[i % nsets], GRAVITY, out=s_outs[i % nsets])

        for i in range(3):
            sstep(i)
        barrier()
        _, sms_max = _timed_graph(torch, dist, world, stream, sstep, args.steps)
        strong = {"value": BATCH * args.steps / (sms_max * 1e-3), "unit": "knots/s", "global_batch": BATCH,
                  "knots_per_gpu": hi - lo, "ms_per_step": sms_max / args.steps,
                  "note": "one 65536-knot batch cut by runtime.shard_bounds, one rank per GPU, no collective, "
                          "max over ranks; compare with the 1-GPU value of the same bench"}

    # ---- end to end through the C-ABI host entry point (pinned host buffers)
    h_in = torch.from_numpy(rng.uniform(-1, 1, (BATCH, 3 * n)).astype(np.float32)).pin_memory()
    h_out = torch.empty((BATCH, 2 * n * n), dtype=torch.float32).pin_memory()
    e2e_steps = max(3, min(args.steps, 50))
    for _ in range(3):
        lib.forward_dynamics_gradient_host(h_in, h_out, GRAVITY)
    barrier()
    sampler.active.set()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        lib.forward_dynamics_gradient_host(h_in, h_out, GRAVITY)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    sampler.active.clear()
    sampler.stop_flag.set()
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * BATCH * e2e_steps / float(t.item())
    checksum = float(h_out[:: 4096].double().abs().sum())

    if rank == 0:
        hbm_peak, sm_max_mhz, which = _peaks()
        achieved_gbs = BYTES_PER_KNOT * BATCH / (kernel_ms * 1e-3) / 1e9
        clocks = sampler.summary()
        fp32_peak = 148 * 128 * 2 * (clocks.get("sm_max_mhz") or sm_max_mhz) * 1e6 / 1e12
        ex = _load_json("executed_flops.json") or {}
        ex_flops = ex.get("flop_per_knot")
        ex_tflops = (ex_flops * BATCH / (kernel_ms * 1e-3) / 1e12) if ex_flops else None
        model_tflops = MODEL_FLOPS_PER_KNOT * BATCH / (kernel_ms * 1e-3) / 1e12
        # Register-port bound: an FP32 instruction that reads three distinct registers without operand reuse issues
        # every second cycle on sm_100 (two register banks, one read port each; measured with tools/probes/rf_ports.cu:
        # 0.50-0.59 instead of 0.90 instructions per cycle per scheduler).  tools/sass_model.py counts the issue cycles
        # one 32-knot tile needs under that rule; a scheduler cannot run the batch faster than (tiles it owns) x that.
        ports = dict(ex.get("register_ports") or {})
        if ports.get("dispatch_cycles_per_tile"):
            clk = (clocks.get("sm_mhz") or sm_max_mhz) * 1e6
            tiles, smsps = BATCH // 32, 148 * 4
            fluid_us = tiles / smsps * ports["dispatch_cycles_per_tile"] / clk * 1e6
            granular_us = -(-tiles // smsps) * ports["dispatch_cycles_per_tile"] / clk * 1e6
            ports.update({"bound_us_fluid": fluid_us, "bound_us_whole_tiles_per_scheduler": granular_us,
                          "frac_of_fluid_bound": fluid_us / (kernel_ms * 1e3),
                          "frac_of_whole_tile_bound": granular_us / (kernel_ms * 1e3)})
        cpu_v, cores, done, cpu_dt = cpu_reference_leg(args.cpu_seconds)
        info = lib.kernel_info("forward_dynamics_gradient")
        traffic = _load_json("dram_traffic.json")
        line = {
            "metric": METRIC, "value": value, "unit": "knots/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "value_is": "weak scaling: 65536 knots per GPU (see strong_scaling for "
                       "the single 65536-knot batch split over the GPUs)", "robot": ROBOT, "batch_per_gpu": BATCH,
                       "gravity": GRAVITY, "parallelism": f"batch-sharded x{world}, no collective",
                       "l2": f"rotating {nsets} input/output buffer sets ({nsets * set_bytes >> 20} MB > 126 MB L2)",
                       "launch": "the K timed steps are one CUDA graph of K kernel nodes, replayed once",
                       "kernel": info, "persistent_blocks": lib.persistent_blocks("forward_dynamics_gradient")},
            "roofline": {"bound": "fp32", "achieved": ex_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": (ex_tflops / fp32_peak) if ex_tflops else None,
                         "traffic": traffic["bytes_per_launch"] if traffic else None,
                         "traffic_source": traffic.get("source") if traffic else None,
                         "kernel": "forward_dynamics_gradient_kernel<float>", "kernel_ms": kernel_ms,
                         "executed_flop_per_knot": ex_flops, "executed_flop_source": ex.get("source"),
                         "peak_source": "148 SMs x 128 FP32 lanes x 2 x max SM clock (MEASURED_PEAKS.json has no FP32 "
                                        "entry; clock from NVML / MEASURED_PEAKS)",
                         "model": {"flop_per_knot": MODEL_FLOPS_PER_KNOT, "tflops": model_tflops,
                                   "note": "dense 6x6 operation count of the reference algorithm (SURVEY 8(d)); the "
                                           "kernel folds structural zeros, so this is a throughput figure in model "
                                           "FLOPs, not a roofline fraction"},
                         "register_ports": ports or None,
                         "hbm": {"achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": achieved_gbs / hbm_peak, "peak_source": which,
                                 "algorithmic_bytes_per_knot": BYTES_PER_KNOT}},
            "cpu_baseline": {"value": cpu_v, "unit": "knots/s", "cores": cores, "kind": "port",
                             "sample": f"{done} knots of the same workload in {cpu_dt:.1f} s, C port of reference "
                                       f"_test.py test_fd_grad (oracle/rbd_oracle.c), {cores} threads"},
            "e2e": {"value": e2e_value, "unit": "knots/s", "h2d_bytes_per_step": BATCH * 3 * n * 4,
                    "d2h_bytes_per_step": BATCH * 2 * n * n * 4, "steps": e2e_steps,
                    "api": "grid_forward_dynamics_gradient_host_f32 (pinned host buffers, chunked 3-stream pipeline)",
                    "note": "PCIe-bound: 31.2 MB per step over the host link", "checksum": checksum},
            "parity": {"max_rel_err_vs_oracle": parity_err, "tol": PARITY_TOL,
                       "sample": "256 knots of each timed output buffer set (64 first + 192 random)"},
            "gpu_launches": args.steps,
            "clocks": clocks,
        }
        if strong is not None:
            line["strong_scaling"] = strong
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=4.0, help="wall time of the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


def _emit(line: dict):
    sys.stdout.flush()
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = os.dup(1)


if __name__ == "__main__":
    sys.exit(main())
