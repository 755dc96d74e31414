#!/usr/bin/env python
"""BASELINE.json configs that are quoted on 8 GPUs, run as they are stated: ONE global batch cut into contiguous
per-rank slices (runtime.shard_bounds), one process per GPU, no collective on the compute path; the optional final
gather of the per-rank results (runtime.gather_results, one NCCL all_gather) is timed separately.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
      tools/bench_sharded.py

Timing: K launches captured in one CUDA graph per rank, CUDA events, barrier + synchronize on both sides, MAX over
ranks; inputs rotate over buffer sets larger than L2 (or at least two).  Each rank also checks a 64-knot sample of
its slice of the outputs (first order, second order and EE Hessian) against the CPU oracle.  One JSON line per case, printed by rank 0.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # (BASELINE config, robot, method, input columns in units of n, GLOBAL batch, oracle method or None)
    (4, "atlas", "forward_dynamics_gradient", 3, 65536, "batch_fd_grad"),
    (4, "atlas", "crba", 1, 65536, "batch_crba"),
    (4, "atlas", "direct_minv", 1, 65536, "batch_minv"),
    (3, "hyq", "forward_dynamics_gradient", 3, 16384, "batch_fd_grad"),
    (3, "hyq", "aba", 3, 16384, "batch_aba"),
    (5, "iiwa14", "idsva_so", 3, 262144, "batch_idsva_so"),
    (5, "iiwa14", "fdsva_so", 3, 262144, "batch_fdsva_so"),
    # EE Hessian: the oracle's second central differences are good to ~1e-5 and RPY derivatives grow like
    # 1/cos(pitch)^2, so the figure reported for it is the MEDIAN over the sampled knots, relative to each knot's scale
    (5, "iiwa14", "end_effector_pose_gradient_hessian", 1, 262144, "batch_ee_pose_hessian"),
    (0, "iiwa14", "forward_dynamics_gradient", 3, 8 * 65536, "batch_fd_grad"),
    (0, "iiwa14", "forward_dynamics_linearization", 3, 8 * 65536, None),
]


def main():
    import numpy as np
    import torch
    import torch.distributed as dist
    from gridcodegenerator_b200.runtime import gather_results, get_library, shard_bounds

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    steps = int(os.environ.get("SHARDED_STEPS", "20"))
    L2 = 126 << 20

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for cfg, robot, method, cols, B, oracle_fn in CASES:
        lib = get_library(robot)
        n = lib.n
        lo, hi = shard_bounds(B, world, rank)
        mine = hi - lo
        rng = np.random.default_rng(100 + rank)

        def call(x, out=None):
            if method == "end_effector_pose_gradient_hessian":
                return lib.end_effector_pose_gradient_hessian(x)[1]
            return getattr(lib, method)(x) if out is None else getattr(lib, method)(x, out=out)

        probe = torch.from_numpy(rng.uniform(-1, 1, (32, cols * n)).astype(np.float32)).cuda()
        out_elems = call(probe).shape[1]
        bytes_per_knot = 4 * (cols * n + out_elems)
        nsets = max(2, min(8, -(-2 * L2 // max(1, mine * bytes_per_knot))))
        xs = [torch.from_numpy(rng.uniform(-1, 1, (mine, cols * n)).astype(np.float32)).cuda() for _ in range(nsets)]
        has_out = method != "end_effector_pose_gradient_hessian"
        outs = [torch.empty((mine, out_elems), dtype=torch.float32, device="cuda") for _ in range(nsets)] if has_out else None

        def step(i):
            return call(xs[i % nsets], outs[i % nsets] if has_out else None)

        for i in range(3):
            res = step(i)
        torch.cuda.synchronize()
        err = None
        if oracle_fn is not None:
            from gridcodegenerator_b200.robot import load_robot
            from oracle import Oracle
            orc = Oracle(load_robot(robot))
            got = step(0)[:64].cpu().numpy().astype(np.float64)
            x64 = xs[0][:64].cpu().numpy().astype(np.float64)
            if cols < 3:
                x64 = np.concatenate([x64, np.zeros((64, (3 - cols) * n))], axis=1)
            ref = getattr(orc, oracle_fn)(x64)
            per_knot = np.abs(got - ref).max(axis=1) / np.maximum(1.0, np.abs(ref).max(axis=1))
            err = float(np.median(per_knot) if oracle_fn == "batch_ee_pose_hessian" else per_knot.max())
        side = torch.cuda.Stream()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            step(0)
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=side):
                for i in range(steps):
                    step(i)
        g.replay()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / steps, err if err is not None else 0.0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        # optional final gather of every rank's result slice (off the compute path)
        gather_ms = None
        if world > 1 and has_out and B * out_elems * 4 <= (4 << 30):
            gather_results(outs[0], B)
            barrier()
            e0.record()
            full = gather_results(outs[0], B)
            e1.record()
            barrier()
            tg = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tg, op=dist.ReduceOp.MAX)
            gather_ms = float(tg[0])
            assert full.shape == (B, out_elems)
            del full
        if rank == 0:
            print(json.dumps({"config": cfg, "robot": robot, "algorithm": method, "global_batch": B, "n_gpus": world,
                              "knots_per_rank": mine, "ms_per_launch_max_over_ranks": ms, "knots_per_s": B / (ms * 1e-3),
                              "algorithmic_bytes_per_knot": bytes_per_knot,
                              "achieved_GBps_total": B * bytes_per_knot / (ms * 1e-3) / 1e9,
                              "oracle_max_rel_err_64_knots_per_rank": float(t[1]) if oracle_fn else None,
                              "oracle_err_statistic": ("median over knots" if oracle_fn == "batch_ee_pose_hessian"
                                                       else "max over knots") if oracle_fn else None,
                              "gather_ms": gather_ms,
                              "gather_GBps_per_rank_out": (B * out_elems * 4 / (gather_ms * 1e-3) / 1e9) if gather_ms else None,
                              "note": f"CUDA-graph replay of {steps} launches per rank, {nsets} rotating buffer sets"}),
                  flush=True)
        del xs, outs, g
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
