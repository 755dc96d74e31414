#!/usr/bin/env python
"""Which host resource caps the multi-GPU END-TO-END rate?  Concurrent pinned-memory copies on 1, 2, 4, 8 GPUs of one
box (one process, one stream per device): aggregate device->host, host->device and both-at-once bandwidth, plus the
PCIe / NUMA topology the driver reports.  The end-to-end metric moves 476 B per knot (84 in, 392 out).

  python tools/pcie_scaling.py            (on a multi-GPU box; prints JSON lines)
"""
import json
import subprocess
import time

import torch

MB = 256


def run(devs, mode):
    bufs = []
    for d in devs:
        torch.cuda.set_device(d)
        g = torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{d}")
        h = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
        g2 = torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{d}")
        h2 = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
        bufs.append((d, g, h, g2, h2, torch.cuda.Stream(device=d), torch.cuda.Stream(device=d)))

    def issue():
        for d, g, h, g2, h2, s1, s2 in bufs:
            torch.cuda.set_device(d)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s1):
                    h.copy_(g, non_blocking=True)
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s2):
                    g2.copy_(h2, non_blocking=True)

    def sync():
        for d, *_ in bufs:
            torch.cuda.synchronize(d)
    issue(); sync()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(4):
            issue()
        sync()
        best = min(best, (time.perf_counter() - t0) / 4)
    nbytes = len(devs) * (MB << 20) * (2 if mode == "both" else 1)
    return nbytes / best / 1e9


def main():
    n = torch.cuda.device_count()
    for cmd in (["nvidia-smi", "topo", "-m"], ["lscpu"], ["numactl", "-H"]):
        try:
            out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=30).stdout
            keep = [l for l in out.splitlines() if cmd[0] != "lscpu" or any(k in l for k in ("Model name", "Socket", "NUMA", "CPU(s):"))]
            print("# " + " ".join(cmd))
            print("\n".join("#   " + l for l in keep[:40]))
        except Exception as e:       # tool not installed
            print(f"# {' '.join(cmd)}: {e}")
    for k in (1, 2, 4, 8):
        if k > n:
            break
        devs = list(range(k))
        row = {"gpus": k}
        for mode in ("d2h", "h2d", "both"):
            row[mode + "_GBps_total"] = round(run(devs, mode), 1)
        row["e2e_bound_Mknots_s"] = round(row["both_GBps_total"] * 1e9 / 476 / 1e6, 1)
        print(json.dumps(row), flush=True)
    if n >= 8:       # second half of the box alone: do the two halves hang from different host bridges?
        row = {"gpus": "4-7"}
        for mode in ("d2h", "both"):
            row[mode + "_GBps_total"] = round(run([4, 5, 6, 7], mode), 1)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
