#!/usr/bin/env python
"""Pinned host -> device copy bandwidth by copy size (what one record, one micro-batch cost on this box's PCIe link).

Alone: one GPU.  Under torchrun (`python -m torch.distributed.run --nproc-per-node 8 tools/probe_h2d.py`): every rank copies to its
own GPU at the same time, and rank 0 prints the per-rank rates and their sum -- what the host can feed to all GPUs at once, which is
what bounds the end-to-end figure of bench.py at 8 GPUs.
"""
import os
import time

import torch


def numa_node_of(dev_index):
    try:
        import pynvml
        pynvml.nvmlInit()
        bdf = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(dev_index)).busId
        bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
        if len(bdf.split(":")[0]) == 8:
            bdf = bdf[4:]
        return int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
    except Exception:
        return None


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    total = 256 << 20
    host = torch.empty(total, dtype=torch.uint8).pin_memory()
    dev = torch.empty(total, dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    sizes = (256 << 10, 1 << 20, 1072812, 4 << 20, 16 << 20, 64 << 20, 256 << 20)
    if rank == 0:
        print(f"{world} rank(s); NUMA node of GPU {local}: {numa_node_of(local)}; host CPUs visible: {len(os.sched_getaffinity(0))}")
    for size in sizes:
        n = total // size
        for rep in range(2):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            with torch.cuda.stream(s):
                for i in range(n):
                    dev[i * size:(i + 1) * size].copy_(host[i * size:(i + 1) * size], non_blocking=True)
            t_issue = time.perf_counter() - t0
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        rate = n * size / dt / 1e9
        if world > 1:
            rates = torch.zeros(world, device="cuda")
            rates[rank] = rate
            dist.all_reduce(rates)
            if rank == 0:
                per = " ".join(f"{r:5.1f}" for r in rates.tolist())
                print(f"copy size {size / 2**20:8.2f} MiB x {n:5d}: per rank [{per}] GB/s, all ranks together {rates.sum().item():6.1f} GB/s")
        else:
            print(f"copy size {size / 2**20:8.2f} MiB x {n:5d}: {rate:6.1f} GB/s  (host issue {t_issue / n * 1e6:6.1f} us per copy)")
    if world == 1:
        # the same record-sized copies spread over several streams (copy engines): does the per-copy gap overlap?
        size = 1072812
        n = total // size
        for k in (1, 2, 3, 4):
            streams = [torch.cuda.Stream() for _ in range(k)]
            for rep in range(2):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for i in range(n):
                    with torch.cuda.stream(streams[i % k]):
                        dev[i * size:(i + 1) * size].copy_(host[i * size:(i + 1) * size], non_blocking=True)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
            print(f"copy size {size / 2**20:8.2f} MiB x {n:5d} over {k} stream(s): {n * size / dt / 1e9:6.1f} GB/s")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
