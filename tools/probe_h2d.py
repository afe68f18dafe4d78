#!/usr/bin/env python
"""Pinned host -> device copy bandwidth by copy size (what one record, one micro-batch cost on this box's PCIe link)."""
import time
import torch

def main():
    total = 256 << 20
    host = torch.empty(total, dtype=torch.uint8).pin_memory()
    dev = torch.empty(total, dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    for size in (256 << 10, 1 << 20, 1072812, 4 << 20, 16 << 20, 64 << 20, 256 << 20):
        n = total // size
        for rep in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with torch.cuda.stream(s):
                for i in range(n):
                    dev[i * size:(i + 1) * size].copy_(host[i * size:(i + 1) * size], non_blocking=True)
            t_issue = time.perf_counter() - t0
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        print(f"copy size {size / 2**20:8.2f} MiB x {n:5d}: {n * size / dt / 1e9:6.1f} GB/s  (host issue {t_issue / n * 1e6:6.1f} us per copy)")

if __name__ == "__main__":
    main()
