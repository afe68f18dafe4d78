#!/usr/bin/env python
"""J1-J3 decode time by batch size (files of bench.py's jpeg workload): what a 64-file micro-batch costs against 256."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import ftf_loader

ftf_loader.load()


def main():
    from flink_tensorflow_b200 import device
    torch.cuda.set_device(0)
    for n in (16, 32, 64, 128, 256):
        files = bench.jpeg_files(np, 7, n)
        jb = device.JpegBatch(0).load(files)
        out = torch.empty(jb.rgb_bytes, dtype=torch.uint8, device="cuda")
        for _ in range(3):
            jb.decode(out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            jb.decode(out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"{n:4d} files: {ms * 1e3:8.1f} us per decode, {ms * 1e3 / n:6.2f} us per file, {n / ms:8.1f} k files/s")


if __name__ == "__main__" and len(sys.argv) == 1:
    main()


def streaming():
    """The same files as records through ftf_submit/ftf_poll on builtin:normalize_label, micro-batches of 64 and of 256."""
    import ctypes as C  # noqa: F401
    ftf = ftf_loader.load()
    lib = ftf._abi.load_library()
    for mb in (64, 128, 256):
        res = bench.streaming_bench(ftf, lib, np, 0, n_records=mb * 40, kind="jpeg_label", max_batch=mb)
        print(f"micro-batches of {mb:3d} JPEG records: {res['records_per_s']:9.0f} records/s, "
              f"p50 {res['latency_us_p50'] / 1e3:.2f} ms, p99 {res['latency_us_p99'] / 1e3:.2f} ms, {res['host_us_in_submit_per_record']} us per record inside ftf_submit "
              f"of {1e6 / res['records_per_s']:.1f} us per record")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "streaming":
    streaming()
