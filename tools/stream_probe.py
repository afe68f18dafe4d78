#!/usr/bin/env python
"""Where does a streaming micro-batch spend its time?  Submits fp32 299x299x3 records through ftf_submit/ftf_poll and logs,
per completed batch, when its results became visible to the host."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
import ftf_loader


def main():
    ftf = ftf_loader.load()
    lib = ftf._abi.load_library()
    from flink_tensorflow_b200.host import FtfTensor, _check
    INC = bench.INC
    batches = int(os.environ.get("BATCHES", "100"))
    depth = int(os.environ.get("DEPTH", "3"))               # full batches kept outstanding
    cfg = ftf.default_config(device=0, max_batch=INC["batch"], max_delay_us=2000, top_k=INC["top_k"], out_h=299, out_w=299,
                             mean=128.0, scale=128.0, num_classes=INC["classes"], seed=INC["seed"])
    r = ftf.Runner("builtin:inception_v3", None, None, cfg)
    rec_bytes = 299 * 299 * 3 * 4
    n_buf = (depth + 1) * INC["batch"]
    p, ring = bench.pinned_array(lib, np, n_buf * rec_bytes, np.float32)
    ring[:] = np.random.default_rng(3).standard_normal(ring.size, dtype=np.float32)
    base = ring.view(np.uint8).ctypes.data
    dims = (C.c_int64 * 3)(299, 299, 3)
    recs = [FtfTensor(ftf.DataType.FLOAT, 3, dims, base + i * rec_bytes, rec_bytes) for i in range(n_buf)]
    keys = (C.c_char_p * 2)(b"top_k_indices", b"top_k_scores")
    tags = (C.c_uint64 * 256)()
    outs = (FtfTensor * 512)()
    n_ready = C.c_int32()
    done = 0
    log = []

    def drain(block):
        nonlocal done
        _check(lib.ftf_poll(r.handle, 1 if block else 0, tags, outs, 256, C.byref(n_ready)), r.handle)
        if n_ready.value:
            done += n_ready.value
            log.append((time.perf_counter(), done))

    n = batches * INC["batch"]
    for i in range(2 * INC["batch"]):
        _check(lib.ftf_submit(r.handle, i, C.byref(recs[i % n_buf])), r.handle)
        if i == 0:
            _check(lib.ftf_poll_select(r.handle, 2, keys), r.handle)
    _check(lib.ftf_flush(r.handle), r.handle)
    while done < 2 * INC["batch"]:
        drain(True)
    done = 0
    log.clear()
    t0 = time.perf_counter()
    t_submit = 0.0
    slow = []
    for i in range(n):
        while i - done >= n_buf - INC["batch"]:
            drain(True)
        ts = time.perf_counter()
        _check(lib.ftf_submit(r.handle, i, C.byref(recs[i % n_buf])), r.handle)
        d_sub = time.perf_counter() - ts
        t_submit += d_sub
        if d_sub > 0.5e-3:
            slow.append((i, round(d_sub * 1e3, 2), round((ts - t0) * 1e3, 1)))
        if i % 16 == 15:
            drain(False)
    _check(lib.ftf_flush(r.handle), r.handle)
    while done < n:
        drain(True)
    dt = time.perf_counter() - t0
    print(f"{n} records in {dt * 1e3:.1f} ms -> {n / dt:.0f} records/s; {dt / batches * 1e3:.3f} ms per batch; "
          f"submit calls took {t_submit / n * 1e6:.1f} us each ({t_submit / batches * 1e3:.3f} ms per batch)")
    gaps = [(log[k][0] - log[k - 1][0]) * 1e3 for k in range(1, len(log))]
    gaps.sort()
    print(f"poll events {len(log)}, gap between result batches: median {gaps[len(gaps) // 2]:.3f} ms, p10 {gaps[len(gaps) // 10]:.3f}, p90 {gaps[len(gaps) * 9 // 10]:.3f}, max {gaps[-1]:.3f}")
    print('slow submits (index, ms, at ms):', slow[:40])
    print('first result batches at ms:', [round((t - t0) * 1e3, 1) for t, _ in log[:6]], 'last:', [round((t - t0) * 1e3, 1) for t, _ in log[-3:]])
    st = r.stats()
    print({k: st[k] for k in st if "latency" in k or "batches" in k or "run" in k})
    r.close()


if __name__ == "__main__":
    main()
