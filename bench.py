#!/usr/bin/env python
"""bench.py -- one JSON line per run (see the round contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload half_plus_two|normalize|inception|jpeg]
  python bench.py --impl reference ...          the reference's CPU path (oracle stand-in) on the host cores
  torchrun --nproc-per-node N bench.py --gpus N one rank per GPU, records sharded, no collective on the data path

A step is one pass of the hot path over one batch of synthetic input:
  half_plus_two  BASELINE config 2: y = 0.5x + 2 over 2^28 fp32 (1 GiB in, 1 GiB out), 8 algorithmic B/element
  normalize      BASELINE config 3: 256 x (480x640 u8 RGB -> 299x299 f32), 1 994 412 algorithmic B/image
  inception      BASELINE config 4: Inception-v3 random-init bf16, batch 64, top-5 labels
`value` is device-resident throughput (inputs already in HBM), `e2e` the same metric through ftf_run with host
buffers (H2D + D2H inside the timed region), `roofline` the dominant kernel against MEASURED_PEAKS.json,
`cpu_baseline` the CPU oracle timed on this box's cores on a bounded sample.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HPT_N = 1 << 28
NORM = dict(batch=256, in_h=480, in_w=640, out_h=299, out_w=299, mean=128.0, scale=128.0)
NORM_BYTES_PER_IMAGE = 480 * 640 * 3 + 299 * 299 * 3 * 4          # 1 994 412
INC = dict(batch=64, side=299, classes=1000, top_k=5, seed=2024)
INC_FLOP_PER_IMAGE = 11.43e9                                       # SURVEY 8d (2*MAC, 94 convs + FC)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region through NVML (the same counters as the
    nvidia-smi clocks line of B200_PROFILING.md, at a 2 ms period so that short timed regions are covered)."""

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None
        self.sm_max = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                  nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=2)

    def summary(self):
        if self.nv is None or not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        nv = self.nv
        sm = sorted(r[0] for r in self.rows)
        bits = 0
        for _, r in self.rows:
            bits |= r
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.sm_max, "reasons": sorted(k for k, v in names.items() if bits & v),
                "samples": len(sm)}


def dist_env():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    return rank, int(os.environ.get("LOCAL_RANK", rank)), world


# ================================================================================================
# reference arm: the reference's CPU implementation of the path (oracle stand-in for TF-Java Session.run)
# ================================================================================================
def torchvision_cpu(threads, unit, seconds=10.0):
    """Strongest CPU stand-in available offline: torchvision's Inception-v3 (same architecture, random init) on oneDNN,
    fp32, all host threads.  An extra beside the oracle number, never the thing the parity tests compare against."""
    try:
        import torch
        import torchvision
        tvm = torchvision.models.inception_v3(weights=None, aux_logits=False, transform_input=False, init_weights=False).eval()
        torch.set_num_threads(threads)
        tx = torch.randn(16, 3, 299, 299)
        with torch.no_grad():
            tvm(tx)
            t0, reps = time.perf_counter(), 0
            while time.perf_counter() - t0 < seconds:
                tvm(tx)
                reps += 1
        return {"value": 16 * reps / (time.perf_counter() - t0), "unit": unit, "cores": threads, "sample": f"{reps} passes over 16 images, fp32"}
    except Exception as ex:  # noqa: BLE001
        return {"unavailable": str(ex)[:200]}


JPEG = {"batch": 256, "h": 480, "w": 640, "quality": 85, "subsampling": "4:2:0", "distinct": 16}


def jpeg_files(np, seed, n=None):
    """Synthetic photo-like 640x480 images (smooth colour fields + sensor-like noise) written as baseline JPEG by Pillow
    (quality 85, 4:2:0, standard Huffman tables, no restart markers: what a camera or `convert` writes by default)."""
    import io
    from PIL import Image
    rng = np.random.default_rng(seed)
    h, w = JPEG["h"], JPEG["w"]
    y, x = np.mgrid[0:h, 0:w]
    files = []
    for _ in range(JPEG["distinct"]):
        p = rng.uniform(0, 6.28, 3)
        img = np.stack([128 + 90 * np.sin(x / rng.uniform(20, 60) + p[c]) * np.cos(y / rng.uniform(15, 50)) for c in range(3)], -1)
        img = np.clip(img + rng.normal(0, 10, (h, w, 3)), 0, 255).astype(np.uint8)
        b = io.BytesIO()
        Image.fromarray(img).save(b, "JPEG", quality=JPEG["quality"])
        files.append(b.getvalue())
    n = JPEG["batch"] if n is None else n
    return [files[i % len(files)] for i in range(n)]


def cpu_jpeg(np, oracle, files, threads, seconds):
    """The oracle's decoder over `files`, one file per task on `threads` host threads (ctypes drops the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(oracle.jpeg_decode, files[:threads]))
        t0, done = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds:
            list(ex.map(oracle.jpeg_decode, files))
            done += len(files)
        return done / (time.perf_counter() - t0)


def cpu_pillow(np, files, threads, seconds):
    """libjpeg-turbo (SIMD) through Pillow on all host threads: the library TF's DecodeJpeg wraps, the strongest CPU stand-in."""
    import io
    from concurrent.futures import ThreadPoolExecutor
    from PIL import Image
    dec = lambda f: np.asarray(Image.open(io.BytesIO(f)).convert("RGB")).shape
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(dec, files[:threads]))
        t0, done = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds:
            list(ex.map(dec, files))
            done += len(files)
        return done / (time.perf_counter() - t0)


def run_reference(args):
    import numpy as np
    import oracle
    rank, _, _ = dist_env()
    if rank != 0:
        return
    threads = oracle.num_threads()
    rng = np.random.default_rng(1234)
    wl = args.workload
    warm = None
    if wl == "half_plus_two":
        n = 1 << 26                                           # bounded sample: 1/4 of the 2^28 workload per step
        x = (rng.standard_normal(n) * 100).astype(np.float32)
        step = lambda: oracle.affine_f32(x, 0.5, 2.0, threads=threads)
        units, unit, metric, sample = 8.0 * n / 1e9, "GB/s", "half_plus_two GB/s", f"2^26 of the 2^28 floats per step"
        config = {"workload": "half_plus_two y=0.5x+2 over 2^28 fp32 (BASELINE config 2)", "sample_elements": n}
        dtype = "f32"
    elif wl == "normalize":
        b = 32
        img = rng.integers(0, 256, (b, NORM["in_h"], NORM["in_w"], 3), dtype=np.uint8)
        step = lambda: oracle.resize_bilinear_norm_u8(img, NORM["out_h"], NORM["out_w"], NORM["mean"], NORM["scale"], threads=threads)
        units, unit, metric, sample = float(b), "images/s", "ImageNormalization images/sec", "32 of the 256 images per step"
        config = {"workload": "ImageNormalization 640x480 -> 299x299 batch 256 (BASELINE config 3)", "sample_images": b}
        dtype = "f32"
    elif wl == "jpeg":
        from concurrent.futures import ThreadPoolExecutor
        files = jpeg_files(np, 1234, 64)
        pool = ThreadPoolExecutor(threads)
        step = lambda: list(pool.map(oracle.jpeg_decode, files))
        units, unit, metric, sample = float(len(files)), "images/s", "DecodeJpeg images/sec", "64 of the 256 files per step"
        config = {"workload": "DecodeJpeg 640x480 baseline JPEG (quality 85, 4:2:0) batch 256", "sample_images": len(files)}
        dtype = "u8"
    else:
        import oracle.inception as oinc                       # noqa: F401  (exists once the conv path lands)
        net = oinc.Net(INC["classes"], INC["seed"])
        b = INC["batch"]                                      # every timed step is the whole batch-64 workload (a few seconds)
        img = rng.standard_normal((b, 299, 299, 3)).astype(np.float32)
        step = lambda: net.forward(img, threads=threads)
        warm = lambda: net.forward(img[:4], threads=threads)  # warm-up only has to generate and cache the weights
        units, unit, metric, sample = float(b), "images/s", "Inception-v3 images/sec", "all 64 images per timed step, fp32 (warm-up steps: 4 images)"
        config = {"workload": "Inception-v3 random-init batch 64 top-5 (BASELINE config 4)", "sample_images": b}
        dtype = "f32"
    for _ in range(args.warmup):
        (warm or step)()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = units * args.steps / dt
    line = {"impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
            "config": config,
            "cpu_baseline": {"value": value, "unit": unit, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "CPU oracle (C restatement of the TF-1.1 op sequence) stands in for TF-Java Session.run, which "
                    "cannot run without a JVM; it omits the reference's per-record JNI/dispatch overhead"}
    if wl == "inception":
        line["cpu_torchvision_onednn"] = torchvision_cpu(threads, unit)
    if wl == "jpeg":
        line["cpu_pillow_libjpeg_turbo"] = {"value": cpu_pillow(np, files, threads, 5.0), "unit": unit, "cores": threads}
    print(json.dumps(line), flush=True)


# ================================================================================================
# our arm
# ================================================================================================
def cuda_time(fn, steps, torch):
    """K back-to-back steps between two events on the current stream; returns total milliseconds."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def pinned_array(lib, np, nbytes, dtype):
    p = lib.ftf_host_alloc(nbytes)
    if not p:
        raise MemoryError("ftf_host_alloc failed")
    return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,)).view(dtype)


def streaming_bench(ftf, lib, np, device, n_records=64 * 40, kind="normalize_label", sync=None, wire=False, offered_rate=None, max_batch=None):
    """Records through ftf_submit/ftf_poll at saturation (micro-batches of 64, 2 ms deadline); latency = submit -> result
    polled.  kind "normalize_label": BASELINE config 5 on one GPU, raw 640x480 RGB records (normalise -> Inception-v3 ->
    top-5 on the device).  kind "inception": config 4's records, normalised 299x299x3 fp32 images (Inception-v3 -> top-5),
    the per-record map() the reference's InceptionModel.label runs.
    wire: every record sits in the pinned ring as a TensorValue wire image (TensorValue.java:142-187) and is decoded in place
    (ftf_tv_decode, no copy) before it is submitted.  offered_rate: records/s the source emits (None = saturation)."""
    from flink_tensorflow_b200.host import _as_ftf, FtfTensor, _check
    max_batch = max_batch or INC["batch"]
    cfg = ftf.default_config(device=device, max_batch=max_batch, max_delay_us=2000, top_k=INC["top_k"], out_h=299, out_w=299,
                             mean=128.0, scale=128.0, num_classes=INC["classes"], seed=INC["seed"])
    n_buf = 4 * max_batch                                   # records stay untouched until polled: a ring of pinned payloads
    if kind == "jpeg_label":
        # BASELINE config 5 with the records the reference's source emits: image FILES (ImageInputFormat.scala:63-80 reads the
        # file, ImageFileTensor wraps its bytes), here 640x480 baseline JPEGs as 0-D STRING tensors in pinned memory
        r = ftf.Runner("builtin:normalize_label", None, None, cfg)
        blobs = [ftf.bytes_to_tensor(f).buffer for f in jpeg_files(np, 3, JPEG["distinct"])]
        stride = (max(b.nbytes for b in blobs) + 255) // 256 * 256
        p, ring = pinned_array(lib, np, n_buf * stride, np.uint8)
        recs = []
        for i in range(n_buf):
            b = blobs[i % len(blobs)]
            ring[i * stride: i * stride + b.nbytes] = b
            recs.append(FtfTensor(ftf.DataType.STRING, 0, None, ring.ctypes.data + i * stride, b.nbytes))
        rec_bytes = int(sum(b.nbytes for b in blobs) / len(blobs))
        pipeline = "640x480 JPEG file -> Huffman/IDCT/colour -> resize+normalise (bf16) -> Inception-v3 -> softmax/top-5"
    elif kind == "normalize_label":
        r = ftf.Runner("builtin:normalize_label", None, None, cfg)
        rec_bytes = 480 * 640 * 3
        p, ring = pinned_array(lib, np, n_buf * rec_bytes, np.uint8)
        ring[:] = np.random.default_rng(3).integers(0, 256, ring.size, dtype=np.uint8)
        dims = (C.c_int64 * 3)(480, 640, 3)
        dtype, pipeline = ftf.DataType.UINT8, "u8 640x480 RGB -> resize+normalise (bf16) -> Inception-v3 -> softmax/top-5"
    else:
        r = ftf.Runner("builtin:inception_v3", None, None, cfg)
        rec_bytes = 299 * 299 * 3 * 4
        p, ring32 = pinned_array(lib, np, n_buf * rec_bytes, np.float32)
        ring32[:] = np.random.default_rng(3).standard_normal(ring32.size, dtype=np.float32)
        ring = ring32.view(np.uint8)
        dims = (C.c_int64 * 3)(299, 299, 3)
        dtype, pipeline = ftf.DataType.FLOAT, "fp32 299x299x3 -> bf16 -> Inception-v3 -> softmax/top-5"
    if kind != "jpeg_label":
        recs = [FtfTensor(dtype, 3, dims, ring.ctypes.data + i * rec_bytes, rec_bytes) for i in range(n_buf)]
    wire_p = None
    if wire:
        # the same payloads as Flink wire images in a second pinned ring: 0x01 | BE header | payload (37 header bytes here)
        wlen = int(lib.ftf_tv_binary_length(3, rec_bytes))
        wstride = (wlen + 255) // 256 * 256
        wire_p, wring = pinned_array(lib, np, n_buf * wstride, np.uint8)
        written = C.c_int64()
        for i in range(n_buf):
            _check(lib.ftf_tv_encode(C.byref(recs[i]), wring.ctypes.data + i * wstride, wlen, C.byref(written)), r.handle)
            assert written.value == wlen
        wire_ptrs = [wring.ctypes.data + i * wstride for i in range(n_buf)]
        dims_buf = (C.c_int64 * 32)()
        decoded = FtfTensor()

    def submit(i):
        slot = i % n_buf
        while True:
            if wire:                                           # parse the wire image where it lies; the payload is not copied
                _check(lib.ftf_tv_decode(wire_ptrs[slot], wlen, C.byref(decoded), dims_buf), r.handle)
                rc = lib.ftf_submit(r.handle, i, C.byref(decoded))
            else:
                t_in = time.perf_counter()
                rc = lib.ftf_submit(r.handle, i, C.byref(recs[slot]))
                in_submit[0] += time.perf_counter() - t_in
            if rc != 8:                                        # RESOURCE_EXHAUSTED: every slot holds undelivered results
                _check(rc, r.handle)
                return
            drain(True)                                        # take what is ready, then offer the record again
    keys = (C.c_char_p * 2)(b"top_k_indices", b"top_k_scores")
    tags = (C.c_uint64 * 256)()
    outs = (FtfTensor * 512)()
    n_ready = C.c_int32()
    done = 0

    def drain(block):
        nonlocal done
        _check(lib.ftf_poll(r.handle, 1 if block else 0, tags, outs, 256, C.byref(n_ready)), r.handle)
        for q in range(n_ready.value):                         # the host reads every result: top-1 index of each record
            checksum[0] += C.cast(outs[2 * q].data, C.POINTER(C.c_int32))[0]
        done += n_ready.value
    checksum = [0]
    in_submit = [0.0]                                        # seconds the source thread spends inside ftf_submit
    # warm-up (graph capture, allocations)
    n_warm = 4 * INC["batch"]                                # every slot has run once before the clock starts
    for i in range(n_warm):
        while i - done >= n_buf - INC["batch"]:
            drain(True)
        submit(i)
        if i == 0:
            _check(lib.ftf_poll_select(r.handle, 2, keys), r.handle)
    _check(lib.ftf_flush(r.handle), r.handle)
    while done < n_warm:
        drain(True)
    done = 0
    checksum[0] = 0
    in_submit[0] = 0.0
    if sync:
        sync()
    r2 = r.stats()
    t0 = time.perf_counter()
    period = 1.0 / offered_rate if offered_rate else 0.0
    for i in range(n_records):
        while i - done >= n_buf - INC["batch"]:                # back-pressure: never overwrite an unpolled record
            drain(True)
        if period:                                             # a source that emits at a fixed rate: poll while waiting
            due = t0 + i * period
            while time.perf_counter() < due:
                drain(False)
        submit(i)
        if i % 16 == 15:
            drain(False)
    _check(lib.ftf_flush(r.handle), r.handle)
    while done < n_records:
        drain(True)
    dt = time.perf_counter() - t0
    st = r.stats()
    r.close()
    lib.ftf_host_free(p)
    if wire_p:
        lib.ftf_host_free(wire_p)
    return {"records_per_s": n_records / dt, "offered_records_per_s": offered_rate,
            "record_format": "TensorValue wire image, decoded in place" if wire else ("JPEG file as a 0-D STRING tensor" if kind == "jpeg_label" else "raw tensor"), "records": n_records, "latency_us_p50": st["queue_latency_us_p50"],
            "latency_us_p99": st["queue_latency_us_p99"], "max_batch": max_batch, "max_delay_us": 2000,
            "h2d_bytes_per_record": rec_bytes, "d2h_bytes_per_record": 2 * INC["top_k"] * 4, "pipeline": pipeline, "seconds": dt,
            "api": "ftf_submit/ftf_poll, pinned record payloads", "launches": int(st["kernel_launches"] - r2["kernel_launches"]),
            "checksum": int(checksum[0]), "host_us_in_submit_per_record": None if wire else round(in_submit[0] / n_records * 1e6, 2)}


def blocking_two_subtasks(ftf, lib, np, device, seconds=1.5):
    """Two task threads, each with its own label runner on the same GPU (two Flink slots per GPU), each calling the blocking
    ftf_run over a stacked [64,299,299,3] fp32 tensor in pinned memory: one thread's upload overlaps the other's kernels."""
    import threading
    from flink_tensorflow_b200.host import _as_ftf, FtfTensor, _check
    nb = INC["batch"] * 299 * 299 * 3 * 4
    runners, bufs, counts = [], [], [0, 0]
    for t in range(2):
        cfg = ftf.default_config(device=device, max_batch=INC["batch"], top_k=INC["top_k"], num_classes=INC["classes"], seed=INC["seed"])
        runners.append(ftf.Runner("builtin:inception_v3", None, None, cfg))
        p, hx = pinned_array(lib, np, nb, np.float32)
        hx[:] = np.random.default_rng(11 + t).standard_normal(nb // 4, dtype=np.float32)
        bufs.append((p, hx))
    stop = threading.Event()
    go = threading.Barrier(3)

    def task(t):
        r, (p, hx) = runners[t], bufs[t]
        t_in, keep = _as_ftf(ftf.DataType.FLOAT, (INC["batch"], 299, 299, 3), hx.view(np.uint8))
        ik, ok = (C.c_char_p * 1)(b"inputs"), (C.c_char_p * 1)(b"top_k_indices")
        tin, tout = (FtfTensor * 1)(t_in), (FtfTensor * 1)()
        for _ in range(2):
            _check(lib.ftf_run(r.handle, 1, ik, tin, 1, ok, tout), r.handle)
        go.wait()
        while not stop.is_set():
            _check(lib.ftf_run(r.handle, 1, ik, tin, 1, ok, tout), r.handle)
            counts[t] += 1
    th = [threading.Thread(target=task, args=(t,)) for t in range(2)]
    for x in th:
        x.start()
    go.wait()
    t0 = time.perf_counter()
    time.sleep(seconds)
    stop.set()
    for x in th:
        x.join()
    dt = time.perf_counter() - t0
    for r in runners:
        r.close()
    for p, _ in bufs:
        lib.ftf_host_free(p)
    return {"value": sum(counts) * INC["batch"] / dt, "unit": "images/s", "calls": sum(counts), "seconds": dt,
            "api": "ftf_run (C ABI) from two task threads with one runner each on one GPU (two Flink slots per GPU)"}


def sub_leg(name, torch, device, np, lib, pk, rank, local_rank, min_ms=60.0):
    """A compact device-resident measurement of one of the HBM-bound kernels (BASELINE configs 2 and 3), run inside the
    default bench so that the driver's record holds it: value, ms per step, roofline fraction against the measured copy
    rate, DRAM traffic from profiles/traffic.json, clocks sampled during the timed region (it lasts >= min_ms so that the
    2 ms sampler sees >= 10 samples)."""
    if name == "half_plus_two":
        g = torch.Generator(device="cuda").manual_seed(1234 + rank)
        x = torch.randn(HPT_N, device="cuda", generator=g) * 100
        y = torch.empty_like(x)
        step = lambda: device.affine_f32(x, 0.5, 2.0, out=y)
        units, unit, algo = 8.0 * HPT_N / 1e9, "GB/s", 8.0 * HPT_N
        what = "y = 0.5x + 2 over 2^28 fp32, 8 algorithmic B/element (BASELINE config 2)"
    else:
        b = NORM["batch"]
        g = torch.Generator(device="cuda").manual_seed(42 + rank)
        img = torch.randint(0, 256, (b, NORM["in_h"], NORM["in_w"], 3), dtype=torch.uint8, device="cuda", generator=g)
        bf16 = name == "normalize_bf16"                 # the output the label chain really uses (K2 in front of Inception-v3)
        out = torch.empty((b, NORM["out_h"], NORM["out_w"], 3), dtype=torch.bfloat16 if bf16 else torch.float32, device="cuda")
        step = lambda: device.resize_norm(img, NORM["out_h"], NORM["out_w"], NORM["mean"], NORM["scale"], out=out)
        per_image = NORM["in_h"] * NORM["in_w"] * 3 + NORM["out_h"] * NORM["out_w"] * 3 * (2 if bf16 else 4)
        units, unit, algo = float(b), "images/s", float(per_image) * b
        what = ("256 x (480x640 u8 RGB -> 299x299 bf16), 1 458 006 algorithmic B/image (config 3's kernel as the label chain runs it)" if bf16
                else "256 x (480x640 u8 RGB -> 299x299 f32), 1 994 412 algorithmic B/image (BASELINE config 3)")
    for _ in range(5):
        step()
    ms1 = cuda_time(step, 10, torch) / 10
    steps = max(20, int(min_ms / max(ms1, 1e-3)) + 1)
    with ClockSampler(local_rank) as clk:
        ms = cuda_time(step, steps, torch) / steps
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(name, {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    achieved = algo / (ms / 1e3) / 1e9
    return {"workload": what, "value": units / (ms / 1e3), "unit": unit, "ms_per_step": ms, "steps": steps,
            "l2": "tensors of one step exceed L2, no flush",
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"],
                         "frac_of_nominal_8TBps": achieved / 8000.0, "traffic": traffic, "peak_source": pk["src"]},
            "clocks": clk.summary()}


def per_record_leg(ftf, lib, np, local_rank, lab, seconds=2.0):
    """The reference's real calling pattern (BASELINE config 1): ONE record per model call -- one tf.Example per
    regress_x_to_y call (FTT/ml/RegressITCase.scala:54) and one image per InceptionModel.label call
    (EX/inception/inception.scala:39-46) -- through the blocking ftf_run, host buffers in, result read by the host."""
    from flink_tensorflow_b200.host import _as_ftf, FtfTensor, _check
    out = {}
    r = ftf.Runner(os.path.join(ROOT, "models", "half_plus_two"), "regress_x_to_y", None, ftf.default_config(device=local_rank))
    rng = np.random.default_rng(1234)
    xs = rng.uniform(-1000, 1000, 4096).astype(np.float32)
    recs = [ftf.messages_to_tensor([ftf.example(x=float(v))]) for v in xs]
    ik, ok = (C.c_char_p * 1)(b"inputs"), (C.c_char_p * 1)(b"outputs")
    tout = (FtfTensor * 1)()
    prepared = []
    for tv in recs:
        t, keep = _as_ftf(tv.data_type, tv.shape, tv.buffer)
        prepared.append(((FtfTensor * 1)(t), keep))
    lat, wrong, t0, i = [], 0, time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        tin, _ = prepared[i % len(prepared)]
        a = time.perf_counter()
        _check(lib.ftf_run(r.handle, 1, ik, tin, 1, ok, tout), r.handle)
        y = C.cast(tout[0].data, C.POINTER(C.c_float))[0]
        lat.append(time.perf_counter() - a)
        wrong += y != np.float32(np.float32(0.5) * xs[i % len(xs)]) + np.float32(2.0)
        i += 1
    dt = time.perf_counter() - t0
    r.close()
    lat = np.sort(np.array(lat)) * 1e6
    out["half_plus_two"] = {"calls_per_s": i / dt, "latency_us_p50": float(lat[len(lat) // 2]), "latency_us_p99": float(lat[int(len(lat) * 0.99)]),
                            "calls": i, "wrong_answers": int(wrong), "api": "one serialized tf.Example per ftf_run on regress_x_to_y (parse, K1 on the mapped pinned buffers: no copies)"}
    # one image per label call on the batch-64 runner (a batch of one replays the 8-image graph bucket)
    nb = 299 * 299 * 3 * 4
    p, hx = pinned_array(lib, np, nb, np.float32)
    hx[:] = rng.standard_normal(nb // 4, dtype=np.float32)
    t_in, keep = _as_ftf(ftf.DataType.FLOAT, (1, 299, 299, 3), hx.view(np.uint8))
    tin, ok2 = (FtfTensor * 1)(t_in), (C.c_char_p * 1)(b"top_k_indices")
    for _ in range(3):
        _check(lib.ftf_run(lab.handle, 1, ik, tin, 1, ok2, tout), lab.handle)
    lat, t0, i = [], time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        a = time.perf_counter()
        _check(lib.ftf_run(lab.handle, 1, ik, tin, 1, ok2, tout), lab.handle)
        _ = C.cast(tout[0].data, C.POINTER(C.c_int32))[0]
        lat.append(time.perf_counter() - a)
        i += 1
    dt = time.perf_counter() - t0
    lib.ftf_host_free(p)
    lat = np.sort(np.array(lat)) * 1e6
    out["inception_label"] = {"calls_per_s": i / dt, "latency_us_p50": float(lat[len(lat) // 2]), "latency_us_p99": float(lat[int(len(lat) * 0.99)]),
                              "calls": i, "api": "one fp32 299x299x3 image per ftf_run on builtin:inception_v3 (upload 1.07 MB, network, top-5 back)"}
    return out


def per_record_cpu(ftf, np, oracle, seconds=2.0):
    """The CPU oracle called the way the reference calls TF: one serialized Example per call (STRING tensor -> ParseExample ->
    a*x + b), one thread, plus one image per call through the oracle's Inception-v3 on all host threads."""
    rng = np.random.default_rng(1234)
    xs = rng.uniform(-1000, 1000, 4096).astype(np.float32)
    recs = [oracle.string_tensor_encode([ftf.example(x=float(v))]) for v in xs]
    t0, i = time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        rc, y = oracle.regress(recs[i % len(recs)], 1, "x", 0.5, 2.0)
        i += 1
    dt = time.perf_counter() - t0
    out = {"half_plus_two": {"calls_per_s": i / dt, "cores": 1, "kind": "port",
                             "sample": f"{i} one-Example calls of the C oracle through ctypes (no JNI, no TF executor dispatch: "
                                       "faster than the reference's Session.run could be)"}}
    import oracle.inception as oinc
    net = oinc.Net(INC["classes"], INC["seed"], 0)
    img = rng.standard_normal((1, 299, 299, 3)).astype(np.float32)
    threads = oracle.num_threads()
    net.forward(img, threads=threads)
    t0, i = time.perf_counter(), 0
    while time.perf_counter() - t0 < 2 * seconds:
        net.forward(img, threads=threads)
        i += 1
    out["inception_label"] = {"calls_per_s": i / (time.perf_counter() - t0), "cores": threads, "kind": "port",
                              "sample": f"{i} one-image calls of the fp32 C oracle, all host threads"}
    return out


def outputs_digest(ftf, lib, np, torch, lab):
    """Top-5 and logits of ONE common seeded batch (the same on every rank), as 8 int64 words of two SHA-256 digests."""
    import hashlib
    from flink_tensorflow_b200.host import _check
    bsz = INC["batch"]
    g = torch.Generator(device="cuda").manual_seed(4242)                  # not rank-dependent
    img = torch.randn((bsz, 299, 299, 3), device="cuda", generator=g).to(torch.bfloat16)
    logits = torch.empty((bsz, INC["classes"]), dtype=torch.float32, device="cuda")
    idx = torch.empty((bsz, INC["top_k"]), dtype=torch.int32, device="cuda")
    score = torch.empty((bsz, INC["top_k"]), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    _check(lib.ftf_label_dev(lab.handle, img.data_ptr(), ftf._abi.BFLOAT16, bsz, logits.data_ptr(), None, idx.data_ptr(), score.data_ptr()), lab.handle)
    torch.cuda.synchronize()
    words = []
    for t in (idx, logits):
        d = hashlib.sha256(t.cpu().numpy().tobytes()).digest()
        words += [int.from_bytes(d[8 * k: 8 * k + 8], "little", signed=True) for k in range(4)]
    return words



def bind_to_gpu_numa_node(torch, device):
    """Runs this rank (and so its pinned staging buffers, which are placed on first touch) on the CPUs of the NUMA node its
    GPU hangs off (the one shared resource of the record-sharded path is host memory / PCIe).  With eight ranks
    feeding 44 GB/s each, records that sit on the other socket cross the inter-socket link on their way to the GPU.
    Returns what it did, for the bench line."""
    if os.environ.get("FTF_BENCH_NO_NUMA"):
        return "off (FTF_BENCH_NO_NUMA)"
    try:
        pr = torch.cuda.get_device_properties(device)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read())
        if node < 0:
            return "unbound (numa_node = -1: the box exposes no NUMA topology for this GPU)"
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "no local CPUs allowed"
        os.sched_setaffinity(0, cpus)
        return f"bound to node {node}, {len(cpus)} cpus"
    except Exception as e:  # noqa: BLE001  (no sysfs entry, container without the attribute: run unbound)
        return f"unavailable ({type(e).__name__})"


def run_ours(args):
    import numpy as np
    import torch
    import ftf_loader
    ftf = ftf_loader.load()
    from flink_tensorflow_b200 import device
    lib = ftf._abi.load_library()
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the runner has no CPU path")
    torch.cuda.set_device(local_rank)
    # one rank alone has the host to itself (and the CPU baseline below wants every core): bind only when ranks share it
    numa = bind_to_gpu_numa_node(torch, local_rank) if world > 1 else "not bound (single rank)"
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pk = peaks()
    wl = args.workload
    K, Wm = args.steps, max(3, args.warmup)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from flink_tensorflow_b200 import sharding

    def max_over_ranks(ms):
        return sharding.max_over_ranks(ms, device="cuda")

    from flink_tensorflow_b200.host import _check as _chk
    launches0 = lib.ftf_launch_count()
    ext = None
    extra = {}
    if wl == "half_plus_two":
        n = HPT_N
        g = torch.Generator(device="cuda").manual_seed(1234 + rank)
        x = torch.randn(n, device="cuda", generator=g) * 100
        y = torch.empty_like(x)
        step = lambda: device.affine_f32(x, 0.5, 2.0, out=y)
        units_per_step, unit, metric, dtype = 8.0 * n / 1e9, "GB/s", "half_plus_two GB/s", "f32"
        algo_bytes = 8.0 * n
        config = {"workload": "half_plus_two y=0.5x+2 over 2^28 fp32 per GPU (BASELINE config 2)", "elements": n,
                  "l2": "inputs (1 GiB) larger than L2, no flush", "parallelism": f"dp{world} (records sharded, no collective)"}
        bound, peak, peak_unit = "hbm", pk["hbm"], "GB/s"
    elif wl == "normalize":
        b = NORM["batch"]
        g = torch.Generator(device="cuda").manual_seed(42 + rank)
        img = torch.randint(0, 256, (b, NORM["in_h"], NORM["in_w"], 3), dtype=torch.uint8, device="cuda", generator=g)
        out = torch.empty((b, NORM["out_h"], NORM["out_w"], 3), dtype=torch.float32, device="cuda")
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        step = lambda: device.resize_norm(img, NORM["out_h"], NORM["out_w"], NORM["mean"], NORM["scale"], out=out)
        units_per_step, unit, metric, dtype = float(b), "images/s", "ImageNormalization images/sec", "f32"
        algo_bytes = float(NORM_BYTES_PER_IMAGE) * b
        config = {"workload": "ImageNormalization 640x480 u8 RGB -> ResizeBilinear 299x299 + mean/scale, batch 256 per GPU "
                              "(BASELINE config 3)", "l2": "input+output 511 MB per step, larger than L2",
                  "parallelism": f"dp{world} (records sharded, no collective)", **NORM}
        bound, peak, peak_unit = "hbm", pk["hbm"], "GB/s"
    elif wl == "jpeg":
        files = jpeg_files(np, 1234 + rank)
        jb = device.JpegBatch(local_rank)
        jb.load(files)
        rgb = torch.empty(jb.rgb_bytes, dtype=torch.uint8, device="cuda")
        step = lambda: jb.decode(out=rgb)
        step(); jb.status()
        units_per_step, unit, metric, dtype = float(len(files)), "images/s", "DecodeJpeg images/sec", "u8"
        file_bytes = sum(len(f) for f in files)
        algo_bytes = float(file_bytes + jb.rgb_bytes)
        config = {"workload": "DecodeJpeg: 640x480 baseline JPEG files (quality 85, 4:2:0, no restart markers) -> RGB u8, batch 256 "
                              "per GPU (the node in front of BASELINE config 3's resize)", "file_bytes_per_image": file_bytes // len(files),
                  "segments": jb.segments, "l2": "coefficients + planes + RGB of one step (about 590 MB) exceed L2; no flush",
                  "parallelism": f"dp{world} (records sharded, no collective)", **JPEG}
        bound, peak, peak_unit = "hbm", pk["hbm"], "GB/s"
    else:
        bsz = INC["batch"]
        cfg = ftf.default_config(device=local_rank, max_batch=bsz, top_k=INC["top_k"], num_classes=INC["classes"], seed=INC["seed"])
        lab = ftf.Runner("builtin:inception_v3", None, None, cfg)
        ext = torch.cuda.ExternalStream(lib.ftf_runner_stream(lab.handle))
        g = torch.Generator(device="cuda").manual_seed(7 + rank)
        img = torch.randn((bsz, 299, 299, 3), device="cuda", generator=g).to(torch.bfloat16)
        logits = torch.empty((bsz, INC["classes"]), dtype=torch.float32, device="cuda")
        probs = torch.empty_like(logits)
        idx = torch.empty((bsz, INC["top_k"]), dtype=torch.int32, device="cuda")
        score = torch.empty((bsz, INC["top_k"]), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()

        def step():
            _chk(lib.ftf_label_dev(lab.handle, img.data_ptr(), ftf._abi.BFLOAT16, bsz, logits.data_ptr(), probs.data_ptr(),
                                     idx.data_ptr(), score.data_ptr()), lab.handle)
        units_per_step, unit, metric, dtype = float(bsz), "images/s", "Inception-v3 images/sec", "bf16"
        algo_bytes = None
        config = {"workload": "Inception-v3 random-init labeling, bf16 (fp32 accumulate), batch 64 per GPU, top-5 "
                              "(BASELINE config 4)", "l2": "activations of one step (about 1 GB written) exceed L2; no flush",
                  "parallelism": f"dp{world} (records sharded, no collective)", **INC}
        bound, peak, peak_unit = "tensor", pk["tf_sustained"], "TFLOP/s"

    # ---- device-resident timing ------------------------------------------------------------------
    for _ in range(Wm):
        step()
    barrier()
    launches0 = lib.ftf_launch_count()          # kernels of the timed steps only (not model set-up, plan-time timing or warm-up)
    # FTF_BENCH_CUDA_PROFILER=1: bracket the timed region with cudaProfilerStart/Stop, so that
    # `ncu --profile-from-start off` lists exactly the launches of the timed steps (not the plan-time tuning launches)
    prof_range = os.environ.get("FTF_BENCH_CUDA_PROFILER") == "1"
    if prof_range:
        torch.cuda.cudart().cudaProfilerStart()
    with ClockSampler(local_rank) as clk:
        if ext is not None:
            with torch.cuda.stream(ext):
                ms_total = cuda_time(step, K, torch)
        else:
            ms_total = cuda_time(step, K, torch)
        barrier()
    if prof_range:
        torch.cuda.cudart().cudaProfilerStop()
    value, ms_step = sharding.aggregate_throughput(units_per_step, ms_total / K, device="cuda")   # all ranks' units over the slowest rank's time
    if wl == "inception":
        achieved = INC_FLOP_PER_IMAGE * INC["batch"] / (ms_step / 1e3) / 1e12     # algorithmic TFLOP/s of the whole step
    else:
        achieved = algo_bytes / (ms_step / 1e3) / 1e9
    gpu_launches = int(lib.ftf_launch_count() - launches0)

    if wl == "inception":
        # Every rank labels ONE common seeded batch: do N GPUs give the single GPU's outputs?  (Each rank timed its own
        # kernel candidates; all kernel families produce the same bits, tests/test_gpu_conv.py.)  SHA-256 of the top-5
        # indices and of the logits, compared with rank 0's.
        words = outputs_digest(ftf, lib, np, torch, lab)
        mine = torch.tensor(words, dtype=torch.int64, device="cuda")
        if world > 1:
            allw = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allw, mine)
        else:
            allw = [mine]
        extra["outputs_match_rank0"] = bool(all(torch.equal(w[:4], allw[0][:4]) for w in allw))
        extra["logits_bits_match_rank0"] = bool(all(torch.equal(w[4:], allw[0][4:]) for w in allw))
        extra["outputs_digest"] = {"batch": "64 images, torch CUDA generator seed 4242 on every rank", "top5_sha256_head": "%016x" % (words[0] & (2**64 - 1)),
                                   "logits_sha256_head": "%016x" % (words[4] & (2**64 - 1)), "ranks_compared": world}

    # ---- end to end through ftf_run with host buffers ----------------------------------------------
    e2e_steps = max(1, min(K, 5))
    if wl == "half_plus_two":
        r = ftf.Runner(os.path.join(ROOT, "models", "half_plus_two"), "serving_default", None,
                       ftf.default_config(device=local_rank))
        p, hx = pinned_array(lib, np, 4 * HPT_N, np.float32)
        hx[:] = np.random.default_rng(5 + rank).standard_normal(HPT_N, dtype=np.float32)
        tv = ftf.TensorValue.__new__(ftf.TensorValue)
        tv.data_type, tv.shape, tv.buffer = ftf.DataType.FLOAT, (HPT_N, 1), hx.view(np.uint8)
        in_key, out_key = "x", "y"
        h2d = d2h = 4 * HPT_N
    elif wl == "jpeg":
        # the files through the public call: STRING tensor of JPEG files -> decode -> resize/normalise -> Inception-v3 -> top-5
        r = ftf.Runner("builtin:normalize_label", None, None,
                       ftf.default_config(device=local_rank, max_batch=64, top_k=INC["top_k"], num_classes=INC["classes"], seed=INC["seed"],
                                          out_h=299, out_w=299, mean=128.0, scale=128.0))
        tv = ftf.messages_to_tensor(files)
        p, hx = pinned_array(lib, np, tv.buffer.nbytes, np.uint8)
        hx[:] = tv.buffer
        tv.buffer = hx
        in_key, out_key = "inputs", "top_k_indices"
        h2d, d2h = int(tv.buffer.nbytes), len(files) * INC["top_k"] * 4             # only the fetched output is downloaded
    elif wl == "inception":
        r = lab
        nb = INC["batch"] * 299 * 299 * 3 * 4
        p, hx = pinned_array(lib, np, nb, np.float32)
        hx[:] = np.random.default_rng(5 + rank).standard_normal(nb // 4, dtype=np.float32)
        tv = ftf.TensorValue.__new__(ftf.TensorValue)
        tv.data_type, tv.shape, tv.buffer = ftf.DataType.FLOAT, (INC["batch"], 299, 299, 3), hx.view(np.uint8)
        in_key, out_key = "inputs", "top_k_indices"
        h2d, d2h = nb, INC["batch"] * INC["top_k"] * 4                               # only the fetched output is downloaded
    else:
        r = ftf.Runner("builtin:image_normalization", None, None,
                       ftf.default_config(device=local_rank, out_h=NORM["out_h"], out_w=NORM["out_w"], mean=NORM["mean"],
                                          scale=NORM["scale"]))
        nb = NORM["batch"] * NORM["in_h"] * NORM["in_w"] * 3
        p, hx = pinned_array(lib, np, nb, np.uint8)
        hx[:] = np.random.default_rng(5 + rank).integers(0, 256, nb, dtype=np.uint8)
        tv = ftf.TensorValue.__new__(ftf.TensorValue)
        tv.data_type, tv.shape, tv.buffer = ftf.DataType.UINT8, (NORM["batch"], NORM["in_h"], NORM["in_w"], 3), hx
        in_key, out_key = "inputs", "outputs"
        h2d, d2h = nb, NORM["batch"] * NORM["out_h"] * NORM["out_w"] * 3 * 4

    from flink_tensorflow_b200.host import _as_ftf, FtfTensor, _check
    t_in, keep = _as_ftf(tv.data_type, tv.shape, tv.buffer)
    ikeys, okeys = (C.c_char_p * 1)(in_key.encode()), (C.c_char_p * 1)(out_key.encode())
    tin, tout = (FtfTensor * 1)(t_in), (FtfTensor * 1)()

    def e2e_step():
        _check(lib.ftf_run(r.handle, 1, ikeys, tin, 1, okeys, tout), r.handle)
        # the result (runner-owned pinned memory) is read by the host: first and last element
        res = np.ctypeslib.as_array(C.cast(tout[0].data, C.POINTER(C.c_int32 if wl in ("inception", "jpeg") else C.c_float)),
                                    shape=(tout[0].nbytes // 4,))
        return float(res[0]) + float(res[-1])
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_steps
    e2e_value = units_per_step * world / (e2e_ms / 1e3)
    st = r.stats()
    e2e_launches = st["kernel_launches"]
    if wl == "inception" and world == 1 and not args.quick:
        extra["per_record"] = per_record_leg(ftf, lib, np, local_rank, lab)
    r.close()
    lib.ftf_host_free(p)

    e2e = {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "gpu_launches": int(e2e_launches),
           "api": "ftf_run (C ABI) with pinned host buffers, chunked H2D/kernel/D2H pipeline"}
    if wl == "jpeg":
        # the per-record path with files as records (what the reference's source emits), next to the blocking call
        sres = streaming_bench(ftf, lib, np, local_rank, n_records=INC["batch"] * (100 if args.quick else 400), kind="jpeg_label", sync=barrier)
        barrier()
        extra["streaming_files"] = sres
        e2e["api"] = ("ftf_run (C ABI) on builtin:normalize_label with a STRING tensor of 256 JPEG files in host memory: header parse, "
                      "upload of the entropy-coded bytes, J1-J3, resize/normalise, Inception-v3, top-5 read back")
    if wl == "inception":
        # The reference labels images record by record inside a map() (InceptionModel.label); the drop-in for that call
        # is ftf_submit/ftf_poll, which micro-batches 64 records, uploads each as it arrives and overlaps the upload of
        # the next batch with the kernels of the current one.  That is the end-to-end figure; the blocking one-call
        # ftf_run over a stacked [64,299,299,3] tensor (upload, then compute, then download) is kept beside it.
        n_b = max(K, 100)
        sres = streaming_bench(ftf, lib, np, local_rank, n_records=INC["batch"] * n_b, kind="inception", sync=barrier)
        barrier()
        s_ms = max_over_ranks(sres["seconds"] * 1e3)
        extra["e2e_blocking_run"] = e2e
        if world == 1 and not args.quick:
            extra["e2e_blocking_run_two_subtasks"] = blocking_two_subtasks(ftf, lib, np, local_rank)
        e2e = {"value": sres["records"] * world / (s_ms / 1e3), "unit": unit,
               "h2d_bytes_per_step": INC["batch"] * sres["h2d_bytes_per_record"], "d2h_bytes_per_step": INC["batch"] * sres["d2h_bytes_per_record"],
               "ms_per_step": s_ms / n_b, "steps": n_b, "gpu_launches": sres["launches"],
               "latency_us_p50": sres["latency_us_p50"], "latency_us_p99": sres["latency_us_p99"],
               "api": "ftf_submit/ftf_poll (C ABI): one fp32 299x299x3 record per call from pinned host memory, micro-batches of 64, "
                      "top-5 indices and scores read back per record"}
    if wl == "inception" and not args.quick:
        # BASELINE config 5, every rank on its own GPU with its own source thread: >= 10^5 records per rank at saturation --
        # (a) the record the reference's source really emits, a 640x480 JPEG FILE (ImageInputFormat.scala:63-80), decoded on the
        # GPU; (b) raw 640x480 u8 records as TensorValue wire images; then (c) a source emitting (b) at half the saturation
        # rate (latency at a sustainable load).  Whole-job figures: records of all ranks over the slowest rank's time,
        # latencies as the worst rank's.
        def whole_job(res):
            secs = max_over_ranks(res["seconds"])
            res = dict(res)
            res["records_per_s_this_rank"] = res.pop("records_per_s")
            res["records_per_s"] = res["records"] * world / secs
            res["seconds"] = secs
            res["latency_us_p50"] = max_over_ranks(res["latency_us_p50"])
            res["latency_us_p99"] = max_over_ranks(res["latency_us_p99"])
            res["ranks"] = world
            return res
        jp = streaming_bench(ftf, lib, np, local_rank, n_records=INC["batch"] * 1600, kind="jpeg_label", sync=barrier)
        barrier()
        extra["streaming_jpeg"] = whole_job(jp)
        sat = streaming_bench(ftf, lib, np, local_rank, n_records=INC["batch"] * 1600, wire=True, sync=barrier)
        barrier()
        half = streaming_bench(ftf, lib, np, local_rank, n_records=INC["batch"] * 800, wire=True, offered_rate=0.5 * sat["records_per_s"], sync=barrier)
        barrier()
        extra["streaming"] = whole_job(sat)
        extra["streaming_half_rate"] = whole_job(half)
        # configs 2 and 3 in the same run (rank-local kernels; reported from rank 0)
        extra["half_plus_two"] = sub_leg("half_plus_two", torch, device, np, lib, pk, rank, local_rank)
        extra["normalize"] = sub_leg("normalize", torch, device, np, lib, pk, rank, local_rank)
        extra["normalize_bf16"] = sub_leg("normalize_bf16", torch, device, np, lib, pk, rank, local_rank)
        barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- CPU baseline on a bounded sample (rank 0, N=1 only) -----------------------------------------
    cpu = None
    if world == 1 and not args.quick:
        import oracle                                        # the checker is loaded for this leg only, after every GPU measurement
        threads = oracle.num_threads()
        rng = np.random.default_rng(1234)
        if wl == "half_plus_two":
            n = 1 << 26
            cx = (rng.standard_normal(n) * 100).astype(np.float32)
            oracle.affine_f32(cx, 0.5, 2.0, threads=threads)
            t0, reps = time.perf_counter(), 0
            while time.perf_counter() - t0 < 5.0:
                oracle.affine_f32(cx, 0.5, 2.0, threads=threads)
                reps += 1
            cpu = {"value": 8.0 * n * reps / 1e9 / (time.perf_counter() - t0), "unit": unit, "cores": threads, "kind": "port",
                   "sample": f"{reps} passes over 2^26 floats (1/4 of the workload), all host threads"}
        elif wl == "jpeg":
            cpu = {"value": cpu_jpeg(np, oracle, files[:64], threads, 8.0), "unit": unit, "cores": threads, "kind": "port",
                   "sample": "passes over 64 of the 256 files for 8 s, C oracle, one file per task on all host threads"}
            extra["cpu_pillow_libjpeg_turbo"] = {"value": cpu_pillow(np, files[:64], threads, 5.0), "unit": unit, "cores": threads,
                                                 "note": "libjpeg-turbo (SIMD) through Pillow, the library TF's DecodeJpeg wraps"}
        elif wl == "inception":
            import oracle.inception as oinc
            net = oinc.Net(INC["classes"], INC["seed"], 0)
            b = 4
            cimg = rng.standard_normal((b, 299, 299, 3)).astype(np.float32)
            t0, reps = time.perf_counter(), 0
            while time.perf_counter() - t0 < 10.0:
                net.forward(cimg, threads=threads)
                reps += 1
            cpu = {"value": b * reps / (time.perf_counter() - t0), "unit": unit, "cores": threads, "kind": "port",
                   "sample": f"{reps} passes over 4 images (1/16 of the batch), fp32 C oracle, all host threads"}
            extra["cpu_torchvision_onednn"] = torchvision_cpu(threads, unit)
            if "per_record" in extra:
                extra["per_record"]["cpu"] = per_record_cpu(ftf, np, oracle)
        else:
            b = 32
            cimg = rng.integers(0, 256, (b, NORM["in_h"], NORM["in_w"], 3), dtype=np.uint8)
            oracle.resize_bilinear_norm_u8(cimg, NORM["out_h"], NORM["out_w"], NORM["mean"], NORM["scale"], threads=threads)
            t0, reps = time.perf_counter(), 0
            while time.perf_counter() - t0 < 5.0:
                oracle.resize_bilinear_norm_u8(cimg, NORM["out_h"], NORM["out_w"], NORM["mean"], NORM["scale"], threads=threads)
                reps += 1
            cpu = {"value": b * reps / (time.perf_counter() - t0), "unit": unit, "cores": threads, "kind": "port",
                   "sample": f"{reps} passes over 32 images (1/8 of the batch), all host threads"}

    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(wl, {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
            "data": "synthetic", "config": config,
            "roofline": {"bound": bound, "achieved": achieved, "peak": peak, "unit": peak_unit, "frac": achieved / peak,
                         "traffic": traffic, "peak_source": pk["src"] + (" (sustained)" if wl == "inception" else ""),
                         **({"frac_of_nominal_8TBps": achieved / 8000.0} if wl != "inception" else
                            {"frac_of_burst": achieved / pk["tf_burst"], "scope": "whole step (all kernels), algorithmic FLOPs"}),
                         "kernel_ms": ms_step,
                         **({"scope": "whole four-kernel step, file bytes in + RGB bytes out; the dominant kernel (Huffman, 3/4 of the step) "
                                      "is bound by issue slots and its serial correction chain, not by bytes (profiles/r01_jpeg_notes.md)"}
                            if wl == "jpeg" else {})},
            "e2e": e2e,
            "gpu_launches": gpu_launches, "clocks": clk.summary(), "cpu_baseline": cpu, "host_numa_binding": numa}
    line.update(extra)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quick", action="store_true", help="development A/B runs: skip the CPU baseline and the streaming leg")
    ap.add_argument("--workload", default="inception", choices=["half_plus_two", "normalize", "inception", "jpeg"])
    args = ap.parse_args()
    # stdout carries exactly one JSON line: whatever libraries print at the C level on the way (NCCL's version banner
    # under NCCL_DEBUG=VERSION, for one) is sent to stderr, and print() gets the real stdout back
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
