"""The label model in the configuration bench.py measures -- micro-batches of 64, kernels chosen by the plan-time timing --
against the CPU oracle, through every way a record reaches it: device buffers (ftf_label_dev), one blocking call
(ftf_run) and the per-record path (ftf_submit / ftf_poll) with payloads in pinned memory, in pageable memory and as
TensorValue wire images decoded in place.  Needs a B200 (-m gpu).

Reference: the per-record map() around InceptionModel.label (EX/inception/inception.scala:39-46,
EX/inception/InceptionModel.scala:58-67,81-89; FT/streaming/package.scala:23-34)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BATCH = 64


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _images(seed, n):
    return np.random.default_rng(seed).standard_normal((n, 299, 299, 3)).astype(np.float32)


def _check_against_oracle(oracle, logits, top5, want):
    """logits within 1e-2 of the oracle's scale for every image; top-5 indices equal wherever the oracle's own order of its
    six best logits is decided by more than the error the bf16 tolerance allows -- and that must be nearly every image."""
    scale = np.abs(want).max()
    err = np.abs(logits - want).max(axis=1)
    assert err.max() <= 1e-2 * scale, (err.max(), scale)
    wi, _ = oracle.top_labels(oracle.softmax_rows(want), want.shape[1], 5)
    top6 = -np.sort(-want, axis=1)[:, :6]
    margin = (top6[:, :-1] - top6[:, 1:]).min(axis=1)
    decided = margin > 4.0 * err.max()
    assert decided.sum() >= int(0.9 * len(want)), (int(decided.sum()), float(err.max()), np.sort(margin)[:8])
    assert np.array_equal(top5[decided], wi[decided])
    # the others: same five labels or a swap of near-tied neighbours -- every index the runner names is among the oracle's six best
    wi6 = np.argsort(-want, axis=1, kind="stable")[:, :6]
    for i in np.nonzero(~decided)[0]:
        assert set(top5[i]) <= set(wi6[i]), i


@pytest.fixture(scope="module")
def net64(ftf):
    """builtin:inception_v3 exactly as bench.py builds it: max_batch 64, top-5, seed 2024, plan-time tuning at its default."""
    r = ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=BATCH, top_k=5, num_classes=1000, seed=2024))
    yield r
    r.close()


@pytest.fixture(scope="module")
def oracle_logits64(oracle):
    import oracle.inception as oi
    x = _images(7, BATCH)
    return x, oi.Net(1000, 2024, mode=1).forward(x)


def test_batch_64_tuned_plan_matches_oracle_on_every_image(ftf, oracle, net64, oracle_logits64):
    """The benched plan (tile shapes, kernel families and SM shares chosen for 64 images) is the one compared here."""
    import torch
    x, want = oracle_logits64
    lib = ftf._abi.load_library()
    # device buffers, bf16 images (bench.py's device-resident arm) and fp32 images (read in place by the first convolution)
    xd = torch.from_numpy(x).cuda()
    outs = {}
    for name, t, code in (("bf16", xd.to(torch.bfloat16), ftf._abi.BFLOAT16), ("f32", xd, ftf._abi.FLOAT)):
        logits = torch.empty((BATCH, 1000), dtype=torch.float32, device="cuda")
        probs = torch.empty_like(logits)
        idx = torch.empty((BATCH, 5), dtype=torch.int32, device="cuda")
        score = torch.empty((BATCH, 5), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()                           # the runner's stream does not wait for torch's
        rc = lib.ftf_label_dev(net64.handle, t.data_ptr(), code, BATCH, logits.data_ptr(), probs.data_ptr(), idx.data_ptr(), score.data_ptr())
        assert rc == 0, lib.ftf_last_error(net64.handle)
        torch.cuda.synchronize()
        outs[name] = (logits.cpu().numpy(), idx.cpu().numpy(), probs.cpu().numpy(), score.cpu().numpy())
    assert np.array_equal(_bits(outs["bf16"][0]), _bits(outs["f32"][0]))          # rounding at load == a cast pass in front
    logits, idx, probs, score = outs["bf16"]
    _check_against_oracle(oracle, logits, idx, want)
    assert np.allclose(probs, oracle.softmax_rows(logits), rtol=1e-5, atol=1e-12)
    assert np.array_equal(score, np.take_along_axis(probs, idx, 1))
    # the blocking call over host memory gives the same bits
    res = net64.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices"])
    assert np.array_equal(_bits(res["logits"].to_array()), _bits(logits))
    assert np.array_equal(res["top_k_indices"].to_array(), idx)


def test_a_second_runner_takes_the_first_ones_kernels(ftf, net64, oracle_logits64):
    """Two subtasks of one process must return the same bits for the same record: the second runner reuses the kernel
    choices the first one timed (no FTF_CONV_TUNE=0 crutch)."""
    x, _ = oracle_logits64
    a = net64.run({"inputs": ftf.TensorValue.from_array(x[:BATCH])}, ["logits"])["logits"].to_array()
    r2 = ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=BATCH, top_k=5, num_classes=1000, seed=2024))
    try:
        b = r2.run({"inputs": ftf.TensorValue.from_array(x[:BATCH])}, ["logits"])["logits"].to_array()
    finally:
        r2.close()
    assert np.array_equal(_bits(a), _bits(b))


def test_short_batches_replay_the_next_bucket_with_the_same_bits(ftf, net64, oracle_logits64):
    """Batches of 1, 5 and 13 images run the graph of the next multiple of 8; a record's logits do not depend on it."""
    x, _ = oracle_logits64
    full = net64.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits"])["logits"].to_array()
    for lo, n in ((0, 1), (3, 5), (20, 13), (40, 24)):
        part = net64.run({"inputs": ftf.TensorValue.from_array(x[lo:lo + n])}, ["logits"])["logits"].to_array()
        assert np.array_equal(_bits(part), _bits(full[lo:lo + n])), (lo, n)


def _stream_records(ftf, runner, recs, how):
    """Submits every record (how = "pinned" | "pageable" | "wire"), polls everything; returns [(tag, [logits, top5])]."""
    from flink_tensorflow_b200.host import FtfTensor, _check
    lib = ftf._abi.load_library()
    n, rec_bytes = len(recs), recs[0].nbytes
    keep, tensors = [], []
    dims = (C.c_int64 * recs[0].ndim)(*recs[0].shape)
    code = ftf.DataType.of(recs[0])
    if how == "pinned":
        p = lib.ftf_host_alloc(n * rec_bytes)
        assert p
        ring = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n * rec_bytes,))
        for i, a in enumerate(recs):
            ring[i * rec_bytes:(i + 1) * rec_bytes] = a.reshape(-1).view(np.uint8)
            tensors.append(FtfTensor(code, a.ndim, dims, p + i * rec_bytes, rec_bytes))
        free = lambda: lib.ftf_host_free(p)
    elif how == "mixed":
        # runs of records back to back in one pinned allocation (uploaded by one merged copy), records in a second allocation in
        # reverse order (never mergeable), single-record allocations, and pageable payloads in between (staged by the runner)
        pa, pb = lib.ftf_host_alloc(n * rec_bytes), lib.ftf_host_alloc(n * rec_bytes)
        assert pa and pb
        singles = []
        for i, a in enumerate(recs):
            raw = np.ascontiguousarray(a).reshape(-1).view(np.uint8)
            if i % 7 == 0:
                bcopy = raw.copy()
                keep.append(bcopy)
                addr = bcopy.ctypes.data
            elif i % 7 in (1, 2, 3):
                addr = pa + i * rec_bytes
            elif i % 7 in (4, 5):
                addr = pb + (n - 1 - i) * rec_bytes
            else:
                addr = lib.ftf_host_alloc(rec_bytes)
                assert addr
                singles.append(addr)
            if i % 7:
                np.ctypeslib.as_array(C.cast(addr, C.POINTER(C.c_uint8)), shape=(rec_bytes,))[:] = raw
            tensors.append(FtfTensor(code, a.ndim, dims, addr, rec_bytes))
        free = lambda: [lib.ftf_host_free(q) for q in [pa, pb] + singles]
    elif how == "pageable":
        for a in recs:
            b = np.ascontiguousarray(a).reshape(-1).view(np.uint8).copy()
            keep.append(b)
            tensors.append(FtfTensor(code, a.ndim, dims, b.ctypes.data, rec_bytes))
        free = lambda: None
    else:   # TensorValue wire images (TensorValue.java:142-187) in pinned memory, decoded where they lie
        wlen = int(lib.ftf_tv_binary_length(recs[0].ndim, rec_bytes))
        stride = (wlen + 255) // 256 * 256
        p = lib.ftf_host_alloc(n * stride)
        assert p
        written = C.c_int64()
        for i, a in enumerate(recs):
            src = np.ascontiguousarray(a).reshape(-1).view(np.uint8)
            t = FtfTensor(code, a.ndim, dims, src.ctypes.data, rec_bytes)
            _check(lib.ftf_tv_encode(C.byref(t), p + i * stride, wlen, C.byref(written)))
            dbuf = (C.c_int64 * 32)()
            d = FtfTensor()
            _check(lib.ftf_tv_decode(p + i * stride, wlen, C.byref(d), dbuf))
            assert d.data >= p + i * stride and d.data + d.nbytes <= p + i * stride + wlen        # a view, not a copy
            keep.append(dbuf)
            tensors.append(d)
        free = lambda: lib.ftf_host_free(p)
    got = []
    try:
        selected = False
        for i, t in enumerate(tensors):
            while True:
                rc = lib.ftf_submit(runner.handle, 1000 + i, C.byref(t))
                if rc != ftf._abi.RESOURCE_EXHAUSTED:
                    _check(rc, runner.handle)
                    break
                got += runner.poll(block=True)
            if not selected:
                runner.poll_select(["logits", "top_k_indices"])
                selected = True
            if i % 50 == 49:
                got += runner.poll()
        runner.flush()
        while len(got) < n:
            got += runner.poll(block=True)
    finally:
        free()
    return got


@pytest.mark.parametrize("how", ["pinned", "pageable", "wire", "mixed"])
def test_label_records_through_submit_poll(ftf, oracle, net64, how):
    """builtin:inception_v3 behind the 1-in/1-out map: three full micro-batches and a short one, every record's logits and
    top-5 bit-equal to the blocking call on the same records, in submission order; a sample checked against the oracle."""
    import oracle.inception as oi
    n = 3 * BATCH + 17
    x = _images(100 + len(how), n)
    want = net64.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices"])
    wl, wi = want["logits"].to_array(), want["top_k_indices"].to_array()
    got = _stream_records(ftf, net64, [x[i] for i in range(n)], how)
    assert [t for t, _ in got] == list(range(1000, 1000 + n))
    gl = np.stack([v[0].to_array() for _, v in got])
    gi = np.stack([v[1].to_array() for _, v in got])
    assert np.array_equal(_bits(gl), _bits(wl)) and np.array_equal(gi, wi)
    # the short last batch and one record of each full one against the oracle itself
    pick = [0, BATCH + 5, 2 * BATCH + 63, n - 1]
    ref = oi.Net(1000, 2024, mode=1).forward(x[pick])
    assert np.abs(gl[pick] - ref).max() <= 1e-2 * np.abs(ref).max()


def test_raw_rgb_records_through_submit_poll_on_normalize_label(ftf, oracle):
    """builtin:normalize_label (BASELINE config 5's chain: u8 640x480 -> resize/normalise -> Inception-v3 -> top-5) behind the
    per-record path, micro-batches of 64: bit-equal to the blocking call, in order; a sample against the oracle chain."""
    import oracle.inception as oi
    cfg = ftf.default_config(max_batch=BATCH, top_k=5, out_h=299, out_w=299, mean=128.0, scale=128.0, num_classes=1000, seed=2024)
    r = ftf.Runner("builtin:normalize_label", None, None, cfg)
    try:
        n = 3 * BATCH + 9
        img = np.random.default_rng(55).integers(0, 256, (n, 480, 640, 3), dtype=np.uint8)
        want = r.run({"inputs": ftf.TensorValue.from_array(img)}, ["logits", "top_k_indices"])
        wl, wi = want["logits"].to_array(), want["top_k_indices"].to_array()
        for how in ("pinned", "wire", "pageable"):
            got = _stream_records(ftf, r, [img[i] for i in range(n)], how)
            assert [t for t, _ in got] == list(range(1000, 1000 + n)), how
            gl = np.stack([v[0].to_array() for _, v in got])
            gi = np.stack([v[1].to_array() for _, v in got])
            assert np.array_equal(_bits(gl), _bits(wl)) and np.array_equal(gi, wi), how
        pick = [0, BATCH + 1, n - 1]
        norm = oracle.resize_bilinear_norm_u8(img[pick], 299, 299, 128.0, 128.0, threads=8)
        ref = oi.Net(1000, 2024, mode=1).forward(norm)
        assert np.abs(wl[pick] - ref).max() <= 1e-2 * np.abs(ref).max()
        assert np.array_equal(wi[pick], oracle.top_labels(oracle.softmax_rows(ref), 1000, 5)[0])
    finally:
        r.close()


def test_outputs_selected_after_launch_are_still_delivered(ftf, net64):
    """Only the selected outputs travel back with a batch; selecting another one while a batch is in flight fetches it at
    poll time from the device copy."""
    x = _images(9, 3)
    want = net64.run({"inputs": ftf.TensorValue.from_array(x)}, ["outputs", "top_k_scores"])
    net64.poll_select(["logits"])
    for i in range(3):
        net64.submit(i, ftf.TensorValue.from_array(x[i]))
    net64.flush()                                          # launched with "logits" selected
    net64.poll_select(["outputs", "top_k_scores"])
    got = []
    while len(got) < 3:
        got += net64.poll(block=True)
    assert np.array_equal(_bits(np.stack([v[0].to_array() for _, v in got])), _bits(want["outputs"].to_array()))
    assert np.array_equal(_bits(np.stack([v[1].to_array() for _, v in got])), _bits(want["top_k_scores"].to_array()))


def test_trained_weights_are_read_from_a_weight_directory(ftf, tmp_path):
    """ftf_create("inception_v3:<dir>"): the label model with weights from a TensorBundle directory (where the reference
    loads its frozen graph file, InceptionModel.scala:51-55).  A seeded runner exports its weights; a runner created from
    that directory with ANOTHER seed returns the same bits (the file decides, not the seed); labels.txt is picked up; a
    damaged data file, a missing tensor and a misshapen tensor are refused."""
    import os
    import shutil
    d = str(tmp_path / "inception_v3_weights")
    x = _images(21, 3)
    cfg = ftf.default_config(max_batch=4, top_k=5, num_classes=1000, seed=77)
    a = ftf.Runner("builtin:inception_v3", None, None, cfg)
    try:
        want = a.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices"])
        a.export_weights(d)
    finally:
        a.close()
    assert os.path.getsize(os.path.join(d, "variables", "variables.data-00000-of-00001")) > 90e6      # 23.8 M fp32 parameters
    # the exported directory is a TensorBundle in the half_plus_two layout's format: the index carries the bundle magic
    idx = open(os.path.join(d, "variables", "variables.index"), "rb").read()
    assert idx[-8:] == bytes([0x57, 0xfb, 0x80, 0x8b, 0x24, 0x75, 0x47, 0xdb]) and b"Mixed_7c.branch_pool/weights" in idx and b"fc/biases" in idx
    with open(os.path.join(d, "labels.txt"), "w") as f:
        f.write("\n".join(f"label {i}" for i in range(1000)) + "\n")
    m = ftf.InceptionModel(model_dir=d, max_batch=4, top_k=5, num_classes=1000, seed=12345)
    m.open()
    try:
        assert m.labels[7] == "label 7" and len(m.labels) == 1000
        got = m.session.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices"])
        assert np.array_equal(_bits(got["logits"].to_array()), _bits(want["logits"].to_array()))
        assert np.array_equal(got["top_k_indices"].to_array(), want["top_k_indices"].to_array())
        with pytest.raises(RuntimeError, match="loaded from a weight file"):
            m.session.export_weights(str(tmp_path / "again"))
    finally:
        m.close()
    # the fp32 mode reads the same file
    r32 = ftf.Runner("inception_v3:" + d, None, None, ftf.default_config(max_batch=1, top_k=5, compute_dtype=ftf.DataType.FLOAT))
    try:
        l32 = r32.run({"inputs": ftf.TensorValue.from_array(x[:1])}, ["logits"])["logits"].to_array()
        assert np.abs(l32 - want["logits"].to_array()[:1]).max() <= 1e-2 * np.abs(l32).max()
    finally:
        r32.close()
    # one flipped byte in the data file: the crc32c of that tensor no longer matches
    bad = str(tmp_path / "bad")
    shutil.copytree(d, bad)
    p = os.path.join(bad, "variables", "variables.data-00000-of-00001")
    with open(p, "r+b") as f:
        f.seek(1 << 20)
        b = f.read(1)
        f.seek(1 << 20)
        f.write(bytes([b[0] ^ 0x40]))
    with pytest.raises(ValueError, match="crc32c mismatch"):
        ftf.Runner("inception_v3:" + bad, None, None, cfg)
    with pytest.raises(KeyError):
        ftf.Runner("inception_v3:" + str(tmp_path / "nothing_here"), None, None, cfg)
    with pytest.raises(ValueError, match="expected FLOAT"):                                            # 1001 classes: fc has the wrong shape
        ftf.Runner("inception_v3:" + d, None, None, ftf.default_config(max_batch=1, num_classes=1001))


def test_kernel_choices_travel_between_processes_through_the_plan_file(ftf, tmp_path):
    """FTF_CONV_PLAN_FILE: the first process times its candidates and writes its choices, the second one reads them and
    times nothing -- same kernels (a profiler run replays the plan of the plain run; every rank can be handed one file)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    plan = str(tmp_path / "plan.txt")
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import ftf_loader; ftf = ftf_loader.load();"
            "r = ftf.Runner('builtin:inception_v3', None, None, ftf.default_config(max_batch=8, top_k=5, seed=2024));"
            "x = np.random.default_rng(3).standard_normal((2, 299, 299, 3)).astype(np.float32);"
            "print('DIGEST', r.run({'inputs': ftf.TensorValue.from_array(x)}, ['logits'])['logits'].to_array().tobytes().hex()[:64]); r.close()") % root
    env = dict(os.environ, FTF_CONV_PLAN_FILE=plan, FTF_CONV_VERBOSE="1")
    a = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert a.returncode == 0, a.stderr[-2000:]
    assert ", timed)" in a.stderr and os.path.getsize(plan) > 0
    n_lines = sum(1 for _ in open(plan))
    b = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert b.returncode == 0, b.stderr[-2000:]
    assert ", timed)" not in b.stderr and "choice of an earlier runner" in b.stderr and "refine:" not in b.stderr
    assert sum(1 for _ in open(plan)) == n_lines                      # nothing new was learnt
    assert a.stdout.split("DIGEST")[1].strip() == b.stdout.split("DIGEST")[1].strip()
