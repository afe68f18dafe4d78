"""Parity of the CUDA path with the CPU oracle, through the C ABI.  Needs a B200 (-m gpu)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev(ftf):
    import torch
    assert torch.cuda.is_available()
    from flink_tensorflow_b200 import device
    return device


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# ---- K1 -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 3, 7, 8, 31, 257, 4099, 1 << 20, (1 << 22) + 5])
def test_affine_bit_exact(dev, oracle, n):
    import torch
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 100).astype(np.float32)
    y = dev.affine_f32(torch.from_numpy(x).cuda(), 0.5, 2.0).cpu().numpy()
    assert np.array_equal(_bits(y), _bits(oracle.affine_f32(x, 0.5, 2.0)))


def test_affine_edge_values_and_generic_coefficients(dev, oracle):
    import torch
    edge = np.array([0.0, -0.0, 1e-45, -1e-45, 1.1754942e-38, np.inf, -np.inf, np.nan, 3.4028235e38, -3.4028235e38,
                     1.1754944e-38, 2.3509887e-38, -4.0, 4.0], np.float32)
    rng = np.random.default_rng(11)
    x = np.concatenate([edge, rng.standard_normal(5000).astype(np.float32) * 1e3,
                        rng.integers(0, 2**32, 5000, dtype=np.uint32).view(np.float32)])   # every bit pattern class
    for a, b in [(0.5, 2.0), (0.5, 3.0), (1.0 / 3.0, 0.1), (-7.25, 1e-30), (3.4e38, -3.4e38)]:
        y = dev.affine_f32(torch.from_numpy(x).cuda(), a, b).cpu().numpy()
        want = oracle.affine_f32(x, a, b)
        nan = np.isnan(want)
        assert np.array_equal(nan, np.isnan(y))
        assert np.array_equal(_bits(y)[~nan], _bits(want)[~nan])


def test_affine_unaligned_views(dev, oracle):
    import torch
    x = (np.random.default_rng(12).standard_normal(10007) * 10).astype(np.float32)
    xd = torch.from_numpy(x).cuda()
    for off in (1, 2, 3, 4, 5):                          # 4/8/12/16/20-byte offsets: scalar, 128- and 256-bit paths
        src = xd[off:]
        out = torch.empty(x.size + 8, dtype=torch.float32, device="cuda")[off: off + src.numel()]
        y = dev.affine_f32(src, 0.5, 2.0, out=out).cpu().numpy()
        assert np.array_equal(_bits(y), _bits(oracle.affine_f32(x[off:], 0.5, 2.0)))


def test_affine_full_size_properties(dev):
    """BASELINE config 2 size (2^28 floats): exactness of 0.5*x+2 through size-independent properties."""
    import torch
    n = 1 << 28
    g = torch.Generator(device="cuda").manual_seed(1234)
    x = torch.randn(n, device="cuda", generator=g) * 100
    y = dev.affine_f32(x, 0.5, 2.0)
    # 0.5*x is exact, so y == fl(x/2 + 2): torch's own fp32 ops give the same single rounding
    assert torch.equal(y, x * 0.5 + 2.0)
    # linearity in exact arithmetic: with a power-of-two slope and zero intercept the kernel is a bit shift
    z = dev.affine_f32(x, 0.5, 0.0)
    assert torch.equal(z + z, x)
    # a checksum over blocks equals the checksum of a second, independent pass
    assert torch.equal(dev.affine_f32(x, 0.5, 2.0).view(torch.int32).sum(), y.view(torch.int32).sum())


# ---- K2 -----------------------------------------------------------------------------------------
def _images(rng, n, h, w):
    img = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    if n >= 2:   # one gradient and one checkerboard image
        img[0] = (np.add.outer(np.arange(h), np.arange(w))[:, :, None] * np.array([1, 2, 3])) % 256
        img[1] = ((np.add.outer(np.arange(h), np.arange(w)) % 2) * 255)[:, :, None]
    return img


@pytest.mark.parametrize("shape", [
    (3, 480, 640, 224, 224, 117.0, 1.0),      # the reference's constants (ImageNormalization.scala:51-54)
    (3, 480, 640, 299, 299, 128.0, 128.0),    # BASELINE config 3
    (2, 480, 640, 299, 299, 117.0, 3.0),      # scale that is not a power of two: true division
    (2, 64, 48, 128, 96, 0.0, 1.0),           # exact 2x upscale
    (2, 37, 53, 37, 53, 1.5, 0.25),           # identity size, unaligned row pitch (direct kernel)
    (1, 1, 1, 5, 7, 10.0, 2.0),               # 1x1 source
    (2, 720, 1280, 299, 299, 117.0, 1.0),     # HD source, 3840-byte rows (15 TMA chunks)
    (1, 2000, 16, 4, 4, 0.0, 1.0),            # 500x vertical downscale
    (1, 16, 3000, 8, 3, 0.0, 1.0),            # 1000x horizontal downscale
    (2, 100, 100, 1, 1, 3.0, 7.0),            # single output pixel
    (1, 33, 65, 700, 900, 117.0, 1.0),        # large upscale, output rows of 2700 elements
])
def test_resize_norm_bit_exact(dev, oracle, shape):
    import torch
    n, h, w, oh, ow, mean, scale = shape
    img = _images(np.random.default_rng(h * w + oh), n, h, w)
    got = dev.resize_norm(torch.from_numpy(img).cuda(), oh, ow, mean, scale).cpu().numpy()
    want = oracle.resize_bilinear_norm_u8(img, oh, ow, mean, scale, threads=8)
    assert got.shape == want.shape
    assert np.array_equal(_bits(got), _bits(want))


def test_resize_norm_bf16_is_rounded_f32(dev, oracle):
    import torch
    img = _images(np.random.default_rng(21), 3, 480, 640)
    got = dev.resize_norm(torch.from_numpy(img).cuda(), 299, 299, 128.0, 128.0, out_dtype=torch.bfloat16)
    want = oracle.bf16_bits(oracle.resize_bilinear_norm_u8(img, 299, 299, 128.0, 128.0, threads=8))
    assert np.array_equal(got.view(torch.int16).cpu().numpy().view(np.uint16), want)


def test_resize_norm_unaligned_base_pointer(dev, oracle):
    import torch
    img = _images(np.random.default_rng(22), 2, 48, 64)
    flat = torch.empty(img.size + 16, dtype=torch.uint8, device="cuda")
    view = flat[3: 3 + img.size].view(2, 48, 64, 3)      # base not 16-byte aligned -> direct kernel
    view.copy_(torch.from_numpy(img))
    got = dev.resize_norm(view, 31, 17, 117.0, 1.0).cpu().numpy()
    assert np.array_equal(_bits(got), _bits(oracle.resize_bilinear_norm_u8(img, 31, 17, 117.0, 1.0)))


def test_resize_norm_full_batch_properties(dev, oracle):
    """BASELINE config 3 size (batch 256): batch independence + a sampled oracle check."""
    import torch
    g = torch.Generator(device="cuda").manual_seed(42)
    img = torch.randint(0, 256, (256, 480, 640, 3), dtype=torch.uint8, device="cuda", generator=g)
    out = dev.resize_norm(img, 299, 299, 128.0, 128.0)
    for i in (0, 17, 255):                                # record i alone == record i inside the batch
        assert torch.equal(dev.resize_norm(img[i:i + 1].contiguous(), 299, 299, 128.0, 128.0)[0], out[i])
    pick = [3, 128, 254]
    want = oracle.resize_bilinear_norm_u8(img[pick].cpu().numpy(), 299, 299, 128.0, 128.0, threads=8)
    assert np.array_equal(_bits(out[pick].cpu().numpy()), _bits(want))
    const = torch.full((4, 480, 640, 3), 77, dtype=torch.uint8, device="cuda")     # constant image stays constant
    assert torch.all(dev.resize_norm(const, 299, 299, 128.0, 128.0) == (77.0 - 128.0) / 128.0)


# ---- K7 -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("m,n,labels,k", [(1, 1008, 1008, 3), (64, 1000, 1000, 5), (7, 1001, 1000, 5), (3, 10, 4, 6),
                                          (5, 1, 1, 1), (2, 5000, 5000, 32), (300, 17, 17, 17)])
def test_top_labels_exact(dev, oracle, m, n, labels, k):
    import torch
    rng = np.random.default_rng(m * n + k)
    x = rng.standard_normal((m, n)).astype(np.float32)
    x[:, ::7] = x[:, :1]                                  # plenty of exact ties
    _, idx, sc = dev.softmax_topk(torch.from_numpy(x).cuda(), k, n_labels=labels, apply_softmax=False)
    oi, osc = oracle.top_labels(x, labels, k)
    assert np.array_equal(idx.cpu().numpy(), oi)
    assert np.array_equal(_bits(sc.cpu().numpy()), _bits(osc))


def test_top_labels_all_equal_and_infinities(dev, oracle):
    import torch
    x = np.zeros((2, 50), np.float32)
    x[1, 10] = np.inf; x[1, 20] = -np.inf; x[1, 30] = np.inf
    _, idx, _ = dev.softmax_topk(torch.from_numpy(x).cuda(), 4, apply_softmax=False)
    assert idx.cpu().numpy().tolist() == [[0, 1, 2, 3], [10, 30, 0, 1]]
    assert np.array_equal(idx.cpu().numpy(), oracle.top_labels(x, 50, 4)[0])


def test_softmax_then_top5(dev, oracle):
    import torch
    rng = np.random.default_rng(31)
    x = (rng.standard_normal((64, 1000)) * 3).astype(np.float32)
    probs, idx, sc = dev.softmax_topk(torch.from_numpy(x).cuda(), 5)
    p = probs.cpu().numpy()
    want = oracle.softmax_rows(x)
    # fp32 tolerance: exp() differs by <= 2 ulp between libm and CUDA, the row sum is reduced in another order
    assert np.allclose(p, want, rtol=2e-6, atol=1e-30)
    oi, _ = oracle.top_labels(want, 1000, 5)
    assert np.array_equal(idx.cpu().numpy(), oi)          # top-5 label indices bit-exact
    # the kernel's own selection is exactly the stable top-k of the probabilities it wrote
    si, ss = oracle.top_labels(p, 1000, 5)
    assert np.array_equal(idx.cpu().numpy(), si) and np.array_equal(_bits(sc.cpu().numpy()), _bits(ss))


# ---- through ModelFunction / ftf_run ------------------------------------------------------------------
def test_regress_itcase(ftf, golden, model_dir):
    """RegressITCase.scala:40-68: one Example per call through a flatMap, float == on the result."""
    model = ftf.HalfPlusTwo(model_dir)

    def fun(rec, m):
        ex, expected = rec
        with m.regress_x_to_y(ftf.messages_to_tensor([ex])) as y:
            assert y.shape == (1, 1) and y.data_type == ftf.DataType.FLOAT
            actual = y.to_array().reshape(-1)[0]          # .taggedAs[1D].as[Array[Float]](0)
            assert actual == np.float32(expected)
            return float(actual)
    recs = [(ftf.example(x=x), y) for x, y in golden["half_plus_two"]["known_answers"]["regress_x_to_y"]]
    assert list(ftf.map_with_model(recs, model, fun)) == [2.0, 2.5, 3.0, 3.5]


def test_half_plus_two_signatures(ftf, oracle, model_dir):
    model = ftf.HalfPlusTwo(model_dir)
    model.open()
    try:
        rng = np.random.default_rng(1234)
        xs = rng.uniform(-1000, 1000, 3001).astype(np.float32)
        t = ftf.messages_to_tensor([ftf.example(x=float(v)) for v in xs])
        with model.regress_x_to_y(t) as y:
            assert y.shape == (3001, 1)
            assert np.array_equal(_bits(y.to_array().ravel()), _bits(oracle.affine_f32(xs, 0.5, 2.0)))
        with model.regress_x_to_y2(t) as y2:
            assert np.array_equal(_bits(y2.to_array().ravel()), _bits(oracle.affine_f32(xs, 0.5, 3.0)))
        with model.serving_default(ftf.TensorValue.from_array(xs.reshape(-1, 1))) as y:
            assert y.shape == (3001, 1)
            assert np.array_equal(_bits(y.to_array().ravel()), _bits(oracle.affine_f32(xs, 0.5, 2.0)))
        # the session's signature as ModelFunction sees it
        info = model.session("regress_x_to_y").signature_info()
        assert info == {"method_name": "tensorflow/serving/regress", "inputs": {"inputs": "tf_example:0"},
                        "outputs": {"outputs": "y:0"}}
        # empty batch
        with model.serving_default(ftf.TensorValue.from_array(np.zeros((0, 1), np.float32))) as y:
            assert y.shape == (0, 1)
        # missing required feature -> IllegalArgumentException
        bad = ftf.messages_to_tensor([ftf.example(z=1.0)])
        with pytest.raises(ValueError, match="Feature: x"):
            with model.regress_x_to_y(bad):
                pass
        with pytest.raises(ValueError, match="FLOAT"):
            with model.serving_default(ftf.TensorValue.from_array(np.zeros((2, 1), np.float64))):
                pass
    finally:
        model.close()


def test_create_errors(ftf, model_dir, tmp_path):
    with pytest.raises(KeyError, match="signature 'nope'"):
        ftf.Runner(model_dir, "nope")
    with pytest.raises(KeyError, match="SavedModel not found"):
        ftf.Runner(str(tmp_path), "regress_x_to_y")
    with pytest.raises(ValueError, match="does not match"):           # ModelFunction.scala:37
        ftf.Runner(model_dir, "regress_x_to_y", "tensorflow/serving/predict")
    with pytest.raises(ValueError, match="out of range"):
        ftf.Runner(model_dir, "regress_x_to_y", None, ftf.default_config(device=99))
    r = ftf.Runner(model_dir, "regress_x_to_y", "tensorflow/serving/regress")
    with pytest.raises(KeyError, match="no output named"):
        r.run({"inputs": ftf.messages_to_tensor([ftf.example(x=1.0)])}, ["scores"])
    with pytest.raises(ValueError, match="An input tensor named inputs must be provided"):
        r.run({"x": ftf.messages_to_tensor([ftf.example(x=1.0)])}, ["outputs"])
    r.close()
    r.close()                                                          # idempotent


def test_image_normalization_model(ftf, oracle):
    m = ftf.ImageNormalization()                                       # reference constants: 224x224, 117, 1
    m.open()
    try:
        img = _images(np.random.default_rng(41), 5, 480, 640)
        with m.normalize(ftf.TensorValue.from_array(img)) as out:
            assert out.shape == (5, 224, 224, 3) and out.data_type == ftf.DataType.FLOAT
            want = oracle.resize_bilinear_norm_u8(img, 224, 224, 117.0, 1.0, threads=8)
            assert np.array_equal(_bits(out.to_array()), _bits(want))
        with m.normalize(ftf.TensorValue.from_array(img[0])) as out:   # one HWC record -> ExpandDims -> [1,H,W,3]
            assert out.shape == (1, 224, 224, 3)
            assert np.array_equal(_bits(out.to_array()[0]), _bits(want[0]))
        with pytest.raises(ValueError, match="UINT8"):
            with m.normalize(ftf.TensorValue.from_array(img.astype(np.float32))):
                pass
    finally:
        m.close()


def test_large_run_is_chunked_and_pinned_input(ftf, oracle, model_dir):
    """> 32 MiB of input crosses several pipeline chunks; pinned caller memory skips the staging copy."""
    import ctypes as C
    lib = ftf._abi.load_library()
    r = ftf.Runner(model_dir, "serving_default")
    n = (24 << 20) + 13
    x = (np.random.default_rng(51).standard_normal(n) * 50).astype(np.float32)
    want = oracle.affine_f32(x, 0.5, 2.0, threads=8)
    got = r.run({"x": ftf.TensorValue.from_array(x.reshape(-1, 1))}, ["y"])["y"]
    assert got.shape == (n, 1) and np.array_equal(_bits(got.to_array().ravel()), _bits(want))
    s = r.stats()
    assert s["h2d_bytes"] == 4 * n and s["d2h_bytes"] == 4 * n and s["batches"] >= 3 and s["kernel_launches"] >= 3
    p = lib.ftf_host_alloc(4 * n)
    assert p
    pinned = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n,))
    pinned[:] = x
    tv = ftf.TensorValue.__new__(ftf.TensorValue)
    tv.data_type, tv.shape, tv.buffer = ftf.DataType.FLOAT, (n, 1), pinned.view(np.uint8)
    got = r.run({"x": tv}, ["y"])["y"]
    assert np.array_equal(_bits(got.to_array().ravel()), _bits(want))
    r.close()
    lib.ftf_host_free(p)


def test_micro_batched_records_keep_order_and_values(ftf, oracle, model_dir):
    """submit/poll behind a 1-in/1-out map: results in submission order, independent of batch composition."""
    cfg = ftf.default_config(max_batch=16, max_delay_us=500)
    r = ftf.Runner(model_dir, "regress_x_to_y", None, cfg)
    xs = np.random.default_rng(61).uniform(-10, 10, 103).astype(np.float32)
    got = {}
    for i, v in enumerate(xs):
        rec = ftf.messages_to_tensor([ftf.example(x=float(v))])
        while True:
            try:
                r.submit(i, rec)
                break
            except RuntimeError as e:                                   # RESOURCE_EXHAUSTED on a slow box: every slot is
                assert "undelivered" in str(e)                          # waiting to be polled; take results, offer again
                got.update(r.poll(block=True))
        if i % 5 == 0:
            for tag, tv in r.poll():
                got[tag] = tv
    r.flush()                                                           # snapshotState()/close(): drain
    while len(got) < len(xs):
        res = r.poll(block=True)
        assert res
        tags = [t for t, _ in res]
        assert tags == sorted(tags)
        got.update(res)
    assert list(got) == list(range(len(xs)))                            # submission order
    y = np.array([got[i].to_array().ravel()[0] for i in range(len(xs))], np.float32)
    assert np.array_equal(_bits(y), _bits(oracle.affine_f32(xs, 0.5, 2.0)))
    assert got[0].shape == (1,)
    s = r.stats()
    assert s["records"] == len(xs) and s["batches"] >= 7 and s["queue_latency_us_p99"] >= s["queue_latency_us_p50"] > 0
    with pytest.raises(ValueError, match="Feature: x"):
        r.submit(999, ftf.messages_to_tensor([ftf.example(q=1.0)]))
    r.close()


def test_micro_batch_deadline_and_images(ftf, oracle):
    import time
    cfg = ftf.default_config(max_batch=8, max_delay_us=100000, out_h=32, out_w=24, mean=117.0, scale=1.0)   # 100 ms: a busy box must not close the batch early
    r = ftf.Runner("builtin:image_normalization", None, None, cfg)
    img = _images(np.random.default_rng(71), 11, 60, 80)
    for i in range(3):
        r.submit(100 + i, ftf.TensorValue.from_array(img[i]))
    assert r.poll() == []                                              # batch still open, deadline not reached
    time.sleep(0.15)
    res = []
    while len(res) < 3:                                                # deadline passed: poll launches the short batch
        res += r.poll()
    for i in range(3, 11):
        r.submit(100 + i, ftf.TensorValue.from_array(img[i]))          # 8 records: launched by max_batch
    r.flush()
    while len(res) < 11:
        res += r.poll(block=True)
    assert [t for t, _ in res] == list(range(100, 111))
    want = oracle.resize_bilinear_norm_u8(img, 32, 24, 117.0, 1.0)
    for (tag, tv), w in zip(res, want):
        assert tv.shape == (32, 24, 3) and np.array_equal(_bits(tv.to_array()), _bits(w))
    r.close()


def test_mini_cluster_subtasks(ftf, oracle, model_dir):
    """4 subtasks (threads), one model instance each -- the multi-subtask run the reference never tests
    (both ITCases pin parallelism = 1).  Outputs must not depend on the parallelism."""
    import torch
    xs = np.random.default_rng(81).uniform(-100, 100, 64).astype(np.float32)
    recs = [ftf.example(x=float(v)) for v in xs]

    def fun(ex, m):
        with m.regress_x_to_y(ftf.messages_to_tensor([ex])) as y:
            return float(y.to_array().ravel()[0])
    n_gpus = torch.cuda.device_count()
    outs = [ftf.MiniCluster(p, n_gpus).run(recs, lambda dev: ftf.HalfPlusTwo(model_dir, device=dev), fun)
            for p in (1, 4)]
    assert outs[0] == outs[1] == oracle.affine_f32(xs, 0.5, 2.0).tolist()


@pytest.mark.parametrize("n", [1, 2, 4095, 4096, 4097, 9000])
def test_small_affine_calls_run_on_the_pinned_buffers_themselves(ftf, oracle, model_dir, n):
    """Up to 16 KB a blocking affine call makes no copies: K1 reads the pinned input and writes the pinned output (mapped host
    memory).  Same bits either side of that limit, edge values included, and repeated calls do not see one another's data."""
    model = ftf.HalfPlusTwo(model_dir)
    model.open()
    try:
        rng = np.random.default_rng(n)
        for rep in range(3):
            xs = rng.uniform(-1e6, 1e6, n).astype(np.float32)
            xs[:min(n, 6)] = np.array([np.nan, np.inf, -np.inf, 1e-42, -0.0, 3.4e38], np.float32)[:min(n, 6)]
            with model.serving_default(ftf.TensorValue.from_array(xs.reshape(-1, 1))) as y:
                assert y.shape == (n, 1)
                got, want = y.to_array().ravel(), oracle.affine_f32(xs, 0.5, 2.0)
                nan = np.isnan(want)
                assert np.array_equal(nan, np.isnan(got)) and np.array_equal(_bits(got)[~nan], _bits(want)[~nan])
    finally:
        model.close()
