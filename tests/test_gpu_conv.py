"""The tensor-core convolution (K3) and the Inception-v3 label path against the CPU oracle.  Needs a B200."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def dev(ftf):
    from flink_tensorflow_b200 import device
    return device


def _bf16(oracle, a):
    return oracle.bf16_round(np.asarray(a, np.float32))


def _to_bf16_cuda(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda().to(torch.bfloat16)


# every distinct kind of convolution Inception-v3 has, at small batch: (n, h, w, cin, cout, kh, kw, stride, ph, pw)
CONV_SHAPES = [
    (2, 17, 17, 64, 64, 1, 1, 1, 0, 0),        # 1x1, K = one 64-channel chunk
    (1, 35, 35, 192, 208, 1, 1, 1, 0, 0),      # Mixed_5b fused 1x1s: N = 208
    (2, 19, 19, 32, 32, 3, 3, 1, 0, 0),        # Conv2d_2a: Cin = 32 (half-filled K chunk), VALID
    (2, 19, 19, 32, 64, 3, 3, 1, 1, 1),        # Conv2d_2b: padded
    (1, 23, 23, 80, 192, 3, 3, 1, 0, 0),       # Conv2d_4a: Cin = 80 (1.25 chunks)
    (2, 35, 35, 48, 64, 5, 5, 1, 2, 2),        # 5x5, Cin = 48
    (2, 35, 35, 96, 96, 3, 3, 1, 1, 1),
    (2, 35, 35, 288, 384, 3, 3, 2, 0, 0),      # Mixed_6a: stride 2, N = 384 (two N tiles)
    (2, 17, 17, 128, 128, 1, 7, 1, 0, 3),      # 1x7
    (2, 17, 17, 160, 192, 7, 1, 1, 3, 0),      # 7x1, Cin = 160
    (3, 17, 17, 192, 320, 3, 3, 2, 0, 0),      # Mixed_7a: stride 2 to 8x8
    (3, 8, 8, 384, 384, 1, 3, 1, 0, 1),        # 1x3
    (3, 8, 8, 448, 384, 3, 3, 1, 1, 1),
    (2, 8, 8, 1280, 1344, 1, 1, 1, 0, 0),      # Mixed_7b fused 1x1s: N = 1344 (six N tiles)
    (5, 1, 1, 2048, 1008, 1, 1, 1, 0, 0),      # fc as a 1x1 convolution
    (1, 9, 9, 64, 16, 3, 3, 1, 1, 1),          # smallest N tile
]


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_tc_matches_oracle(dev, oracle, shape):
    import torch
    import oracle.inception as oi
    n, h, w, cin, cout, kh, kw, s, ph, pw = shape
    rng = np.random.default_rng(sum((i + 1) * v for i, v in enumerate(shape)))
    x = _bf16(oracle, rng.standard_normal((n, h, w, cin)))
    wt = _bf16(oracle, rng.standard_normal((cout, kh, kw, cin)) * np.sqrt(2.0 / (kh * kw * cin)))
    b = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    want = oi.conv2d_nhwc(x, wt, b, stride=s, pad=(ph, pw), relu=True)            # fp32 accumulate, unrounded
    xd, wd, bd = _to_bf16_cuda(x), _to_bf16_cuda(wt), torch.from_numpy(b).cuda()
    # fp32 output: only the accumulation order differs from the oracle -> 1e-4 relative to the output scale
    got32 = dev.conv2d_bf16(xd, wd, bd, s, (ph, pw), relu=True, out_dtype=torch.float32).cpu().numpy()
    assert got32.shape == want.shape
    scale = np.abs(want).max() + 1e-6
    assert np.abs(got32 - want).max() <= 1e-4 * scale, np.abs(got32 - want).max() / scale
    # bf16 output: the same values rounded once (RNE) -> within one bf16 ulp (2^-8 relative) of the oracle
    got16 = dev.conv2d_bf16(xd, wd, bd, s, (ph, pw), relu=True).float().cpu().numpy()
    assert np.all(np.abs(got16 - want) <= np.abs(want) * 2.0 ** -8 + 1e-4 * scale)
    # no bias, no ReLU (the pool-branch pre-activation)
    raw = dev.conv2d_bf16(xd, wd, None, s, (ph, pw), relu=False, out_dtype=torch.float32).cpu().numpy()
    want_raw = oi.conv2d_nhwc(x, wt, b, stride=s, pad=(ph, pw), relu=False, add_bias=False)
    assert np.abs(raw - want_raw).max() <= 1e-4 * (np.abs(want_raw).max() + 1e-6)


def test_conv_tc_writes_a_concat_slice(dev, oracle):
    import torch
    import oracle.inception as oi
    rng = np.random.default_rng(5)
    x = _bf16(oracle, rng.standard_normal((2, 17, 17, 64)))
    wt = _bf16(oracle, rng.standard_normal((32, 3, 3, 64)) * 0.06)
    b = rng.uniform(-0.1, 0.1, 32).astype(np.float32)
    out = torch.full((2, 17, 17, 96), 7.0, dtype=torch.bfloat16, device="cuda")
    dev.conv2d_bf16(_to_bf16_cuda(x), _to_bf16_cuda(wt), torch.from_numpy(b).cuda(), 1, (1, 1), out=out, out_off=48)
    got = out.float().cpu().numpy()
    want = oi.conv2d_nhwc(x, wt, b, pad=(1, 1))
    assert np.all(got[..., :48] == 7.0) and np.all(got[..., 80:] == 7.0)          # neighbours untouched
    assert np.all(np.abs(got[..., 48:80] - want) <= np.abs(want) * 2.0 ** -8 + 1e-4 * np.abs(want).max())


def test_conv_tc_is_batch_independent(dev, oracle):
    """A record's result must not depend on what else is in the micro-batch."""
    import torch
    rng = np.random.default_rng(6)
    x = _to_bf16_cuda(rng.standard_normal((5, 17, 17, 128)))
    wt = _to_bf16_cuda(rng.standard_normal((192, 1, 7, 128)) * 0.05)
    b = torch.from_numpy(rng.uniform(-0.1, 0.1, 192).astype(np.float32)).cuda()
    full = dev.conv2d_bf16(x, wt, b, 1, (0, 3))
    for i in (0, 2, 4):
        one = dev.conv2d_bf16(x[i:i + 1].contiguous(), wt, b, 1, (0, 3))
        assert torch.equal(one[0], full[i])


def test_conv_variants_agree(dev, monkeypatch):
    """The plan may pick either kernel (pixel-major, also as a CTA pair / weight-major), with or without the halo window.  All of them
    accumulate K in the same order (channel chunk, filter tap, 16-element step): the halo window changes nothing in a
    kernel's result, and the two kernels agree to fp32 accumulation rounding."""
    import torch
    rng = np.random.default_rng(21)
    for (n, h, w, cin, cout, kh, kw, ph, pw) in [(2, 35, 35, 96, 96, 3, 3, 1, 1), (2, 17, 17, 160, 192, 7, 1, 3, 0),
                                                   (2, 19, 19, 32, 64, 3, 3, 1, 1), (3, 8, 8, 448, 384, 3, 3, 1, 1)]:
        x = _to_bf16_cuda(rng.standard_normal((n, h, w, cin)))
        wt = _to_bf16_cuda(rng.standard_normal((cout, kh, kw, cin)) * np.sqrt(2.0 / (kh * kw * cin)))
        b = torch.from_numpy(rng.uniform(-0.1, 0.1, cout).astype(np.float32)).cuda()
        outs = {}
        for kind in ("pm", "wm", "pm2", "pmw"):
            for no_halo in ("1", "0"):
                if kind in ("pm2", "pmw") and no_halo == ("0" if kind == "pm2" else "1"):
                    continue                                   # the CTA-pair kernel has no halo variant, the wide one is nothing else
                monkeypatch.setenv("FTF_CONV_KIND", kind)
                monkeypatch.setenv("FTF_NO_HALO", no_halo)
                monkeypatch.setenv("FTF_CONV_TUNE", "0")
                outs[(kind, no_halo)] = dev.conv2d_bf16(x, wt, b, 1, (ph, pw), out_dtype=torch.float32).clone()
        for k in ("FTF_CONV_KIND", "FTF_NO_HALO", "FTF_CONV_TUNE"):
            monkeypatch.delenv(k)
        scale = float(outs[("pm", "1")].abs().max())
        diffs = {k: float((v - outs[("pm", "1")]).abs().max()) / scale for k, v in outs.items()}
        assert torch.equal(outs[("pm", "0")], outs[("pm", "1")]), diffs
        assert torch.equal(outs[("wm", "0")], outs[("wm", "1")]), diffs
        assert torch.equal(outs[("pm2", "1")], outs[("pm", "1")]), diffs      # cta_group::2: same K order, same bits
        assert torch.equal(outs[("pmw", "0")], outs[("pm", "1")]), diffs      # halo window with one wide accumulator set
        assert diffs[("wm", "1")] <= 1e-5, diffs


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_pixel_major_and_weight_major_give_the_same_bits(dev, monkeypatch, shape):
    """Every kernel family walks K in the same order and the tensor core's fp32 accumulation does not depend on which
    operand sits on the M side: whichever family the plan-time timing picks on a given GPU, a record's result is the same.
    (This is what lets subtasks on different GPUs, which time their candidates independently, agree bit for bit.)"""
    import torch
    n, h, w, cin, cout, kh, kw, s, ph, pw = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    x = torch.randn((n, h, w, cin), device="cuda", generator=g).to(torch.bfloat16)
    wt = (torch.randn((cout, kh, kw, cin), device="cuda", generator=g) * (2.0 / (kh * kw * cin)) ** 0.5).to(torch.bfloat16)
    b = torch.rand(cout, device="cuda", generator=g) * 0.2 - 0.1
    outs = {}
    monkeypatch.setenv("FTF_CONV_TUNE", "0")
    for kind in ("pm", "wm"):
        monkeypatch.setenv("FTF_CONV_KIND", kind)
        try:
            outs[kind] = dev.conv2d_bf16(x, wt, b, s, (ph, pw), out_dtype=torch.float32).clone()
        except ValueError:
            pytest.skip(f"no {kind} plan for this shape")
    assert torch.equal(outs["pm"], outs["wm"]), float((outs["pm"] - outs["wm"]).abs().max())


def test_pool_kernels_match_oracle(dev, oracle):
    """K4 max pool, K5 pool branch, K6 global average pool on bf16 tensors."""
    import torch
    import oracle.inception as oi
    rng = np.random.default_rng(12)
    x = _bf16(oracle, rng.standard_normal((3, 17, 17, 96)))
    xd = _to_bf16_cuda(x)
    # K4: exact (max of bf16 values), into a concat slice
    out = torch.zeros((3, 8, 8, 128), dtype=torch.bfloat16, device="cuda")
    dev.maxpool3x3s2(xd, out=out, out_off=32)
    win = np.lib.stride_tricks.sliding_window_view(x, (3, 3), axis=(1, 2))[:, ::2, ::2]      # [n,8,8,c,3,3]
    assert np.array_equal(out.float().cpu().numpy()[..., 32:], win.max(axis=(-1, -2)))
    assert torch.all(out[..., :32] == 0)
    # K5: relu(avgpool3x3(conv1x1(x)) / 9 + b), fp32 until the final bf16 rounding
    w = _bf16(oracle, rng.standard_normal((64, 96)) * 0.1)
    b = rng.uniform(-0.1, 0.1, 64).astype(np.float32)
    got = dev.pool_branch(xd, _to_bf16_cuda(w), torch.from_numpy(b).cuda()).float().cpu().numpy()
    pre = oi.conv2d_nhwc(x, w.reshape(64, 1, 1, 96), b, relu=False, add_bias=False)
    padded = np.pad(pre, ((0, 0), (1, 1), (1, 1), (0, 0)))
    pooled = np.lib.stride_tricks.sliding_window_view(padded, (3, 3), axis=(1, 2)).sum(axis=(-1, -2), dtype=np.float32)
    want = np.maximum(pooled * np.float32(1.0 / 9.0) + b, 0.0)
    assert np.all(np.abs(got - want) <= np.abs(want) * 2.0 ** -8 + 1e-4 * np.abs(want).max())
    # K6: fp32 mean of the pixels, rounded to bf16
    g = dev.global_avgpool(xd).float().cpu().numpy()
    want = x.reshape(3, -1, 96).mean(axis=1, dtype=np.float32)
    assert np.all(np.abs(g - want) <= np.abs(want) * 2.0 ** -8 + 1e-6)


@pytest.mark.parametrize("shape", [(2, 147, 147, 64), (2, 71, 71, 192), (1, 35, 35, 288), (2, 17, 17, 768), (3, 9, 13, 40), (1, 3, 3, 32), (2, 20, 6, 64)])
def test_maxpool_exact_on_every_tile_shape(dev, oracle, shape):
    """K4 on the Inception pool shapes (TMA-tiled kernel: 64- and 32-channel chunks, tiles that overhang the output grid) and on
    a channel count that falls back to the one-thread-per-output kernel; exact, written into a concat slice."""
    import torch
    rng = np.random.default_rng(sum(shape))
    n, H, W, C = shape
    x = _bf16(oracle, rng.standard_normal(shape))
    OH, OW = (H - 3) // 2 + 1, (W - 3) // 2 + 1
    out = torch.full((n, OH, OW, C + 16), -7.0, dtype=torch.bfloat16, device="cuda")
    dev.maxpool3x3s2(_to_bf16_cuda(x), out=out, out_off=8)
    win = np.lib.stride_tricks.sliding_window_view(x, (3, 3), axis=(1, 2))[:, ::2, ::2]
    got = out.float().cpu().numpy()
    assert np.array_equal(got[..., 8:8 + C], win.max(axis=(-1, -2)))
    assert np.all(got[..., :8] == -7.0) and np.all(got[..., 8 + C:] == -7.0)


@pytest.mark.parametrize("shape", [(1, 1, 1, 32), (1, 3, 5, 32), (2, 35, 35, 32), (1, 8, 8, 192), (2, 17, 17, 64), (1, 13, 9, 64), (2, 6, 7, 48), (1, 40, 40, 32)])
def test_pool_branch_sums_taps_in_row_major_order(dev, oracle, shape):
    """K5 alone, bit for bit: with identity 1x1 weights the pre-activation IS the input, and the kernel must add the nine taps
    row-major in fp32 (zero for taps outside the image), times 1/9, plus bias, ReLU, one bf16 rounding -- whatever the strip
    of rows a thread walks (strips of 6 or 4 rows; heights that are no multiple of either)."""
    import torch
    rng = np.random.default_rng(sum(shape))
    n, H, W, C = shape
    x = _bf16(oracle, rng.standard_normal(shape))
    b = rng.uniform(-0.1, 0.1, C).astype(np.float32)
    got = dev.pool_branch(_to_bf16_cuda(x), _to_bf16_cuda(np.eye(C, dtype=np.float32)), torch.from_numpy(b).cuda())
    padded = np.pad(x, ((0, 0), (1, 1), (1, 1), (0, 0)))
    s = np.zeros(shape, np.float32)
    for ky in range(3):
        for kx in range(3):
            s = s + padded[:, ky:ky + H, kx:kx + W, :]
    want = np.maximum(s * np.float32(1.0 / 9.0) + b, np.float32(0.0))
    want_bf16 = torch.from_numpy(want).to(torch.bfloat16)
    assert torch.equal(got.cpu(), want_bf16)


@pytest.fixture(scope="module")
def label_runner(ftf):
    r = ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=4, top_k=5, num_classes=1000, seed=2024))
    yield r
    r.close()


def _activation(ftf, r, name, n):
    lib = ftf._abi.load_library()
    dims = (C.c_int32 * 4)()
    assert lib.ftf_debug_activation(r.handle, name.encode(), n, None, 0, dims) == 0, lib.ftf_last_error(r.handle)
    out = np.empty(tuple(dims), np.float32)
    assert lib.ftf_debug_activation(r.handle, name.encode(), n, out.ctypes.data, out.size, dims) == 0
    return out


def test_inception_blocks_follow_the_oracle(ftf, oracle, label_runner):
    """Layer by layer against the bf16-emulating oracle: each tap within bf16 rounding noise of the oracle's."""
    import oracle.inception as oi
    x = np.random.default_rng(7).standard_normal((2, 299, 299, 3)).astype(np.float32)
    out = label_runner.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits"])["logits"].to_array()
    net = oi.Net(1000, 2024, mode=1)
    worst = {}
    logits = None
    for name in ["Conv2d_1a_3x3", "Conv2d_2a_3x3", "Conv2d_2b_3x3", "pool1", "Conv2d_3b_1x1", "Conv2d_4a_3x3", "pool2",
                 "Mixed_5b", "Mixed_5d", "Mixed_6a", "Mixed_6c", "Mixed_6e", "Mixed_7a", "Mixed_7c", "gap"]:
        logits, want = net.forward(x, tap=name)
        got = _activation(ftf, label_runner, name, 2)
        assert got.shape == want.shape, name
        # activations are bf16: a different accumulation order can flip a rounding (1 ulp = 2^-8 relative) and the
        # flips accumulate slowly with depth; bound the error relative to the tensor's scale
        err = np.abs(got - want).max() / (np.abs(want).max() + 1e-6)
        worst[name] = float(err)
        assert err <= 2e-2, (name, err, worst)
    assert np.abs(out - logits).max() <= 1e-2 * np.abs(logits).max(), worst


def test_inception_logits_and_top5(ftf, oracle, label_runner):
    """north_star: logits within 1e-2 relative in bf16, top-5 label indices bit-exact."""
    import oracle.inception as oi
    x = np.random.default_rng(7).standard_normal((2, 299, 299, 3)).astype(np.float32)
    res = label_runner.run({"inputs": ftf.TensorValue.from_array(x)}, ["outputs", "logits", "top_k_indices", "top_k_scores"])
    logits = res["logits"].to_array()
    assert logits.shape == (2, 1000) and res["outputs"].shape == (2, 1000) and res["top_k_indices"].shape == (2, 5)
    want = oi.Net(1000, 2024, mode=1).forward(x)
    assert np.abs(logits - want).max() <= 1e-2 * np.abs(want).max()
    wi, _ = oracle.top_labels(oracle.softmax_rows(want), 1000, 5)
    assert np.array_equal(res["top_k_indices"].to_array(), wi)
    # committed golden vectors: torchvision's fp32 logits for the same seeded images and weights
    g = np.load(os.path.join(HERE, "golden", "inception_v3.npz"))
    assert np.abs(logits - g["logits_torchvision"]).max() <= 1e-2 * np.abs(g["logits_torchvision"]).max()
    assert np.abs(want - g["logits_oracle_bf16"]).max() == 0.0                     # the oracle is deterministic
    gi = np.argsort(-g["logits_torchvision"], axis=1, kind="stable")[:, :5]
    assert np.array_equal(res["top_k_indices"].to_array(), gi)
    # probabilities: softmax of the logits the runner returned
    p = res["outputs"].to_array()
    assert np.allclose(p, oracle.softmax_rows(logits), rtol=1e-5, atol=1e-12)
    assert np.array_equal(res["top_k_scores"].to_array(), np.take_along_axis(p, res["top_k_indices"].to_array(), 1))


@pytest.mark.parametrize("kind", ["pm", "pm2", "pmw", "wm"])
def test_whole_network_on_one_kernel_family(ftf, oracle, monkeypatch, kind):
    """The network built from ONE kernel family (one-CTA pixel-major, CTA-pair pixel-major, weight-major), so that every
    layer shape -- the multi-segment fused 1x1 groups and the fp32 pool-branch segments included -- runs through that
    family's epilogue and shape logic, not only the layers the plan-time timing happens to give it."""
    import oracle.inception as oi
    monkeypatch.setenv("FTF_CONV_KIND", kind)
    monkeypatch.setenv("FTF_CONV_TUNE", "0")
    r = ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=2, top_k=5, num_classes=1000, seed=2024))
    try:
        x = np.random.default_rng(7).standard_normal((2, 299, 299, 3)).astype(np.float32)
        res = r.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices"])
    finally:
        r.close()
    want = oi.Net(1000, 2024, mode=1).forward(x)
    logits = res["logits"].to_array()
    assert np.abs(logits - want).max() <= 1e-2 * np.abs(want).max()
    wi, _ = oracle.top_labels(oracle.softmax_rows(want), 1000, 5)
    assert np.array_equal(res["top_k_indices"].to_array(), wi)


def test_two_label_runners_share_one_gpu(ftf):
    """Two Flink subtasks on one GPU: two runners created and driven concurrently from two threads give the bits a single
    runner gives.  Kernel choices are timed once per process and layer (under a lock) and taken over by every later runner,
    so all three runners here run the same kernels whatever the timing noise."""
    import threading
    x = np.random.default_rng(11).standard_normal((4, 299, 299, 3)).astype(np.float32)
    cfg = ftf.default_config(max_batch=4, top_k=5, num_classes=1000, seed=2024)
    ref = ftf.Runner("builtin:inception_v3", None, None, cfg)
    try:
        want = ref.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits"])["logits"].to_array()
    finally:
        ref.close()
    results, errors = [None, None], []

    def subtask(i):
        try:
            r = ftf.Runner("builtin:inception_v3", None, None, cfg)
            try:
                for _ in range(5):
                    results[i] = r.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits"])["logits"].to_array()
            finally:
                r.close()
        except Exception as ex:  # noqa: BLE001
            errors.append(ex)

    threads = [threading.Thread(target=subtask, args=(i,)) for i in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    assert np.array_equal(results[0], want) and np.array_equal(results[1], want)


def test_label_results_do_not_depend_on_batch(ftf, label_runner):
    x = np.random.default_rng(8).standard_normal((4, 299, 299, 3)).astype(np.float32)
    full = label_runner.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits"])["logits"].to_array()
    for i in (0, 3):
        one = label_runner.run({"inputs": ftf.TensorValue.from_array(x[i:i + 1])}, ["logits"])["logits"].to_array()
        assert np.array_equal(one[0], full[i])


def test_inception_model_api_and_text_labels(ftf, oracle):
    """InceptionModel.label + toTextLabels (InceptionModel.scala:67,81-89) through the mirror API."""
    labels = [f"label{i}" for i in range(1000)]
    m = ftf.InceptionModel(labels=labels, max_batch=2, top_k=3, seed=2024)
    m.open()
    try:
        x = np.random.default_rng(7).standard_normal((1, 299, 299, 3)).astype(np.float32)
        with m.label(ftf.TensorValue.from_array(x)) as t:
            assert t.shape == (1, 1000)
            got = m.to_text_labels(t)                      # default take = 3
            probs = t.to_array()
        assert len(got) == 1 and len(got[0].labels) == 3
        oi_, os_ = oracle.top_labels(probs, 1000, 3)
        assert [l for _, l in got[0].labels] == [labels[i] for i in oi_[0]]
        assert [s for s, _ in got[0].labels] == os_[0].tolist()
        with pytest.raises(ValueError, match="FLOAT tensor shaped"):
            with m.label(ftf.TensorValue.from_array(np.zeros((1, 224, 224, 3), np.float32))):
                pass
    finally:
        m.close()


def test_normalize_label_chain(ftf, oracle):
    """Raw RGB records -> K2 (bf16) -> Inception-v3 -> top-5 on one GPU; equals the two operators run apart."""
    import oracle.inception as oi
    cfg = ftf.default_config(max_batch=2, top_k=5, out_h=299, out_w=299, mean=128.0, scale=128.0, seed=2024)
    r = ftf.Runner("builtin:normalize_label", None, None, cfg)
    img = np.random.default_rng(9).integers(0, 256, (2, 480, 640, 3), dtype=np.uint8)
    res = r.run({"inputs": ftf.TensorValue.from_array(img)}, ["logits", "top_k_indices"])
    norm = oracle.resize_bilinear_norm_u8(img, 299, 299, 128.0, 128.0, threads=8)
    want = oi.Net(1000, 2024, mode=1).forward(norm)
    assert np.abs(res["logits"].to_array() - want).max() <= 1e-2 * np.abs(want).max()
    assert np.array_equal(res["top_k_indices"].to_array(), oracle.top_labels(oracle.softmax_rows(want), 1000, 5)[0])
    r.close()


def test_fp32_compute_mode_matches_the_fp32_oracle(ftf, oracle):
    """compute_dtype = FTF_FLOAT (north_star: logits within 1e-4 in fp32): fp32 activations, weights and accumulation --
    the precision the reference's Session.run computes in (InceptionModel.scala:58-67).  Checked against the fp32 oracle
    (mode 0) and the committed torchvision logits, layer taps included."""
    import oracle.inception as oi
    g = np.load(os.path.join(HERE, "golden", "inception_v3.npz"))
    x = np.random.default_rng(int(g["image_seed"])).standard_normal((2, 299, 299, 3)).astype(np.float32)
    cfg = ftf.default_config(max_batch=2, top_k=5, num_classes=1000, seed=int(g["seed"]), compute_dtype=ftf.DataType.FLOAT)
    r = ftf.Runner("builtin:inception_v3", None, None, cfg)
    try:
        res = r.run({"inputs": ftf.TensorValue.from_array(x)}, ["logits", "top_k_indices", "outputs"])
        logits = res["logits"].to_array()
        net = oi.Net(1000, int(g["seed"]), mode=0)
        want = net.forward(x)
        assert np.abs(logits - want).max() <= 1e-4 * np.abs(want).max(), np.abs(logits - want).max() / np.abs(want).max()
        tv = g["logits_torchvision"]
        assert np.abs(logits - tv).max() <= 1e-4 * np.abs(tv).max()
        assert np.array_equal(res["top_k_indices"].to_array(), np.argsort(-want, axis=1, kind="stable")[:, :5])
        assert np.allclose(res["outputs"].to_array(), oracle.softmax_rows(logits), rtol=1e-5, atol=1e-12)
        for name in ("Conv2d_1a_3x3", "Conv2d_4a_3x3", "Mixed_5b", "Mixed_6a", "Mixed_6e", "Mixed_7c", "gap"):
            _, tap = net.forward(x, tap=name)
            got = _activation(ftf, r, name, 2)
            assert got.shape == tap.shape and np.abs(got - tap).max() <= 1e-4 * (np.abs(tap).max() + 1e-6), name
        # a bf16 image batch has no meaning for this network
        with pytest.raises(ValueError):
            ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=1, compute_dtype=ftf.DataType.INT32))
    finally:
        r.close()
    # the chained runner normalises to fp32 in this mode
    cfg2 = ftf.default_config(max_batch=1, top_k=5, out_h=299, out_w=299, mean=128.0, scale=128.0, seed=2024, compute_dtype=ftf.DataType.FLOAT)
    r2 = ftf.Runner("builtin:normalize_label", None, None, cfg2)
    try:
        img = np.random.default_rng(9).integers(0, 256, (1, 480, 640, 3), dtype=np.uint8)
        got = r2.run({"inputs": ftf.TensorValue.from_array(img)}, ["logits"])["logits"].to_array()
        norm = oracle.resize_bilinear_norm_u8(img, 299, 299, 128.0, 128.0, threads=8)
        want = oi.Net(1000, 2024, mode=0).forward(norm)
        assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()
    finally:
        r2.close()


def test_early_layers_in_groups_of_images_give_the_same_bits(ftf, monkeypatch):
    """FTF_EARLY_GROUP=<n> (a measurement switch): the layers ahead of Mixed_5b run group by group of n images
    instead of layer by layer over the batch.  Same kernels on the same tensors: logits bit for bit those of the default order,
    also for a batch that ends inside a group."""
    x = np.random.default_rng(77).standard_normal((20, 299, 299, 3)).astype(np.float32)
    outs = []
    for group in (None, "8"):
        if group is None:
            monkeypatch.delenv("FTF_EARLY_GROUP", raising=False)
        else:
            monkeypatch.setenv("FTF_EARLY_GROUP", group)
        r = ftf.Runner("builtin:inception_v3", None, None, ftf.default_config(max_batch=24, top_k=5, num_classes=1000, seed=2024))
        try:
            outs.append([r.run({"inputs": ftf.TensorValue.from_array(x[:n])}, ["logits"])["logits"].to_array() for n in (20, 11, 3)])
        finally:
            r.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)
