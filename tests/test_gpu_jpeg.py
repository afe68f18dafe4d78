"""DecodeJpeg on the GPU (J1-J3, csrc/jpeg.cu) through the C ABI: bit-exact against the libjpeg-turbo fixtures and the
oracle, alone and in front of the normalisation and labeling models.  Needs a B200 (-m gpu)."""
import io
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dev(ftf):
    import torch
    assert torch.cuda.is_available()
    from flink_tensorflow_b200 import device
    return device


@pytest.fixture(scope="module")
def cases():
    z = np.load(os.path.join(ROOT, "tests", "golden", "jpeg", "cases.npz"))
    names = sorted(k[:-4] for k in z.files if k.endswith(".jpg") and not k.startswith("refuse_"))
    return {n: (z[n + ".jpg"].tobytes(), z[n + ".rgb"]) for n in names}


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _photo(rng, h=480, w=640, noise=10.0):
    y, x = np.mgrid[0:h, 0:w]
    p = rng.uniform(0, 6.28, 3)
    img = np.stack([128 + 90 * np.sin(x / rng.uniform(20, 60) + p[c]) * np.cos(y / rng.uniform(15, 50)) for c in range(3)], -1)
    return np.clip(img + rng.normal(0, noise, (h, w, 3)), 0, 255).astype(np.uint8)


def _jpeg(arr, **kw):
    from PIL import Image
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def test_every_fixture_in_one_ragged_batch(dev, cases):
    """All fixtures (4:4:4 / 4:2:2 / 4:2:0 / grey, odd sizes down to 1x1, optimised tables, restart markers) as ONE batch:
    every image equals what libjpeg-turbo decoded."""
    import torch
    names = list(cases)
    b = dev.JpegBatch()
    flat = b.load([cases[n][0] for n in names]).decode()
    assert b.status() == [0] * len(names)
    assert b.segments > len(names)                      # restart intervals became extra segments
    for n, img in zip(names, b.images(flat)):
        assert np.array_equal(img.cpu().numpy(), cases[n][1]), n
    # a second load on the same object (smaller, other tables) and two decodes of one load
    flat2 = b.load([cases["gray_48x64"][0], cases["444_q98"][0]]).decode()
    flat3 = b.decode()
    b.status()
    torch.cuda.synchronize()
    assert torch.equal(flat2, flat3)
    assert np.array_equal(b.images(flat2)[1].cpu().numpy(), cases["444_q98"][1])
    b.close()


@pytest.mark.parametrize("kw", [dict(quality=85), dict(quality=92, subsampling=0), dict(quality=60, subsampling=1),
                                dict(quality=75, restart_marker_rows=1), dict(quality=90, optimize=True)])
def test_photo_sized_batch_matches_oracle(dev, oracle, kw):
    rng = np.random.default_rng(77)
    files = [_jpeg(_photo(rng), **kw) for _ in range(12)]
    imgs = dev.decode_jpeg(files)
    for f, g in zip(files, imgs):
        assert np.array_equal(g.cpu().numpy(), oracle.jpeg_decode(f))


def test_full_batch_is_independent_of_batch_composition(dev, oracle):
    """The bench's size (256 files of 640x480): an image decodes to the same bytes whatever batch it travels in (alone,
    in a batch of 7, in the full batch), and a sample of the full batch equals the oracle."""
    import torch
    rng = np.random.default_rng(55)
    distinct = [_jpeg(_photo(rng), quality=85) for _ in range(8)] + [_jpeg(_photo(rng), quality=70, restart_marker_rows=2)]
    files = [distinct[i % len(distinct)] for i in range(256)]
    b = dev.JpegBatch()
    flat = b.load(files).decode()
    assert b.status() == [0] * 256
    full = [im.clone() for im in b.images(flat)]
    for i in range(len(distinct), 256):                       # repeats of a file give the same image
        assert torch.equal(full[i], full[i % len(distinct)])
    for i in (0, 3, 8):
        assert np.array_equal(full[i].cpu().numpy(), oracle.jpeg_decode(files[i]))
    alone = dev.decode_jpeg([files[5]])[0]
    assert torch.equal(alone, full[5])
    seven = dev.decode_jpeg(files[2:9])
    for k, im in enumerate(seven):
        assert torch.equal(im, full[2 + k])
    b.close()


def test_corrupt_entropy_data_is_reported_per_image(dev, oracle):
    rng = np.random.default_rng(3)
    good = [_jpeg(_photo(rng, 96, 128), quality=80, optimize=True) for _ in range(3)]
    i = good[1].index(b"\xff\xda")
    start = i + 2 + int.from_bytes(good[1][i + 2:i + 4], "big")
    bad = good[1][:start + 50] + b"\xfe" * 200 + good[1][start + 250:]
    b = dev.JpegBatch()
    flat = b.load([good[0], bad, good[2]]).decode()
    try:
        st = b.status()
    except ValueError as e:
        assert "image 1" in str(e)
        st = b.last_status
    # an optimised table has no all-ones code, so the run of 0xFE bytes cannot decode
    assert st == [0, 3, 0]
    ims = b.images(flat)
    assert np.array_equal(ims[0].cpu().numpy(), oracle.jpeg_decode(good[0]))
    assert np.array_equal(ims[2].cpu().numpy(), oracle.jpeg_decode(good[2]))
    with pytest.raises(NotImplementedError):
        b.load([_jpeg(_photo(rng, 32, 32), progressive=True)])
    with pytest.raises(ValueError, match="image 1"):
        b.load([good[0], b"\xff\xd8 nonsense"])
    b.close()


def test_damaged_files_decode_to_completion_like_the_oracle(dev, oracle):
    """Random damage to the entropy-coded bytes of small files, one ragged batch: the call returns, images the device
    accepts equal the oracle's decode of the same bytes, and the others are flagged."""
    rng = np.random.default_rng(101)
    files = []
    for t in range(48):
        f = bytearray(_jpeg(_photo(rng, 40 + 8 * (t % 5), 56 + 8 * (t % 3)), quality=int(rng.integers(40, 96)),
                            subsampling=t % 3, optimize=bool(t & 1), **({"restart_marker_blocks": 4} if t % 4 == 0 else {})))
        sos = f.index(b"\xff\xda") + 14
        for _ in range(int(rng.integers(1, 5))):
            f[int(rng.integers(sos, len(f) - 2))] = int(rng.integers(0, 256))
        files.append(bytes(f))
    b = dev.JpegBatch()
    flat = b.load(files).decode()
    try:
        st = b.status()
    except ValueError:
        st = b.last_status
    assert set(st) <= {0, 3}
    checked = 0
    for f, s, im in zip(files, st, b.images(flat)):
        try:
            want = oracle.jpeg_decode(f)
        except ValueError:
            continue                                   # the strict oracle gives up where libjpeg would only warn
        if s == 0:
            assert np.array_equal(im.cpu().numpy(), want)
            checked += 1
    assert checked >= 5
    b.close()


def test_truncated_file_is_refused_like_tf_with_default_attributes(dev, oracle):
    """The reference's graph leaves try_recover_truncated at false: a file that ends inside its scan is invalid JPEG data,
    for the oracle and for the product; one that lost only its end-of-image marker too.  A premature end-of-image marker
    is libjpeg's warning case: the rest of the image decodes from zero bits."""
    rng = np.random.default_rng(8)
    f = _jpeg(_photo(rng, 80, 112), quality=80)
    for cut in (f[: len(f) * 2 // 3], f[:-2]):
        with pytest.raises(ValueError):
            oracle.jpeg_decode(cut)
        with pytest.raises(ValueError, match="ends inside the scan"):
            dev.decode_jpeg([cut])
    early = f[: len(f) * 2 // 3] + b"\xff\xd9"
    assert np.array_equal(dev.decode_jpeg([early])[0].cpu().numpy(), oracle.jpeg_decode(early))


def test_normalization_model_takes_image_files(ftf, oracle):
    """The reference's own call: ImageFileTensor (file bytes as a 0-D STRING tensor) -> normalize -> [1,224,224,3]
    (EX/inception/ImageInputFormat.scala:63-80, ImageNormalization.scala:56-65), constants 224/117/1."""
    rng = np.random.default_rng(19)
    files = [_jpeg(_photo(rng), quality=85) for _ in range(5)]
    m = ftf.ImageNormalization()
    m.open()
    try:
        with m.normalize(ftf.bytes_to_tensor(files[0])) as out:
            assert out.shape == (1, 224, 224, 3) and out.data_type == ftf.DataType.FLOAT
            want0 = oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(files[0])[None], 224, 224, 117.0, 1.0)
            assert np.array_equal(_bits(out.to_array()), _bits(want0))
        with m.normalize(ftf.messages_to_tensor(files)) as out:             # a vector of files: one batch
            want = np.concatenate([oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(f)[None], 224, 224, 117.0, 1.0) for f in files])
            assert np.array_equal(_bits(out.to_array()), _bits(want))
        ragged = [_jpeg(_photo(rng, h, w), quality=80, subsampling=s) for (h, w, s) in ((100, 150, 2), (33, 47, 0), (240, 320, 1))]
        with m.normalize(ftf.messages_to_tensor(ragged)) as out:            # files of different sizes
            want = np.concatenate([oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(f)[None], 224, 224, 117.0, 1.0) for f in ragged])
            assert np.array_equal(_bits(out.to_array()), _bits(want))
        with pytest.raises(ValueError, match="DecodeJpeg"):
            with m.normalize(ftf.bytes_to_tensor(b"GIF89a not a jpeg")):
                pass
        img = np.random.default_rng(2).integers(0, 256, (2, 60, 80, 3), dtype=np.uint8)
        with m.normalize(ftf.TensorValue.from_array(img)) as out:           # raw RGB still works on the same session
            assert np.array_equal(_bits(out.to_array()), _bits(oracle.resize_bilinear_norm_u8(img, 224, 224, 117.0, 1.0)))
    finally:
        m.close()


def test_camera_sized_file(ftf, dev, oracle):
    """A 3-megapixel file (several waves of chunks per Huffman CTA, the direct resize kernel behind it) against the oracle."""
    rng = np.random.default_rng(31)
    f = _jpeg(_photo(rng, 1500, 2000, noise=14.0), quality=90)
    want = oracle.jpeg_decode(f)
    assert np.array_equal(dev.decode_jpeg([f])[0].cpu().numpy(), want)
    m = ftf.ImageNormalization()
    m.open()
    try:
        with m.normalize(ftf.bytes_to_tensor(f)) as out:
            assert np.array_equal(_bits(out.to_array()), _bits(oracle.resize_bilinear_norm_u8(want[None], 224, 224, 117.0, 1.0)))
    finally:
        m.close()


def test_file_records_through_submit_poll(ftf, oracle):
    """One JPEG file per record through the micro-batched path: order kept, values equal the oracle's."""
    rng = np.random.default_rng(23)
    files = [_jpeg(_photo(rng, 120, 160), quality=80) for _ in range(11)]
    cfg = ftf.default_config(max_batch=4, max_delay_us=500, out_h=32, out_w=24, mean=117.0, scale=1.0)
    r = ftf.Runner("builtin:image_normalization", None, None, cfg)
    for i, f in enumerate(files):
        r.submit(500 + i, ftf.bytes_to_tensor(f))
    with pytest.raises(NotImplementedError):
        r.submit(999, ftf.bytes_to_tensor(_jpeg(_photo(rng, 32, 32), progressive=True)))   # refused, joins no batch
    r.flush()
    res = []
    while len(res) < len(files):
        res += r.poll(block=True)
    assert [t for t, _ in res] == list(range(500, 500 + len(files)))
    for (_, tv), f in zip(res, files):
        want = oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(f)[None], 32, 24, 117.0, 1.0)[0]
        assert np.array_equal(_bits(tv.to_array()), _bits(want))
    r.close()


def test_one_corrupt_file_fails_alone_in_its_micro_batch(ftf, oracle):
    """A record whose file turns out corrupt on the device fails by itself (the reference fails that record's map() call
    only); the records batched and polled with it keep their results, in order."""
    rng = np.random.default_rng(31)
    files = [_jpeg(_photo(rng, 96, 128), quality=80, optimize=True) for _ in range(6)]
    i = files[2].index(b"\xff\xda")
    start = i + 2 + int.from_bytes(files[2][i + 2:i + 4], "big")
    files[2] = files[2][:start + 50] + b"\xfe" * 200 + files[2][start + 250:]     # no all-ones code in an optimised table
    cfg = ftf.default_config(max_batch=4, max_delay_us=500, out_h=32, out_w=24, mean=117.0, scale=1.0)
    r = ftf.Runner("builtin:image_normalization", None, None, cfg)
    try:
        for k, f in enumerate(files):
            r.submit(70 + k, ftf.bytes_to_tensor(f))
        r.flush()
        res = []
        while len(res) < len(files):
            res += r.poll(block=True)                     # must not raise: the bad record is an entry, not a failed call
        assert [t for t, _ in res] == list(range(70, 76))
        for k, (_, tv) in enumerate(res):
            if k == 2:
                assert isinstance(tv, ValueError) and "invalid JPEG data" in str(tv)
                continue
            want = oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(files[k])[None], 32, 24, 117.0, 1.0)[0]
            assert np.array_equal(_bits(tv.to_array()), _bits(want)), k
        # the C-level view of the same thing: ftf_poll returns OK, ftf_poll_status names the record
        import ctypes as C
        lib = ftf._abi.load_library()
        r.submit(1, ftf.bytes_to_tensor(files[2]))
        r.submit(2, ftf.bytes_to_tensor(files[0]))
        r.flush()
        from flink_tensorflow_b200.host import FtfTensor
        tags, outs, n = (C.c_uint64 * 8)(), (FtfTensor * 8)(), C.c_int32()
        assert lib.ftf_poll(r.handle, 1, tags, outs, 8, C.byref(n)) == 0 and n.value == 2
        st, msg = (C.c_int32 * 8)(), C.c_char_p()
        assert lib.ftf_poll_status(r.handle, st, 8, C.byref(msg)) == ftf._abi.INVALID_ARGUMENT
        assert list(st[:2]) == [ftf._abi.INVALID_ARGUMENT, 0] and b"image 0" in msg.value
    finally:
        r.close()


def test_forged_giant_header_is_refused_before_anything_is_sized(dev, cases):
    """TF refuses frames of 2^29 or more output bytes; a 65535x65535 header must not become a multi-gigabyte allocation."""
    f = bytearray(cases["444_q98"][0])
    i = f.index(b"\xff\xc0")
    f[i + 5:i + 9] = b"\xff\xff\xff\xff"                  # height, width
    b = dev.JpegBatch()
    with pytest.raises(ValueError, match="too large"):
        b.load([bytes(f)])
    b.close()


def test_refused_file_leaves_the_open_micro_batch_as_it_was(ftf, oracle, cases):
    """A file that ends inside its scan is refused at submit (try_recover_truncated is false in the reference's graph) and
    joins no batch; what it had reserved in the open batch is given back, so its neighbours decode and resize as one
    uniform batch and come back in order."""
    good = cases["420_48x64_q75"][0]
    cfg = ftf.default_config(max_batch=8, max_delay_us=500, out_h=20, out_w=28, mean=117.0, scale=1.0)
    r = ftf.Runner("builtin:image_normalization", None, None, cfg)
    try:
        r.submit(1, ftf.bytes_to_tensor(good))
        with pytest.raises(ValueError, match="ends inside the scan"):
            r.submit(2, ftf.bytes_to_tensor(good[: len(good) * 2 // 3]))
        r.submit(3, ftf.bytes_to_tensor(good))
        r.flush()
        res = r.poll(block=True)
        assert [t for t, _ in res] == [1, 3]
        want = oracle.resize_bilinear_norm_u8(cases["420_48x64_q75"][1][None], 20, 28, 117.0, 1.0)[0]
        for _, tv in res:
            assert np.array_equal(_bits(tv.to_array()), _bits(want))
    finally:
        r.close()


def test_files_and_decoded_pixels_label_alike(ftf, oracle):
    """normalize_label fed JPEG files gives exactly what it gives when fed the oracle-decoded pixels of those files."""
    rng = np.random.default_rng(29)
    files = [_jpeg(_photo(rng), quality=85) for _ in range(3)]
    cfg = ftf.default_config(max_batch=4, top_k=5, out_h=299, out_w=299, mean=128.0, scale=128.0, seed=2024)
    r = ftf.Runner("builtin:normalize_label", None, None, cfg)
    a = r.run({"inputs": ftf.messages_to_tensor(files)}, ["logits", "top_k_indices"])
    pix = np.stack([oracle.jpeg_decode(f) for f in files])
    b = r.run({"inputs": ftf.TensorValue.from_array(pix)}, ["logits", "top_k_indices"])
    assert np.array_equal(_bits(a["logits"].to_array()), _bits(b["logits"].to_array()))
    assert np.array_equal(a["top_k_indices"].to_array(), b["top_k_indices"].to_array())
    r.close()


def test_file_records_in_pinned_memory_are_staged_behind_the_callers_back(ftf, oracle, cases):
    """JPEG records whose bytes lie in ftf_host_alloc memory are unstuffed by the slot's worker threads while the caller submits
    the next ones (pageable payloads are staged inside ftf_submit).  Both kinds in one micro-batch, a file with restart markers,
    one that does not end with its EOI marker (never deferred), many batches back to back: order kept, values equal the oracle's."""
    from flink_tensorflow_b200.host import FtfTensor, _check
    lib = ftf._abi.load_library()
    rng = np.random.default_rng(31)
    base = [_jpeg(_photo(rng, 120, 160), quality=80) for _ in range(5)]
    base += [cases["420_rst_rows"][0], cases["444_rst_blocks"][0]]          # restart intervals: several segments per file
    base.append(base[1] + b"\x00\x00")          # trailing bytes behind EOI: staged in place
    files = [base[i % len(base)] for i in range(90)]
    blobs = [ftf.bytes_to_tensor(f).buffer for f in files]
    stride = (max(b.nbytes for b in blobs) + 255) // 256 * 256
    p = lib.ftf_host_alloc(len(files) * stride)
    assert p
    ring = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(len(files) * stride,))
    cfg = ftf.default_config(max_batch=16, max_delay_us=100000, out_h=32, out_w=24, mean=117.0, scale=1.0)
    r = ftf.Runner("builtin:image_normalization", None, None, cfg)
    res = []
    try:
        for i, b in enumerate(blobs):
            if i % 5 == 4:                      # a pageable payload between the pinned ones
                t = FtfTensor(ftf.DataType.STRING, 0, None, b.ctypes.data, b.nbytes)
            else:
                ring[i * stride:i * stride + b.nbytes] = b
                t = FtfTensor(ftf.DataType.STRING, 0, None, p + i * stride, b.nbytes)
            while True:
                rc = lib.ftf_submit(r.handle, 700 + i, C.byref(t))
                if rc != ftf._abi.RESOURCE_EXHAUSTED:
                    _check(rc, r.handle)
                    break
                res += r.poll(block=True)
        r.flush()
        while len(res) < len(files):
            res += r.poll(block=True)
    finally:
        r.close()
        lib.ftf_host_free(p)
    assert [t for t, _ in res] == list(range(700, 700 + len(files)))
    want = {}
    for (_, tv), f in zip(res, files):
        if f not in want:
            want[f] = oracle.resize_bilinear_norm_u8(oracle.jpeg_decode(f)[None], 32, 24, 117.0, 1.0)[0]
        assert np.array_equal(_bits(tv.to_array()), _bits(want[f]))
