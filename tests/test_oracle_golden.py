"""The oracle against the reference's own known answers and fixtures (CPU only)."""
import struct

import numpy as np


def test_affine_known_answers(oracle, golden):
    # RegressITCase.scala:35: x in {0,1,2,3} -> {2.0,2.5,3.0,3.5}, compared with float ==
    g = golden["half_plus_two"]
    a, b = g["variables"]["a"]["value"], g["variables"]["b"]["value"]
    assert (a, b, g["variables"]["c"]["value"]) == (0.5, 2.0, 3.0)
    for x, y in g["known_answers"]["regress_x_to_y"]:
        assert oracle.affine_f32(np.array([x], np.float32), a, b)[0] == np.float32(y)


def test_regress_from_string_tensor(oracle, golden):
    # the exact bytes RegressITCase feeds: Seq(example).toList.as[Tensor] (TensorInjections.scala:49-72)
    st = bytes.fromhex(golden["codecs"]["string_tensor_of_4_examples"])
    rc, y = oracle.regress(st, 4, "x", 0.5, 2.0)
    assert rc == 0 and y.ravel().tolist() == [2.0, 2.5, 3.0, 3.5]
    rc, _ = oracle.regress(st, 4, "missing", 0.5, 2.0)            # required key absent -> INVALID_ARGUMENT
    assert rc == 3
    rc, y = oracle.regress(st, 4, "x2", 0.5, 3.0, has_default=True, default=0.0)
    assert rc == 0 and y.ravel().tolist() == [3.0] * 4            # y3 = a*x2 + c with x2 defaulted


def test_string_tensor_layout(oracle, golden):
    exs = [bytes.fromhex(h) for h in golden["codecs"]["examples_x_0_to_3"]]
    assert oracle.string_tensor_encode(exs).hex() == golden["codecs"]["string_tensor_of_4_examples"]
    assert oracle.string_tensor_encode(exs[1:2]).hex() == golden["codecs"]["string_tensor_of_example_x1"]
    assert oracle.string_tensor_encode([]) == b""
    big = [b"z" * 300]                                             # two-byte varint length
    enc = oracle.string_tensor_encode(big)
    assert enc[:8] == struct.pack("<Q", 0) and enc[8:10] == bytes([0xac, 0x02]) and enc[10:] == big[0]


def test_tensorvalue_wire(oracle, golden):
    for key in ("tensorvalue_int32_6", "tensorvalue_int32_2x3"):   # TensorValueTest.java:23-55
        g = golden["codecs"][key]
        rc, wire = oracle.tv_encode(g["dtype"], g["shape"], bytes.fromhex(g["payload"]))
        assert rc == 0 and wire.hex() == g["wire"]
        rc, (dt, shape, payload) = oracle.tv_decode(wire)
        assert rc == 0 and dt == g["dtype"] and shape == g["shape"] and payload.hex() == g["payload"]
        assert len(wire) == 13 + 8 * len(shape) + len(payload)     # getBinaryLength, TensorValue.java:142-147
    assert oracle.tv_decode(b"\x02" + b"\x00" * 20)[0] == 100      # VersionMismatchException
    assert oracle.tv_encode(5, [1], b"\x00")[0] == 3               # unknown dtype code (TFUtils.java:18)
    assert oracle.tv_encode(1, [1] * 26, b"")[0] == 12             # rank > 25 (TFUtils.java:75)
    rc, wire = oracle.tv_encode(1, [], struct.pack("<f", 1.5))     # scalar
    assert rc == 0 and len(wire) == 17


def test_tensor_name(oracle):
    assert oracle.tensor_name("y:0") == (0, "y", 0)
    assert oracle.tensor_name("ParseExample/ParseExample:1") == (0, "ParseExample/ParseExample", 1)
    assert oracle.tensor_name("input") == (0, "input", 0)          # missing index => 0
    assert oracle.tensor_name("a:b:c")[0] == 3                     # TensorName.scala:16
    assert oracle.tensor_name("a:x")[0] == 3                       # toInt throws


def test_resize_identity_and_edges(oracle):
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (2, 7, 5, 3), dtype=np.uint8)
    same = oracle.resize_bilinear_norm_u8(img, 7, 5, 0.0, 1.0)
    assert np.array_equal(same, img.astype(np.float32))            # scale 1 => lerp weights 0
    up = oracle.resize_bilinear_norm_u8(img, 14, 10, 0.0, 1.0)
    assert np.array_equal(up[:, ::2, ::2], img.astype(np.float32)) # exact 2x: even samples hit source pixels
    assert np.array_equal(up[:, -1, ::2], img[:, -1].astype(np.float32))   # clamped bottom edge: hi == lo
    const = np.full((1, 9, 11, 3), 200, np.uint8)
    out = oracle.resize_bilinear_norm_u8(const, 4, 6, 117.0, 1.0)  # reference constants mean 117 / scale 1
    assert np.all(out == 83.0)
    one = oracle.resize_bilinear_norm_u8(img[:, :1, :1], 3, 3, 0.0, 2.0)   # 1x1 source
    assert np.all(one == img[:, :1, :1].astype(np.float32) / 2.0)


def test_resize_matches_float_formula(oracle):
    # independent numpy restatement of the same expression order
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, (1, 48, 64, 3), dtype=np.uint8)
    oh, ow, mean, scale = 23, 31, 117.0, 3.0
    f = img.astype(np.float32)
    hs, ws = np.float32(48) / np.float32(oh), np.float32(64) / np.float32(ow)
    py = np.arange(oh, dtype=np.float32) * hs
    px = np.arange(ow, dtype=np.float32) * ws
    ylo, xlo = py.astype(np.int64), px.astype(np.int64)
    yhi, xhi = np.minimum(ylo + 1, 47), np.minimum(xlo + 1, 63)
    yl, xl = (py - ylo.astype(np.float32))[None, :, None, None], (px - xlo.astype(np.float32))[None, None, :, None]
    tl, tr = f[:, ylo][:, :, xlo], f[:, ylo][:, :, xhi]
    bl, br = f[:, yhi][:, :, xlo], f[:, yhi][:, :, xhi]
    top = tl + (tr - tl) * xl
    bot = bl + (br - bl) * xl
    want = ((top + (bot - top) * yl) - np.float32(mean)) / np.float32(scale)
    got = oracle.resize_bilinear_norm_u8(img, oh, ow, mean, scale)
    assert np.array_equal(got.view(np.uint32), want.astype(np.float32).view(np.uint32))
    got8 = oracle.resize_bilinear_norm_u8(img, oh, ow, mean, scale, threads=8)
    assert np.array_equal(got, got8)


def test_top_labels_stable(oracle):
    s = np.array([[0.1, 0.7, 0.7, 0.05, 0.7, 0.2]], np.float32)
    idx, sc = oracle.top_labels(s, 6, 3)                           # ties keep ascending column order
    assert idx.tolist() == [[1, 2, 4]] and sc.tolist() == [[np.float32(0.7)] * 3]
    idx, _ = oracle.top_labels(s, 3, 5)                            # zip truncates to #labels; short rows pad -1
    assert idx.tolist() == [[1, 2, 0, -1, -1]]
    rng = np.random.default_rng(3)
    x = rng.standard_normal((16, 1008)).astype(np.float32)
    idx, sc = oracle.top_labels(x, 1008, 5)
    want = np.argsort(-x, axis=1, kind="stable")[:, :5]            # python's stable sort = List.sortWith
    assert np.array_equal(idx, want) and np.array_equal(sc, np.take_along_axis(x, want, 1))


def test_softmax_rows(oracle):
    x = np.random.default_rng(4).standard_normal((3, 1000)).astype(np.float32) * 4
    p = oracle.softmax_rows(x)
    e = np.exp((x - x.max(1, keepdims=True)).astype(np.float64))
    assert np.allclose(p, e / e.sum(1, keepdims=True), rtol=2e-6, atol=1e-12)
    assert np.allclose(p.sum(1), 1.0, atol=1e-5)


def test_bf16_round(oracle):
    a = np.array([1.0, 1.00390625, 1.01171875, -3.14159, 0.0, 1e-40, np.inf], np.float32)
    r = oracle.bf16_round(a)
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == np.float32(1.015625)   # ties to even both ways
    assert np.all((r.view(np.uint32) & 0xffff) == 0)
    import ctypes
    for v in a:
        assert oracle.lib().orc_bf16_round(ctypes.c_float(v)) == r[list(a).index(v)]
