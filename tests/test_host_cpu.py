"""Host-side logic of the product and the C ABI surface; no GPU needed (no compute calls)."""
import ctypes as C
import json
import os
import re
import struct

import numpy as np
import pytest


def test_abi_exports_every_declared_symbol(ftf):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "ftf.h")).read()
    declared = set(re.findall(r"\b(ftf_[a-z0-9_]+)\s*\(", header))
    declared -= {"ftf_runner", "ftf_tensor", "ftf_config", "ftf_stats"}
    lib = ftf._abi.load_library()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/ftf.h but not exported"
    assert declared == set(ftf._abi.SYMBOLS), declared ^ set(ftf._abi.SYMBOLS)
    assert b"sm_100a" in lib.ftf_version()


def test_no_device_is_a_loud_error(ftf, model_dir):
    import torch
    if torch.cuda.is_available():
        pytest.skip("only meaningful on a box without a GPU")
    with pytest.raises(RuntimeError, match="no CPU path"):
        ftf.Runner(model_dir, "regress_x_to_y")


def _describe(ftf, path):
    lib = ftf._abi.load_library()
    need = C.c_int64()
    rc = lib.ftf_savedmodel_describe(path.encode(), None, None, 0, C.byref(need))
    assert rc == 0, lib.ftf_last_error(None)
    buf = C.create_string_buffer(need.value)
    assert lib.ftf_savedmodel_describe(path.encode(), None, buf, need.value, C.byref(need)) == 0
    return json.loads(buf.value)


@pytest.mark.parametrize("which", ["own", "reference"])
def test_savedmodel_reader_matches_reference_metagraph(ftf, golden, model_dir, which):
    path = model_dir if which == "own" else "/root/reference/models/half_plus_two"
    if not os.path.isdir(path):
        pytest.skip("reference checkout not present on this box")
    d, g = _describe(ftf, path), golden["half_plus_two"]
    assert d["tags"] == g["tags"]
    assert set(d["signatures"]) == set(g["signatures"])
    for key, sig in g["signatures"].items():
        got = d["signatures"][key]
        assert got["method_name"] == sig["method_name"]
        assert got["inputs"] == sig["inputs"] and got["outputs"] == sig["outputs"]
    for name, v in g["variables"].items():
        assert d["variables"][name]["value"] == v["value"] and d["variables"][name]["offset"] == v["offset"]
    aff = {k: v["affine"] for k, v in d["signatures"].items()}
    assert (aff["regress_x_to_y"]["a"], aff["regress_x_to_y"]["b"]) == (0.5, 2.0)
    assert aff["regress_x_to_y"]["examples"] and aff["regress_x_to_y"]["feature_key"] == "x"
    assert not aff["regress_x_to_y"]["has_default"]
    assert (aff["regress_x_to_y2"]["b"], aff["regress_x2_to_y3"]["b"]) == (3.0, 3.0)
    assert not aff["serving_default"]["examples"] and not aff["regress_x2_to_y3"]["examples"]


def test_savedmodel_errors(ftf, tmp_path, model_dir):
    lib = ftf._abi.load_library()
    need = C.c_int64()
    assert lib.ftf_savedmodel_describe(str(tmp_path).encode(), None, None, 0, C.byref(need)) == ftf._abi.NOT_FOUND
    assert lib.ftf_savedmodel_describe(model_dir.encode(), b"train", None, 0, C.byref(need)) == ftf._abi.NOT_FOUND
    # corrupt the bundle: a flipped payload bit must fail the crc check, a bad magic must be rejected
    import shutil
    bad = tmp_path / "m"
    shutil.copytree(model_dir, bad)
    data = bad / "variables" / "variables.data-00000-of-00001"
    raw = bytearray(data.read_bytes()); raw[0] ^= 1; data.write_bytes(bytes(raw))
    d = _describe(ftf, str(bad))
    assert "value" not in d["variables"]["a"] and d["variables"]["b"]["value"] == 2.0
    assert "affine" not in d["signatures"]["regress_x_to_y"]
    idx = bad / "variables" / "variables.index"
    raw = bytearray(idx.read_bytes()); raw[-1] ^= 0xff; idx.write_bytes(bytes(raw))
    assert lib.ftf_savedmodel_describe(str(bad).encode(), None, None, 0, C.byref(need)) == ftf._abi.INVALID_ARGUMENT


def test_tensorvalue_wire_roundtrip_and_golden(ftf, oracle, golden):
    # TensorValueTest.java:23-35: int[2][3] -> shape (2,3), row-major flat buffer
    tv = ftf.TensorValue.from_array(np.array([[1, 2, 3], [4, 5, 6]], np.int32))
    assert tv.shape == (2, 3) and tv.num_dimensions() == 2
    flat = tv.buffer.view(np.int32)
    assert flat[1] == 2 and flat[3] == 4
    assert tv.write().hex() == golden["codecs"]["tensorvalue_int32_2x3"]["wire"]
    # TensorValueTest.java:39-56: 6 int32 -> 1-D, round trip
    tv = ftf.TensorValue.from_array(np.arange(1, 7, dtype=np.int32))
    wire = tv.write()
    assert wire.hex() == golden["codecs"]["tensorvalue_int32_6"]["wire"]
    assert len(wire) == tv.get_binary_length()
    back = ftf.TensorValue.read(wire)
    assert back.shape == (6,) and back.data_type == ftf.DataType.INT32
    assert back.to_array().tolist() == [1, 2, 3, 4, 5, 6]
    # product codec == oracle codec on random tensors of every dtype and rank 0..4
    rng = np.random.default_rng(5)
    for dt in (np.float32, np.float64, np.int32, np.int64, np.bool_, np.uint8):
        for rank in range(5):
            shape = tuple(int(v) for v in rng.integers(0, 4, rank))
            arr = rng.integers(0, 2, shape).astype(dt)
            tv = ftf.TensorValue.from_array(arr)
            rc, want = oracle.tv_encode(tv.data_type, list(shape), arr.tobytes())
            if dt is np.uint8:                      # not a TFUtils.java code: the oracle (== reference) refuses it,
                assert rc == 3                      # the product carries it as an extension for raw RGB records
                want = tv.write()
            assert rc in (0, 3) and tv.write() == want
            rt = ftf.TensorValue.read(want)
            assert rt.shape == shape and rt.buffer.tobytes() == arr.tobytes()


def test_tensorvalue_wire_errors(ftf):
    good = ftf.TensorValue.from_array(np.zeros(3, np.float32)).write()
    with pytest.raises(ftf.VersionMismatchException, match="incompatible tensor value"):
        ftf.TensorValue.read(b"\x02" + good[1:])
    with pytest.raises(ValueError):
        ftf.TensorValue.read(good[:1] + struct.pack(">i", 5) + good[5:])      # dtype code 5 unknown
    with pytest.raises(NotImplementedError, match="rank > 25"):
        ftf.TensorValue.read(good[:5] + struct.pack(">i", 26) + good[9:])
    with pytest.raises(IndexError):
        ftf.TensorValue.read(good[:-1])                                         # truncated payload
    with pytest.raises(NotImplementedError):
        ftf.TensorValue(1, (1,) * 26, b"\0\0\0\0").write()


def test_tensor_name(ftf, oracle):
    for s in ["y:0", "ParseExample/ParseExample:1", "input", "x:12", "op:"]:
        t = ftf.TensorName(s)
        rc, op, idx = oracle.tensor_name(s)
        assert rc == 0 and (t.name, t.index) == (op, idx)
    assert str(ftf.TensorName("input")) == "input:0"
    for s in ["a:b:c", "a:x", "a:1:2"]:
        with pytest.raises(ValueError, match="not a valid tensor name"):
            ftf.TensorName(s)
        assert oracle.tensor_name(s)[0] == 3


def test_string_tensor_and_example_codec(ftf, oracle, golden):
    exs = [ftf.example(x=float(v)) for v in range(4)]
    assert [e.hex() for e in exs] == golden["codecs"]["examples_x_0_to_3"]
    t = ftf.messages_to_tensor(exs)
    assert t.data_type == ftf.DataType.STRING and t.shape == (4,)
    assert t.buffer.tobytes().hex() == golden["codecs"]["string_tensor_of_4_examples"]
    assert ftf.messages_to_tensor(exs[1:2]).buffer.tobytes().hex() == golden["codecs"]["string_tensor_of_example_x1"]
    rng = np.random.default_rng(6)
    items = [rng.bytes(int(n)) for n in rng.integers(0, 700, 50)]            # > the reference's 10 000-byte scratch
    assert ftf.messages_to_tensor(items).buffer.tobytes() == oracle.string_tensor_encode(items)


def test_parse_examples_host(ftf, oracle):
    lib = ftf._abi.load_library()
    rng = np.random.default_rng(7)
    xs = rng.uniform(-1000, 1000, 257).astype(np.float32)
    msgs = [ftf.example(x=float(v), other=[1.0, 2.0]) for v in xs]
    st = ftf.messages_to_tensor(msgs).buffer
    out = np.empty(len(xs), np.float32)
    assert lib.ftf_parse_examples_f32(st.ctypes.data, st.nbytes, len(xs), b"x", 0, 0.0, out.ctypes.data) == 0
    assert np.array_equal(out, xs)
    for m, v in zip(msgs[:8], xs):
        assert oracle.parse_example_dense_f32(m, "x") == (0, v)
    # missing required key / default / wrong arity / malformed
    assert lib.ftf_parse_examples_f32(st.ctypes.data, st.nbytes, len(xs), b"x2", 0, 0.0, out.ctypes.data) == 3
    assert b"x2" in lib.ftf_last_error(None)
    assert lib.ftf_parse_examples_f32(st.ctypes.data, st.nbytes, len(xs), b"x2", 1, 7.5, out.ctypes.data) == 0
    assert np.all(out == 7.5)
    assert lib.ftf_parse_examples_f32(st.ctypes.data, st.nbytes, len(xs), b"other", 0, 0.0, out.ctypes.data) == 3
    assert oracle.parse_example_dense_f32(msgs[0], "other")[0] == 3
    trunc = ftf.messages_to_tensor([msgs[0][:-3]]).buffer
    assert lib.ftf_parse_examples_f32(trunc.ctypes.data, trunc.nbytes, 1, b"x", 0, 0.0, out.ctypes.data) == 3
    assert oracle.parse_example_dense_f32(msgs[0][:-3], "x")[0] == 3
    # unpacked float encoding (wire type 5) is accepted as well
    unpacked = bytes.fromhex("0a0e0a0c0a0178120712050d0000803f")
    assert oracle.parse_example_dense_f32(unpacked, "x") == (0, 1.0)
    u = ftf.messages_to_tensor([unpacked]).buffer
    assert lib.ftf_parse_examples_f32(u.ctypes.data, u.nbytes, 1, b"x", 0, 0.0, out.ctypes.data) == 0 and out[0] == 1.0


def test_model_function_requires(ftf):
    with pytest.raises(ValueError, match="a session must be provided"):
        ftf.ModelFunction(None, ftf.RegressionMethod())

    class FakeSession:
        def signature_info(self):
            return {"method_name": "tensorflow/serving/regress", "inputs": {"inputs": "tf_example:0"},
                    "outputs": {"outputs": "y:0"}}

        def run(self, fed, fetch):
            return {"outputs": fed["inputs"]}
    with pytest.raises(ValueError, match="requirement failed"):                 # ModelFunction.scala:37
        ftf.ModelFunction(FakeSession(), ftf.LabelMethod())
    f = ftf.ModelFunction(FakeSession(), ftf.RegressionMethod())
    with pytest.raises(ValueError, match="Input must be provided"):
        with f(None):
            pass

    class WrongKey(ftf.RegressionMethod):
        def inputs(self, i):
            return {"examples": i}
    with pytest.raises(ValueError, match="An input tensor named inputs must be provided"):
        with ftf.ModelFunction(FakeSession(), WrongKey())(1):
            pass
    with f(42) as y:
        assert y == 42


def test_tensorflow_model_lifecycle(ftf, model_dir):
    m = ftf.HalfPlusTwo(model_dir)
    m.close()                                            # close before open is a no-op (GenericModel.scala:34-43)
    m.open()
    with pytest.raises(RuntimeError, match="checkState"):   # TensorFlowModel.scala:40
        m.open()
    m.close()
    m.close()
    with pytest.raises(ValueError, match="a session must be provided"):
        m.regress_x_to_y
