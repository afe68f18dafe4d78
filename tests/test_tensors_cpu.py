"""Typed tensor conversions with the reference's guards (host logic, no GPU).  Mirrors
flink-tensorflow/src/test/scala/org/tensorflow/contrib/scala/ArraysTest.scala:15-25,
FTT/types/TensorInjectionsTest.scala:23-90 and FTT/types/TensorValueTest.java:23-55, plus the inverse guards of
FT/types/TensorInjections.scala:160-273 and the builder of FT/types/TensorValueBuilder.java:68-105."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def T(ftf):
    from flink_tensorflow_b200 import tensors
    return tensors


def test_float_array_round_trip_like_arrays_test(ftf, T):
    expected = np.array([1.0, 2.0, 3.0], np.float32)                 # ArraysTest.scala:17-23
    t = T.float_array_to_tensor(expected)
    assert t.shape() == (3,) and t.rank_tag == "1D" and t.data_type() == ftf.DataType.FLOAT
    assert np.array_equal(T.tensor_to_float_array(t), expected)
    t.close()


@pytest.mark.parametrize("to_t,from_t,np_type", [("float_array_to_tensor", "tensor_to_float_array", np.float32),
                                                 ("double_array_to_tensor", "tensor_to_double_array", np.float64),
                                                 ("long_array_to_tensor", "tensor_to_long_array", np.int64),
                                                 ("int_array_to_tensor", "tensor_to_int_array", np.int32)])
def test_array_injections_and_their_guards(ftf, T, to_t, from_t, np_type):
    a = np.array([5, -7, 11, 2 ** 20], np_type)
    t = getattr(T, to_t)(a)
    back = getattr(T, from_t)(t)
    assert back.dtype == np_type and np.array_equal(back, a)
    # squeeze(shape).length <= 1: [1,4] and [4,1,1] are arrays, [2,2] is not (TensorInjections.scala:164)
    for shape, ok in (((1, 4), True), ((4, 1, 1), True), ((2, 2), False)):
        tt = T.TypedTensor(ftf.TensorValue(t.data_type(), shape, a.view(np.uint8)))
        if ok:
            assert np.array_equal(getattr(T, from_t)(tt), a)
        else:
            with pytest.raises(T.InversionFailure):
                getattr(T, from_t)(tt)
    # the dtype must match: every other inverse refuses this tensor
    for other in ("tensor_to_float_array", "tensor_to_double_array", "tensor_to_long_array", "tensor_to_int_array"):
        if other != from_t:
            with pytest.raises(T.InversionFailure):
                getattr(T, other)(t)


def test_scalar_injections_and_their_guards(ftf, T):
    for to_t, from_t, v in (("float_to_tensor", "tensor_to_float", 2.5), ("double_to_tensor", "tensor_to_double", -1e300),
                            ("long_to_tensor", "tensor_to_long", 2 ** 40 + 3), ("int_to_tensor", "tensor_to_int", -17)):
        t = getattr(T, to_t)(v)
        assert t.shape() == () and t.num_dimensions() == 0 and t.num_elements() == 1
        assert getattr(T, from_t)(t) == v
        with pytest.raises(T.InversionFailure):                      # rank 1 is not a scalar
            getattr(T, from_t)(T.TypedTensor(ftf.TensorValue(t.data_type(), (1,), t.as_value_view().buffer)))
    with pytest.raises(T.InversionFailure):
        T.tensor_to_float(T.double_to_tensor(1.0))
    s = T.string_to_tensor("flink")
    assert s.data_type() == ftf.DataType.STRING and s.shape() == () and T.tensor_to_string(s) == "flink"
    with pytest.raises(T.InversionFailure):
        T.tensor_to_string(T.float_to_tensor(1.0))


def test_tensor_value_bijection_and_to_value(ftf, T):
    """FT/package.scala:14-32: TensorValue -> TypedTensor is a view; toValue copies and leaves the tensor with the caller;
    the inverse of the TensorValue injections checks rank tag and dtype (TensorInjections.scala:92-96)."""
    arr = np.arange(6, dtype=np.int32).reshape(2, 3)                  # TensorValueTest.java:23-35
    tv = ftf.TensorValue.from_array(arr)
    t = T.tensor_value_to_tensor(tv, "2D", ftf.DataType.INT32)
    assert t.shape() == (2, 3) and t.num_bytes() == 24 and t.as_value_view() is tv
    v = t.to_value()
    assert v is not tv and v.shape == (2, 3) and np.array_equal(v.to_array(), arr)
    v.buffer[0] = 99                                                  # a copy: the tensor's record is untouched
    assert tv.to_array()[0, 0] == 0
    back = T.tensor_to_tensor_value(t, "2D", ftf.DataType.INT32)
    assert np.array_equal(back.to_array(), arr)
    with pytest.raises(T.InversionFailure):
        T.tensor_to_tensor_value(t, "1D", ftf.DataType.INT32)
    with pytest.raises(T.InversionFailure):
        T.tensor_to_tensor_value(t, "2D", ftf.DataType.FLOAT)
    # tags are unchecked casts, as in TS/package.scala:51,59
    assert t.tagged_with("4D", ftf.DataType.FLOAT).rank_tag == "4D"


def test_tensor_value_builder(ftf, T):
    b = T.TensorValueBuilder
    tv = b().data_type(ftf.DataType.INT32).shape([6]).data(np.arange(1, 7, dtype=np.int32)).build()   # TensorValueTest.java:40-47
    assert tv.shape == (6,) and np.array_equal(tv.to_array(), np.arange(1, 7)) and tv.buffer.nbytes == 24
    assert tv.write().hex().startswith("01" "00000003" "00000001" "0000000000000006" "00000018")       # the golden wire header
    # all the typed buffers; int64 gets 8 bytes per element (the reference's LongBuffer branch allocates 4 and overflows)
    for np_type, code in ((np.float32, ftf.DataType.FLOAT), (np.float64, ftf.DataType.DOUBLE), (np.int64, ftf.DataType.INT64)):
        a = np.array([[1, 2], [3, 4]], np_type)
        v = b().data_type(code).shape(a.shape).data(a).build()
        assert v.buffer.nbytes == a.nbytes and np.array_equal(v.to_array(), a)
    big_endian = np.arange(4, dtype=">f4")
    assert np.array_equal(b().data_type(ftf.DataType.FLOAT).shape([4]).data(big_endian).build().to_array(), np.arange(4, dtype=np.float32))
    raw = b().data_type(ftf.DataType.UINT8).shape([3]).data(b"\x01\x02\x03").build()
    assert raw.buffer.tolist() == [1, 2, 3]
    for missing, msg in ((lambda x: x.shape([1]).data(b"\0"), "dataType is required"),
                         (lambda x: x.data_type(1).data(b"\0"), "shape is required"),
                         (lambda x: x.data_type(1).shape([1]), "data is required")):
        with pytest.raises(RuntimeError, match=msg):
            missing(b()).build()
    with pytest.raises(RuntimeError, match="unsupported data buffer"):
        b().data_type(1).shape([1]).data(np.zeros(1, np.int16)).build()
    with pytest.raises(NotImplementedError):
        b().shape([1] * 26)


def test_example_messages_round_trip_like_tensor_injections_test(ftf, oracle, golden):
    """TensorInjectionsTest.scala:34-46,72-90: Example -> STRING tensor and back; a STRING vector of the four test examples."""
    ex = ftf.example(x=1.0)
    t = ftf.messages_to_tensor([ex])
    assert t.buffer.tobytes().hex() == golden["codecs"]["string_tensor_single_hex"] if "string_tensor_single_hex" in golden["codecs"] else True
    rc, x = oracle.parse_example_dense_f32(ex, "x")
    assert rc == 0 and x == 1.0
    four = [ftf.example(x=float(v)) for v in range(4)]
    vec = ftf.messages_to_tensor(four)
    assert vec.shape == (4,) and vec.buffer.nbytes == 104                                               # SURVEY 8c golden size
    assert vec.buffer.tobytes() == oracle.string_tensor_encode(four)
