"""Typed tensor conversions of the reference's marshalling layer, with its guards (host side, numpy only).

What each piece mirrors (paths relative to /root/reference; FT = flink-tensorflow/src/main/scala/org/apache/flink/contrib/
tensorflow, TS = flink-tensorflow/src/main/scala/org/tensorflow/contrib/scala):

  TypedTensor, tagged_with / tagged_as      TS/package.scala:13-60   (phantom rank/element tags: unchecked casts)
  TensorValueBuilder                        FT/types/TensorValueBuilder.java:68-105
  tensor_value_to_tensor, TypedTensor.to_value   FT/package.scala:14-32   (bijection TensorValue <-> TypedTensor, .toValue)
  TensorValue2TensorInjections.invert       FT/types/TensorInjections.scala:86-153  (rank tag and dtype must match)
  *_array_to_tensor / tensor_to_*_array     FT/types/TensorInjections.scala:160-211, TS/Arrays.scala:17-60
                                            (inverse guarded by dtype and squeeze(shape).length <= 1, TFUtils.java:34-45)
  *_to_tensor / tensor_to_* scalars         FT/types/TensorInjections.scala:218-273 (inverse: dtype and rank 0)

In the reference a `Tensor` is a native TF-Java object and every conversion is a JNI copy.  Here a tensor is a host view
that ftf_run / ftf_submit read in place (include/ftf.h: ftf_tensor), so `TypedTensor` wraps a TensorValue without copying and
`to_value()` is the one copy the reference's `.toValue` makes ("does not take ownership", FT/package.scala:29).

An inverse whose guard fails raises InversionFailure (bijection's `attemptWhen` returns a Failure(InversionFailure)).
Two reference bugs are not reproduced: the LongBuffer branch of TensorValueBuilder.build sizes 4 bytes per element
(TensorValueBuilder.java:96) and overflows; messages2Tensor's fixed 10 000-byte scratch (TensorInjections.scala:54).
"""
import numpy as np

from .host import DataType, TensorValue


class InversionFailure(ValueError):
    """com.twitter.bijection.InversionFailure: the value is not in the image of the injection."""


RANK_TAGS = {"0D": 0, "1D": 1, "2D": 2, "3D": 3, "4D": 4, "5D": 5}      # TS/Rank.scala:19-44


def squeeze(shape):
    """TFUtils.squeeze (FT/util/TFUtils.java:34-45): the dimensions that are not 1."""
    return tuple(int(d) for d in shape if int(d) != 1)


class TypedTensor:
    """A tensor tagged with a rank K and an element type V.  The tags are phantom: `tagged_with` is an unchecked cast, as
    `taggedWith` / `taggedAs` are in TS/package.scala:51,59 (the reference's "2D" ExampleTensor is really [N])."""

    def __init__(self, value, rank=None, dtype=None):
        if not isinstance(value, TensorValue):
            raise TypeError("a TypedTensor wraps a TensorValue")
        self._value, self.rank_tag, self.dtype_tag = value, rank, dtype

    # the Tensor surface the reference uses (TensorFlow-Java 1.1: dataType, shape, numDimensions, numElements, numBytes)
    def data_type(self):
        return self._value.data_type

    def shape(self):
        return self._value.shape

    def num_dimensions(self):
        return len(self._value.shape)

    def num_elements(self):
        return int(np.prod(self._value.shape, dtype=np.int64)) if self._value.shape else 1

    def num_bytes(self):
        return int(self._value.buffer.nbytes)

    def tagged_with(self, rank, dtype):
        return TypedTensor(self._value, rank, dtype)

    tagged_as = tagged_with

    def as_value_view(self):
        """The wrapped record, uncopied: what is handed to Runner.run / ModelFunction."""
        return self._value

    def to_value(self):
        """RichTensor.toValue (FT/package.scala:24-32): a serializable copy; the tensor stays the caller's."""
        return TensorValue(self._value.data_type, self._value.shape, self._value.buffer.copy())

    def close(self):
        self._value = None


def tensor_value_to_tensor(value, rank=None, dtype=None):
    """FT/package.scala:14-19: TensorValue -> TypedTensor[K,V] (value.toTensor.taggedWith[K,V])."""
    return TypedTensor(value, rank, dtype)


def tensor_to_tensor_value(t, rank, dtype):
    """TensorValue2TensorInjections.invert (TensorInjections.scala:86-153): only when the tensor's rank is the tag K and its
    dtype is V's."""
    want_rank = RANK_TAGS.get(rank, rank)
    if t.num_dimensions() != want_rank or t.data_type() != dtype:
        raise InversionFailure(f"tensor of rank {t.num_dimensions()} and dtype {t.data_type()} is not a TensorValue[{rank},{dtype}]")
    return t.to_value()


# ---- TensorValueBuilder ---------------------------------------------------------------------------
class TensorValueBuilder:
    """TensorValueBuilder.java:68-105: dataType + shape + a typed buffer -> TensorValue over a native-order byte buffer."""
    _TYPED = (np.int32, np.float32, np.float64, np.int64)

    def __init__(self):
        self._dtype = self._shape = self._buffer = None

    def data_type(self, code):
        self._dtype = int(code)
        return self

    def shape(self, shape):
        self._shape = tuple(int(d) for d in shape)
        if len(self._shape) > 25:                                   # TFUtils.shapeOf: Tuple0..Tuple25 (TFUtils.java:75)
            raise NotImplementedError("rank > 25")
        return self

    def data(self, buffer):
        self._buffer = buffer
        return self

    def build(self):
        if self._dtype is None:
            raise RuntimeError("dataType is required")              # IllegalStateException
        if self._shape is None:
            raise RuntimeError("shape is required")
        if self._buffer is None:
            raise RuntimeError("data is required")
        buf = self._buffer
        if isinstance(buf, (bytes, bytearray, memoryview)):
            data = np.frombuffer(bytes(buf), np.uint8)              # a ByteBuffer in native order is taken as it is
        elif isinstance(buf, np.ndarray) and buf.dtype == np.uint8:
            data = np.ascontiguousarray(buf).reshape(-1)
        elif isinstance(buf, np.ndarray) and any(buf.dtype.newbyteorder("=") == np.dtype(t) for t in self._TYPED):
            if not buf.dtype.isnative:                              # a typed buffer of the other byte order is converted
                buf = buf.astype(buf.dtype.newbyteorder("="))
            data = np.ascontiguousarray(buf).reshape(-1).view(np.uint8)      # itemsize bytes per element, int64 included
        else:
            raise RuntimeError("unsupported data buffer")
        self._buffer = None
        return TensorValue(self._dtype, self._shape, data.copy())


# ---- arrays <-> 1-D tensors -----------------------------------------------------------------------
def _array_to_tensor(arr, np_type, code):
    a = np.ascontiguousarray(np.asarray(arr, dtype=np_type))
    if a.ndim != 1:
        raise ValueError("a 1-D array is expected")
    return TypedTensor(TensorValue(code, a.shape, a.view(np.uint8)), "1D", code)


def _tensor_to_array(t, np_type, code):
    if t.data_type() != code or len(squeeze(t.shape())) > 1:        # TensorInjections.scala:164
        raise InversionFailure(f"tensor of dtype {t.data_type()} and shape {t.shape()} is not an array of dtype {code}")
    return t.as_value_view().buffer.view(np_type).reshape(-1).copy()


def float_array_to_tensor(a):
    return _array_to_tensor(a, np.float32, DataType.FLOAT)


def double_array_to_tensor(a):
    return _array_to_tensor(a, np.float64, DataType.DOUBLE)


def long_array_to_tensor(a):
    return _array_to_tensor(a, np.int64, DataType.INT64)


def int_array_to_tensor(a):
    return _array_to_tensor(a, np.int32, DataType.INT32)


def tensor_to_float_array(t):
    return _tensor_to_array(t, np.float32, DataType.FLOAT)


def tensor_to_double_array(t):
    return _tensor_to_array(t, np.float64, DataType.DOUBLE)


def tensor_to_long_array(t):
    return _tensor_to_array(t, np.int64, DataType.INT64)


def tensor_to_int_array(t):
    return _tensor_to_array(t, np.int32, DataType.INT32)


# ---- scalars <-> 0-D tensors ----------------------------------------------------------------------
def _scalar_to_tensor(v, np_type, code):
    a = np.asarray(v, dtype=np_type).reshape(())
    return TypedTensor(TensorValue(code, (), a.reshape(1).view(np.uint8)), "0D", code)


def _tensor_to_scalar(t, np_type, code):
    if t.data_type() != code or t.num_dimensions() != 0:            # TensorInjections.scala:223,236,249,262,271
        raise InversionFailure(f"tensor of dtype {t.data_type()} and rank {t.num_dimensions()} is not a scalar of dtype {code}")
    return t.as_value_view().buffer.view(np_type)[0].item()


def float_to_tensor(v):
    return _scalar_to_tensor(v, np.float32, DataType.FLOAT)


def double_to_tensor(v):
    return _scalar_to_tensor(v, np.float64, DataType.DOUBLE)


def long_to_tensor(v):
    return _scalar_to_tensor(v, np.int64, DataType.INT64)


def int_to_tensor(v):
    return _scalar_to_tensor(v, np.int32, DataType.INT32)


def tensor_to_float(t):
    return _tensor_to_scalar(t, np.float32, DataType.FLOAT)


def tensor_to_double(t):
    return _tensor_to_scalar(t, np.float64, DataType.DOUBLE)


def tensor_to_long(t):
    return _tensor_to_scalar(t, np.int64, DataType.INT64)


def tensor_to_int(t):
    return _tensor_to_scalar(t, np.int32, DataType.INT32)


def string_to_tensor(s):
    """string2Tensor (TensorInjections.scala:218-226): the string's bytes as a 0-D STRING tensor."""
    from .host import bytes_to_tensor
    return TypedTensor(bytes_to_tensor(s.encode() if isinstance(s, str) else bytes(s)), "0D", DataType.STRING)


def tensor_to_string(t):
    if t.data_type() != DataType.STRING or t.num_dimensions() != 0:
        raise InversionFailure("not a 0-D STRING tensor")
    buf = t.as_value_view().buffer
    # TF-1.x scalar STRING layout: one u64 offset (0), then varint length | bytes (TensorInjections.scala:60-64)
    pos, n, shift = 8, 0, 0
    while True:
        c = int(buf[pos]); pos += 1
        n |= (c & 0x7f) << shift
        shift += 7
        if not c & 0x80:
            break
    return bytes(buf[pos:pos + n]).decode()
