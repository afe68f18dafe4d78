"""Host-side mirror of the reference's operator API for the model-invocation path, over the C ABI.

Same names, argument meaning and error behaviour as the Scala/Java classes they mirror (paths relative to
/root/reference; FT = flink-tensorflow/src/main/scala/org/apache/flink/contrib/tensorflow, EX =
flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples):

  TensorValue            FT/types/TensorValue.java:48-264
  TensorName             FT/types/TensorName.scala:14-19
  GraphMethod & co       FT/graphs/GraphMethod.scala:8-36, FT/ml/signatures/RegressionMethod.scala:18-36,
                         EX/inception/ImageNormalization.scala:14-32, EX/inception/InceptionModel.scala:19-39
  ModelFunction          FT/models/ModelFunction.scala:18-78
  RichModel/TensorFlowModel/GenericModel   FT/models/Model.scala:24-44, FT/models/savedmodel/TensorFlowModel.scala,
                         FT/models/generic/GenericModel.scala:22-43
  map_with_model         FT/streaming/package.scala:23-34

Exceptions follow TF-Java's status mapping: INVALID_ARGUMENT -> ValueError (IllegalArgumentException),
FAILED_PRECONDITION/RESOURCE_EXHAUSTED -> RuntimeError (IllegalStateException), OUT_OF_RANGE -> IndexError,
UNIMPLEMENTED -> NotImplementedError (UnsupportedOperationException), NOT_FOUND -> KeyError.
"""
import contextlib
import ctypes as C
import threading

import numpy as np

from . import _abi
from ._abi import FtfConfig, FtfStats, FtfTensor


class VersionMismatchException(Exception):
    """TensorValue.java:152-154"""


_ERRORS = {
    _abi.INVALID_ARGUMENT: ValueError, _abi.NOT_FOUND: KeyError, _abi.RESOURCE_EXHAUSTED: RuntimeError,
    _abi.FAILED_PRECONDITION: RuntimeError, _abi.OUT_OF_RANGE: IndexError, _abi.UNIMPLEMENTED: NotImplementedError,
    _abi.INTERNAL: RuntimeError, _abi.VERSION_MISMATCH: VersionMismatchException,
}


def _check(rc, runner=None):
    if rc == _abi.OK:
        return
    lib = _abi.load_library()
    msg = lib.ftf_last_error(runner)
    raise _ERRORS.get(rc, RuntimeError)((msg or b"").decode("utf-8", "replace") + f" [status {rc}]")


# ---- dtype codes: FT/util/TFUtils.java:10-32 ------------------------------------------------------
class DataType:
    FLOAT, DOUBLE, INT32, UINT8, STRING, INT64, BOOL, BFLOAT16 = 1, 2, 3, 4, 7, 9, 10, 14
    _NP = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 9: np.int64, 10: np.bool_}

    @staticmethod
    def of(arr):
        for code, t in DataType._NP.items():
            if arr.dtype == np.dtype(t):
                return code
        raise ValueError("dataType")  # TFUtils.java:18


class TensorName:
    """TensorName.apply: "op:idx" -> (name, index); missing idx => 0; >1 colon => ValueError."""

    def __init__(self, s):
        lib = _abi.load_library()
        buf = C.create_string_buffer(len(s.encode()) + 1)
        idx = C.c_int32()
        _check(lib.ftf_tensor_name_parse(s.encode(), buf, len(buf), C.byref(idx)))
        self.name, self.index = buf.value.decode(), idx.value

    def __str__(self):
        return f"{self.name}:{self.index}"


def _as_ftf(dtype, shape, buf):
    """(FtfTensor, keepalive) over a numpy uint8/typed buffer."""
    dims = (C.c_int64 * max(1, len(shape)))(*shape)
    t = FtfTensor(dtype, len(shape), dims, buf.ctypes.data if buf.size else None, buf.nbytes)
    return t, (dims, buf)


class TensorValue:
    """Serializable tensor record: dtype code, shape tuple, immutable native-order payload."""

    def __init__(self, data_type, shape, buffer):
        self.data_type = int(data_type)
        self.shape = tuple(int(d) for d in shape)
        self.buffer = np.ascontiguousarray(np.frombuffer(bytes(buffer), dtype=np.uint8)) \
            if not isinstance(buffer, np.ndarray) else np.ascontiguousarray(buffer).view(np.uint8).reshape(-1)

    @staticmethod
    def from_array(arr):
        arr = np.asarray(arr)
        shape = arr.shape                                  # ascontiguousarray promotes rank 0 to rank 1
        flat = np.ascontiguousarray(arr).reshape(-1)
        return TensorValue(DataType.of(arr), shape, flat.view(np.uint8))

    def num_dimensions(self):
        return len(self.shape)

    def to_array(self):
        return self.buffer.view(DataType._NP[self.data_type]).reshape(self.shape)

    def get_binary_length(self):  # TensorValue.java:142-147
        return _abi.load_library().ftf_tv_binary_length(len(self.shape), self.buffer.nbytes)

    def write(self):  # TensorValue.java:168-187
        lib = _abi.load_library()
        t, keep = _as_ftf(self.data_type, self.shape, self.buffer)
        out = np.empty(self.get_binary_length(), np.uint8)
        n = C.c_int64()
        _check(lib.ftf_tv_encode(C.byref(t), out.ctypes.data, out.nbytes, C.byref(n)))
        return out[: n.value].tobytes()

    @staticmethod
    def read(wire):  # TensorValue.java:150-165
        lib = _abi.load_library()
        raw = np.frombuffer(wire, dtype=np.uint8)
        t = FtfTensor()
        dims = (C.c_int64 * 25)()
        _check(lib.ftf_tv_decode(raw.ctypes.data if raw.size else None, raw.nbytes, C.byref(t), dims))
        off = (t.data or 0) - raw.ctypes.data if t.nbytes else 0
        return TensorValue(t.dtype, [dims[i] for i in range(t.rank)], raw[off: off + t.nbytes].copy())


def messages_to_tensor(messages):
    """List[bytes] -> STRING tensor [N] TensorValue (TensorInjections.scala:49-72)."""
    lib = _abi.load_library()
    n = len(messages)
    arr = (C.c_char_p * max(1, n))(*messages)
    lens = (C.c_int64 * max(1, n))(*[len(m) for m in messages])
    need = lib.ftf_string_tensor_encode(n, arr, lens, None, 0)
    out = np.empty(max(need, 1), np.uint8)
    got = lib.ftf_string_tensor_encode(n, arr, lens, out.ctypes.data, out.nbytes)
    if got < 0:
        _check(-got)
    return TensorValue(DataType.STRING, (n,), out[:got])


def bytes_to_tensor(data):
    """One byte string -> 0-D STRING tensor (TS/ByteStrings.scala:22-25: Tensor.create(bytes)); this is how the reference
    wraps an image file into its ImageFileTensor (EX/inception/ImageInputFormat.scala:63-80)."""
    tv = messages_to_tensor([bytes(data)])
    return TensorValue(DataType.STRING, (), tv.buffer)


def example(**features):
    """tf.Example with float_list features (FTT/util/TestData.scala:22-34), serialized by hand."""
    def ld(field, payload):
        out, n = bytearray([field << 3 | 2]), len(payload)
        while n >= 0x80:
            out.append(n & 0x7f | 0x80)
            n >>= 7
        out.append(n)
        return bytes(out) + payload
    entries = b""
    for key, values in features.items():
        vals = np.atleast_1d(np.asarray(values, np.float32))
        feature = ld(2, ld(1, vals.tobytes()))            # Feature.float_list{ value (packed) }
        entries += ld(1, ld(1, key.encode()) + ld(2, feature))
    return ld(1, entries)


# ---- graph methods ------------------------------------------------------------------------------
class GraphMethod:
    name = None

    def inputs(self, i):
        raise NotImplementedError

    def outputs(self, o):
        raise NotImplementedError


class RegressionMethod(GraphMethod):
    name = "tensorflow/serving/regress"      # SignatureConstants.java:52

    def inputs(self, i):
        return {"inputs": i}                 # REGRESS_INPUTS

    def outputs(self, o):
        return o["outputs"]                  # REGRESS_OUTPUTS


class PredictMethod(GraphMethod):
    """tensorflow/serving/predict with one input and one output (serving_default of half_plus_two)."""
    name = "tensorflow/serving/predict"

    def __init__(self, in_key="x", out_key="y"):
        self.in_key, self.out_key = in_key, out_key

    def inputs(self, i):
        return {self.in_key: i}

    def outputs(self, o):
        return o[self.out_key]


class ImageNormalizationMethod(GraphMethod):
    name = "inception/normalize"

    def inputs(self, i):
        return {"inputs": i}

    def outputs(self, o):
        return o["outputs"]


class LabelMethod(GraphMethod):
    name = "inception/label"

    def inputs(self, i):
        return {"inputs": i}

    def outputs(self, o):
        return o["outputs"]


# ---- runner handle ------------------------------------------------------------------------------
def default_config(**kw):
    cfg = FtfConfig()
    _abi.load_library().ftf_default_config(C.byref(cfg))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError(f"unknown config field {k}")
        setattr(cfg, k, v)
    return cfg


class Runner:
    """One ftf_runner (== one TF Session bound to one signature) on one GPU."""

    def __init__(self, model, signature_key=None, method_name=None, config=None):
        self._lib = _abi.load_library()
        self._h = C.c_void_p()
        cfg = config if config is not None else default_config()
        _check(self._lib.ftf_create(model.encode(), signature_key.encode() if signature_key else None,
                                    method_name.encode() if method_name else None, C.byref(cfg), C.byref(self._h)))
        self.config = cfg

    @property
    def handle(self):
        return self._h

    def signature_info(self):
        g = lambda k, i: (self._lib.ftf_signature_info(self._h, k, i) or b"").decode()
        ins, outs, i = {}, {}, 0
        while self._lib.ftf_signature_info(self._h, 1, i) is not None:
            ins[g(1, i)] = g(2, i)
            i += 1
        i = 0
        while self._lib.ftf_signature_info(self._h, 3, i) is not None:
            outs[g(3, i)] = g(4, i)
            i += 1
        return {"method_name": g(0, 0), "inputs": ins, "outputs": outs}

    def run(self, inputs, fetch):
        """inputs: {key: TensorValue}; fetch: [key] -> {key: TensorValue (copied out of runner memory)}."""
        if self._h is None or not self._h.value:
            raise ValueError("a session must be provided")
        n_in = len(inputs)
        keys = (C.c_char_p * max(1, n_in))(*[k.encode() for k in inputs])
        tens = (FtfTensor * max(1, n_in))()
        keep = []
        for j, tv in enumerate(inputs.values()):
            t, k = _as_ftf(tv.data_type, tv.shape, tv.buffer)
            tens[j] = t
            keep.append(k)
        okeys = (C.c_char_p * max(1, len(fetch)))(*[k.encode() for k in fetch])
        outs = (FtfTensor * max(1, len(fetch)))()
        _check(self._lib.ftf_run(self._h, n_in, keys, tens, len(fetch), okeys, outs), self._h)
        result = {}
        for j, k in enumerate(fetch):
            o = outs[j]
            shape = [o.dims[i] for i in range(o.rank)]
            data = np.ctypeslib.as_array(C.cast(o.data, C.POINTER(C.c_uint8)), shape=(o.nbytes,)).copy() \
                if o.nbytes else np.empty(0, np.uint8)
            result[k] = TensorValue(o.dtype, shape, data)
        self._lib.ftf_release(self._h, outs, len(fetch))     # ManagedResource release, ModelFunction.scala:68-69
        return result

    # micro-batched per-record path
    def poll_select(self, keys):
        arr = (C.c_char_p * len(keys))(*[k.encode() for k in keys])
        _check(self._lib.ftf_poll_select(self._h, len(keys), arr), self._h)
        self._nsel = len(keys)

    def submit(self, tag, tv):
        t, keep = _as_ftf(tv.data_type, tv.shape, tv.buffer)
        _check(self._lib.ftf_submit(self._h, tag, C.byref(t)), self._h)

    def poll(self, block=False, cap=256):
        """[(tag, TensorValue or [TensorValue])] of the records that are ready, in submission order.  A record that failed
        on its own (a JPEG file found corrupt on the device) comes back as (tag, exception): it fails alone, like one
        map() call of the reference, and the records polled with it keep their results."""
        nsel = getattr(self, "_nsel", 1)
        tags = (C.c_uint64 * cap)()
        outs = (FtfTensor * (cap * nsel))()
        n = C.c_int32()
        _check(self._lib.ftf_poll(self._h, 1 if block else 0, tags, outs, cap, C.byref(n)), self._h)
        status = (C.c_int32 * max(1, n.value))()
        msg = C.c_char_p()
        any_bad = self._lib.ftf_poll_status(self._h, status, n.value, C.byref(msg)) != _abi.OK
        res = []
        for i in range(n.value):
            if any_bad and status[i] != _abi.OK:
                res.append((tags[i], _ERRORS.get(status[i], RuntimeError)(
                    f"record {tags[i]}: DecodeJpeg: invalid JPEG data (first bad record of this poll: "
                    f"{(msg.value or b'').decode('utf-8', 'replace')}) [status {status[i]}]")))
                continue
            vals = []
            for q in range(nsel):
                o = outs[i * nsel + q]
                data = np.ctypeslib.as_array(C.cast(o.data, C.POINTER(C.c_uint8)), shape=(o.nbytes,)).copy()
                vals.append(TensorValue(o.dtype, [o.dims[d] for d in range(o.rank)], data))
            res.append((tags[i], vals[0] if nsel == 1 else vals))
        return res

    def flush(self):
        _check(self._lib.ftf_flush(self._h), self._h)

    def export_weights(self, directory):
        _check(self._lib.ftf_label_export_weights(self._h, str(directory).encode()), self._h)

    def stats(self):
        s = FtfStats()
        _check(self._lib.ftf_stats_get(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in FtfStats._fields_}

    def close(self):
        if self._h is not None and self._h.value:
            self._lib.ftf_destroy(self._h)
        self._h = C.c_void_p()


# ---- ModelFunction ------------------------------------------------------------------------------
class ModelFunction:
    """ModelFunction(session, signatureDef)(method): apply(in) -> context manager over the output."""

    def __init__(self, session, method):
        if session is None:
            raise ValueError("requirement failed: a session must be provided")
        if method is None:
            raise ValueError("requirement failed: a signatureDef must be provided")
        info = session.signature_info()
        if method.name != info["method_name"]:                       # ModelFunction.scala:37
            raise ValueError("requirement failed")
        self._session, self._method, self._info = session, method, info

    @contextlib.contextmanager
    def __call__(self, value):
        if value is None:
            raise ValueError("requirement failed: Input must be provided")
        inputs = self._method.inputs(value)
        for key in self._info["inputs"]:                              # ModelFunction.scala:51-53
            if key not in inputs:
                raise ValueError(f"An input tensor named {key} must be provided for this computation.")
        fed = {k: inputs[k] for k in self._info["inputs"]}
        outs = self._session.run(fed, list(self._info["outputs"]))
        yield self._method.outputs(outs)

    apply = __call__


# ---- models -------------------------------------------------------------------------------------
class RichModel:
    def open(self):
        raise NotImplementedError

    def close(self):
        raise NotImplementedError


class TensorFlowModel(RichModel):
    """SavedModel-backed model: TensorFlowModel.load(path, tags) + signatureDef(name)."""

    def __init__(self, model_path, *tags, device=0):
        self.model_path, self.tags, self.device = str(model_path), tags or ("serve",), device
        self._sessions = None

    def open(self):
        if self._sessions is not None:                                # TensorFlowModel.scala:40
            raise RuntimeError("checkState failed: model is already open")
        self._sessions = {}

    def close(self):                                                  # null-safe, idempotent
        if self._sessions:
            for s in self._sessions.values():
                s.close()
        self._sessions = None

    def session(self, signature_key, method_name=None):
        if self._sessions is None:
            raise ValueError("requirement failed: a session must be provided")
        if signature_key not in self._sessions:
            self._sessions[signature_key] = Runner(self.model_path, signature_key, method_name,
                                                   default_config(device=self.device))
        return self._sessions[signature_key]

    def model_function(self, signature_key, method):
        return ModelFunction(self.session(signature_key), method)


class HalfPlusTwo(TensorFlowModel):
    """FTT/ml/HalfPlusTwo.scala:15-28"""

    @property
    def regress_x_to_y(self):
        return self.model_function("regress_x_to_y", RegressionMethod())

    @property
    def regress_x_to_y2(self):
        return self.model_function("regress_x_to_y2", RegressionMethod())

    @property
    def serving_default(self):
        return self.model_function("serving_default", PredictMethod("x", "y"))


class GenericModel(RichModel):
    """Builtin-graph model: one runner, opened in open(), closed in close() (GenericModel.scala:22-43)."""
    kind = None
    method = None

    def __init__(self, **config):
        self._config = config
        self._session = None

    def open(self):
        self._session = Runner(self.kind, None, None, default_config(**self._config))

    def close(self):
        if self._session is not None:
            self._session.close()
            self._session = None

    @property
    def session(self):
        return self._session


class ImageNormalization(GenericModel):
    """EX/inception/ImageNormalization.scala:40-85 on raw RGB; constants default to the reference's."""
    kind = "builtin:image_normalization"

    @property
    def normalize(self):
        return ModelFunction(self._session, ImageNormalizationMethod())


class LabeledImage:
    def __init__(self, labels):
        self.labels = labels  # [(score, label)] sorted by score descending

    def __repr__(self):
        return f"LabeledImage({self.labels})"


class InceptionModel(GenericModel):
    """EX/inception/InceptionModel.scala:47-90; `labels` is the label file's lines."""
    kind = "builtin:inception_v3"

    def __init__(self, labels=None, model_dir=None, **config):
        """labels: the label file's lines.  model_dir: a directory with trained weights (variables/*, see include/ftf.h)
        and, optionally, labels.txt -- read like imagenet_comp_graph_label_strings.txt (InceptionModel.scala:54-55,
        GraphUtils.readAllLines) when `labels` is not given.  Without model_dir: seeded random-init weights."""
        super().__init__(**config)
        if model_dir is not None:
            import os
            self.kind = "inception_v3:" + str(model_dir)
            lf = os.path.join(str(model_dir), "labels.txt")
            if labels is None and os.path.exists(lf):
                with open(lf, encoding="utf-8") as f:
                    labels = [line.rstrip("\n") for line in f]
        self.labels = labels

    @property
    def label(self):
        return ModelFunction(self._session, LabelMethod())

    def to_text_labels(self, t, take=3):
        """RichLabelTensor.toTextLabels (InceptionModel.scala:81-89) through K7 on the device."""
        if t.num_dimensions() != 2:
            raise ValueError("requirement failed: expected a [M N] shaped tensor")
        from . import device
        scores = t.to_array()
        labels = self.labels if self.labels is not None else [str(i) for i in range(scores.shape[1])]
        idx, sc = device.top_labels(scores, len(labels), take)
        return [LabeledImage([(float(s), labels[i]) for i, s in zip(ri, rs) if i >= 0]) for ri, rs in zip(idx, sc)]


# ---- operator glue ------------------------------------------------------------------------------
def map_with_model(stream, model, fun):
    """DataStream.mapWithModel (FT/streaming/package.scala:23-34): open -> map* -> close on one 'task thread'."""
    if fun is None:
        raise TypeError("Map function must not be null.")
    model.open()
    try:
        for rec in stream:
            yield fun(rec, model)
    finally:
        model.close()


class MiniCluster:
    """N subtasks, one thread and one model instance each, subtask i -> GPU i % n_gpus: the stand-in for
    LocalFlinkMiniCluster (FTT/util/FlinkTestBase.scala:45-57).  Records are dealt round-robin (rebalance)."""

    def __init__(self, parallelism, n_gpus=1):
        self.parallelism, self.n_gpus = parallelism, max(1, n_gpus)

    def run(self, records, model_factory, fun):
        parts = [records[i::self.parallelism] for i in range(self.parallelism)]
        results = [None] * self.parallelism
        errors = []

        def task(i):
            try:
                model = model_factory(i % self.n_gpus)
                results[i] = list(map_with_model(parts[i], model, fun))
            except Exception as e:  # task failure -> job failure
                errors.append(e)

        threads = [threading.Thread(target=task, args=(i,)) for i in range(self.parallelism)]
        [t.start() for t in threads]
        [t.join() for t in threads]
        if errors:
            raise errors[0]
        out = [None] * len(records)
        for i, part in enumerate(results):
            out[i::self.parallelism] = part
        return out
