#!/usr/bin/env python
"""Regenerates tests/golden/*.json from the reference's own fixtures and tests.  Runs only where
/root/reference exists (the build container); the JSON it writes is committed and travels to the GPU box.

Sources (relative to /root/reference):
  models/half_plus_two/{saved_model.pb,variables/*}   decoded with tensorboard's TF proto stubs (independent of
                                                      the product's own reader in csrc/wire.cpp)
  flink-tensorflow/src/test/scala/.../ml/RegressITCase.scala:35          x -> y known answers
  flink-tensorflow/src/test/scala/.../types/TensorValueTest.java:23-55   int32 layouts
  flink-tensorflow/src/test/scala/.../util/TestData.scala:15-34          Example{x: 0..3}
"""
import json
import os
import struct

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def half_plus_two():
    from tensorboard.compat.proto import meta_graph_pb2
    data = open(os.path.join(REF, "models/half_plus_two/saved_model.pb"), "rb").read()

    def varint(b, i):
        r = s = 0
        while True:
            c = b[i]; i += 1; r |= (c & 0x7f) << s; s += 7
            if c < 0x80:
                return r, i
    i, mg = 0, None
    while i < len(data):
        tag, i = varint(data, i)
        if tag & 7 == 0:
            _, i = varint(data, i)
        else:
            ln, i = varint(data, i)
            if tag >> 3 == 2 and mg is None:
                mg = meta_graph_pb2.MetaGraphDef()
                mg.ParseFromString(data[i:i + ln])
            i += ln
    sigs = {}
    for key, s in mg.signature_def.items():
        info = lambda m: {k: {"name": v.name, "dtype": v.dtype, "shape": [d.size for d in v.tensor_shape.dim]}
                          for k, v in m.items()}
        sigs[key] = {"method_name": s.method_name, "inputs": info(s.inputs), "outputs": info(s.outputs)}
    raw = open(os.path.join(REF, "models/half_plus_two/variables/variables.data-00000-of-00001"), "rb").read()
    values = struct.unpack("<3f", raw)
    return {
        "tags": list(mg.meta_info_def.tags),
        "signatures": sigs,
        "variables": {"a": {"offset": 0, "value": values[0]}, "b": {"offset": 4, "value": values[1]},
                      "c": {"offset": 8, "value": values[2]}},
        # RegressITCase.scala:35
        "known_answers": {"regress_x_to_y": [[0.0, 2.0], [1.0, 2.5], [2.0, 3.0], [3.0, 3.5]]},
    }


def codecs():
    def example_x(x):  # Example{features{feature{"x": float_list[x]}}}, canonical protobuf encoding
        fl = b"\x0a\x04" + struct.pack("<f", x)
        feat = b"\x12" + bytes([len(fl)]) + fl
        entry = b"\x0a\x01x" + b"\x12" + bytes([len(feat)]) + feat
        feats = b"\x0a" + bytes([len(entry)]) + entry
        return b"\x0a" + bytes([len(feats)]) + feats
    exs = [example_x(float(v)) for v in range(4)]                      # TestData.scala:15-17
    table = b"".join(struct.pack("<Q", sum(1 + len(e) for e in exs[:i])) for i in range(4))
    batch = table + b"".join(bytes([len(e)]) + e for e in exs)
    one = struct.pack("<Q", 0) + bytes([len(exs[1])]) + exs[1]
    payload = struct.pack("<6i", 1, 2, 3, 4, 5, 6)                     # TensorValueTest.java:40
    wire = b"\x01" + struct.pack(">ii", 3, 1) + struct.pack(">q", 6) + struct.pack(">i", 24) + payload
    wire23 = b"\x01" + struct.pack(">ii", 3, 2) + struct.pack(">qq", 2, 3) + struct.pack(">i", 24) + payload
    return {
        "examples_x_0_to_3": [e.hex() for e in exs],
        "string_tensor_of_example_x1": one.hex(),
        "string_tensor_of_4_examples": batch.hex(),
        "tensorvalue_int32_6": {"dtype": 3, "shape": [6], "payload": payload.hex(), "wire": wire.hex()},
        "tensorvalue_int32_2x3": {"dtype": 3, "shape": [2, 3], "payload": payload.hex(), "wire": wire23.hex()},
    }


if __name__ == "__main__":
    with open(os.path.join(HERE, "half_plus_two.json"), "w") as f:
        json.dump(half_plus_two(), f, indent=1, sort_keys=True)
    with open(os.path.join(HERE, "codecs.json"), "w") as f:
        json.dump(codecs(), f, indent=1, sort_keys=True)
    print("wrote golden fixtures")
