#!/usr/bin/env python
"""Writes models/half_plus_two/ in the SavedModel directory layout the reference loads
(/root/reference/models/half_plus_two: saved_model.pb + variables/variables.{index,data-00000-of-00001} +
assets/), with a hand-rolled protobuf / TensorBundle encoder -- no TensorFlow, no copy of the reference's files.

Graph (same tensor names and signatures as the reference fixture, decoded in SURVEY.md 8a-H):
  tf_example -> ParseExample(dense "x" required, "x2" default [0.0]) -> x, x2
  y = a*x + b,  y2 = a*x + c,  y3 = a*x2 + c      with a=0.5, b=2.0, c=3.0
"""
import os
import struct
import sys


def varint(n):
    n &= (1 << 64) - 1
    out = bytearray()
    while n >= 0x80:
        out.append(n & 0x7f | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out)


def fv(field, n):            # varint field
    return varint(field << 3) + varint(n)


def fb(field, payload):      # length-delimited field
    if isinstance(payload, str):
        payload = payload.encode()
    return varint(field << 3 | 2) + varint(len(payload)) + payload


def f32(field, x):
    return varint(field << 3 | 5) + struct.pack("<f", x)


DT_FLOAT, DT_INT32, DT_STRING, DT_INT64 = 1, 3, 7, 9


def shape(dims):
    return b"".join(fb(2, fv(1, d)) for d in dims)


def attr(key, value):
    return fb(5, fb(1, key) + fb(2, value))


def tensor_f32(dims, vals):
    return fv(1, DT_FLOAT) + fb(2, shape(dims)) + b"".join(f32(5, v) for v in vals)


def tensor_str(dims, vals):
    return fv(1, DT_STRING) + fb(2, shape(dims)) + b"".join(fb(8, v) for v in vals)


def tensor_i32(dims, vals):
    return fv(1, DT_INT32) + fb(2, shape(dims)) + b"".join(fv(7, v) for v in vals)


def node(name, op, inputs=(), attrs=()):
    return fb(1, fb(1, name) + fb(2, op) + b"".join(fb(3, i) for i in inputs) + b"".join(attrs))


def const(name, dtype, tensor):
    return node(name, "Const", (), [attr("dtype", fv(6, dtype)), attr("value", fb(8, tensor))])


def variable(name, init):
    return [
        const(name + "/initial_value", DT_FLOAT, tensor_f32([], [init])),
        node(name, "VariableV2", (), [attr("dtype", fv(6, DT_FLOAT)), attr("shape", fb(7, b"")),
                                      attr("container", fb(2, "")), attr("shared_name", fb(2, ""))]),
        node(name + "/read", "Identity", [name], [attr("T", fv(6, DT_FLOAT))]),
    ]


def graph_def():
    T = attr("T", fv(6, DT_FLOAT))
    n = []
    for name, v in (("a", 0.5), ("b", 2.0), ("c", 3.0)):
        n += variable(name, v)
    n.append(node("tf_example", "Placeholder", (), [attr("dtype", fv(6, DT_STRING)), attr("shape", fb(7, fv(3, 1)))]))
    n.append(const("ParseExample/Const", DT_FLOAT, tensor_f32([0], [])))
    n.append(const("ParseExample/key_x2", DT_FLOAT, tensor_f32([1], [0.0])))
    n.append(const("ParseExample/Reshape/shape", DT_INT32, tensor_i32([1], [1])))
    n.append(node("ParseExample/Reshape", "Reshape", ["ParseExample/key_x2", "ParseExample/Reshape/shape"],
                  [T, attr("Tshape", fv(6, DT_INT32))]))
    n.append(const("ParseExample/ParseExample/names", DT_STRING, tensor_str([0], [])))
    n.append(const("ParseExample/ParseExample/dense_keys_0", DT_STRING, tensor_str([], [b"x"])))
    n.append(const("ParseExample/ParseExample/dense_keys_1", DT_STRING, tensor_str([], [b"x2"])))
    n.append(node("ParseExample/ParseExample", "ParseExample",
                  ["tf_example", "ParseExample/ParseExample/names", "ParseExample/ParseExample/dense_keys_0",
                   "ParseExample/ParseExample/dense_keys_1", "ParseExample/Const", "ParseExample/Reshape"],
                  [attr("Nsparse", fv(3, 0)), attr("Ndense", fv(3, 2)),
                   attr("Tdense", fb(1, fv(6, DT_FLOAT) + fv(6, DT_FLOAT))),
                   attr("dense_shapes", fb(1, fb(7, shape([1])) + fb(7, shape([1]))))]))
    n.append(node("x", "Identity", ["ParseExample/ParseExample"], [T]))
    n.append(node("Mul", "Mul", ["a/read", "x"], [T]))
    n.append(node("y", "Add", ["Mul", "b/read"], [T]))
    n.append(node("Mul_1", "Mul", ["a/read", "x"], [T]))
    n.append(node("y2", "Add", ["Mul_1", "c/read"], [T]))
    n.append(node("x2", "Identity", ["ParseExample/ParseExample:1"], [T]))
    n.append(node("Mul_2", "Mul", ["a/read", "x2"], [T]))
    n.append(node("y3", "Add", ["Mul_2", "c/read"], [T]))
    return b"".join(n) + fb(4, fv(1, 21))      # versions.producer = 21


def tensor_info(name, dtype, dims):
    return fb(1, name) + fv(2, dtype) + fb(3, shape(dims))


def signature(method, inputs, outputs):
    body = b""
    for key, ti in inputs:
        body += fb(1, fb(1, key) + fb(2, ti))
    for key, ti in outputs:
        body += fb(2, fb(1, key) + fb(2, ti))
    return body + fb(3, method)


def meta_graph():
    ex = tensor_info("tf_example:0", DT_STRING, [])
    f = lambda name: tensor_info(name, DT_FLOAT, [-1, 1])
    REG, CLS, PRED = "tensorflow/serving/regress", "tensorflow/serving/classify", "tensorflow/serving/predict"
    sigs = {
        "regress_x_to_y": signature(REG, [("inputs", ex)], [("outputs", f("y:0"))]),
        "regress_x_to_y2": signature(REG, [("inputs", ex)], [("outputs", f("y2:0"))]),
        "regress_x2_to_y3": signature(REG, [("inputs", f("x2:0"))], [("outputs", f("y3:0"))]),
        "classify_x_to_y": signature(CLS, [("inputs", ex)], [("scores", f("y:0"))]),
        "serving_default": signature(PRED, [("x", f("x:0"))], [("y", f("y:0"))]),
    }
    meta_info = fb(4, "serve") + fb(5, "flink-tensorflow_b200 fixture")
    body = fb(1, meta_info) + fb(2, graph_def())
    for key, sig in sigs.items():
        body += fb(5, fb(1, key) + fb(2, sig))
    return body


# ---- TensorBundle (LevelDB-format table) --------------------------------------------------------
def crc32c(data):
    tbl = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        tbl.append(c)
    c = 0xFFFFFFFF
    for b in data:
        c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def mask(c):
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def block(entries):
    out, prev = bytearray(), b""
    for key, val in entries:
        shared = 0
        while shared < min(len(prev), len(key)) and prev[shared] == key[shared]:
            shared += 1
        if not out:
            shared = 0
        out += varint(shared) + varint(len(key) - shared) + varint(len(val)) + key[shared:] + val
        prev = key
    out += struct.pack("<II", 0, 1)          # one restart point at 0
    return bytes(out)


def with_trailer(blk):
    return blk + b"\x00" + struct.pack("<I", mask(crc32c(blk + b"\x00")))


def bundle(variables):
    data = b"".join(struct.pack("<f", v) for _, v in variables)
    entries = [(b"", fv(1, 1) + fb(3, fv(1, 1)))]                       # BundleHeaderProto
    off = 0
    for name, v in variables:
        payload = struct.pack("<f", v)
        e = fv(1, DT_FLOAT) + fb(2, b"")
        if off:
            e += fv(4, off)
        e += fv(5, 4) + varint(6 << 3 | 5) + struct.pack("<I", mask(crc32c(payload)))
        entries.append((name.encode(), e))
        off += 4
    data_block = block(entries)
    table = with_trailer(data_block)
    meta_off = len(table)
    meta_block = block([])
    table += with_trailer(meta_block)
    index_off = len(table)
    last_key = entries[-1][0]
    sep = last_key[:-1] + bytes([last_key[-1] + 1])                       # a key >= every key in the block
    index_block = block([(sep, varint(0) + varint(len(data_block)))])
    table += with_trailer(index_block)
    footer = varint(meta_off) + varint(len(meta_block)) + varint(index_off) + varint(len(index_block))
    footer += b"\x00" * (40 - len(footer)) + bytes.fromhex("57fb808b247547db")
    return table + footer, data


def main(out_dir):
    os.makedirs(os.path.join(out_dir, "variables"), exist_ok=True)
    os.makedirs(os.path.join(out_dir, "assets"), exist_ok=True)
    with open(os.path.join(out_dir, "saved_model.pb"), "wb") as f:
        f.write(fv(1, 1) + fb(2, meta_graph()))
    index, data = bundle([("a", 0.5), ("b", 2.0), ("c", 3.0)])
    with open(os.path.join(out_dir, "variables", "variables.index"), "wb") as f:
        f.write(index)
    with open(os.path.join(out_dir, "variables", "variables.data-00000-of-00001"), "wb") as f:
        f.write(data)
    with open(os.path.join(out_dir, "assets", "foo.txt"), "w") as f:
        f.write("asset-file-contents")


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "..", "models", "half_plus_two"))
