"""Generates tests/golden/msg_wire.json: SecureMultiplication__Msg wire bytes produced by the
google.protobuf runtime (a third-party implementation of the format protobuf-c emits for
the reference, src/protobuf/secure_multiplication{,_32}.proto) for fixed vectors.  Run in the
build container:  python tests/golden/make_wire_golden.py
"""
import json
import os

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

HERE = os.path.dirname(os.path.abspath(__file__))


def msg_class(bits):
    """message msg { repeated uintW vector = 1 [packed=true]; required uintW value = 2; }"""
    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "secure_multiplication_%d.proto" % bits
    fd.package = "secure_multiplication%d" % bits
    fd.syntax = "proto2"
    m = fd.message_type.add()
    m.name = "msg"
    ty = F.TYPE_UINT64 if bits == 64 else F.TYPE_UINT32
    f1 = m.field.add()
    f1.name, f1.number, f1.label, f1.type = "vector", 1, F.LABEL_REPEATED, ty
    f1.options.packed = True
    f2 = m.field.add()
    f2.name, f2.number, f2.label, f2.type = "value", 2, F.LABEL_REQUIRED, ty
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName(fd.package + ".msg"))


def cases(bits):
    rng = np.random.default_rng(20170401 + bits)
    top = (1 << bits) - 1
    edge = [0, 1, 127, 128, 16383, 16384, (1 << 21) - 1, 1 << 21, (1 << 28) - 1, 1 << 28, top, top - 1, 1 << (bits - 1)]
    if bits == 64:
        edge += [(1 << 35) - 1, 1 << 35, (1 << 56) - 1, 1 << 56, (1 << 63) - 1]
    out = [([], 0), ([], top), ([0], 5), (edge, 300), (edge[::-1], 1 << (bits - 1))]
    for n in (1, 2, 17, 130, 1000):
        out.append(([int(v) for v in rng.integers(0, top, n, dtype=np.uint64, endpoint=True)], int(rng.integers(0, top, dtype=np.uint64))))
    # mixed magnitudes: every varint length occurs
    sh = rng.integers(0, bits, 257)
    out.append(([int(v) >> int(s) for v, s in zip(rng.integers(0, top, 257, dtype=np.uint64, endpoint=True), sh)], 77))
    return out


def main():
    doc = {"comment": "google.protobuf %s SerializeToString of msg{vector, value}" % __import__("google.protobuf").protobuf.__version__,
           "cases": []}
    for bits in (64, 32):
        M = msg_class(bits)
        for vec, value in cases(bits):
            m = M()
            m.vector.extend(vec)
            m.value = value
            wire = m.SerializeToString()
            back = M.FromString(wire)
            assert list(back.vector) == vec and back.value == value
            doc["cases"].append({"bits": bits, "vector": [str(v) for v in vec], "value": str(value), "wire_hex": wire.hex()})
    with open(os.path.join(HERE, "msg_wire.json"), "w") as f:
        json.dump(doc, f)
    print("wrote", len(doc["cases"]), "cases")


if __name__ == "__main__":
    main()
