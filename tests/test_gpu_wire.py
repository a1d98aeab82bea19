"""GPU parity of the wire-format kernels: p1_msg_pack / p1_msg_unpack
against the oracle's restatement of the proto2 encoding (pinned to google.protobuf bytes in
tests/golden/msg_wire.json) and p1_share_bits against src/input.c:41."""
import ctypes
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as po
from test_gpu_parity import mkctx, vp, dev_array, dev_zeros, to_host

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
c_vp = ctypes.c_void_p


def mixed_words(rng, shape, bits):
    """random words of every varint length"""
    v = rng.integers(0, 2 ** 64 - 1, shape, dtype=np.uint64, endpoint=True)
    sh = rng.integers(0, 64, shape).astype(np.uint64)
    v = v >> sh
    return (v & np.uint64((1 << bits) - 1)).astype(po.word_dtype(bits))


def gpu_pack(lib, ctx, vecs, values, n, ld):
    count = len(values)
    stride = int(lib.p1_msg_max_bytes(ctx, n))
    vd = dev_array(lib, ctx, vecs)
    out = dev_zeros(lib, ctx, count * stride)
    sizes = dev_zeros(lib, ctx, 8 * count)
    lib.check(lib.p1_msg_pack(ctx, vd, ld, n, count, vp(values), out, stride, sizes))
    sz = to_host(lib, ctx, sizes, count, np.uint64)
    raw = to_host(lib, ctx, out, (count, stride), np.uint8)
    for p in (vd, sizes):
        lib.check(lib.p1_dev_free(ctx, p))
    return raw, sz, out, stride


@pytest.mark.parametrize("bits", [64, 32])
@pytest.mark.parametrize("n", [0, 1, 7, 2048, 2049, 10000])
@pytest.mark.parametrize("kind", ["mixed", "full"])
def test_msg_pack_unpack_parity(p1lib, oracle, bits, n, kind):
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(100 + n + bits)
    dt, count = po.word_dtype(bits), 3
    ld = max(4, (n + 255) // 256 * 256)
    vecs = np.zeros((count, ld), dt)
    if kind == "mixed":
        vecs[:, :n] = mixed_words(rng, (count, n), bits)
    else:   # masked ring words: uniformly random, almost always the longest encoding
        vecs[:, :n] = rng.integers(0, 2 ** 64 - 1, (count, n), dtype=np.uint64, endpoint=True).astype(dt)
    values = rng.integers(0, 2 ** 64 - 1, count, dtype=np.uint64, endpoint=True)
    values[0] = 0
    raw, sz, out, stride = gpu_pack(p1lib, ctx, vecs, values, n, ld)
    for p in range(count):
        want = oracle.msg_pack(vecs[p, :n], int(values[p]) & ((1 << bits) - 1), bits)
        assert int(sz[p]) == len(want)
        assert np.array_equal(raw[p, :len(want)], want), "message %d differs" % p
    # and back, from the device bytes
    back = dev_zeros(p1lib, ctx, max(vecs.nbytes, 16))
    vals = dev_zeros(p1lib, ctx, 8 * count)
    status = dev_zeros(p1lib, ctx, 4 * count)
    p1lib.check(p1lib.p1_msg_unpack(ctx, out, stride, vp(sz), count, n, back, ld, vals, status))
    assert not to_host(p1lib, ctx, status, count, np.int32).any()
    got = to_host(p1lib, ctx, back, (count, ld), dt)
    assert np.array_equal(got[:, :n], vecs[:, :n])
    assert np.array_equal(to_host(p1lib, ctx, vals, count, np.uint64), values & np.uint64((1 << bits) - 1))
    p1lib.p1_ctx_destroy(ctx)


def test_msg_unpack_google_protobuf_golden(p1lib):
    """bytes serialised by the google.protobuf runtime decode to the golden vectors"""
    with open(os.path.join(ROOT, "tests", "golden", "msg_wire.json")) as f:
        cases = json.load(f)["cases"]
    for bits in (64, 32):
        ctx = mkctx(p1lib, bits, bits)
        for c in [c for c in cases if c["bits"] == bits]:
            vec = np.array([int(v) for v in c["vector"]], dtype=po.word_dtype(bits))
            wire = np.frombuffer(bytes.fromhex(c["wire_hex"]), dtype=np.uint8)
            stride = (len(wire) + 15) // 16 * 16
            buf = np.zeros(stride, np.uint8)
            buf[:len(wire)] = wire
            ind = dev_array(p1lib, ctx, buf)
            n, ld = len(vec), max(4, (len(vec) + 3) // 4 * 4)
            back = dev_zeros(p1lib, ctx, 8 * ld + 16)
            vals, status = dev_zeros(p1lib, ctx, 8), dev_zeros(p1lib, ctx, 4)
            sizes = np.array([len(wire)], np.uint64)
            p1lib.check(p1lib.p1_msg_unpack(ctx, ind, stride, vp(sizes), 1, n, back, ld, vals, status))
            assert int(to_host(p1lib, ctx, status, 1, np.int32)[0]) == 0
            assert np.array_equal(to_host(p1lib, ctx, back, ld, po.word_dtype(bits))[:n], vec)
            assert int(to_host(p1lib, ctx, vals, 1, np.uint64)[0]) == int(c["value"])
            # the same vector packed on the GPU gives the golden bytes
            raw, sz, _, _ = gpu_pack(p1lib, ctx, np.concatenate([vec, np.zeros(ld - n, vec.dtype)])[None, :],
                                     np.array([int(c["value"])], np.uint64), n, ld)
            assert int(sz[0]) == len(wire) and np.array_equal(raw[0, :len(wire)], wire)
        p1lib.p1_ctx_destroy(ctx)


def test_msg_unpack_flags_malformed(p1lib, oracle):
    from linreg_mpc_b200 import ffi
    ctx = mkctx(p1lib, 64, 64)
    rng = np.random.default_rng(5)
    n = 3000
    vec = mixed_words(rng, n, 64)
    good = oracle.msg_pack(vec, 77, 64)
    stride = (len(good) + 64 + 15) // 16 * 16

    def status_of(wire, n_expect=n):
        buf = np.zeros(stride, np.uint8)
        buf[:len(wire)] = wire
        ind = dev_array(p1lib, ctx, buf)
        back = dev_zeros(p1lib, ctx, 8 * (n + 8))
        vals, status = dev_zeros(p1lib, ctx, 8), dev_zeros(p1lib, ctx, 4)
        p1lib.check(p1lib.p1_msg_unpack(ctx, ind, stride, vp(np.array([len(wire)], np.uint64)), 1, n_expect, back, n + 8, vals, status))
        return int(to_host(p1lib, ctx, status, 1, np.int32)[0])

    assert status_of(good) == 0
    assert status_of(good, n - 1) & ffi.P1_MSG_BAD_COUNT
    assert status_of(good[:-1]) & ffi.P1_MSG_BAD_HEADER
    bad = good.copy(); bad[0] = 0x12
    assert status_of(bad) != 0
    bad = good.copy(); bad[1] += 1                     # payload one byte longer: swallows the trailer tag
    assert status_of(bad) != 0
    assert status_of(np.concatenate([good, [0]]).astype(np.uint8)) & ffi.P1_MSG_BAD_HEADER
    bad = good.copy(); bad[10:22] |= 0x80              # an element of more than 10 bytes
    assert status_of(bad) != 0
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("b1,b2", [(64, 64), (32, 32), (64, 32)])
def test_share_bits_parity(p1lib, oracle, b1, b2):
    ctx = mkctx(p1lib, b1, b2)
    rng = np.random.default_rng(3)
    n = 1237
    shares = rng.integers(0, 2 ** 64 - 1, n, dtype=np.uint64, endpoint=True).astype(po.word_dtype(b2))
    want = oracle.share_bits(shares, b2)
    sel = np.zeros(n * b2, np.uint8)
    p1lib.check(p1lib.p1_share_bits(ctx, vp(shares), n, vp(sel)))           # host in, host out
    assert np.array_equal(sel, want)
    sd = dev_array(p1lib, ctx, shares)
    od = dev_zeros(p1lib, ctx, n * b2)
    p1lib.check(p1lib.p1_share_bits(ctx, sd, n, od))                        # device in, device out
    assert np.array_equal(to_host(p1lib, ctx, od, n * b2, np.uint8), want)
    p1lib.p1_ctx_destroy(ctx)
