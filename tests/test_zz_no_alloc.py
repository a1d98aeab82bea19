"""p1_msg_pack_static: the wire-format pack for hosts that must not allocate on the GPU.

Who needs it: a provider process whose own arena already owns all of the GPU's memory (there is
nothing left for the driver to hand out when a message is due), the per-peer threads of --use_ot
mode (cudaMalloc / cudaFree wait for the whole device, so one thread's allocation holds up every
other peer's session on that GPU), and a party run under a memory cap that could refuse a request
in the middle of the protocol.  The entry point keeps its tile tables in a fixed 768 KB array that
is part of libp1gpu.so's device image, so it asks the driver for nothing.

What is checked, through the C-ABI only: the messages and their sizes are byte-for-byte those of
p1_msg_pack (the tolerance of every integer test here: none), on two calls in a row that reuse the
shared array with different shapes - the second call is smaller and of the other varint mix, so the
tables still hold the first call's entries when it starts; a call beyond the stated maximum and a
host `values` array are refused with a message and leave the output untouched.

This file is named so that pytest collects it last: `pytest -x` stops at the first failure, and a
failure here must not hide the parity tests.  It passes alone in a fresh process (the first call
then also does the per-device address look-up) and after the rest of the suite."""
import numpy as np
import pytest

from oracle import pyoracle as po
from test_gpu_parity import mkctx, vp, dev_array, dev_zeros, to_host

pytestmark = pytest.mark.gpu
MAX_TILES = 65536      # P1_MSG_PACK_STATIC_MAX_TILES (include/p1gpu.h)
TILE_WORDS = 2048


def _words(rng, shape, bits, mixed):
    v = rng.integers(0, 2 ** 64 - 1, shape, dtype=np.uint64, endpoint=True)
    if mixed:   # every varint length, not only the longest
        v = v >> rng.integers(0, 64, shape).astype(np.uint64)
    return (v & np.uint64((1 << bits) - 1)).astype(po.word_dtype(bits))


def _pack_both(lib, ctx, bits, n, count, rng, mixed):
    """one batch through p1_msg_pack and through p1_msg_pack_static; returns (bytes, sizes) of each"""
    ld = max(4, (n + 255) // 256 * 256)
    vecs = np.zeros((count, ld), po.word_dtype(bits))
    vecs[:, :n] = _words(rng, (count, n), bits, mixed)
    values = rng.integers(0, 2 ** 64 - 1, count, dtype=np.uint64, endpoint=True)
    values[0] = 0
    stride = int(lib.p1_msg_max_bytes(ctx, n))
    vd = dev_array(lib, ctx, vecs)
    vals_dev = dev_array(lib, ctx, values)
    res = []
    for static in (False, True):
        out = dev_zeros(lib, ctx, count * stride)
        sizes = dev_zeros(lib, ctx, 8 * count)
        if static:
            lib.check(lib.p1_msg_pack_static(ctx, vd, ld, n, count, vals_dev, out, stride, sizes))
        else:
            lib.check(lib.p1_msg_pack(ctx, vd, ld, n, count, vp(values), out, stride, sizes))
        res.append((to_host(lib, ctx, out, (count, stride), np.uint8), to_host(lib, ctx, sizes, count, np.uint64)))
        for p in (out, sizes):
            lib.check(lib.p1_dev_free(ctx, p))
    for p in (vd, vals_dev):
        lib.check(lib.p1_dev_free(ctx, p))
    return res


@pytest.mark.parametrize("bits", [64, 32])
def test_msg_pack_static_matches_msg_pack_twice_in_a_row(p1lib, bits):
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(20170401 + bits)
    # first call: several tiles per message, a ragged last tile; second call: fewer, shorter messages of
    # the other mix (stale table entries of the first call lie beside and behind the ones it writes);
    # third: empty vectors (no tiles at all, only header-less trailers)
    for n, count, mixed in ((3 * TILE_WORDS + 77, 5, True), (TILE_WORDS + 1, 2, False), (0, 3, True)):
        (want, want_sz), (got, got_sz) = _pack_both(p1lib, ctx, bits, n, count, rng, mixed)
        assert np.array_equal(got_sz, want_sz)
        assert np.array_equal(got, want), "n=%d count=%d: bytes differ from p1_msg_pack" % (n, count)
        assert want_sz.min() >= 2 and (n == 0 or want_sz.min() > n)   # the reference run did produce messages
    p1lib.p1_ctx_destroy(ctx)


def test_msg_pack_static_refuses_what_it_cannot_hold(p1lib):
    ctx = mkctx(p1lib, 64, 64)
    n, count = 64, 2
    stride = int(p1lib.p1_msg_max_bytes(ctx, n))
    vecs = np.arange(count * 256, dtype=np.uint64).reshape(count, 256)
    values = np.array([1, 2], np.uint64)
    vd, vals_dev = dev_array(p1lib, ctx, vecs), dev_array(p1lib, ctx, values)
    out, sizes = dev_zeros(p1lib, ctx, count * stride), dev_zeros(p1lib, ctx, 8 * count)
    # one tile entry more than the array holds: refused on the sizes alone, before any pointer is used
    big_n = (MAX_TILES // count) * TILE_WORDS + 1
    big_stride = int(p1lib.p1_msg_max_bytes(ctx, big_n))
    assert p1lib.p1_msg_pack_static(ctx, vd, big_n + 255, big_n, count, vals_dev, out, big_stride, sizes) == 1
    assert b"built-in scratch" in p1lib.p1_last_error()
    # values on the host would need a staging copy
    assert p1lib.p1_msg_pack_static(ctx, vd, 256, n, count, vp(values), out, stride, sizes) == 1
    assert b"on the device" in p1lib.p1_last_error()
    assert not to_host(p1lib, ctx, out, count * stride, np.uint8).any()
    assert not to_host(p1lib, ctx, sizes, count, np.uint64).any()
    # the largest shape it accepts per message count of 1 is not exercised here (1 GB of words); the call
    # right at a small shape still works after the refusals
    p1lib.check(p1lib.p1_msg_pack_static(ctx, vd, 256, n, count, vals_dev, out, stride, sizes))
    assert to_host(p1lib, ctx, sizes, count, np.uint64).min() > n
    for p in (vd, vals_dev, out, sizes):
        p1lib.check(p1lib.p1_dev_free(ctx, p))
    p1lib.p1_ctx_destroy(ctx)
