"""GPU parity of the text ingestion: p1_parse_doubles against glibc's
strtod token by token (oracle.parse_doubles = what fscanf("%lf") does in src/linear.c:46,82),
and the parsed device buffer fed straight into p1_encode."""
import ctypes
import math
import struct

import numpy as np
import pytest

from oracle import pyoracle as po
from test_gpu_parity import mkctx, vp, dev_zeros, to_host, download

pytestmark = pytest.mark.gpu
c_vp = ctypes.c_void_p
c_u64 = ctypes.c_uint64


def gpu_parse(lib, ctx, text, cap=1 << 16, max_tokens=None):
    buf = np.frombuffer(bytes(text), dtype=np.uint8)
    mt = (len(buf) // 2 + 1) if max_tokens is None else max_tokens
    out = dev_zeros(lib, ctx, 8 * max(mt, 1))
    n, nd = c_u64(), c_u64()
    deferred = np.zeros(max(cap, 1), np.uint64)
    lib.check(lib.p1_parse_doubles(ctx, vp(buf), len(buf), out, mt, ctypes.byref(n), ctypes.byref(nd), vp(deferred), cap))
    vals = to_host(lib, ctx, out, min(int(n.value), mt), np.float64)
    return vals, int(n.value), deferred[:min(int(nd.value), cap)].copy(), int(nd.value), out


def token_offsets(text):
    b = np.frombuffer(bytes(text), dtype=np.uint8)
    sp = np.isin(b, np.frombuffer(b" \t\n\v\f\r", dtype=np.uint8))
    prev = np.concatenate([[True], sp[:-1]])
    return np.flatnonzero(~sp & prev)


def check_against_strtod(oracle, text, vals, n, deferred, allow_long_mantissa=()):
    want, whole = oracle.parse_doubles(text)
    assert n == len(want)
    offs = token_offsets(text)
    dset = set(int(o) for o in deferred)
    assert dset <= set(int(o) for o in offs)
    is_def = np.array([int(o) in dset for o in offs])
    # every token strtod does not take whole (or that is inf/nan/hex) must be deferred
    assert np.all(is_def[~whole]), "tokens outside the decimal grammar were converted"
    extra = np.flatnonzero(is_def & whole)
    assert set(extra.tolist()) <= set(allow_long_mantissa), "plain decimals deferred: %s" % extra[:10]
    ok = ~is_def
    assert np.array_equal(vals[ok].view(np.uint64), want[ok].view(np.uint64)), "conversion differs from strtod"
    assert np.all(np.isnan(vals[is_def]))


def random_doubles(rng, count):
    bits = rng.integers(0, 2 ** 63, count, dtype=np.uint64) | (rng.integers(0, 2, count, dtype=np.uint64) << np.uint64(63))
    v = bits.view(np.float64)
    return v[np.isfinite(v)]


def test_parse_formats_match_strtod(p1lib, oracle):
    ctx = mkctx(p1lib)
    rng = np.random.default_rng(11)
    toks = []
    for x in random_doubles(rng, 3000):
        toks += [repr(float(x)), "%.17g" % x, "%.12g" % x, "%e" % x, "%.19e" % x]
    for x in rng.standard_normal(3000):
        toks += [repr(float(x)), "%.12g" % x, "%f" % x, "%.6f" % (x * 1e4)]
    toks += [str(int(v)) for v in rng.integers(-10 ** 18, 10 ** 18, 500)]
    toks += ["0", "-0", "0.0", "-0.0", "+1", ".5", "5.", "-.5e1", "1E5", "1e+05", "1e-05", "00012.500", "0.000",
             "1e400", "-1e400", "1e-400", "4.9e-324", "2.4703282292062327e-324", "2.4703282292062328e-324",
             "2.2250738585072014e-308", "2.2250738585072011e-308", "1.7976931348623157e308", "1.7976931348623159e308",
             "9007199254740993", "9007199254740992", "9007199254740995", "18014398509481985", "1e22", "1e23", "8.5e22",
             "0.1", "0.3", "1234567890123456789", "12345678901234567890", "123456789012345678901234567890",
             "0.10000000000000000555111512312578270211815834045410156250", "1" + "0" * 40, "0." + "0" * 40 + "1",
             "100000000000000000000000e-23", "1e0000000000000000000001", "7" * 300]
    n_plain = len(toks)
    toks += ["inf", "-inf", "nan", "0x1p3", "0x10", "localhost:1234", "1.5abc", "e5", ".", "-", "+.e1", "1e", "1e+", "--1", "1..2",
             "1," "2"]
    order = rng.permutation(len(toks))
    seps = [" ", "\n", "\t", "  ", " \n", "\r\n"]
    text = "".join(toks[i] + seps[rng.integers(len(seps))] for i in order).encode()
    vals, n, deferred, nd, _ = gpu_parse(p1lib, ctx, text)
    assert nd == len(deferred)
    # which plain tokens may be deferred: mantissas of more than 19 digits whose truncation is ambiguous
    inv = np.argsort(order)
    long_ok = [int(inv[i]) for i, t in enumerate(toks[:n_plain]) if sum(ch.isdigit() for ch in t.split("e")[0].split("E")[0]) > 19]
    check_against_strtod(oracle, text, vals, n, deferred, allow_long_mantissa=long_ok)
    assert nd <= (len(toks) - n_plain) + len(long_ok)
    p1lib.p1_ctx_destroy(ctx)


def test_parse_edges(p1lib, oracle):
    ctx = mkctx(p1lib)
    for text in [b"", b"   \n\t ", b"1", b" 1", b"1 ", b"1.5\n2.5", b"3" * 40 + b" 4"]:
        vals, n, deferred, nd, _ = gpu_parse(p1lib, ctx, text, cap=4)
        want, whole = oracle.parse_doubles(text)
        assert n == len(want)
        if n:
            check_against_strtod(oracle, text, vals, n, deferred, allow_long_mantissa=range(n))
    # counted but not stored beyond max_tokens; deferred list capped
    text = b"1 2 3 x y z 4 5"
    vals, n, deferred, nd, _ = gpu_parse(p1lib, ctx, text, cap=2, max_tokens=4)
    assert n == 8 and nd == 3 and len(deferred) == 2 and len(vals) == 4
    assert list(vals[:3]) == [1.0, 2.0, 3.0] and math.isnan(vals[3])
    p1lib.p1_ctx_destroy(ctx)


def test_parse_large_buffer_many_tiles(p1lib, oracle):
    """> 4096 tiles (two scan partitions), tokens straddling tile boundaries"""
    ctx = mkctx(p1lib)
    rng = np.random.default_rng(12)
    x = rng.standard_normal(2_200_000)
    text = " ".join(map(repr, x.tolist())).encode()
    assert len(text) > 4096 * 8192
    vals, n, deferred, nd, _ = gpu_parse(p1lib, ctx, text)
    assert n == len(x) and nd == 0
    assert np.array_equal(vals.view(np.uint64), x.view(np.uint64))      # repr round-trips, strtod is correctly rounded
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("b1,b2,prec", [(64, 64, 56), (32, 32, 24)])
def test_instance_text_to_encoded_words(p1lib, oracle, b1, b2, prec, tmp_path):
    """the X / y part of a reference instance file -> p1_parse_doubles -> p1_encode (device pointers)
    gives the words read_matrix / read_vector produce (src/linear.c:27-102)"""
    ctx = mkctx(p1lib, b1, b2)
    n, d = 700, 9
    X, y, _ = po.generate_lin_regression(n, d, seed=5)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, [0, 3, 6])
    raw = open(path, "rb").read()
    # config.c consumes the header lines (n d p, endpoints); the data part starts at the "n d" line of X
    lines = raw.split(b"\n")
    start = len(b"\n".join(lines[:1 + 2 + 3])) + 1
    text = raw[start:]
    vals, ntok, deferred, nd, out = gpu_parse(p1lib, ctx, text)
    assert nd == 0 and int(vals[0]) == n and int(vals[1]) == d and int(vals[2 + n * d]) == n
    normalizer = oracle.normalizer(prec, n)
    data = c_vp()
    Xp = c_vp(out.value + 8 * 2)
    yp = c_vp(out.value + 8 * (2 + n * d + 1))
    p1lib.check(p1lib.p1_encode(ctx, Xp, yp, n, d, normalizer, prec, ctypes.byref(data)))
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    Xt = np.array(text.split()[2:2 + n * d], dtype=np.float64)          # numpy's parser is correctly rounded too
    yt = np.array(text.split()[3 + n * d:3 + n * d + n], dtype=np.float64)
    assert np.array_equal(Xw.ravel(), oracle.encode(Xt, normalizer, prec, b1, b2))
    assert np.array_equal(yw, oracle.encode(yt, normalizer, prec, b1, b2))
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)
