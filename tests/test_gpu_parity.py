"""GPU suite (-m gpu): every kernel of libp1gpu.so through the C-ABI against the oracle,
bit-exact, plus whole phase-1 runs against the committed reference golden vectors."""
import ctypes

import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

c_vp = ctypes.c_void_p
WIDTHS = [(64, 64, 56, 0), (32, 32, 24, 0), (64, 32, 56, 24)]


def mkctx(lib, b1=64, b2=64):
    ctx = c_vp()
    lib.check(lib.p1_ctx_create(0, b1, b2, ctypes.byref(ctx)))
    return ctx


def vp(a):
    return a.ctypes.data_as(c_vp) if a is not None else None


def encode(lib, ctx, X, y, normalizer, precision):
    out = c_vp()
    lib.check(lib.p1_encode(ctx, vp(X), vp(y), X.shape[0], X.shape[1], normalizer, precision, ctypes.byref(out)))
    return out


def from_words(lib, ctx, Xw, yw):
    out = c_vp()
    lib.check(lib.p1_data_from_words(ctx, vp(Xw), vp(yw), Xw.shape[0], Xw.shape[1], ctypes.byref(out)))
    return out


def download(lib, ctx, data, n, d, bits, with_y=True):
    dt = po.word_dtype(bits)
    Xw = np.empty((n, d), dt)
    yw = np.empty(n, dt) if with_y else None
    lib.check(lib.p1_data_download(ctx, data, vp(Xw), vp(yw)))
    return Xw, yw


def dev_array(lib, ctx, host):
    p = c_vp()
    lib.check(lib.p1_dev_alloc(ctx, max(host.nbytes, 8), ctypes.byref(p)))
    lib.check(lib.p1_copy_h2d(ctx, p, vp(host), host.nbytes))
    return p


def dev_zeros(lib, ctx, nbytes):
    p = c_vp()
    lib.check(lib.p1_dev_alloc(ctx, max(nbytes, 8), ctypes.byref(p)))
    lib.check(lib.p1_dev_zero(ctx, p, nbytes))
    return p


def to_host(lib, ctx, p, shape, dtype):
    out = np.empty(shape, dtype)
    lib.check(lib.p1_copy_d2h(ctx, vp(out), p, out.nbytes))
    return out


# ---------------------------------------------------------------------------
@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
def test_encode_parity(p1lib, oracle, b1, b2, prec, prec2):
    rng = np.random.default_rng(1)
    n, d = 777, 37  # ragged: not multiples of the 32x32 tiles
    X = rng.standard_normal((n, d))
    y = rng.standard_normal(n) * 3
    # x86 cast edge cases (src/fixed.c:3-5 compiled by gcc): out of range, NaN, +-0, denormals
    X[0, :8] = [1e30, -1e30, float("nan"), float("inf"), float("-inf"), 0.0, -0.0, 5e-324]
    X[1, :4] = [2.0 ** 31, -2.0 ** 31, 2.0 ** 63, -2.0 ** 63]
    y[:3] = [1e300, -1e300, float("nan")]
    nz = oracle.normalizer(prec, n)
    ctx = mkctx(p1lib, b1, b2)
    data = encode(p1lib, ctx, X, y, nz, prec)
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    assert np.array_equal(Xw, oracle.encode(X, nz, prec, b1, b2))
    assert np.array_equal(yw, oracle.encode(y, nz, prec, b1, b2))
    # the same from device-resident doubles
    import torch
    Xd = torch.from_numpy(X).cuda()
    out = c_vp()
    p1lib.check(p1lib.p1_encode(ctx, c_vp(Xd.data_ptr()), None, n, d, nz, prec, ctypes.byref(out)))
    Xw2, _ = download(p1lib, ctx, out, n, d, b1, with_y=False)
    assert np.array_equal(Xw2, Xw)
    p1lib.p1_data_free(out)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


def _check_local_block(p1lib, oracle, ctx, data, Xw, yw, lo, hi, owns_target, bits, algo):
    n, d = Xw.shape
    nA = d * (d + 1) // 2
    L = int(p1lib.p1_acc_len(d))
    acc = dev_zeros(p1lib, ctx, 8 * L)
    p1lib.check(p1lib.p1_local_block(ctx, data, lo, hi, int(owns_target), acc, algo))
    got = to_host(p1lib, ctx, acc, L, np.uint64)
    mask = (1 << bits) - 1
    want = np.zeros(nA + d, np.uint64)
    for i in range(lo, hi):
        for j in range(lo, i):
            want[i * (i + 1) // 2 + j] = oracle.dot(Xw[:, i], Xw[:, j], bits)
        if owns_target:
            want[nA + i] = oracle.dot(yw, Xw[:, i], bits)
    off = np.ones(nA + d, bool)
    off[[i * (i + 1) // 2 + i for i in range(d)]] = False  # diagonal slots belong to the FP64 path
    assert np.array_equal(got[:nA + d][off] & np.uint64(mask), want[off])
    # diagonal statistics: exact sum of squares of the owned columns
    for c in range(lo, hi):
        col = Xw[:, c].astype(np.int64 if bits == 64 else np.int32).astype(object)
        S = sum(int(v) * int(v) for v in col)
        viol = sum(1 for v in col if abs(int(v)) >= 2 ** 26)
        if viol == 0:
            assert (int(got[nA + 2 * d + c]) << 32) + int(got[nA + d + c]) == S
        assert int(got[nA + 3 * d + c]) == viol
    p1lib.check(p1lib.p1_dev_free(ctx, acc))


@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
def test_local_block_imad_parity(p1lib, oracle, b1, b2, prec, prec2):
    n, d = 1500, 70  # two column tiles, ragged rows
    X, y, _ = po.generate_lin_regression(n, d, seed=2)
    nz = oracle.normalizer(prec, n)
    ctx = mkctx(p1lib, b1, b2)
    data = encode(p1lib, ctx, X, y, nz, prec)
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, b1, 1)       # single provider: whole Gram + X^T y
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 20, 47, False, b1, 1)    # middle provider
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 47, 70, True, b1, 1)     # last provider
    p1lib.p1_data_free(data)
    # full-range ring words (the wide 64x64 IMAD path)
    rng = np.random.default_rng(4)
    dt = po.word_dtype(b1)
    Xf = rng.integers(0, 2 ** 63, (600, 9), dtype=np.uint64).astype(dt) * dt(3)
    yf = rng.integers(0, 2 ** 63, 600, dtype=np.uint64).astype(dt) * dt(5)
    data = from_words(p1lib, ctx, Xf, yf)
    _check_local_block(p1lib, oracle, ctx, data, Xf, yf, 0, 9, True, b1, 1)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


def _signed_words(rng, n, d, bits, mag_bits):
    """ring words that are signed integers of at most mag_bits significant bits (full range if >= bits)"""
    dt = po.word_dtype(bits)
    if mag_bits >= bits:
        return rng.integers(0, 2 ** bits, (n, d), dtype=np.uint64).astype(dt)
    lim = 2 ** (mag_bits - 1)
    sdt = np.int64 if bits == 64 else np.int32
    a = rng.integers(-lim, lim, (n, d)).astype(sdt)
    a[0, 0], a[1 % n, 0] = lim - 1, -lim       # extremes of the range
    return a.view(dt)


@pytest.mark.parametrize("bits,mag_bits", [(64, 7), (64, 15), (64, 20), (64, 27), (64, 40), (64, 64),
                                           (32, 6), (32, 14), (32, 22), (32, 32)])
def test_local_block_tcgen05_digits(p1lib, oracle, bits, mag_bits):
    """tcgen05 kind::i8 Gram (src/phase1.c:14-20,569-571) for every digit count L = 1..W/8:
    operands of growing magnitude up to full-range ring words, with the target row."""
    rng = np.random.default_rng(100 + mag_bits)
    n, d = 1100, 70
    Xw = _signed_words(rng, n, d, bits, mag_bits)
    yw = _signed_words(rng, n, 1, bits, mag_bits)[:, 0].copy()
    ctx = mkctx(p1lib, bits, bits)
    data = from_words(p1lib, ctx, Xw, yw)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, bits, 2)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 3, 41, False, bits, 2)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
def test_local_block_tcgen05_shapes(p1lib, oracle, b1, b2, prec, prec2):
    """several 128x64 output tiles, ragged column/row counts, multi-chunk drains and row
    slices (forced small through the tuning knobs), on encoded synthetic data"""
    n, d = 5000, 300
    X, y, _ = po.generate_lin_regression(n, d, seed=3)
    nz = oracle.normalizer(prec, n)
    ctx = mkctx(p1lib, b1, b2)
    data = encode(p1lib, ctx, X, y, nz, prec)
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, b1, 2)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_max_slices", 2))
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_chunk_blocks", 3))   # 20 blocks per CTA -> 7 drains
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, b1, 2)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 100, 229, False, b1, 2)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 171, 300, True, b1, 2)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("d,owns", [(16, False), (100, True), (128, False), (128, True), (200, True), (255, True), (256, False)])
def test_local_block_tcgen05_one_digit_whole_block(p1lib, oracle, bits, d, owns):
    """one-digit words (|word| <= 127: what the 32-bit configurations hold) in a block of at most 256 columns
    take the kernel in which every CTA multiplies the WHOLE block over its rows (k_gram_whole: one tile load per
    K-block, two MMAs into 384 TMEM columns, straight red.add epilogue).  Against the oracle, for one and two row
    tiles, ragged rows, with the row slices and the s32 chunks forced small, and against the tiled kernel."""
    rng = np.random.default_rng(7 * d + bits)
    n = 3001
    Xw = _signed_words(rng, n, d, bits, 7)
    yw = _signed_words(rng, n, 1, bits, 7)[:, 0].copy()
    Xw[2, d - 1] = 127                                                   # the largest one-digit magnitude
    ctx = mkctx(p1lib, bits, bits)
    data = from_words(p1lib, ctx, Xw, yw)
    L = ctypes.c_int(0)
    p1lib.check(p1lib.p1_gram_digits(ctx, data, 0, d, int(owns), ctypes.byref(L)))
    assert L.value == 1
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, owns, bits, 2)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_max_slices", 3))
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_chunk_blocks", 3))   # 8 K-blocks per CTA -> 3 chunks
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, owns, bits, 2)
    if d > 40:
        _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 5, d - 3, False, bits, 2)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_whole", 0))
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, owns, bits, 2)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits,d", [(32, 40), (64, 20), (32, 90)])
def test_local_block_tcgen05_whole_block_narrow_many_rows(p1lib, oracle, bits, d):
    """narrow one-digit blocks over enough rows that a pipeline stage of k_gram_whole holds several K-blocks
    (one TMA box of up to 24 KB; the last stage of the last CTA runs past the rows and is zero-filled)"""
    rng = np.random.default_rng(11 * d + bits)
    n = 148 * 128 * 5 + 777
    Xw = _signed_words(rng, n, d, bits, 7)
    yw = _signed_words(rng, n, 1, bits, 7)[:, 0].copy()
    ctx = mkctx(p1lib, bits, bits)
    data = from_words(p1lib, ctx, Xw, yw)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, bits, 2)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"gram_tc_chunk_blocks", 2))
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 1, d - 2, False, bits, 2)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


def _limb_split_launches(p1lib, ctx):
    from linreg_mpc_b200 import ffi
    ms, cnt = ctypes.c_double(), ctypes.c_uint64()
    p1lib.check(p1lib.p1_ctx_profile_read(ctx, ffi.P1_K_GRAM_LIMBS, ctypes.byref(ms), ctypes.byref(cnt)))
    return cnt.value


@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
@pytest.mark.parametrize("n,d", [(4503, 64), (1030, 500)])
def test_encode_async_pipeline_and_fallbacks(p1lib, oracle, b1, b2, prec, prec2, n, d):
    """the cp.async encode kernel (even d, 16-byte aligned source; the default) against the oracle, with the x86
    cast edge cases, ragged row tiles and a ragged last column tile; the same input at an 8-byte aligned
    address and with the pipeline switched off must take the register-staged kernels and give the same words
    and the same statistics (through the exact diagonal)"""
    import torch
    rng = np.random.default_rng(12)
    X = rng.standard_normal((n, d))
    y = rng.standard_normal(n)
    X[0, :8] = [1e30, -1e30, float("nan"), float("inf"), float("-inf"), 0.0, -0.0, 5e-324]
    X[n - 1, d - 4:] = [2.0 ** 31, -2.0 ** 31, 2.0 ** 63, -2.0 ** 63]
    nz = oracle.normalizer(prec, n)
    want = oracle.encode(X, nz, prec, b1, b2)
    flat = torch.zeros(n * d + 1, dtype=torch.float64, device="cuda")
    flat[1:] = torch.from_numpy(X).cuda().reshape(-1)
    Xal = torch.from_numpy(X).cuda()
    assert Xal.data_ptr() % 16 == 0 and flat[1:].data_ptr() % 16 == 8
    L = int(p1lib.p1_acc_len(d))
    accs = []
    for name, ptr_, async_ in (("pipeline", Xal.data_ptr(), 1), ("unaligned", flat[1:].data_ptr(), 1), ("off", Xal.data_ptr(), 0), ("host", None, 1)):
        ctx = mkctx(p1lib, b1, b2)
        p1lib.check(p1lib.p1_ctx_set_option(ctx, b"encode_async", async_))
        out = c_vp()
        p1lib.check(p1lib.p1_encode(ctx, c_vp(ptr_) if ptr_ else vp(X), vp(y), n, d, nz, prec, ctypes.byref(out)))
        Xw, yw = download(p1lib, ctx, out, n, d, b1)
        assert np.array_equal(Xw, want), name
        assert np.array_equal(yw, oracle.encode(y, nz, prec, b1, b2)), name
        acc = dev_zeros(p1lib, ctx, 8 * L)
        p1lib.check(p1lib.p1_local_block(ctx, out, 0, d, 1, acc, 1))
        accs.append(to_host(p1lib, ctx, acc, L, np.uint64))
        p1lib.check(p1lib.p1_dev_free(ctx, acc))
        p1lib.p1_data_free(out)
        p1lib.p1_ctx_destroy(ctx)
    nA = d * (d + 1) // 2
    with np.errstate(over="ignore"):
        for a in accs[1:]:
            assert np.array_equal(a[:nA + d], accs[0][:nA + d])                        # ring words
            assert np.array_equal(a[nA + 3 * d:], accs[0][nA + 3 * d:])                # |word| >= 2^26 counts
            # sum of squares = (hi << 32) + lo; how it splits into the two rows depends on the flush grouping
            sq = lambda v: (v[nA + 2 * d:nA + 3 * d] << np.uint64(32)) + v[nA + d:nA + 2 * d]
            assert np.array_equal(sq(a), sq(accs[0]))


@pytest.mark.parametrize("d,async_", [(2, 1), (2, 0), (3, 1)])
def test_encode_more_rows_than_one_grid(p1lib, oracle, d, async_):
    """a device-resident matrix of more than 2^24 rows is encoded in row slabs (grid.y holds <= 65535 tiles)"""
    import torch
    n = (1 << 24) + 1000
    g = torch.Generator(device="cuda"); g.manual_seed(31)
    Xd = torch.randn((n, d), dtype=torch.float64, device="cuda", generator=g)
    yd = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
    nz = oracle.normalizer(56, n)
    ctx = mkctx(p1lib, 64, 64)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"encode_async", async_))
    out = c_vp()
    p1lib.check(p1lib.p1_encode(ctx, c_vp(Xd.data_ptr()), c_vp(yd.data_ptr()), n, d, nz, 56, ctypes.byref(out)))
    Xw, yw = download(p1lib, ctx, out, n, d, 64)
    assert np.array_equal(Xw, oracle.encode(Xd.cpu().numpy(), nz, 56, 64, 64))
    assert np.array_equal(yw, oracle.encode(yd.cpu().numpy(), nz, 56, 64, 64))
    p1lib.p1_data_free(out)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("async_", [1, 0])
@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
@pytest.mark.parametrize("n,d", [(5000, 300), (70001, 37)])
def test_encode_with_digit_planes(p1lib, oracle, b1, b2, prec, prec2, n, d, async_):
    """option encode_planes: the encode pass also writes the digit planes of the tcgen05 Gram.
    Words, statistics and every Gram block must equal the two-pass result, and the Gram must
    not run its own digit split (except where the planes cannot serve it)."""
    X, y, _ = po.generate_lin_regression(n, d, seed=5)
    nz = oracle.normalizer(prec, n)
    ctx = mkctx(p1lib, b1, b2)
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"encode_planes", 1))
    p1lib.check(p1lib.p1_ctx_set_option(ctx, b"encode_async", async_))
    p1lib.check(p1lib.p1_ctx_profile(ctx, 1))
    data = encode(p1lib, ctx, X, y, nz, prec)
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    assert np.array_equal(Xw, oracle.encode(X, nz, prec, b1, b2))
    assert np.array_equal(yw, oracle.encode(y, nz, prec, b1, b2))
    digits = ctypes.c_int()
    p1lib.check(p1lib.p1_gram_digits(ctx, data, 0, d, 1, ctypes.byref(digits)))
    planes_serve = digits.value <= 4                      # more digits than the encode pass provided: split again
    base = _limb_split_launches(p1lib, ctx)
    lo, hi = d // 3, d - d // 4
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, b1, 2)        # single provider
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, lo, hi, False, b1, 2)     # middle provider
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, hi, d, True, b1, 2)       # last provider
    if planes_serve:
        assert _limb_split_launches(p1lib, ctx) == base
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, hi, True, b1, 2)       # target with a non-adjacent block: own split
    assert _limb_split_launches(p1lib, ctx) > base
    p1lib.p1_data_free(data)
    # values far beyond the planes' digit budget: the Gram falls back to its own split
    data = encode(p1lib, ctx, X * 4096.0, y, nz, prec)
    Xw, yw = download(p1lib, ctx, data, n, d, b1)
    _check_local_block(p1lib, oracle, ctx, data, Xw, yw, 0, d, True, b1, 2)
    p1lib.p1_data_free(data)
    # no target vector
    out = c_vp()
    p1lib.check(p1lib.p1_encode(ctx, vp(X), None, n, d, nz, prec, ctypes.byref(out)))
    Xw, _ = download(p1lib, ctx, out, n, d, b1, with_y=False)
    _check_local_block(p1lib, oracle, ctx, out, Xw, None, 0, d, False, b1, 2)
    p1lib.p1_data_free(out)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("b1,b2,prec", [(64, 64, 56), (32, 32, 24)])
def test_encode_with_digit_planes_staged_host_input(p1lib, oracle, b1, b2, prec):
    """host rows larger than one staging chunk (256 MB): the chunks of the encode pass land at row offsets
    > 0 of the column store and of the digit planes; the Gram from the planes must equal the Gram that
    splits the stored words itself (verified against the oracle in the tests above), word for word."""
    n, d = 300_001, 128                      # 307 MB of doubles: chunks of 262144 + 37857 rows
    rng = np.random.default_rng(8)
    X = rng.standard_normal((n, d))
    y = rng.standard_normal(n)
    nz = oracle.normalizer(prec, n)
    L = int(p1lib.p1_acc_len(d))
    accs = []
    for planes in (1, 0):
        ctx = mkctx(p1lib, b1, b2)
        p1lib.check(p1lib.p1_ctx_set_option(ctx, b"encode_planes", planes))
        p1lib.check(p1lib.p1_ctx_profile(ctx, 1))
        data = encode(p1lib, ctx, X, y, nz, prec)
        if planes:
            Xw, yw = download(p1lib, ctx, data, n, d, b1)
            assert np.array_equal(Xw[::997], oracle.encode(X[::997], nz, prec, b1, b2))
            assert np.array_equal(Xw[-300:], oracle.encode(X[-300:], nz, prec, b1, b2))
            assert np.array_equal(yw, oracle.encode(y, nz, prec, b1, b2))
        acc = dev_zeros(p1lib, ctx, 8 * L)
        p1lib.check(p1lib.p1_local_block(ctx, data, 0, d, 1, acc, 2))
        accs.append(to_host(p1lib, ctx, acc, L, np.uint64))
        assert (_limb_split_launches(p1lib, ctx) == 0) == bool(planes)
        if planes:
            i, j = 77, 5
            assert int(accs[0][i * (i + 1) // 2 + j]) & ((1 << b1) - 1) == oracle.dot(Xw[:, i], Xw[:, j], b1)
        p1lib.check(p1lib.p1_dev_free(ctx, acc))
        p1lib.p1_data_free(data)
        p1lib.p1_ctx_destroy(ctx)
    assert np.array_equal(accs[0], accs[1])


@pytest.mark.parametrize("bits", [64, 32])
def test_diagonal_exact_and_sequential(p1lib, oracle, bits):
    b2 = bits
    prec = 56 if bits == 64 else 24
    ctx = mkctx(p1lib, bits, b2)
    rng = np.random.default_rng(6)
    n, d = 3000, 6
    dt = po.word_dtype(bits)
    sdt = np.int64 if bits == 64 else np.int32
    lim = 2 ** 20 if bits == 64 else 2 ** 14   # n * lim^2 stays below 2^53
    Xs = rng.integers(-lim, lim, (n, d)).astype(sdt)
    if bits == 64:
        Xs[:, 1] *= 2 ** 12      # |word| up to 2^32 >= 2^26: the exact shortcut must be refused
        Xs[:, 4] = rng.integers(2 ** 25, 2 ** 26, n)  # no single violation but sum of squares > 2^53
    Xw = Xs.view(dt)
    data = from_words(p1lib, ctx, Xw, None)
    L = int(p1lib.p1_acc_len(d)); nA = d * (d + 1) // 2
    acc = dev_zeros(p1lib, ctx, 8 * L)
    p1lib.check(p1lib.p1_local_block(ctx, data, 0, d, 0, acc, 1))
    need = (ctypes.c_int * d)()
    A, B = c_vp(), c_vp()
    st = p1lib.p1_finalize(ctx, acc, d, 0, d, prec, 0, float(d), None, need, ctypes.byref(A), ctypes.byref(B))
    want_need = [0, 1, 0, 0, 1, 0] if bits == 64 else [0] * d
    assert list(need) == want_need
    xy = np.full(d, np.nan)
    if any(want_need):
        assert st == 1  # refuses to guess: the sequential sums are required
        xy = np.zeros(d)
        # chained in two row shards to exercise the continuation
        h1 = from_words(p1lib, ctx, np.ascontiguousarray(Xw[:1234]), None)
        h2 = from_words(p1lib, ctx, np.ascontiguousarray(Xw[1234:]), None)
        p1lib.check(p1lib.p1_diag_sequential(ctx, h1, 0, d, prec, vp(xy)))
        p1lib.check(p1lib.p1_diag_sequential(ctx, h2, 0, d, prec, vp(xy)))
        p1lib.p1_data_free(h1); p1lib.p1_data_free(h2)
        p1lib.check(p1lib.p1_finalize(ctx, acc, d, 0, d, prec, 0, float(d), vp(xy), None, ctypes.byref(A), ctypes.byref(B)))
    else:
        assert st == 0
    resA = np.ctypeslib.as_array(ctypes.cast(A, ctypes.POINTER(ctypes.c_uint64 if bits == 64 else ctypes.c_uint32)), (nA,))
    for c in range(d):
        assert int(resA[c * (c + 1) // 2 + c]) == oracle.diag(Xw[:, c], prec, float(d), bits), c
    # the sequential kernel equals the literal loop on every column, also where the shortcut applies
    xy2 = np.zeros(d)
    p1lib.check(p1lib.p1_diag_sequential(ctx, data, 0, d, prec, vp(xy2)))
    for c in range(d):
        acc_d = ctypes.c_double(0.0)
        oracle.lib.p1o_diag_accumulate(vp(np.ascontiguousarray(Xw[:, c])), 1, n, prec, bits, ctypes.byref(acc_d))
        assert xy2[c] == acc_d.value
    p1lib.free(A); p1lib.free(B)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [64, 32])
@pytest.mark.parametrize("n_total,row0,n_rows", [(1000, 0, 1000), (1001, 0, 1001), (1003, 0, 1003), (4096, 1024, 2048),
                                                 (999, 333, 666), (130, 0, 130), (1, 0, 1)])
def test_ti_generate_parity(p1lib, oracle, bits, n_total, row0, n_rows):
    ctx = mkctx(p1lib, bits, bits)
    dt = po.word_dtype(bits)
    q0, count = 5, 4
    ld = (n_rows + 255) // 256 * 256
    x = dev_zeros(p1lib, ctx, count * ld * dt().nbytes)
    y = dev_zeros(p1lib, ctx, count * ld * dt().nbytes)
    xy = dev_zeros(p1lib, ctx, 8 * count)
    r = dev_zeros(p1lib, ctx, 8 * count)
    p1lib.check(p1lib.p1_ti_generate(ctx, po.TI_KEY, q0, count, n_total, row0, n_rows, x, y, ld, xy, r))
    xh = to_host(p1lib, ctx, x, (count, ld), dt)
    yh = to_host(p1lib, ctx, y, (count, ld), dt)
    xyh = to_host(p1lib, ctx, xy, count, np.uint64)
    rh = to_host(p1lib, ctx, r, count, np.uint64)
    # values-only call (seeded mode on the TI side) gives the same dot products
    xy2 = dev_zeros(p1lib, ctx, 8 * count)
    p1lib.check(p1lib.p1_ti_generate(ctx, po.TI_KEY, q0, count, n_total, row0, n_rows, None, None, ld, xy2, None))
    assert np.array_equal(to_host(p1lib, ctx, xy2, count, np.uint64), xyh)
    mask = (1 << bits) - 1
    for p in range(count):
        xo, yo, va, vb = oracle.ti_pair(po.TI_KEY, q0 + p, n_total, bits)
        assert np.array_equal(xh[p, :n_rows], xo[row0:row0 + n_rows]), p
        assert np.array_equal(yh[p, :n_rows], yo[row0:row0 + n_rows]), p
        assert not xh[p, n_rows:].any()  # padding untouched
        assert int(rh[p]) == vb
        assert int(xyh[p]) & mask == oracle.dot(xo[row0:row0 + n_rows], yo[row0:row0 + n_rows], bits)
        if row0 == 0 and n_rows == n_total:
            assert (int(xyh[p]) - int(rh[p])) & mask == va
    for p_ in (x, y, xy, r, xy2):
        p1lib.check(p1lib.p1_dev_free(ctx, p_))
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [64, 32])
def test_party_kernels_parity(p1lib, oracle, bits):
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(8)
    dt = po.word_dtype(bits)
    n, d, count = 2051, 7, 5  # ragged n exercises the scalar tail
    Xw = rng.integers(0, 2 ** 63, (n, d), dtype=np.uint64).astype(dt)
    yw = rng.integers(0, 2 ** 63, n, dtype=np.uint64).astype(dt)
    data = from_words(p1lib, ctx, Xw, yw)
    ld = (n + 255) // 256 * 256
    T = rng.integers(0, 2 ** 63, (count, ld), dtype=np.uint64).astype(dt)
    IN = rng.integers(0, 2 ** 63, (count, ld), dtype=np.uint64).astype(dt)
    cols = np.array([0, 3, 7, 6, 3], np.int32)  # 7 = target vector
    slots = np.array([2, 9, 4, 2, 11], np.int64)  # slot 2 hit twice: accumulation
    colv = lambda c: yw if c == d else Xw[:, c]
    Td, INd = dev_array(p1lib, ctx, T), dev_array(p1lib, ctx, IN)
    out = dev_zeros(p1lib, ctx, T.nbytes)
    acc = dev_zeros(p1lib, ctx, 8 * 16)
    mask = (1 << bits) - 1
    # b: out = col + T
    p1lib.check(p1lib.p1_party_b_mask(ctx, data, vp(cols), count, Td, ld, out, ld))
    oh = to_host(p1lib, ctx, out, (count, ld), dt)
    for p in range(count):
        assert np.array_equal(oh[p, :n], colv(cols[p]) + T[p, :n])
    # a: out = col - T, acc += <in, T>
    p1lib.check(p1lib.p1_party_a_step(ctx, data, vp(cols), vp(slots), count, Td, ld, INd, ld, out, ld, acc))
    oh = to_host(p1lib, ctx, out, (count, ld), dt)
    want = np.zeros(16, np.uint64)
    for p in range(count):
        assert np.array_equal(oh[p, :n], colv(cols[p]) - T[p, :n])
        want[slots[p]] += np.uint64(oracle.dot(IN[p, :n], T[p, :n], bits))
    got = to_host(p1lib, ctx, acc, 16, np.uint64)
    assert np.array_equal(got & np.uint64(mask), want & np.uint64(mask))
    # b: acc += <in, col>
    p1lib.check(p1lib.p1_party_b_share(ctx, data, vp(cols), vp(slots), count, INd, ld, acc))
    for p in range(count):
        want[slots[p]] += np.uint64(oracle.dot(IN[p, :n], colv(cols[p]), bits))
    got = to_host(p1lib, ctx, acc, 16, np.uint64)
    assert np.array_equal(got & np.uint64(mask), want & np.uint64(mask))
    # acc -= value
    vals = rng.integers(0, 2 ** 63, count, dtype=np.uint64)
    vd = dev_array(p1lib, ctx, vals)
    p1lib.check(p1lib.p1_acc_sub_values(ctx, acc, vp(slots), vd, count))
    for p in range(count):
        want[slots[p]] -= vals[p]
    got = to_host(p1lib, ctx, acc, 16, np.uint64)
    assert np.array_equal(got & np.uint64(mask), want & np.uint64(mask))
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [64, 32])
@pytest.mark.parametrize("n_total,row0,n_rows", [(9000, 0, 9000), (9001, 0, 9001), (9003, 0, 9003), (20480, 4096, 8192),
                                                 (9999, 3333, 6666), (130, 0, 130), (1, 0, 1), (70000, 512, 66000)])
def test_masking_fused_parity(p1lib, oracle, bits, n_total, row0, n_rows):
    """the one-kernel masking loop (p1_masking_fused): both party messages, the masks when kept, <x,y>, r
    and both share sums equal the oracle's TI stream + party algebra (src/phase1.c:173-224,270-278), for
    every phase of a pair's mask inside the 16-byte AES blocks (odd pairs, odd n, row shards)"""
    _masking_fused_case(p1lib, oracle, bits, n_total, row0, n_rows, 7)


@pytest.mark.parametrize("bits", [64, 32])
def test_masking_fused_counter_crosses_2_32_blocks(p1lib, oracle, bits):
    """a launch whose pairs straddle a multiple of 2^32 keystream blocks: pairs before it run under the key folded
    for the launch's first high counter word, the pair that contains it through the general loop, pairs after it
    under the key folded for the next one (k_fused.cu: key / keyb)"""
    n_total = 70000
    per_pair = (2 * n_total + 1) * (bits // 8)               # keystream bytes per pair
    q_cross = (16 << 32) // per_pair                         # the pair that contains block 2^32
    _masking_fused_case(p1lib, oracle, bits, n_total, 512, 66000, q_cross - 2)


def _masking_fused_case(p1lib, oracle, bits, n_total, row0, n_rows, q0):
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(80 + n_total)
    dt = po.word_dtype(bits)
    d, count = 5, 6
    Xw = rng.integers(0, 2 ** 63, (n_rows, d), dtype=np.uint64).astype(dt)
    yw = rng.integers(0, 2 ** 63, n_rows, dtype=np.uint64).astype(dt)
    data = from_words(p1lib, ctx, Xw, yw)
    colv = lambda c: yw if c == d else Xw[:, c]
    ld = (n_rows + 255) // 256 * 256
    cols_a = np.array([5, 4, 4, 3, 2, 5], np.int32)      # 5 = the target vector
    cols_b = np.array([0, 1, 2, 0, 1, 3], np.int32)
    slots_a = np.array([0, 1, 2, 3, 1, 5], np.int64)      # slot 1 is hit twice: accumulation
    slots_b = np.array([8, 9, 10, 11, 12, 9], np.int64)
    mask = (1 << bits) - 1
    for keep in (False, True):
        bufs = [dev_zeros(p1lib, ctx, count * ld * dt().nbytes) for _ in range(4)]
        mb, ma, x, y = bufs
        xy = dev_zeros(p1lib, ctx, 8 * count)
        r = dev_zeros(p1lib, ctx, 8 * count)
        acc = dev_zeros(p1lib, ctx, 8 * 16)
        p1lib.check(p1lib.p1_masking_fused(ctx, po.TI_KEY, q0, count, n_total, row0, data, vp(cols_a), vp(cols_b),
                                           vp(slots_a), vp(slots_b), mb, ma, ld, x if keep else None, y if keep else None,
                                           xy, r, acc))
        mbh, mah = to_host(p1lib, ctx, mb, (count, ld), dt), to_host(p1lib, ctx, ma, (count, ld), dt)
        xh, yh = to_host(p1lib, ctx, x, (count, ld), dt), to_host(p1lib, ctx, y, (count, ld), dt)
        xyh, rh = to_host(p1lib, ctx, xy, count, np.uint64), to_host(p1lib, ctx, r, count, np.uint64)
        acch = to_host(p1lib, ctx, acc, 16, np.uint64)
        want = np.zeros(16, np.uint64)
        for p in range(count):
            xo, yo, va, vb = oracle.ti_pair(po.TI_KEY, q0 + p, n_total, bits)
            xo, yo = xo[row0:row0 + n_rows], yo[row0:row0 + n_rows]
            a, b = colv(cols_a[p]), colv(cols_b[p])
            assert np.array_equal(mbh[p, :n_rows], b + xo), p
            assert np.array_equal(mah[p, :n_rows], a - yo), p
            assert not mbh[p, n_rows:].any() and not mah[p, n_rows:].any()     # padding untouched
            if keep:
                assert np.array_equal(xh[p, :n_rows], xo) and np.array_equal(yh[p, :n_rows], yo)
            else:
                assert not xh.any() and not yh.any()
            assert int(rh[p]) == vb
            assert int(xyh[p]) & mask == oracle.dot(xo, yo, bits)
            want[slots_a[p]] += np.uint64(oracle.dot(b + xo, yo, bits))
            want[slots_b[p]] += np.uint64(oracle.dot(a - yo, b, bits))
        assert np.array_equal(acch & np.uint64(mask), want & np.uint64(mask))
        for p_ in bufs + [xy, r, acc]:
            p1lib.check(p1lib.p1_dev_free(ctx, p_))
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [64, 32])
@pytest.mark.parametrize("n", [64, 37, 515])
def test_cot_inner_product_parity(p1lib, oracle, bits, n):
    """one OT-based inner product: both shares and both wire messages equal the oracle's"""
    from linreg_mpc_b200 import phase1
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(10)
    dt = po.word_dtype(bits)
    Xw = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64).astype(dt) * dt(3)
    data = from_words(p1lib, ctx, Xw, None)
    s0, s1, delta = phase1.derive_cot_session(po.OT_KEY, 2, 4)
    chosen = np.ascontiguousarray(np.where(phase1._bits128(delta)[:, None], s1, s0))
    snd, rcv = c_vp(), c_vp()
    p1lib.check(p1lib.p1_cot_sender_new(ctx, delta.tobytes(), vp(chosen), ctypes.byref(snd)))
    p1lib.check(p1lib.p1_cot_receiver_new(ctx, vp(s0), vp(s1), ctypes.byref(rcv)))
    ub, cb = int(p1lib.p1_cot_u_bytes(ctx, n)), int(p1lib.p1_cot_c_bytes(ctx, n))
    u = dev_zeros(p1lib, ctx, ub); c = dev_zeros(p1lib, ctx, cb)
    acc = dev_zeros(p1lib, ctx, 16)
    ss = oracle.lib.p1o_ot_session_new(po.OT_KEY, 2, 4)
    want_s = want_r = 0
    mask = (1 << bits) - 1
    for rep in range(2):  # session cursor advances
        p1lib.check(p1lib.p1_cot_receiver_extend(ctx, rcv, data, 1, n, 0, u))
        p1lib.check(p1lib.p1_cot_sender_step(ctx, snd, data, 0, n, 0, u, c, acc, 0))
        p1lib.check(p1lib.p1_cot_receiver_finish(ctx, rcv, c, acc, 1))
        os_, or_, uo, co = oracle.ot_inner_product(ss, Xw[:, 0], Xw[:, 1], bits, want_wire=True)
        assert np.array_equal(to_host(p1lib, ctx, u, ub, np.uint8), uo)
        assert np.array_equal(to_host(p1lib, ctx, c, cb, np.uint8)[:co.size], co)
        want_s = (want_s + os_) & mask; want_r = (want_r + or_) & mask
        got = to_host(p1lib, ctx, acc, 2, np.uint64)
        assert int(got[0]) & mask == want_s and int(got[1]) & mask == want_r
    assert (want_s + want_r) & mask == (2 * oracle.dot(Xw[:, 0], Xw[:, 1], bits)) & mask
    oracle.lib.p1o_ot_session_free(ss)
    p1lib.p1_cot_free(snd); p1lib.p1_cot_free(rcv)
    p1lib.p1_data_free(data)
    p1lib.p1_ctx_destroy(ctx)


@pytest.mark.parametrize("bits", [64, 32])
def test_cot_row_shards_and_long_session(p1lib, oracle, bits):
    """OT inner products over row shards with row0 > 0 (as the ranks of a row-sharded box hold them) on a
    session whose PRG block cursor and OT counter cross 2^32 inside the inner product: every shard's U / C
    bytes are the slices of the oracle's wire messages and the shard sums add up to the oracle's shares;
    a second session built from the same seeds and moved with p1_cot_advance lands on the same words"""
    from linreg_mpc_b200 import phase1
    ctx = mkctx(p1lib, bits, bits)
    rng = np.random.default_rng(12)
    dt = po.word_dtype(bits)
    w = bits // 8
    n = 3000
    per_tile = 128 // bits
    Xw = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64).astype(dt) * dt(5)
    cuts = [0, 1024, 1024 + 768, n]            # multiples of 256 rows, the last shard ragged
    shards = [from_words(p1lib, ctx, np.ascontiguousarray(Xw[a:b]), None) for a, b in zip(cuts[:-1], cuts[1:])]
    s0, s1, delta = phase1.derive_cot_session(po.OT_KEY, 3, 2)
    chosen = np.ascontiguousarray(np.where(phase1._bits128(delta)[:, None], s1, s0))
    tiles_total = (n * bits + 127) // 128
    pos0, base0 = 2 ** 32 - tiles_total // 2, 2 ** 32 - 777      # both cross 2^32 inside the first inner product
    ss = oracle.lib.p1o_ot_session_new(po.OT_KEY, 3, 2)
    oracle.lib.p1o_ot_session_set_cursor(ss, pos0, base0)
    mask = (1 << bits) - 1

    def sessions(skip):
        snd, rcv = c_vp(), c_vp()
        p1lib.check(p1lib.p1_cot_sender_new(ctx, delta.tobytes(), vp(chosen), ctypes.byref(snd)))
        p1lib.check(p1lib.p1_cot_receiver_new(ctx, vp(s0), vp(s1), ctypes.byref(rcv)))
        for s_ in (snd, rcv):
            p1lib.check(p1lib.p1_cot_set_cursor(s_, pos0, base0))
            p1lib.check(p1lib.p1_cot_advance(s_, n, skip))
        return snd, rcv

    per_shard = [sessions(0) for _ in shards]      # every rank has its own session objects, all in step
    acc = dev_zeros(p1lib, ctx, 8 * 4)
    want = []
    for rep in range(2):
        os_, or_, uo, co = oracle.ot_inner_product(ss, Xw[:, 0], Xw[:, 1], bits, want_wire=True)
        want.append((os_, or_))
        p1lib.check(p1lib.p1_dev_zero(ctx, acc, 32))
        for (a, b), data, (snd, rcv) in zip(zip(cuts[:-1], cuts[1:]), shards, per_shard):
            rows = b - a
            ub, cb = int(p1lib.p1_cot_u_bytes(ctx, rows)), int(p1lib.p1_cot_c_bytes(ctx, rows))
            u = dev_zeros(p1lib, ctx, ub); c = dev_zeros(p1lib, ctx, cb)
            p1lib.check(p1lib.p1_cot_receiver_extend(ctx, rcv, data, 1, n, a, u))
            p1lib.check(p1lib.p1_cot_sender_step(ctx, snd, data, 0, n, a, u, c, acc, 0))
            p1lib.check(p1lib.p1_cot_receiver_finish(ctx, rcv, c, acc, 1))
            t0 = a // per_tile
            assert np.array_equal(to_host(p1lib, ctx, u, ub, np.uint8), uo[2048 * t0:2048 * t0 + ub])
            assert np.array_equal(to_host(p1lib, ctx, c, cb, np.uint8)[:rows * bits * w], co[a * bits * w:b * bits * w])
            p1lib.check(p1lib.p1_dev_free(ctx, u)); p1lib.check(p1lib.p1_dev_free(ctx, c))
        got = to_host(p1lib, ctx, acc, 2, np.uint64)
        assert int(got[0]) & mask == os_ and int(got[1]) & mask == or_, rep
    # a session that skipped the first inner product computes the second one identically
    snd, rcv = sessions(1)
    full = from_words(p1lib, ctx, Xw, None)
    ub, cb = int(p1lib.p1_cot_u_bytes(ctx, n)), int(p1lib.p1_cot_c_bytes(ctx, n))
    u = dev_zeros(p1lib, ctx, ub); c = dev_zeros(p1lib, ctx, cb)
    p1lib.check(p1lib.p1_dev_zero(ctx, acc, 32))
    p1lib.check(p1lib.p1_cot_receiver_extend(ctx, rcv, full, 1, n, 0, u))
    p1lib.check(p1lib.p1_cot_sender_step(ctx, snd, full, 0, n, 0, u, c, acc, 0))
    p1lib.check(p1lib.p1_cot_receiver_finish(ctx, rcv, c, acc, 1))
    got = to_host(p1lib, ctx, acc, 2, np.uint64)
    assert (int(got[0]) & mask, int(got[1]) & mask) == want[1]
    for s_ in [snd, rcv] + [x for pr in per_shard for x in pr]:
        p1lib.p1_cot_free(s_)
    oracle.lib.p1o_ot_session_free(ss)
    for dd in shards + [full]:
        p1lib.p1_data_free(dd)
    p1lib.p1_ctx_destroy(ctx)


# ---------------------------------------------------------------------------
def _box(p1lib, c, **kw):
    from linreg_mpc_b200 import phase1
    return phase1.Phase1Box(p1lib, c["n"], c["d"], c["col_start"], c["precision"], c["precision_p2"],
                            c["bits_p1"], c["bits_p2"], **kw)


def test_whole_phase1_matches_reference_golden(p1lib, golden):
    """every share word of every party equals the reference's own phase1.c output
    (README example + synthetic instances; masking and OT; 64/64, 32/32, 64/32)"""
    for c in golden["cases"]:
        box = _box(p1lib, c, batch_pairs=3)
        if c["use_ot"]:
            got = box.run_ot(c["X"], c["y"], golden["ot_key"])
        else:
            got = box.run_masking(c["X"], c["y"], golden["ti_key"])               # one-kernel masking loop
            assert np.array_equal(got, c["shares"]), ("masking", c["instance"], c["bits_p1"], c["bits_p2"])
            got = box.run_masking(c["X"], c["y"], golden["ti_key"], fused=False)  # TI / party kernels one by one
            assert np.array_equal(got, c["shares"]), ("masking, split kernels", c["instance"], c["bits_p1"], c["bits_p2"])
            got = box.run_masking(c["X"], c["y"], golden["ti_key"], seeded=True)  # CSP-seeded masks: same words
        assert np.array_equal(got, c["shares"]), (c["instance"], c["bits_p1"], c["bits_p2"], c["use_ot"])
        box.close()


@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
def test_whole_phase1_random_instance(p1lib, oracle, b1, b2, prec, prec2):
    n, d, cs = 5000, 24, [0, 7, 15]
    X, y, _ = po.generate_lin_regression(n, d, seed=21)
    want = oracle.run(X, y, cs, prec, prec2, b1, b2, False, threads=8)
    from linreg_mpc_b200 import phase1
    box = phase1.Phase1Box(p1lib, n, d, cs, prec, prec2, b1, b2, batch_pairs=16)
    got = box.run_masking(X, y, po.TI_KEY)
    assert np.array_equal(got, want)
    assert box.launches() > 0
    assert np.array_equal(box.run_masking(X, y, po.TI_KEY, fused=False), want)
    # same run with every message serialised and parsed on the GPU, and with provider-expanded masks
    assert np.array_equal(box.run_masking(X, y, po.TI_KEY, wire=True), want)
    assert np.array_equal(box.run_masking(X, y, po.TI_KEY, seeded=True), want)
    box.close()


def test_row_shards_combine_exactly(p1lib, oracle):
    """two row shards (as two ranks would hold them) + wrapping u64 sum == unsharded run"""
    from linreg_mpc_b200 import phase1
    n, d, cs = 3000, 10, [0, 4, 7]
    X, y, _ = po.generate_lin_regression(n, d, seed=31)
    want = oracle.run(X, y, cs, 56, threads=4)
    cut = 1280
    boxes = [phase1.Phase1Box(p1lib, n, d, cs, 56, row0=0, n_rows=cut, batch_pairs=8),
             phase1.Phase1Box(p1lib, n, d, cs, 56, row0=cut, n_rows=n - cut, batch_pairs=8)]
    for b, (lo, hi) in zip(boxes, [(0, cut), (cut, n)]):
        b.reset()
        b.encode(np.ascontiguousarray(X[lo:hi]), np.ascontiguousarray(y[lo:hi]))
        b.local_blocks()
        b.cross_pairs_masking(po.TI_KEY)
        b.sync()
    boxes[0].flat += boxes[1].flat  # what the all-reduce does
    boxes[0].torch.cuda.synchronize()
    boxes[0].apply_values()
    got = boxes[0].finalize()
    assert np.array_equal(got, want)
    for b in boxes:
        b.close()


def test_pool_blocks_come_back_zeroed(p1lib):
    """a block that p1_dev_free keeps for the next request of its size is zeroed first (include/p1gpu.h, P1GPU_POOL_SCRUB
    default 1): what one owner kept in it - masks, keys, messages - is not handed to the next"""
    ctx = mkctx(p1lib)
    nbytes = 3 * 4096 + 16          # a size nothing else in the suite asks for
    junk = np.full(nbytes, 0xA5, np.uint8)
    p = dev_array(p1lib, ctx, junk)
    assert to_host(p1lib, ctx, p, nbytes, np.uint8).all()
    first = p.value
    p1lib.check(p1lib.p1_dev_free(ctx, p))
    q = c_vp()
    p1lib.check(p1lib.p1_dev_alloc(ctx, nbytes, ctypes.byref(q)))
    assert q.value == first, "the pool did not hand the kept block back"
    assert not to_host(p1lib, ctx, q, nbytes, np.uint8).any()
    p1lib.check(p1lib.p1_dev_free(ctx, q))
    p1lib.p1_ctx_destroy(ctx)
