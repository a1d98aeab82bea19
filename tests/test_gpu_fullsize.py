"""GPU suite (-m gpu), BASELINE.json's full sizes: size-independent properties of the
domain instead of a word-for-word oracle run (the CPU oracle would need hours there).

  * share algebra: for every off-diagonal index and every share_b entry, the
    sum over ALL providers' shares of the masking run equals the plain ring Gram of the
    encoded data - the masks cancel exactly mod 2^64.  The plain Gram comes from an
    independent path (one provider, tcgen05 kernel over the whole matrix).
  * sampled pairs and all diagonals of that Gram against the CPU oracle on columns
    downloaded from the device (a second, independent check of the tensor-core kernel).
  * the seeded-mask variant and a different batch size give the same share words.
"""
import ctypes

import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
c_vp = ctypes.c_void_p


def _column(lib, ctx, data, c, n, bits=64):
    ld = ctypes.c_size_t()
    base = lib.p1_data_colmajor_dev(data, ctypes.byref(ld))
    w = bits // 8
    out = np.empty(n, dtype=po.word_dtype(bits))
    lib.check(lib.p1_copy_d2h(ctx, out.ctypes.data_as(c_vp), c_vp(base + c * ld.value * w), n * w))
    return out


@pytest.mark.parametrize("n,d,p", [(1_000_000, 500, 5)])
def test_cfg3_full_size_share_algebra(p1lib, oracle, n, d, p):
    import torch
    from linreg_mpc_b200 import phase1
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(20170403)
    X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
    X /= X.abs().amax(dim=0)
    y = X @ torch.rand(d, dtype=torch.float64, device=dev, generator=g) + 0.1 * torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    torch.cuda.synchronize()
    cs = [k * (d // p) for k in range(p)]
    nA = d * (d + 1) // 2

    # masking protocol, 5 providers
    box = phase1.Phase1Box(p1lib, n, d, cs, 56, batch_pairs=32)
    shares = box.run_masking(X, y, po.TI_KEY)
    assert shares.shape == (p, nA + d)
    total = np.zeros(nA + d, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k in range(p):
            total += shares[k]

    # independent path: one provider owns everything (no masks, tcgen05 Gram over 501 columns)
    one = phase1.Phase1Box(p1lib, n, d, [0], 56, gram_algo=phase1.ffi.P1_GRAM_TCGEN05)
    plain = one.run_masking(X, y, po.TI_KEY)[0]
    assert np.array_equal(total, plain), "sum of the providers' shares differs from the plain ring Gram"

    # the diagonal words sit in exactly one provider (the owner), zero elsewhere (calloc, src/phase1.c:482-483)
    diag_idx = np.array([i * (i + 1) // 2 + i for i in range(d)])
    owners = phase1.owners(np.arange(d), cs)
    for k in range(p):
        others = diag_idx[owners != k]
        assert not shares[k][others].any()

    # sampled words of the plain Gram against the CPU oracle on columns read back from the device
    ctx = one.ctxs[0]
    rng = np.random.default_rng(3)
    cols = {}
    col = lambda c: cols.setdefault(c, _column(p1lib, ctx, one.data, c, n))
    for _ in range(24):
        i, j = sorted(rng.choice(d, size=2, replace=False))[::-1]
        assert int(plain[i * (i + 1) // 2 + j]) == oracle.dot(col(int(i)), col(int(j)))
    for j in rng.choice(d, size=6, replace=False):
        assert int(plain[nA + int(j)]) == oracle.dot(col(d), col(int(j)))          # share_b
    for i in rng.choice(d, size=6, replace=False):
        assert int(plain[diag_idx[int(i)]]) == oracle.diag(col(int(i)), 56, float(d))   # sequential double loop
    one.close()

    # seeded masks and another batching: identical share words
    box2 = phase1.Phase1Box(p1lib, n, d, cs, 56, batch_pairs=19, pipes=1)
    again = box2.run_masking(X, y, po.TI_KEY, seeded=True)
    assert np.array_equal(again, shares)
    box2.close()
    box.close()


def test_cfg4_32bit_ring_full_rows(p1lib, oracle):
    """n = 10M rows in the 32-bit ring (config 4, one provider): tcgen05 vs the CUDA-core kernel on
    every word, sampled words vs the oracle."""
    import torch
    from linreg_mpc_b200 import phase1
    n, d = 10_000_000, 64
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(20170404)
    X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
    X /= X.abs().amax(dim=0)
    y = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    torch.cuda.synchronize()
    res = []
    for algo in (phase1.ffi.P1_GRAM_TCGEN05, phase1.ffi.P1_GRAM_IMAD):
        box = phase1.Phase1Box(p1lib, n, d, [0], 30, 0, 32, 32, gram_algo=algo)
        res.append(box.run_masking(X, y, po.TI_KEY)[0])
        if algo == phase1.ffi.P1_GRAM_IMAD:
            c3, c9 = _column(p1lib, box.ctxs[0], box.data, 3, n, 32), _column(p1lib, box.ctxs[0], box.data, 9, n, 32)
            assert int(res[-1][9 * 10 // 2 + 3]) == oracle.dot(c9, c3, 32)
        box.close()
    assert np.array_equal(res[0], res[1])


# ---------------------------------------------------------------------------
# BASELINE.json configs as written (parity holes at BASELINE sizes)

def _synthetic(n, d, seed, dev):
    """experiments/generate_tests.py:159-169 on the device: X ~ N(0,1) scaled to max |.| = 1 per column,
    beta ~ U(0,1), y = X beta + N(0, 0.1^2)"""
    import torch
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
    X /= X.abs().amax(dim=0)
    y = X @ torch.rand(d, dtype=torch.float64, device=dev, generator=g) + 0.1 * torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    torch.cuda.synchronize()
    return X, y


def _reference_shares(oracle, X, y, cs, prec, prec2=0, b1=64, b2=64, use_ot=False):
    """the reference's own phase1.c (oracle/_ref) on the instance; the C port when _ref is not built"""
    import os
    import tempfile
    d = X.shape[1]
    if po.RefLib.available(b1, b2):
        path = tempfile.mktemp(prefix="p1full_", suffix=".in")
        oracle.write_instance(path, X, y, cs)
        try:
            return po.RefLib(b1, b2).run(path, d, len(cs), prec, prec2, use_ot)[0], "reference"
        finally:
            os.unlink(path)
    return oracle.run(X, y, cs, prec, prec2, b1, b2, use_ot, threads=os.cpu_count() or 1), "port"


def test_cfg2_word_for_word_vs_reference(p1lib, oracle):
    """BASELINE config 2 (n = 100k, d = 100, 3 providers, 64-bit, masking): every one of the 3 x 5150
    share words against the reference's phase1.c compiled unmodified ("every word for cfg1-2"),
    through the fused masking loop and through the split TI / party kernels; blocks of 33-34 columns take
    the tcgen05 Gram"""
    from linreg_mpc_b200 import phase1
    n, d, p = 100_000, 100, 3
    X, y, _ = po.generate_lin_regression(n, d, seed=20170402)
    cs = po.even_col_start(d, p)
    want, kind = _reference_shares(oracle, X, y, cs, 56)
    box = phase1.Phase1Box(p1lib, n, d, cs, 56, batch_pairs=64)
    p1lib.check(p1lib.p1_ctx_profile(box.ctxs[0], 1))
    got = box.run_masking(X, y, po.TI_KEY)
    ms, launches = ctypes.c_double(), ctypes.c_uint64()
    p1lib.check(p1lib.p1_ctx_profile_read(box.ctxs[0], phase1.ffi.P1_K_GRAM_TC, ctypes.byref(ms), ctypes.byref(launches)))
    assert launches.value == p, "the local blocks did not run on the tensor cores"
    assert got.shape == (p, d * (d + 1) // 2 + d)
    assert np.array_equal(got, want), "fused masking loop vs %s" % kind
    assert np.array_equal(box.run_masking(X, y, po.TI_KEY, fused=False), want), "split kernels vs %s" % kind
    assert np.array_equal(box.run_masking(X, y, po.TI_KEY, seeded=True), want), "seeded masks vs %s" % kind
    box.close()


def test_use_ot_100k_word_for_word_vs_reference(p1lib, oracle):
    """--use_ot at n = 100k, d = 12, 3 providers (3.07e8 correlated OTs, every session role of
    src/phase1.c:392, both target passes :415-438) word for word against the reference's OT branch, on one
    pipe and on two (p1_cot_advance), and as two row shards with row0 > 0"""
    from linreg_mpc_b200 import phase1
    n, d, p = 100_000, 12, 3
    X, y, _ = po.generate_lin_regression(n, d, seed=20170405)
    cs = po.even_col_start(d, p)
    want, kind = _reference_shares(oracle, X, y, cs, 56, use_ot=True)
    for pipes in (1, 2):
        box = phase1.Phase1Box(p1lib, n, d, cs, 56, pipes=pipes)
        got = box.run_ot(X, y, po.OT_KEY)
        box.close()
        assert np.array_equal(got, want), "use_ot, %d pipe(s), vs %s" % (pipes, kind)
    cut = 40_192      # a multiple of 256 rows
    boxes = [phase1.Phase1Box(p1lib, n, d, cs, 56, row0=0, n_rows=cut), phase1.Phase1Box(p1lib, n, d, cs, 56, row0=cut, n_rows=n - cut)]
    for b, (lo, hi) in zip(boxes, [(0, cut), (cut, n)]):
        b.reset()
        b.encode(np.ascontiguousarray(X[lo:hi]), np.ascontiguousarray(y[lo:hi]))
        b.local_blocks()
        b.cross_pairs_ot(po.OT_KEY)
        b.sync()
    boxes[0].flat += boxes[1].flat      # the ring all-reduce of two ranks
    boxes[0].torch.cuda.synchronize()
    assert np.array_equal(boxes[0].finalize(), want), "use_ot over two row shards vs %s" % kind
    for b in boxes:
        b.close()


@pytest.mark.parametrize("b1,b2,prec,prec2", [(64, 64, 56, 0), (32, 32, 24, 0), (64, 32, 56, 24)])
def test_whole_protocol_on_tensor_cores_vs_reference(p1lib, oracle, b1, b2, prec, prec2):
    """a whole masking run whose local blocks are wide and long enough for P1_GRAM_AUTO to take the tcgen05
    kernel (20 columns >= 16, n >= 4096), every word against the reference build"""
    from linreg_mpc_b200 import phase1
    n, d, p = 8192 + 77, 60, 3
    X, y, _ = po.generate_lin_regression(n, d, seed=77)
    cs = po.even_col_start(d, p)
    want, kind = _reference_shares(oracle, X, y, cs, prec, prec2, b1, b2)
    box = phase1.Phase1Box(p1lib, n, d, cs, prec, prec2, b1, b2, batch_pairs=48)
    p1lib.check(p1lib.p1_ctx_profile(box.ctxs[0], 1))
    got = box.run_masking(X, y, po.TI_KEY)
    ms, launches = ctypes.c_double(), ctypes.c_uint64()
    p1lib.check(p1lib.p1_ctx_profile_read(box.ctxs[0], phase1.ffi.P1_K_GRAM_TC, ctypes.byref(ms), ctypes.byref(launches)))
    box.close()
    assert launches.value == p
    assert np.array_equal(got, want), kind


def test_cfg5_as_written(p1lib, oracle):
    """BASELINE config 5 as written: n = 1M, d = 1000, 20 providers, 64-bit phase 1 / 32-bit phase 2
    (--prec_phase2 = 30): sum invariant on all 501 500 words against the one-provider tcgen05 Gram, sampled
    words against oracle.dot / oracle.diag, the 64 -> 32 shift (src/phase1.c:609-635) on every word"""
    import torch
    from linreg_mpc_b200 import phase1
    from tools.parity_check import check_run
    n, d, p = 1_000_000, 1000, 20
    X, y = _synthetic(n, d, 20170406, torch.device("cuda", 0))
    cs = po.even_col_start(d, p)
    box = phase1.Phase1Box(p1lib, n, d, cs, 56, 30, 64, 32, batch_pairs=32)
    shares = box.run_masking(X, y, po.TI_KEY)
    assert box.sched.num_cross == 475_950
    assert int(shares.max()) < 2 ** 32
    res = check_run(p1lib, oracle, box, shares, X, y, samples=24, diag_samples=6, ti_key=po.TI_KEY, cross_samples=24)
    box.close()
    assert res["ok"] and res["words_checked"] == 500_500 and res["downconvert_all_words"], res
    # both providers' masked words of 24 cross pairs, up to q = 475 949 (110 multiples of 2^32 keystream blocks in)
    assert res["sampled_cross_pairs"] == 24 and res["sampled_cross_pair_words_match"], res
    assert max(res["sampled_cross_pair_2p32_block_epochs"]) == 110, res


def test_cfg4_as_written(p1lib, oracle):
    """BASELINE config 4: n = 10M, d = 200, 32-bit ring (precision 30).  One provider: tcgen05 and the
    CUDA-core kernel agree on all 20 300 words and sampled words match the oracle; five providers (16 160
    cross pairs): sum invariant on all words"""
    import torch
    from linreg_mpc_b200 import phase1
    from tools.parity_check import check_run, column
    n, d = 10_000_000, 200
    X, y = _synthetic(n, d, 20170407, torch.device("cuda", 0))
    res = []
    for algo in (phase1.ffi.P1_GRAM_IMAD, phase1.ffi.P1_GRAM_TCGEN05):
        box = phase1.Phase1Box(p1lib, n, d, [0], 30, 0, 32, 32, gram_algo=algo)
        res.append(box.run_masking(X, y, po.TI_KEY)[0])
        if algo == phase1.ffi.P1_GRAM_IMAD:
            rng = np.random.default_rng(4)
            cols = {}
            col = lambda c: cols.setdefault(int(c), column(p1lib, box.ctxs[0], box.data, int(c), n, 32))
            for _ in range(8):
                i, j = sorted(rng.choice(d, size=2, replace=False))[::-1]
                assert int(res[-1][int(i) * (int(i) + 1) // 2 + int(j)]) == oracle.dot(col(i), col(j), 32)
            assert int(res[-1][d * (d + 1) // 2 + 17]) == oracle.dot(col(d), col(17), 32)
            assert int(res[-1][5 * 6 // 2 + 5]) == oracle.diag(col(5), 30, float(d), 32)
        box.close()
    assert np.array_equal(res[0], res[1])
    box = phase1.Phase1Box(p1lib, n, d, po.even_col_start(d, 5), 30, 0, 32, 32, batch_pairs=24)
    shares = box.run_masking(X, y, po.TI_KEY)
    out = check_run(p1lib, oracle, box, shares, X, y, samples=8, diag_samples=2, ti_key=po.TI_KEY, cross_samples=10)
    box.close()
    assert out["ok"] and out["sampled_cross_pair_words_match"] and out["sampled_cross_pairs"] == 10, out
