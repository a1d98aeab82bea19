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
