"""GPU suite (-m gpu): row sharding behind the C-ABI (p1_group_*, include/p1gpu.h): every group call
against the single-device call / the oracle, bit-exact.  Device lists: (0, 0) - two contexts and two row
slabs on ONE GPU, so the sharding logic runs on the one-GPU box of the regular suite - and (0, 1),
(0, 1, 2, 3) when the box has the devices (gpurun --gpus N)."""
import ctypes

import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t


def vp(a):
    return a.ctypes.data_as(c_vp) if a is not None else None


def _device_lists():
    import torch
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    out = [(0, 0), (0, 0, 0)]
    if n >= 2:
        out.append((0, 1))
    if n >= 4:
        out.append((0, 1, 2, 3))
    if n >= 8:
        out.append(tuple(range(8)))
    return out


def mkgroup(lib, devs, b1=64, b2=64, nccl=False):
    grp = c_vp()
    arr = (ctypes.c_int * len(devs))(*devs)
    lib.check(lib.p1_group_create(arr, len(devs), b1, b2, ctypes.byref(grp)))
    if nccl:
        lib.check(lib.p1_group_set_option(grp, b"reduce", 1))
    return grp


def d2h(lib, ctx, p, shape, dtype):
    out = np.empty(shape, dtype)
    lib.check(lib.p1_copy_d2h(ctx, vp(out), p, out.nbytes))
    return out


def group_acc_sum(lib, grp, d):
    """sum of the devices' accumulators after p1_group_reduce (device 0 holds it)"""
    lib.check(lib.p1_group_reduce(grp))
    L = int(lib.p1_acc_len(d))
    return d2h(lib, lib.p1_group_ctx(grp, 0), lib.p1_group_acc_dev(grp, 0), L, np.uint64)


@pytest.mark.parametrize("devs", _device_lists())
@pytest.mark.parametrize("b1,b2,prec,prec2", [(64, 64, 56, 0), (32, 32, 24, 0), (64, 32, 56, 24)])
def test_group_encode_local_block_finalize(p1lib, oracle, devs, b1, b2, prec, prec2):
    """p1_group_encode + p1_group_local_block + p1_group_finalize == the oracle's one-provider phase 1
    (the reference's run_party with a single data provider), from host and from device-resident doubles"""
    lib = p1lib
    n, d = (5000 if len(devs) <= 4 else 20000) + 77, 40
    X, y, _ = po.generate_lin_regression(n, d, seed=91)
    want = oracle.run(X, y, [0], prec, prec2, b1, b2, False, threads=4)[0]
    nz = oracle.normalizer(prec, n)
    nA = d * (d + 1) // 2
    wdt = np.uint64 if b2 == 64 else np.uint32
    import torch
    Xd, yd = torch.from_numpy(X).cuda(), torch.from_numpy(y).cuda()
    for src in ("host", "device"):
        grp = mkgroup(lib, devs, b1, b2)
        gd = c_vp()
        xa, ya = (vp(X), vp(y)) if src == "host" else (c_vp(Xd.data_ptr()), c_vp(yd.data_ptr()))
        lib.check(lib.p1_group_encode(grp, xa, ya, n, d, nz, prec, ctypes.byref(gd)))
        # every slab holds exactly its rows of the single-device encoding
        Xe = oracle.encode(X, nz, prec, b1, b2)
        for g in range(len(devs)):
            r0, rows = c_sz(), c_sz()
            lib.p1_group_rows(grp, n, g, ctypes.byref(r0), ctypes.byref(rows))
            Xw = np.empty((rows.value, d), po.word_dtype(b1))
            lib.check(lib.p1_data_download(lib.p1_group_ctx(grp, g), lib.p1_gdata_shard(gd, g), vp(Xw), None))
            assert np.array_equal(Xw, Xe[r0.value:r0.value + rows.value]), (src, g)
        lib.check(lib.p1_group_acc_reset(grp, d))
        lib.check(lib.p1_group_local_block(grp, gd, 0, d, 1, 0))
        A, B = c_vp(), c_vp()
        lib.check(lib.p1_group_finalize(grp, gd, None, 0, d, prec, prec2, float(d), ctypes.byref(A), ctypes.byref(B)))
        got = np.concatenate([np.ctypeslib.as_array(ctypes.cast(A, ctypes.POINTER(np.ctypeslib.as_ctypes_type(wdt))), (nA,)),
                              np.ctypeslib.as_array(ctypes.cast(B, ctypes.POINTER(np.ctypeslib.as_ctypes_type(wdt))), (d,))]).astype(np.uint64)
        lib.free(A)
        lib.free(B)
        assert np.array_equal(got, want), src
        lib.p1_gdata_free(gd)
        lib.p1_group_destroy(grp)


@pytest.mark.parametrize("devs", _device_lists())
@pytest.mark.parametrize("nccl", [False, True])
def test_group_reduce_twice_counts_every_partial_once(p1lib, devs, nccl):
    """reduce -> more pairs -> reduce == one reduce at the end (include/p1gpu.h, p1_group_reduce): the
    peers' partial sums are cleared once they have been added, in both reduce modes; a reduce with nothing
    new in between changes nothing"""
    lib = p1lib
    if nccl and len(set(devs)) != len(devs):
        pytest.skip("NCCL needs distinct devices")
    n, d = 4000 + 13, 12
    X, y, _ = po.generate_lin_regression(n, d, seed=33)
    nz = float(np.sqrt(2.0 ** 56 * n))
    res = []
    for split in (False, True):
        grp = mkgroup(lib, devs, 64, 64, nccl)
        gd = c_vp()
        lib.check(lib.p1_group_encode(grp, vp(X), vp(y), n, d, nz, 56, ctypes.byref(gd)))
        lib.check(lib.p1_group_acc_reset(grp, d))
        lib.check(lib.p1_group_local_block(grp, gd, 0, 5, 0, 0))
        if split:
            first = group_acc_sum(lib, grp, d).copy()
            assert np.array_equal(group_acc_sum(lib, grp, d), first)     # nothing new: unchanged
            for g in range(1, len(devs)):                                 # the peers start again from zero
                L = int(lib.p1_acc_len(d))
                assert not d2h(lib, lib.p1_group_ctx(grp, g), lib.p1_group_acc_dev(grp, g), L, np.uint64).any()
        lib.check(lib.p1_group_local_block(grp, gd, 5, d, 1, 0))
        res.append(group_acc_sum(lib, grp, d).copy())
        lib.p1_gdata_free(gd)
        lib.p1_group_destroy(grp)
    assert res[0].any()
    assert np.array_equal(res[0], res[1])


@pytest.mark.parametrize("devs", _device_lists())
def test_group_sequential_diagonal_chain(p1lib, oracle, devs):
    """words >= 2^26 defeat the exact diagonal shortcut: the literal double loop (src/phase1.c:563-566) is
    chained over the slabs in device order and must equal the oracle's single sequential pass"""
    lib = p1lib
    n, d, prec = (3000 if len(devs) <= 4 else 6000), 6, 56
    rng = np.random.default_rng(5)
    X = rng.standard_normal((n, d)) * 40.0        # |word| ~ 40 * 2^28 / sqrt(3000) >> 2^26
    y = rng.standard_normal(n)
    want = oracle.run(X, y, [0], prec, 0, 64, 64, False, threads=2)[0]
    nz = oracle.normalizer(prec, n)
    grp = mkgroup(lib, devs)
    gd = c_vp()
    lib.check(lib.p1_group_encode(grp, vp(X), vp(y), n, d, nz, prec, ctypes.byref(gd)))
    lib.check(lib.p1_group_acc_reset(grp, d))
    lib.check(lib.p1_group_local_block(grp, gd, 0, d, 1, 0))
    A, B = c_vp(), c_vp()
    lib.check(lib.p1_group_finalize(grp, gd, None, 0, d, prec, 0, float(d), ctypes.byref(A), ctypes.byref(B)))
    nA = d * (d + 1) // 2
    got = np.ctypeslib.as_array(ctypes.cast(A, ctypes.POINTER(ctypes.c_uint64)), (nA,)).copy()
    lib.free(A)
    lib.free(B)
    assert np.array_equal(got, want[:nA])
    lib.p1_gdata_free(gd)
    lib.p1_group_destroy(grp)


@pytest.mark.parametrize("devs", _device_lists())
@pytest.mark.parametrize("bits", [64, 32])
@pytest.mark.parametrize("nccl", [False, True])
def test_group_masking_calls(p1lib, oracle, devs, bits, nccl):
    """TI generation, both parties' steps and the ring reduction over the group == oracle algebra on whole
    vectors; masks and messages gathered from / scattered to device 0 (where the wire format lives)"""
    lib = p1lib
    if nccl and len(set(devs)) != len(devs):
        pytest.skip("NCCL needs distinct devices")
    n, d, count, q0 = (2600 if len(devs) <= 4 else 8 * 512 + 40), 5, 4, 3
    rng = np.random.default_rng(17)
    dt = po.word_dtype(bits)
    w = bits // 8
    colv = lambda c: yw if c == d else Xw[:, c]
    grp = mkgroup(lib, devs, bits, bits, nccl)
    ctx0 = lib.p1_group_ctx(grp, 0)
    # small integers as doubles with precision 0 and normalizer 1: the ring word is the integer
    Xi = rng.integers(-2 ** 20, 2 ** 20, (n, d)).astype(np.float64)
    yi = rng.integers(-2 ** 20, 2 ** 20, n).astype(np.float64)
    gd = c_vp()
    lib.check(lib.p1_group_encode(grp, vp(Xi), vp(yi), n, d, 1.0, 0, ctypes.byref(gd)))
    Xw = Xi.astype(np.int64).astype(dt)
    yw = yi.astype(np.int64).astype(dt)
    ld0 = (n + 255) // 256 * 256
    gv = [c_vp() for _ in range(4)]
    for v in gv:
        lib.check(lib.p1_gvec_alloc(grp, n, count, ctypes.byref(v)))
    x, y, mb, ma = gv
    full = [c_vp() for _ in range(4)]
    for f in full:
        lib.check(lib.p1_dev_alloc(ctx0, count * ld0 * w, ctypes.byref(f)))
        lib.check(lib.p1_dev_zero(ctx0, f, count * ld0 * w))
    xy, r = c_vp(), c_vp()
    lib.check(lib.p1_dev_alloc(ctx0, 8 * count, ctypes.byref(xy)))
    lib.check(lib.p1_dev_alloc(ctx0, 8 * count, ctypes.byref(r)))
    lib.check(lib.p1_group_acc_reset(grp, d))
    lib.check(lib.p1_group_ti_generate(grp, po.TI_KEY, q0, count, n, x, y, xy, r))
    cols_a = np.array([5, 4, 3, 5], np.int32)
    cols_b = np.array([0, 1, 0, 2], np.int32)
    slots_a = np.array([0, 1, 2, 1], np.int64)
    slots_b = np.array([8, 9, 10, 9], np.int64)
    lib.check(lib.p1_group_party_b_mask(grp, gd, vp(cols_b), count, x, mb))
    lib.check(lib.p1_group_party_a_step(grp, gd, vp(cols_a), vp(slots_a), count, y, mb, ma))
    lib.check(lib.p1_group_party_b_share(grp, gd, vp(cols_b), vp(slots_b), count, ma))
    for v, f in zip(gv, full):
        lib.check(lib.p1_gvec_gather(grp, v, count, f, ld0))
    xh, yh, mbh, mah = (d2h(lib, ctx0, f, (count, ld0), dt) for f in full)
    xyh, rh = d2h(lib, ctx0, xy, count, np.uint64), d2h(lib, ctx0, r, count, np.uint64)
    mask = np.uint64((1 << bits) - 1)
    want = np.zeros(int(lib.p1_acc_len(d)), np.uint64)
    with np.errstate(over="ignore"):
        for p in range(count):
            xo, yo, va, vb = oracle.ti_pair(po.TI_KEY, q0 + p, n, bits)
            a, b = colv(cols_a[p]), colv(cols_b[p])
            assert np.array_equal(xh[p, :n], xo) and np.array_equal(yh[p, :n], yo), p
            assert np.array_equal(mbh[p, :n], b + xo) and np.array_equal(mah[p, :n], a - yo), p
            assert int(rh[p]) == vb and (int(xyh[p]) - int(rh[p])) & int(mask) == va
            want[slots_a[p]] += np.uint64(oracle.dot(b + xo, yo, bits))
            want[slots_b[p]] += np.uint64(oracle.dot(a - yo, b, bits))
    got = group_acc_sum(lib, grp, d)
    nA = d * (d + 1) // 2
    assert np.array_equal(got[:nA + d] & mask, want[:nA + d] & mask)
    # scatter is the inverse of gather
    lib.check(lib.p1_gvec_scatter(grp, full[0], ld0, count, gv[3]))
    lib.check(lib.p1_dev_zero(ctx0, full[3], count * ld0 * w))
    lib.check(lib.p1_gvec_gather(grp, gv[3], count, full[3], ld0))
    assert np.array_equal(d2h(lib, ctx0, full[3], (count, ld0), dt)[:, :n], xh[:, :n])
    for f in full + [xy, r]:
        lib.check(lib.p1_dev_free(ctx0, f))
    for v in gv:
        lib.p1_gvec_free(v)
    lib.p1_gdata_free(gd)
    lib.p1_group_destroy(grp)


@pytest.mark.parametrize("devs", _device_lists())
@pytest.mark.parametrize("bits", [64, 32])
def test_group_cot_inner_products(p1lib, oracle, devs, bits):
    """--use_ot over the group: the slices of U and C the devices write into the host buffers are the
    oracle's wire messages, the shares add up over the devices"""
    from linreg_mpc_b200 import phase1
    lib = p1lib
    n, d = (2600 if len(devs) <= 4 else 8 * 512 + 40), 2
    rng = np.random.default_rng(23)
    dt = po.word_dtype(bits)
    Xi = rng.integers(-2 ** 30, 2 ** 30, (n, d)).astype(np.float64)
    Xw = Xi.astype(np.int64).astype(dt)
    grp = mkgroup(lib, devs, bits, bits)
    ctx0 = lib.p1_group_ctx(grp, 0)
    gd = c_vp()
    lib.check(lib.p1_group_encode(grp, vp(Xi), None, n, d, 1.0, 0, ctypes.byref(gd)))
    s0, s1, delta = phase1.derive_cot_session(po.OT_KEY, 2, 4)
    chosen = np.ascontiguousarray(np.where(phase1._bits128(delta)[:, None], s1, s0))
    snd, rcv = c_vp(), c_vp()
    lib.check(lib.p1_group_cot_sender_new(grp, delta.tobytes(), vp(chosen), ctypes.byref(snd)))
    lib.check(lib.p1_group_cot_receiver_new(grp, vp(s0), vp(s1), ctypes.byref(rcv)))
    ub, cb = int(lib.p1_cot_u_bytes(ctx0, n)), int(lib.p1_cot_c_bytes(ctx0, n))
    uh, ch = c_vp(), c_vp()
    lib.check(lib.p1_host_alloc(ctx0, ub, ctypes.byref(uh)))
    lib.check(lib.p1_host_alloc(ctx0, cb, ctypes.byref(ch)))
    u_np = np.ctypeslib.as_array(ctypes.cast(uh, ctypes.POINTER(ctypes.c_uint8)), (ub,))
    c_np = np.ctypeslib.as_array(ctypes.cast(ch, ctypes.POINTER(ctypes.c_uint8)), (cb,))
    ss = oracle.lib.p1o_ot_session_new(po.OT_KEY, 2, 4)
    lib.check(lib.p1_group_acc_reset(grp, d))
    mask = (1 << bits) - 1
    ws = wr = 0
    for rep in range(2):
        lib.check(lib.p1_group_cot_receiver_extend(grp, rcv, gd, 1, uh))
        lib.check(lib.p1_group_cot_sender_step(grp, snd, gd, 0, uh, ch, 0))
        lib.check(lib.p1_group_cot_receiver_finish(grp, rcv, gd, ch, 1))
        os_, or_, uo, co = oracle.ot_inner_product(ss, Xw[:, 0], Xw[:, 1], bits, want_wire=True)
        assert np.array_equal(u_np, uo), rep
        assert np.array_equal(c_np[:co.size], co), rep
        ws, wr = (ws + os_) & mask, (wr + or_) & mask
    got = group_acc_sum(lib, grp, d)
    assert int(got[0]) & mask == ws and int(got[1]) & mask == wr
    oracle.lib.p1o_ot_session_free(ss)
    lib.p1_group_cot_free(snd)
    lib.p1_group_cot_free(rcv)
    lib.check(lib.p1_host_free(ctx0, uh))
    lib.check(lib.p1_host_free(ctx0, ch))
    lib.p1_gdata_free(gd)
    lib.p1_group_destroy(grp)
