#!/usr/bin/env python
"""Isolated timings of single libp1gpu.so calls (tuning aid; bench.py is the judged harness).

  python tools/kbench.py ti     --n 1000000 --pairs 33 [--bits 64] [--opt ti_threads=512]
  python tools/kbench.py fused  --n 1000000 --pairs 128       (the one-kernel masking loop, p1_masking_fused)
  python tools/kbench.py party  --n 1000000 --pairs 33
  python tools/kbench.py probe                                 (dense int8 tcgen05 peak, N = 256 / 192 / 64)
  python tools/kbench.py gram   --n 1000000 --cols 500 --algo tcgen05|imad [--target]
  python tools/kbench.py encode --n 1000000 --cols 500
  python tools/kbench.py cot    --n 65536
  python tools/kbench.py wire   --n 1000000 --pairs 33      (varint pack / unpack of masked vectors)
  python tools/kbench.py text   --n 8000000                  (decimal text -> doubles, n tokens)

Every case prints one JSON line: milliseconds per call (CUDA events on the context's
stream, after warm-up) and the derived rate.  Buffers are far larger than the 126 MB L2
unless --n is made small on purpose.
"""
import argparse
import ctypes
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["ti", "fused", "party", "gram", "encode", "cot", "wire", "text", "probe", "reduce"])
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--pairs", type=int, default=33)
    ap.add_argument("--cols", type=int, default=500)
    ap.add_argument("--bits", type=int, default=64)
    ap.add_argument("--algo", default="tcgen05")
    ap.add_argument("--target", action="store_true")
    ap.add_argument("--precision", type=int, default=56)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--opt", action="append", default=[])
    ap.add_argument("--devices", default="0,1", help="reduce: device list of the group")
    ap.add_argument("--novec", action="store_true", help="ti: values only (no mask vectors stored)")
    args = ap.parse_args()

    import torch
    import linreg_mpc_b200 as pkg
    from linreg_mpc_b200.ffi import c_vp, c_sz, ptr
    lib = pkg.P1Lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    bits2 = 64 if args.bits == 64 else 32
    ctx = c_vp()
    lib.check(lib.p1_ctx_create(0, args.bits, bits2, ctypes.byref(ctx)))
    for o in args.opt:
        k, v = o.split("=")
        lib.check(lib.p1_ctx_set_option(ctx, k.encode(), int(v)))
    stream = torch.cuda.ExternalStream(int(lib.p1_ctx_stream(ctx)), device=dev)
    w = args.bits // 8
    wdt = torch.int64 if w == 8 else torch.int32

    def timeit(fn):
        for _ in range(args.warmup):
            fn()
        lib.check(lib.p1_ctx_sync(ctx))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.reps):
            fn()
        b.record(stream)
        lib.check(lib.p1_ctx_sync(ctx))
        torch.cuda.synchronize()
        return a.elapsed_time(b) / args.reps

    def encoded(n, d, with_y=True):
        g = torch.Generator(device=dev)
        g.manual_seed(1)
        X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
        X /= X.abs().amax(dim=0)
        y = torch.randn(n, dtype=torch.float64, device=dev, generator=g) if with_y else None
        torch.cuda.synchronize()
        out = c_vp()
        norm = math.sqrt(math.pow(2, args.precision) * float(n))
        lib.check(lib.p1_encode(ctx, ptr(X), ptr(y), n, d, norm, args.precision, ctypes.byref(out)))
        lib.check(lib.p1_ctx_sync(ctx))
        return out, X, y, norm

    res = {"what": args.what, "n": args.n, "bits": args.bits, "opts": args.opt}
    n, B = args.n, args.pairs
    ld = (n + 255) // 256 * 256
    if args.what == "ti":
        xb = torch.empty(B * ld, dtype=wdt, device=dev)
        yb = torch.empty(B * ld, dtype=wdt, device=dev)
        xy = torch.zeros(B, dtype=torch.int64, device=dev)
        r = torch.zeros(B, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        key = bytes(range(16))
        xp, yp = (None, None) if args.novec else (ptr(xb), ptr(yb))
        ms = timeit(lambda: lib.check(lib.p1_ti_generate(ctx, key, 0, B, n, 0, n, xp, yp, ld, ptr(xy), ptr(r))))
        ks = B * (2 * n * w + 16)
        res.update(ms=ms, pairs=B, keystream_GBs=ks / ms / 1e6, aes_blocks_per_s=ks / 16 / (ms * 1e-3),
                   clk_per_warp_block=148 * 1.965e9 * (ms * 1e-3) / (ks / 16 / 32))
    elif args.what == "fused":
        data, _, _, _ = encoded(n, 64)
        mb = torch.empty(B * ld, dtype=wdt, device=dev)
        ma = torch.empty(B * ld, dtype=wdt, device=dev)
        xy = torch.zeros(B, dtype=torch.int64, device=dev)
        r = torch.zeros(B, dtype=torch.int64, device=dev)
        acc = torch.zeros(int(lib.p1_acc_len(64)), dtype=torch.int64, device=dev)
        cols_b = torch.arange(B, dtype=torch.int32, device=dev) % 32
        cols_a = torch.full((B,), 40, dtype=torch.int32, device=dev)       # one row i, j sweeps (as a run of the schedule)
        sa = torch.arange(B, dtype=torch.int64, device=dev)
        sb = torch.arange(B, dtype=torch.int64, device=dev) + 300
        torch.cuda.synchronize()
        key = bytes(range(16))
        ms = timeit(lambda: lib.check(lib.p1_masking_fused(ctx, key, 0, B, n, 0, data, ptr(cols_a), ptr(cols_b), ptr(sa), ptr(sb),
                                                           ptr(mb), ptr(ma), ld, None, None, ptr(xy), ptr(r), ptr(acc))))
        ks = B * (2 * n * w + 16)
        res.update(ms=ms, pairs=B, us_per_pair=ms * 1e3 / B, keystream_GBs=ks / ms / 1e6, aes_blocks_per_s=ks / 16 / (ms * 1e-3),
                   hbm_GBs_algorithmic=B * 4 * n * w / ms / 1e6, lds_bound_frac_at_1965MHz=(ks / 16 / (ms * 1e-3)) / (148 * 32 * 1.965e9 / 135))
        lib.p1_data_free(data)
    elif args.what == "reduce":
        # p1_group_reduce (the combination of the row shards' accumulators inside p1_group_finalize): peer reads by
        # device 0 vs ncclAllReduce, for the accumulator of `cols` columns; wall clock per call including the sync,
        # which is what a short step pays
        import time
        devs = [int(x) for x in args.devices.split(",")]
        arr = (ctypes.c_int * len(devs))(*devs)
        res["devices"] = devs
        res["payload_bytes"] = int(lib.p1_acc_len(args.cols)) * 8
        for mode, name in ((0, "p2p"), (1, "nccl")):
            grp = c_vp()
            lib.check(lib.p1_group_create(arr, len(devs), args.bits, bits2, ctypes.byref(grp)))
            lib.check(lib.p1_group_set_option(grp, b"reduce", mode))
            lib.check(lib.p1_group_acc_reset(grp, args.cols))
            # p1_group_reduce is a no-op on accumulators that are already combined: every repetition zeroes them
            # again (p1_group_acc_reset), and the cost of that alone is measured and subtracted
            def loop(reduce, synced):
                t0 = time.perf_counter()
                for _ in range(args.reps):
                    lib.check(lib.p1_group_acc_reset(grp, args.cols))
                    if reduce:
                        lib.check(lib.p1_group_reduce(grp))
                    if synced:
                        lib.check(lib.p1_group_sync(grp))
                lib.check(lib.p1_group_sync(grp))
                return (time.perf_counter() - t0) / args.reps * 1e6
            for _ in range(3):
                loop(True, True)
            base_s, base_b = loop(False, True), loop(False, False)
            res[name + "_us_per_call_synced"] = loop(True, True) - base_s
            res[name + "_us_per_call_back_to_back"] = loop(True, False) - base_b
            res[name + "_reset_only_us"] = {"synced": base_s, "back_to_back": base_b}
            lib.p1_group_destroy(grp)
    elif args.what == "probe":
        for ncols in (256, 192, 128, 64):
            tops, ms = ctypes.c_double(), ctypes.c_double()
            lib.check(lib.p1_probe_tensor_i8(ctx, ncols, 300.0, ctypes.byref(tops), ctypes.byref(ms), None))
            res["n%d_tops" % ncols] = tops.value
    elif args.what == "party":
        data, _, _, _ = encoded(n, 64)
        bufs = [torch.empty(B * ld, dtype=wdt, device=dev).random_() for _ in range(4)]
        acc = torch.zeros(int(lib.p1_acc_len(64)), dtype=torch.int64, device=dev)
        cols_b = torch.arange(B, dtype=torch.int32, device=dev) % 32
        cols_a = torch.full((B,), 40, dtype=torch.int32, device=dev)
        slots = torch.arange(B, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        x, y, mb, ma = (ptr(t) for t in bufs)
        nb = n * w
        ms0 = timeit(lambda: lib.check(lib.p1_party_b_mask(ctx, data, ptr(cols_b), B, x, ld, mb, ld)))
        ms1 = timeit(lambda: lib.check(lib.p1_party_a_step(ctx, data, ptr(cols_a), ptr(slots), B, y, ld, mb, ld, ma, ld, ptr(acc))))
        ms2 = timeit(lambda: lib.check(lib.p1_party_b_share(ctx, data, ptr(cols_b), ptr(slots), B, ma, ld, ptr(acc))))
        res.update(b_mask_ms=ms0, b_mask_GBs=3 * B * nb / ms0 / 1e6, a_step_ms=ms1, a_step_GBs=4 * B * nb / ms1 / 1e6,
                   b_share_ms=ms2, b_share_GBs=2 * B * nb / ms2 / 1e6)
        lib.p1_data_free(data)
    elif args.what == "gram":
        d = args.cols
        data, _, _, _ = encoded(n, d, with_y=args.target)
        acc = torch.zeros(int(lib.p1_acc_len(d)), dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        algo = {"imad": 1, "tcgen05": 2, "auto": 0}[args.algo]
        lib.check(lib.p1_ctx_profile(ctx, 1))
        ms = timeit(lambda: lib.check(lib.p1_local_block(ctx, data, 0, d, int(args.target), ptr(acc), algo)))
        prof = {}
        for k in range(pkg.ffi.P1_K_COUNT):
            t, c = ctypes.c_double(), ctypes.c_uint64()
            lib.check(lib.p1_ctx_profile_read(ctx, k, ctypes.byref(t), ctypes.byref(c)))
            if c.value:
                prof[lib.p1_kernel_class_name(k).decode()] = t.value / c.value
        m = d + int(args.target)
        macs = n * (m * (m - 1) // 2)
        res.update(ms=ms, cols=d, algo=args.algo, ring_GMACs=macs / ms / 1e6, per_kernel_ms=prof)
        lib.p1_data_free(data)
    elif args.what == "encode":
        d = args.cols
        data, X, y, norm = encoded(n, d)
        lib.p1_data_free(data)
        lib.check(lib.p1_ctx_profile(ctx, 1))
        for _ in range(args.reps):
            out = c_vp()
            lib.check(lib.p1_encode(ctx, ptr(X), ptr(y), n, d, norm, args.precision, ctypes.byref(out)))
            lib.check(lib.p1_ctx_sync(ctx))
            lib.p1_data_free(out)
        t, c = ctypes.c_double(), ctypes.c_uint64()
        lib.check(lib.p1_ctx_profile_read(ctx, pkg.ffi.P1_K_ENCODE, ctypes.byref(t), ctypes.byref(c)))
        ms = t.value / args.reps          # k_rows_to_cols + k_vec_encode of one call
        res.update(ms=ms, cols=d, GBs=(8 + w) * (n * d + n) / ms / 1e6, note="kernel time only (events around the launches)")
    elif args.what == "cot":
        from linreg_mpc_b200.phase1 import derive_cot_session, _bits128
        data, _, _, _ = encoded(n, 4)
        seeds0, seeds1, delta = derive_cot_session(bytes(range(16, 32)), 2, 3)
        chosen = np.where(_bits128(delta)[:, None], seeds1, seeds0)
        snd, rcv = c_vp(), c_vp()
        lib.check(lib.p1_cot_sender_new(ctx, delta.tobytes(), ptr(np.ascontiguousarray(chosen)), ctypes.byref(snd)))
        lib.check(lib.p1_cot_receiver_new(ctx, ptr(seeds0), ptr(seeds1), ctypes.byref(rcv)))
        ub = torch.empty(int(lib.p1_cot_u_bytes(ctx, n)), dtype=torch.uint8, device=dev)
        cb = torch.empty(int(lib.p1_cot_c_bytes(ctx, n)), dtype=torch.uint8, device=dev)
        acc = torch.zeros(int(lib.p1_acc_len(4)), dtype=torch.int64, device=dev)
        torch.cuda.synchronize()

        def one():
            lib.check(lib.p1_cot_receiver_extend(ctx, rcv, data, 0, n, 0, ptr(ub)))
            lib.check(lib.p1_cot_sender_step(ctx, snd, data, 1, n, 0, ptr(ub), ptr(cb), ptr(acc), 0))
            lib.check(lib.p1_cot_receiver_finish(ctx, rcv, ptr(cb), ptr(acc), 1))
        ms = timeit(one)
        res.update(ms=ms, cots=n * args.bits, MCOTs_per_s=n * args.bits / ms / 1e3)
        lib.p1_cot_free(snd)
        lib.p1_cot_free(rcv)
        lib.p1_data_free(data)
    elif args.what == "wire":
        vec = torch.empty(B * ld, dtype=wdt, device=dev).random_()
        if w == 8:   # random_() leaves the sign bit clear: set it on half of the words like a real mask
            vec ^= (torch.empty(B * ld, dtype=wdt, device=dev).random_() << 1)
        stride = int(lib.p1_msg_max_bytes(ctx, n))
        out = torch.empty(B * stride, dtype=torch.uint8, device=dev)
        back = torch.empty(B * ld, dtype=wdt, device=dev)
        sizes = torch.zeros(B, dtype=torch.int64, device=dev)
        vals = torch.arange(B, dtype=torch.int64, device=dev)
        vals2 = torch.zeros(B, dtype=torch.int64, device=dev)
        status = torch.zeros(B, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        ms_p = timeit(lambda: lib.check(lib.p1_msg_pack(ctx, ptr(vec), ld, n, B, ptr(vals), ptr(out), stride, ptr(sizes))))
        ms_u = timeit(lambda: lib.check(lib.p1_msg_unpack(ctx, ptr(out), stride, ptr(sizes), B, n, ptr(back), ld, ptr(vals2), ptr(status))))
        torch.cuda.synchronize()
        wire = int(sizes.sum().item())
        ok = bool(torch.equal(back.view(B, ld)[:, :n], vec.view(B, ld)[:, :n])) and not bool(status.any().item())
        res.update(pairs=B, wire_bytes=wire, bytes_per_word=wire / (B * n), round_trip_ok=ok,
                   pack_ms=ms_p, pack_GBs=(B * n * w + wire) / ms_p / 1e6, pack_Gwords_s=B * n / ms_p / 1e6,
                   unpack_ms=ms_u, unpack_GBs=(B * n * w + wire) / ms_u / 1e6, unpack_Gwords_s=B * n / ms_u / 1e6,
                   note="GB/s = algorithmic bytes (words + wire bytes, each once) per second")
    elif args.what == "text":
        rng = np.random.default_rng(1)
        x = rng.standard_normal(n)
        raw = " ".join(map(repr, x.tolist())).encode()
        tb = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        out = torch.empty(n, dtype=torch.float64, device=dev)
        nt, nd = ctypes.c_uint64(), ctypes.c_uint64()
        dl = np.zeros(16, np.uint64)
        torch.cuda.synchronize()
        lib.check(lib.p1_ctx_profile(ctx, 1))
        for _ in range(args.reps):
            lib.check(lib.p1_parse_doubles(ctx, ptr(tb), len(raw), ptr(out), n, ctypes.byref(nt), ctypes.byref(nd), ptr(dl), 16))
        t, c = ctypes.c_double(), ctypes.c_uint64()
        lib.check(lib.p1_ctx_profile_read(ctx, pkg.ffi.P1_K_TEXT, ctypes.byref(t), ctypes.byref(c)))
        ms = t.value / args.reps
        ok = bool(np.array_equal(out.cpu().numpy().view(np.uint64), x.view(np.uint64))) and nt.value == n and nd.value == 0
        res.update(ms=ms, text_bytes=len(raw), tokens=n, exact=ok, text_GBs=len(raw) / ms / 1e6,
                   GBs=(2 * len(raw) + 8 * n) / ms / 1e6, Gtokens_s=n / ms / 1e6,
                   note="kernel time (events around the five launches); GBs = text read twice + doubles written")
    print(json.dumps(res))
    lib.p1_ctx_destroy(ctx)


if __name__ == "__main__":
    main()
