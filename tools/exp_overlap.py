#!/usr/bin/env python
"""Experiment (not a benchmark): where does the cfg3 step lose its overlap?  Times the masking loop
of Phase1Box (n=1M, d=500, 5 providers) in variants that are NOT all correct protocols:
  full      : the real step
  ti_novec  : the TI computes <x,y> and r but does not store the masks (party kernels read stale buffers)
  ti_only   : only the TI launches
  party_only: only the party launches
so that the cost of the mask stores under the co-running HBM-bound kernels can be separated."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import linreg_mpc_b200 as pkg  # noqa: E402
from linreg_mpc_b200.phase1 import Phase1Box  # noqa: E402
from linreg_mpc_b200.ffi import c_vp  # noqa: E402


def loop(box, key, mode, frac=0.25):
    C = int(box.sched.num_cross * frac)
    box._ensure_bufs()
    L = box.lib
    box._fan_out()
    for b, q0 in enumerate(range(0, C, box.batch_pairs)):
        B = min(box.batch_pairs, C - q0)
        pipe = b % len(box.ctxs)
        ctx = box.ctxs[pipe]
        xb, yb, mb, ma = (c_vp(tn.data_ptr()) for tn in box.bufs[pipe])
        xy = c_vp(box.xy.data_ptr() + 8 * q0)
        r = c_vp(box.r.data_ptr() + 8 * q0)
        ca = c_vp(box.cols_a.data_ptr() + 4 * q0)
        cb = c_vp(box.cols_b.data_ptr() + 4 * q0)
        sa = c_vp(box.slot_a.data_ptr() + 8 * q0)
        sb = c_vp(box.slot_b.data_ptr() + 8 * q0)
        acc = c_vp(box.acc.data_ptr())
        args = (key, q0, B, box.n_total, box.row0, box.n_rows)
        if mode in ("full", "ti_only"):
            L.check(L.p1_ti_generate(ctx, *args, xb, yb, box.ld, xy, r))
        elif mode == "ti_novec":
            L.check(L.p1_ti_generate(ctx, *args, None, None, box.ld, xy, r))
        if mode in ("full", "ti_novec", "party_only"):
            L.check(L.p1_party_b_mask(ctx, box.data, cb, B, xb, box.ld, mb, box.ld))
            L.check(L.p1_party_a_step(ctx, box.data, ca, sa, B, yb, box.ld, mb, box.ld, ma, box.ld, acc))
            L.check(L.p1_party_b_share(ctx, box.data, cb, sb, B, ma, box.ld, acc))
    box._fan_in()
    box.sync()
    return C


def main():
    lib = pkg.P1Lib()
    n, d, p = int(os.environ.get("N", "1000000")), 500, 5
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(1)
    X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
    X /= X.abs().amax(dim=0)
    y = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    torch.cuda.synchronize()
    pipes = int(os.environ.get("PIPES", "2"))
    batch = int(os.environ.get("BATCH", "33"))
    box = Phase1Box(lib, n, d, [k * (d // p) for k in range(p)], 56, batch_pairs=batch, pipes=pipes)
    for o in os.environ.get("OPTS", "").split(","):
        if o:
            k, v = o.split("=")
            for c in box.ctxs:
                lib.check(lib.p1_ctx_set_option(c, k.encode(), int(v)))
    box.reset(); box.encode(X, y); box.sync()
    key = bytes(range(16))
    out = {"pipes": pipes, "batch": batch, "opts": os.environ.get("OPTS", "")}
    for mode in ("full", "ti_novec", "ti_only", "party_only", "full"):
        loop(box, key, mode, 0.02)
        t0 = time.perf_counter()
        C = loop(box, key, mode)
        dt = time.perf_counter() - t0
        out[mode + "_ms_per_batch"] = round(dt * 1e3 / (C / batch), 4)
        out[mode + "_us_per_pair"] = round(dt * 1e6 / C, 3)
    print(json.dumps(out))
    box.close()


if __name__ == "__main__":
    main()
