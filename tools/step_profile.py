#!/usr/bin/env python
"""Wall-clock of the phases of one Phase1Box step (sync after each), to find host overheads."""
import os, sys, time, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import linreg_mpc_b200 as pkg
from linreg_mpc_b200.phase1 import Phase1Box
import bench

wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "gram"]
n, d, p = wl["n"], wl["d"], wl["p"]
lib = pkg.P1Lib()
dev = torch.device("cuda", 0)
X = torch.randn((n, d), dtype=torch.float64, device=dev); X /= X.abs().amax(dim=0)
y = torch.randn(n, dtype=torch.float64, device=dev)
box = Phase1Box(lib, n, d, bench.even_col_start(d, p), wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"])
key = bytes(range(16))
for it in range(4):
    t = [time.perf_counter()]
    def mark():
        box.sync(); torch.cuda.synchronize(); t.append(time.perf_counter())
    box.reset(); mark()
    box.encode(X, y); mark()
    box.local_blocks(); mark()
    box.cross_pairs_masking(key); mark()
    box.all_reduce(); box.apply_values(); mark()
    out = box.finalize(); mark()
    names = ["reset", "encode", "local", "cross", "reduce+values", "finalize"]
    print(it, " ".join("%s=%.2fms" % (nm, (b - a) * 1e3) for nm, a, b in zip(names, t, t[1:])), "total=%.2fms" % ((t[-1] - t[0]) * 1e3))
box.close()
