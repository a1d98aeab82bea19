#!/usr/bin/env python
"""Times the drop-in host file (linreg-mpc_b200/host/phase1_gpu.c: run_party / run_trusted_initializer
over libp1gpu.so) against the reference's own phase1.c in the SAME loop-back driver (all parties as
threads of one process, oracle/ref_shims), on one instance file, and checks the shares are identical.
Informational (the loop-back transport is test infrastructure); bench.py is the judged harness.

  python tools/dropin_timing.py --n 100000 --d 20 --providers 3 [--use-ot]
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--d", type=int, default=20)
    ap.add_argument("--providers", type=int, default=3)
    ap.add_argument("--use-ot", action="store_true")
    ap.add_argument("--bits", type=int, default=64)
    args = ap.parse_args()
    prec = 56 if args.bits == 64 else 24
    X, y, _ = po.generate_lin_regression(args.n, args.d, seed=31)
    cs = po.even_col_start(args.d, args.providers)
    path = os.path.join(tempfile.mkdtemp(), "inst.in")
    po.Oracle().write_instance(path, X, y, cs)
    out = {"n": args.n, "d": args.d, "providers": args.providers, "use_ot": args.use_ot, "bits": args.bits,
           "text_bytes": os.path.getsize(path)}
    res = {}
    for name, gpuhost in (("gpu_host", True), ("gpu_host_warm", True), ("reference_phase1_c", False)):
        lib = po.RefLib(args.bits, args.bits, gpuhost=gpuhost)
        t0 = time.perf_counter()
        shares, stats, sent = lib.run(path, args.d, args.providers, prec, 0, use_ot=args.use_ot)
        res[name] = shares
        out[name + "_s"] = time.perf_counter() - t0
        out[name + "_bytes_on_sockets"] = int(sent.sum())
    out["shares_identical"] = bool(np.array_equal(res["gpu_host_warm"], res["reference_phase1_c"]))
    out["speedup_warm"] = out["reference_phase1_c_s"] / out["gpu_host_warm_s"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
