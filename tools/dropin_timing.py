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


def measure(n, d, providers, use_ot=False, bits=64, devices=None, seeded=False):
    """one instance file through the drop-in host file (cold, then warm) and through the reference's phase1.c"""
    prec = 56 if bits == 64 else 24
    X, y, _ = po.generate_lin_regression(n, d, seed=31)
    cs = po.even_col_start(d, providers)
    path = os.path.join(tempfile.mkdtemp(), "inst.in")
    po.Oracle().write_instance(path, X, y, cs)
    if devices:
        os.environ["P1GPU_DEVICES"] = devices
    if seeded:
        os.environ["P1GPU_SEEDED_MASKS"] = "1"      # read by the drop-in host file only (the reference has no such mode)
    out = {"n": n, "d": d, "providers": providers, "use_ot": use_ot, "bits": bits,
           "devices": os.environ.get("P1GPU_DEVICES", os.environ.get("P1GPU_DEVICE", "0")),
           "text_bytes": os.path.getsize(path)}
    res, sent_by = {}, {}
    for name, gpuhost in (("gpu_host", True), ("gpu_host_warm", True), ("reference_phase1_c", False)):
        lib = po.RefLib(bits, bits, gpuhost=gpuhost)
        t0 = time.perf_counter()
        shares, stats, sent = lib.run(path, d, providers, prec, 0, use_ot=use_ot)
        res[name] = shares
        sent_by[name] = sent
        out[name + "_s"] = time.perf_counter() - t0
        out[name + "_bytes_on_sockets"] = int(sent.sum())
    os.remove(path)
    out["seeded_masks"] = bool(seeded)
    if seeded:
        # other masks than the reference's single stream: the shares differ word by word, their sum over the parties
        # (what phase 2 consumes) does not; tests/test_gpu_dropin.py::test_run_party_seeded_masks checks every word
        m = np.uint64((1 << bits) - 1)
        out["share_sums_identical"] = bool(np.array_equal(res["gpu_host_warm"].sum(axis=0) & m, res["reference_phase1_c"].sum(axis=0) & m))
        out["ti_bytes_on_sockets"] = int(sent_by["gpu_host_warm"][0].sum())
        out["reference_ti_bytes_on_sockets"] = int(sent_by["reference_phase1_c"][0].sum())
    out["shares_identical"] = bool(np.array_equal(res["gpu_host_warm"], res["reference_phase1_c"]) and
                                   np.array_equal(res["gpu_host"], res["reference_phase1_c"]))
    out["speedup_warm"] = out["reference_phase1_c_s"] / out["gpu_host_warm_s"]
    out["speedup_cold"] = out["reference_phase1_c_s"] / out["gpu_host_s"]
    # every byte on a socket was copied off a GPU by its sender and onto one by its receiver; the inputs go up as text
    out["h2d_bytes"] = out["text_bytes"] + out["gpu_host_warm_bytes_on_sockets"]
    out["d2h_bytes"] = out["gpu_host_warm_bytes_on_sockets"] + int(res["gpu_host_warm"].size) * (bits // 8)
    out["note"] = ("all parties (TI + providers) as threads of one process over the in-process loop-back transport of oracle/ref_shims; "
                   "cold = first run of the process (CUDA context, library load), warm = second run")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--d", type=int, default=20)
    ap.add_argument("--providers", type=int, default=3)
    ap.add_argument("--use-ot", action="store_true")
    ap.add_argument("--bits", type=int, default=64)
    ap.add_argument("--devices", default=None, help="P1GPU_DEVICES for the drop-in's parties, e.g. 0,1")
    ap.add_argument("--seeded", action="store_true", help="P1GPU_SEEDED_MASKS=1: the TI sends keys and values, the providers expand their masks")
    args = ap.parse_args()
    print(json.dumps(measure(args.n, args.d, args.providers, args.use_ot, args.bits, args.devices, args.seeded)))


if __name__ == "__main__":
    main()
