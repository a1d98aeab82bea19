#!/usr/bin/env python
"""bench.py — phase-1 share time / ring GMAC/s on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            (own arm; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W

One "step" = one whole phase-1 pass over the synthetic instance: fixed-point encode of the
shard's rows, local Gram blocks, every cross pair of the TI-masking protocol (mask
expansion, both parties' mask/share kernels), the ring all-reduce over row shards and the
finalize tail.  Rows are sharded over the N ranks (total work fixed: strong scaling).
`value` times the step with X, y resident in HBM as doubles; `e2e` times the same step
through the same calls from pinned HOST buffers, with the device->host read of the shares.
The CPU legs (cpu_baseline, --impl reference) are the only places that touch oracle/.
"""
import argparse
import ctypes
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "phase1_ring_gmacs"
UNIT = "ring GMAC/s"

WORKLOADS = {
    # name: n, d, providers, bits_p1, bits_p2, precision, precision_p2
    "cfg3": dict(n=1_000_000, d=500, p=5, bits_p1=64, bits_p2=64, precision=56, precision_p2=0),
    "cfg2": dict(n=100_000, d=100, p=3, bits_p1=64, bits_p2=64, precision=56, precision_p2=0),
    "cfg4": dict(n=10_000_000, d=200, p=1, bits_p1=32, bits_p2=32, precision=30, precision_p2=0),
    "cfg5": dict(n=1_000_000, d=1000, p=1, bits_p1=64, bits_p2=32, precision=56, precision_p2=30),
    # the multi-provider variants of configs 4 and 5: 16 160 and 475 950 cross pairs
    "cfg4p5": dict(n=10_000_000, d=200, p=5, bits_p1=32, bits_p2=32, precision=30, precision_p2=0),
    "cfg5p20": dict(n=1_000_000, d=1000, p=20, bits_p1=64, bits_p2=32, precision=56, precision_p2=30),
    "gram": dict(n=1_000_000, d=500, p=1, bits_p1=64, bits_p2=64, precision=56, precision_p2=0),
    "tiny": dict(n=65_536, d=40, p=5, bits_p1=64, bits_p2=64, precision=56, precision_p2=0),
}


def even_col_start(d, p):
    return [k * (d // p) for k in range(p)]  # experiments/generate_tests.py:54-61


def workload_config(name, wl, aggregation):
    return {"workload": "%s: synthetic n=%d d=%d, %d providers (vertical split), %d-bit phase 1, %s, precision %d"
                        % (name, wl["n"], wl["d"], wl["p"], wl["bits_p1"], aggregation, wl["precision"]),
            "n": wl["n"], "d": wl["d"], "providers": wl["p"], "ring_bits": wl["bits_p1"], "aggregation": aggregation}


# ----------------------------------------------------------------------------- clocks

class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.f.close()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as f:
            for line in f:
                t = [x.strip() for x in line.split(",")]
                if len(t) < 8:
                    continue
                try:
                    sm.append(float(t[1]))
                    smax.append(float(t[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, t[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(smax)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# ----------------------------------------------------------------------------- CPU legs

def ring_macs(n, d, col_start, masking=True):
    from linreg_mpc_b200.phase1 import Schedule
    return Schedule(d, col_start).ring_macs(n, masking=masking)


class ReferenceSample:
    """The reference's own phase1.c (oracle/_ref, compiled unmodified against shims) on a
    bounded sample of the workload: same n as the workload when it fits, few columns."""

    def __init__(self, wl, d_sample, n_sample=None):
        from oracle import pyoracle as po
        self.po = po
        self.p = min(wl["p"], d_sample)
        self.n = int(n_sample or min(wl["n"], 1_000_000))
        self.d = d_sample
        self.wl = wl
        self.cs = even_col_start(self.d, self.p)
        kind = "reference" if po.RefLib.available(wl["bits_p1"], wl["bits_p2"]) else "port"
        self.kind = kind
        rng = np.random.default_rng(20170401)
        X = rng.standard_normal((self.n, self.d))
        X /= np.max(np.abs(X), axis=0, keepdims=True)
        y = X @ rng.uniform(0, 1, self.d) + rng.normal(0, 0.1, self.n)
        self.X, self.y = X, y
        self.orc = po.Oracle()
        self.path = None
        self.parse_s = 0.0
        if kind == "reference":
            self.path = tempfile.mktemp(prefix="p1sample_", suffix=".in")
            self.orc.write_instance(self.path, X, y, self.cs)
            self.ref = po.RefLib(wl["bits_p1"], wl["bits_p2"])
            _, _, self.parse_s = self.ref.read_inputs(self.path, self.n, self.d, wl["precision"])
        self.macs = ring_macs(self.n, self.d, self.cs)
        self.threads = self.p + 1 if kind == "reference" else (os.cpu_count() or 1)

    def run(self):
        """returns (seconds of phase-1 work, shares)"""
        wl = self.wl
        if self.kind == "reference":
            shares, stats, _ = self.ref.run(self.path, self.d, self.p, wl["precision"], wl["precision_p2"], False)
            wall = float(np.max(stats[:, 0]))
            # every party's run_party starts with the text parse (src/linear.c:27-102), which is
            # outside the path: subtract one (concurrent) parse
            return max(wall - self.parse_s, 1e-9), shares
        t0 = time.perf_counter()
        shares = self.orc.run(self.X, self.y, self.cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"],
                              False, threads=self.threads)
        return time.perf_counter() - t0, shares

    def describe(self):
        return ("n=%d d=%d providers=%d TI masking, 64-bit: %d of the workload's pairs' worth (%.3g ring MACs); "
                "%s; text parse (%.2f s) excluded" %
                (self.n, self.d, self.p, len(self.cs) and self.d * (self.d + 1) // 2 + self.d, self.macs,
                 "reference phase1.c compiled unmodified + loopback shims, one thread per party"
                 if self.kind == "reference" else "oracle port, %d threads" % self.threads, self.parse_s))

    def close(self):
        if self.path and os.path.exists(self.path):
            os.unlink(self.path)


def run_reference_arm(args, wl, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = max(1, args.steps + args.warmup)
    budget = 150.0 / total                      # seconds per step
    # ~0.2 s per cross pair at n = 1M (measured with this shim build); d=5 -> 14, 10 -> 48, 20 -> 176 cross pairs
    d_s = 5
    for cand, pairs in ((20, 176), (10, 48)):
        if 0.2 * pairs * (min(wl["n"], 1_000_000) / 1e6) <= budget:
            d_s = cand
            break
    smp = ReferenceSample(wl, d_s)
    for _ in range(args.warmup):
        smp.run()
    secs = [smp.run()[0] for _ in range(args.steps)]
    smp.close()
    t = float(np.mean(secs))
    value = smp.macs / t / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": smp.threads, "kind": smp.kind, "sample": smp.describe()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "host_cores": os.cpu_count()}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------- own arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--aggregation", default="masking", choices=["masking", "masking-seeded"])
    ap.add_argument("--gram", default="auto", choices=["auto", "imad", "tcgen05"])
    ap.add_argument("--batch-pairs", type=int, default=0)
    ap.add_argument("--pipes", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="library tuning knob name=value (p1_ctx_set_option), repeatable")
    ap.add_argument("--wire-sample", type=int, default=3300, help="cross pairs to time with GPU (de)serialisation of every message (0 = skip)")
    ap.add_argument("--ot-sample", type=int, default=2, help="inner products of the --use_ot path to time (0 = skip)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    agg = "TI masking protocol" + (" (CSP-seeded masks)" if args.aggregation == "masking-seeded" else "")
    cfg = workload_config(args.workload, wl, agg)
    if args.impl == "reference":
        return run_reference_arm(args, wl, cfg)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # ONE JSON line on stdout: NCCL prints its version banner there at NCCL_DEBUG=VERSION (the redirect
        # below only applies above that level), so that level is lowered; INFO/TRACE output goes to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    import linreg_mpc_b200 as pkg
    from linreg_mpc_b200.phase1 import Phase1Box
    lib = pkg.P1Lib()
    n, d, p = wl["n"], wl["d"], wl["p"]
    cs = even_col_start(d, p)
    w = wl["bits_p1"] // 8
    # row shards: boundaries on multiples of 256 rows (16-byte aligned mask slices, whole OT tiles)
    from linreg_mpc_b200.phase1 import shard_rows
    row0, n_rows = shard_rows(n, world, rank)
    per = shard_rows(n, world, 0)[1]
    assert n_rows > 0, "more ranks than 256-row blocks"

    # synthetic instance per experiments/generate_tests.py:159-169 (X ~ N(0,1), columns scaled to
    # max |.| = 1, beta ~ U(0,1), y = X beta + N(0, 0.1^2)), generated on the device per shard
    g = torch.Generator(device=dev)
    g.manual_seed(20170401 + 7919 * rank)
    Xd = torch.randn((n_rows, d), dtype=torch.float64, device=dev, generator=g)
    cmax = Xd.abs().amax(dim=0)
    if group is not None:
        dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
    Xd /= cmax
    gb = torch.Generator(device=dev)
    gb.manual_seed(20170401)
    beta = torch.rand(d, dtype=torch.float64, device=dev, generator=gb)
    yd = Xd @ beta + 0.1 * torch.randn(n_rows, dtype=torch.float64, device=dev, generator=g)
    torch.cuda.synchronize()

    gram = {"auto": pkg.ffi.P1_GRAM_AUTO, "imad": pkg.ffi.P1_GRAM_IMAD, "tcgen05": pkg.ffi.P1_GRAM_TCGEN05}[args.gram]
    batch = args.batch_pairs or max(1, min(256, (256 << 20) // max(1, n_rows * w)))   # ~256 MB of masks per vector buffer
    box = Phase1Box(lib, n, d, cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"], device=local,
                    row0=row0, n_rows=n_rows, batch_pairs=batch, pipes=args.pipes, gram_algo=gram, group=group)
    for o in args.opt:
        name, val = o.split("=")
        for c in box.ctxs:
            lib.check(lib.p1_ctx_set_option(c, name.encode(), int(val)))
    key = bytes(range(16))
    seeded = args.aggregation == "masking-seeded"
    macs = box.sched.ring_macs(n, masking=True)
    macs_gram_p1 = n * (d * (d + 1) // 2)   # the p = 1 full-Gram figure

    def barrier():
        torch.cuda.synchronize()
        if group is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def step(X, y):
        return box.run_masking(X, y, key, seeded=seeded)

    s0 = torch.cuda.ExternalStream(int(lib.p1_ctx_stream(box.ctxs[0])), device=dev)

    def timed(fn, k):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s0)
        t0 = time.perf_counter()
        for _ in range(k):
            fn()
        b.record(s0)
        barrier()
        wall = time.perf_counter() - t0
        ms = torch.tensor([a.elapsed_time(b), wall * 1e3], dtype=torch.float64, device=dev)
        if group is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]) / k, float(ms[1]) / k

    # ---- warm-up, then the timed region (device-resident inputs)
    shares = None
    for _ in range(max(args.warmup, 1)):
        shares = step(Xd, yd)
    for c in box.ctxs:
        lib.check(lib.p1_ctx_profile(c, 1))
    launches0 = box.launches()
    sampler = ClockSampler(local) if rank == 0 else None
    ms_step, ms_wall = timed(lambda: step(Xd, yd), args.steps)
    clocks = sampler.stop() if sampler else None
    launches = (box.launches() - launches0) // max(args.steps, 1)
    prof = {}
    for k in range(pkg.ffi.P1_K_COUNT):
        tot, cnt = 0.0, 0
        for c in box.ctxs:
            ms, nl = ctypes.c_double(), ctypes.c_uint64()
            lib.check(lib.p1_ctx_profile_read(c, k, ctypes.byref(ms), ctypes.byref(nl)))
            tot += ms.value
            cnt += nl.value
        if cnt:
            prof[lib.p1_kernel_class_name(k).decode()] = {"ms_per_step": tot / args.steps, "launches_per_step": cnt / args.steps,
                                                         "avg_ms": tot / cnt}
    for c in box.ctxs:
        lib.check(lib.p1_ctx_profile(c, 0))

    # ---- roofline of the dominant kernel (algorithmic bytes / flops per launch)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
    # no int8 entry in MEASURED_PEAKS.json: dense int8 = 2x dense bf16 on this part, so 2x the measured bf16 burst
    int8_peak = 2.0 * float(peaks.get("bf16_tflops", 1590.0))
    int8_nominal = 4500.0
    C = box.sched.num_cross
    nb = n_rows * w
    local_macs, tc_ops, limb_bytes = 0, 0, 0
    for k in range(box.sched.num_dp):
        lo, hi = box.sched.col_range(k)
        m = hi - lo + int(k == box.sched.num_dp - 1)
        pairs = m * (m - 1) // 2
        Lk = ctypes.c_int(0)
        lib.check(lib.p1_gram_digits(box.ctxs[0], box.data, lo, hi, int(k == box.sched.num_dp - 1), ctypes.byref(Lk)))
        local_macs += n_rows * pairs
        tc_ops += 2 * Lk.value ** 2 * n_rows * pairs         # int8 ops of the lower-triangle pairs (executed digits)
        limb_bytes += m * n_rows * (w + Lk.value)
    alg_bytes = {  # per step on this rank
        "encode": (8 + w) * (n_rows * d + n_rows),
        "ti_generate": C * (2 * nb) * (1 if not seeded else 2),     # keystream words written
        "party_b_mask": C * 3 * nb,                                  # column + mask read, message written
        "party_a_step": C * 4 * nb,                                  # column, mask, peer message read; message written
        "party_b_share": C * 2 * nb,                                 # peer message + column read
        "gram_imad": (d + 1) * nb,                                   # every owned column read once
        "gram_limbs": limb_bytes,                                    # words read, digit planes written
    }
    kernels = {}
    for name, pr in prof.items():
        ent = dict(pr)
        if alg_bytes.get(name):
            gbs = alg_bytes[name] / (pr["ms_per_step"] * 1e-3) / 1e9
            ent.update(bound="hbm", achieved=gbs, peak=hbm_peak, unit="GB/s", frac=gbs / hbm_peak)
        if name == "gram_tc" and tc_ops:
            tops = tc_ops / (pr["ms_per_step"] * 1e-3) / 1e12
            ent.update(bound="tensor", achieved=tops, peak=int8_peak, unit="TOP/s (int8)", frac=tops / int8_peak,
                       frac_of_nominal_dense_int8=tops / int8_nominal, ring_gmacs=local_macs / (pr["ms_per_step"] * 1e-3) / 1e9,
                       peak_source="2 x measured dense bf16 burst (no int8 entry in MEASURED_PEAKS.json); nominal dense int8 is 4500")
        if name == "ti_generate":
            # the honest bound of the T-table AES: 16 lookups x 10 rounds, minus the 12 + 15 of rounds 1-2 that the
            # CTR structure removes, plus 15 per 8 blocks for the group cache = 135 shared-memory lookups per
            # 16-byte block; one 32-lane LDS per clock per SM
            blocks = C * (2 * nb + 16) / 16.0 * (1 if not seeded else 2)
            sm_clk = (clocks or {}).get("sm_mhz") or 1965.0
            lsu_peak = 148 * 32 * sm_clk * 1e6 / 135.0          # blocks/s at one LDS per clock per SM
            ent.update(lsu_bound={"achieved_blocks_per_s": blocks / (pr["ms_per_step"] * 1e-3), "peak_blocks_per_s": lsu_peak,
                                  "frac": blocks / (pr["ms_per_step"] * 1e-3) / lsu_peak,
                                  "note": "135 LDS per AES block, 128 B/clk/SM shared-memory port, at the sampled SM clock; "
                                          "the kernel's own time includes waiting behind the other pipe's co-resident party kernels"})
        kernels[name] = ent
    dominant = max(prof, key=lambda k: prof[k]["ms_per_step"]) if prof else None
    roofline = None
    if dominant and "achieved" in kernels[dominant]:
        kd = kernels[dominant]
        roofline = {"kernel": dominant, "bound": kd["bound"], "achieved": kd["achieved"], "peak": kd["peak"], "unit": kd["unit"],
                    "frac": kd["frac"], "traffic": None, "peak_source": kd.get("peak_source", peak_src),
                    "avg_launch_ms": kd["avg_ms"]}
        if kd["bound"] == "hbm":
            roofline["algorithmic_bytes_per_launch"] = alg_bytes[dominant] / max(kd["launches_per_step"], 1)
        if "lsu_bound" in kd:
            roofline["lsu_bound"] = kd["lsu_bound"]
        try:   # DRAM bytes per launch of this kernel from the committed ncu --set full capture, if the shape matches
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tr = json.load(f).get(dominant)
            if tr and (dominant != "ti_generate" or (tr["pairs"] == batch and tr["n_rows"] == n_rows and tr["ring_bits"] == wl["bits_p1"])):
                roofline["traffic"] = tr["bytes_per_launch"]
                roofline["traffic_source"] = tr["source"]
        except (OSError, ValueError, KeyError):
            pass

    # ---- end to end: host buffers in, shares out
    e2e = None
    if not args.no_e2e:
        Xh = torch.empty((n_rows, d), dtype=torch.float64, pin_memory=True)
        yh = torch.empty(n_rows, dtype=torch.float64, pin_memory=True)
        Xh.copy_(Xd)
        yh.copy_(yd)
        torch.cuda.synchronize()
        Xn, yn = Xh.numpy(), yh.numpy()
        step(Xn, yn)
        ms_e2e, _ = timed(lambda: step(Xn, yn), args.steps)
        h2d = world * (n_rows * d * 8 + n_rows * 8)
        d2h = world * (p * box.acc_len * 8)
        e2e = {"value": macs / (ms_e2e * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}

    # ---- the same protocol with every message through the protobuf wire format on the GPU (what separate
    # boxes execute per pair, minus the network): a bounded sample of the cross pairs, extrapolated
    wire = None
    if args.wire_sample > 0 and C:
        npairs = min(C, args.wire_sample)

        def wire_step(w):
            box.cross_pairs_masking(key, seeded=seeded, wire=w, max_pairs=npairs)
            box.sync()
        wire_step(True)
        ms_w, _ = timed(lambda: wire_step(True), 1)
        ms_p, _ = timed(lambda: wire_step(False), 1)
        wire = {"pairs": npairs, "ms_with_wire": ms_w, "ms_plain": ms_p, "us_per_pair_with_wire": ms_w * 1e3 / npairs,
                "us_per_pair_plain": ms_p * 1e3 / npairs, "all_cross_pairs_ms_extrapolated": ms_w / npairs * C,
                "messages_per_pair": 4, "note": "p1_msg_pack + p1_msg_unpack around each of the 4 messages of a pair "
                "(src/phase1.c:100-145); the headline step hands messages over on the device"}

    # ---- --use_ot aggregation (BASELINE config 3 names it): a bounded sample of inner products over this
    # rank's rows through the correlated-OT kernels, reported as COT/s with the extrapolation to all cross pairs
    ot = None
    if args.ot_sample > 0 and C:
        from linreg_mpc_b200.phase1 import derive_cot_session, _bits128
        L_, ctx0 = lib, box.ctxs[0]
        s0_, s1_, delta_ = derive_cot_session(bytes(range(16, 32)), 3, 4)
        chosen_ = np.where(_bits128(delta_)[:, None], s1_, s0_)
        snd, rcv = ctypes.c_void_p(), ctypes.c_void_p()
        L_.check(L_.p1_cot_sender_new(ctx0, delta_.tobytes(), pkg.ffi.ptr(np.ascontiguousarray(chosen_)), ctypes.byref(snd)))
        L_.check(L_.p1_cot_receiver_new(ctx0, pkg.ffi.ptr(s0_), pkg.ffi.ptr(s1_), ctypes.byref(rcv)))
        ub = torch.empty(int(L_.p1_cot_u_bytes(ctx0, n_rows)), dtype=torch.uint8, device=dev)
        cb = torch.empty(int(L_.p1_cot_c_bytes(ctx0, n_rows)), dtype=torch.uint8, device=dev)
        acc_ot = torch.zeros(box.acc_len, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()

        def ot_step():
            for t in range(args.ot_sample):
                L_.check(L_.p1_cot_receiver_extend(ctx0, rcv, box.data, d - 1, n, row0, pkg.ffi.ptr(ub)))
                L_.check(L_.p1_cot_sender_step(ctx0, snd, box.data, t % (d // p), n, row0, pkg.ffi.ptr(ub), pkg.ffi.ptr(cb),
                                               pkg.ffi.ptr(acc_ot), t))
                L_.check(L_.p1_cot_receiver_finish(ctx0, rcv, pkg.ffi.ptr(cb), pkg.ffi.ptr(acc_ot), t))
        ot_step()
        ms_ot, _ = timed(ot_step, 2)
        cots = args.ot_sample * n * wl["bits_p1"]
        rate = cots / (ms_ot * 1e-3)
        total_cots = C * n * wl["bits_p1"]
        ot = {"inner_products": args.ot_sample, "cots": cots, "ms": ms_ot, "cot_per_s": rate,
              "wire_bytes_per_cot": 16 + w, "all_cross_pairs_cots": total_cots, "all_cross_pairs_seconds_extrapolated": total_cots / rate,
              "note": "reference protocol needs n*W correlated OTs per cross pair (src/phase1.c:58-65); reported, not the headline"}
        L_.p1_cot_free(snd)
        L_.p1_cot_free(rcv)

    # ---- parity property at full size: the providers' shares sum to the plain ring Gram of the encoded data
    check = None
    if rank == 0 and shares is not None:
        check = {"shares_nonzero": bool(np.any(shares))}

    # ---- CPU baseline (rank 0, N = 1 only): the reference's own phase-1 code on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        smp = ReferenceSample(wl, 10)
        secs, ref_shares = smp.run()
        cpu = {"value": smp.macs / secs / 1e9, "unit": UNIT, "cores": smp.threads, "kind": smp.kind,
               "sample": smp.describe(), "seconds": secs, "host_cores": os.cpu_count()}
        # the same sample through the GPU path: every share word must match the reference's
        sbox = Phase1Box(lib, smp.n, smp.d, smp.cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"],
                         device=local, batch_pairs=8)
        got = sbox.run_masking(smp.X, smp.y, key)
        sbox.close()
        cpu["gpu_shares_match_sample"] = bool(np.array_equal(got, ref_shares))
        smp.close()

    box.close()
    if rank == 0:
        line = {"metric": METRIC, "value": macs / (ms_step * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 1), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "int64" if w == 8 else "int32", "data": "synthetic",
                "config": dict(cfg, rows_per_rank=per, batch_pairs=batch, gram=args.gram,
                               l2="inputs (%.1f GB of columns + %.1f GB of messages per step) exceed the 126 MB L2"
                                  % ((d + 1) * nb / 1e9, C * 4 * nb / 1e9)),
                "ring_macs_per_step": macs, "share_time_ms": ms_step, "wall_ms_per_step": ms_wall,
                "gram_p1_equiv_macs": macs_gram_p1,
                "gpu_launches": int(launches) * world, "kernels": kernels, "roofline": roofline, "clocks": clocks,
                "e2e": e2e, "wire_format_sample": wire, "use_ot_sample": ot, "cpu_baseline": cpu, "check": check}
        print(json.dumps(line))
    if group is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
