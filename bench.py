#!/usr/bin/env python
"""bench.py — phase-1 share time / ring GMAC/s on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            (own arm; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W
  python bench.py --workload cfg3 --aggregation ot --ot-pairs 1000     (BASELINE config 3 as written: --use_ot)
  python bench.py --workload cfg4|cfg5|gram|cfg2|cfg4p5|cfg5p20 ...

One "step" = one whole phase-1 pass over the synthetic instance: fixed-point encode of the
shard's rows, local Gram blocks, every cross pair of the aggregation (TI masking: the fused
masking loop - mask expansion and both parties' steps; --use_ot: the correlated-OT
extension), the ring all-reduce over row shards and the finalize tail.  Rows are sharded
over the N ranks (total work fixed: strong scaling).  `value` times the step with X, y
resident in HBM as doubles; `e2e` times the same step through the same calls from pinned
HOST buffers, with the device->host read of the shares.  `check` (outside the timed region,
every N) proves the shares of the timed configuration: sum invariant on all words against an
independent one-provider Gram, sampled words against the CPU oracle.
The CPU legs (cpu_baseline, --impl reference, the oracle inside `check`) are the only places
that touch oracle/.
"""
import argparse
import ctypes
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
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


# index of BASELINE.json's config a workload instantiates (seed = 20170401 + cfg_index)
CFG_INDEX = {"cfg2": 2, "cfg3": 3, "gram": 3, "cfg4": 4, "cfg4p5": 4, "cfg5": 5, "cfg5p20": 5, "tiny": 0}
GEN_CHUNK = 4096          # rows per generator chunk (a multiple of the 256-row shard alignment)


def synthetic_chunk(cfg_index, chunk, d):
    """Rows [chunk * GEN_CHUNK, (chunk + 1) * GEN_CHUNK) of the instance of SURVEY.md 8d before the column scaling:
    X ~ N(0,1) row-major, then the noise of those rows, from numpy.random.default_rng([20170401 + cfg_index, chunk]) -
    one stream per row chunk, so that any row shard of any N generates exactly its rows of the SAME matrix in parallel
    (a single sequential stream would make rank r draw everything before its rows)."""
    rng = np.random.default_rng([20170401 + cfg_index, chunk])
    return rng.standard_normal((GEN_CHUNK, d)), rng.standard_normal(GEN_CHUNK)


def synthetic_rows(cfg_index, row0, n_rows, d, threads=8):
    """(X, noise) of rows [row0, row0 + n_rows) as float64 arrays (unscaled); row0 must be a multiple of 256"""
    from concurrent.futures import ThreadPoolExecutor
    X = np.empty((n_rows, d))
    e = np.empty(n_rows)
    c0, c1 = row0 // GEN_CHUNK, (row0 + n_rows + GEN_CHUNK - 1) // GEN_CHUNK

    def fill(c):
        xc, ec = synthetic_chunk(cfg_index, c, d)
        lo, hi = max(row0, c * GEN_CHUNK), min(row0 + n_rows, (c + 1) * GEN_CHUNK)
        X[lo - row0:hi - row0] = xc[lo - c * GEN_CHUNK:hi - c * GEN_CHUNK]
        e[lo - row0:hi - row0] = ec[lo - c * GEN_CHUNK:hi - c * GEN_CHUNK]
    with ThreadPoolExecutor(max_workers=threads) as ex:      # numpy's generators release the GIL
        list(ex.map(fill, range(c0, c1)))
    return X, e


def even_col_start(d, p):
    return [k * (d // p) for k in range(p)]  # experiments/generate_tests.py:54-61


# the reference's own published phase-1 rows (AWS EC2 hosts of unrecorded type, one party per host,
# experiments/results/phase1*/): different hardware, a network in between - recorded beside the measurement, never divided by
PUBLISHED = {
    "cfg2": {"what": "n=100000 d=100 p=3, TI masking, 64-bit", "dp_wall_s": 63.89, "ti_cpu_s": 36.47, "dp_cpu_s": 27.96,
             "source": "BASELINE.md 1.1 (experiments/results/phase1/test_LR_100000x100_3_*)", "hardware": "published, different hardware (AWS EC2)"},
    "cfg3": {"what": "n=1000000 d=500 p=5, 64-bit", "dp_wall_s": None, "never_run": True,
             "masking_wall_s_extrapolated": 2.0e4, "use_ot_wall_s_extrapolated": 8.0e5,
             "largest_run_published": "n=1000000 d=200 p=5 masking: 3552 s DP wall; n=1000000 d=50 p=5 --use_ot: 8227 s",
             "source": "BASELINE.md 1.1, 1.3, 1.4 (plots/phase1_time.py:18,39 skips d=500 at n=1M)",
             "hardware": "published, different hardware (AWS EC2)"},
}


def workload_config(name, wl, aggregation):
    return {"workload": "%s: synthetic n=%d d=%d, %d providers (vertical split), %d-bit phase 1, %s, precision %d"
                        % (name, wl["n"], wl["d"], wl["p"], wl["bits_p1"], aggregation, wl["precision"]),
            "data_seed": "numpy.random.default_rng([20170401 + %d, row chunk]) per 4096-row chunk (X row-major, then the chunk's noise), "
                         "beta from default_rng(20170401 + %d): the same matrix for every N; deviation from SURVEY 8d's single stream: "
                         "one stream per chunk so that shards generate in parallel" % (CFG_INDEX.get(name, 0), CFG_INDEX.get(name, 0)),
            "n": wl["n"], "d": wl["d"], "providers": wl["p"], "ring_bits": wl["bits_p1"], "aggregation": aggregation}


# ----------------------------------------------------------------------------- clocks

class ClockSampler:
    """SM clock / throttle reasons while the GPU is under load: NVML polled from a thread every ~10 ms
    (a 200 ms nvidia-smi loop sees nothing of a 15 ms timed region); nvidia-smi as the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, period_s=0.01):
        self.sm, self.power, self.reasons, self.smax = [], [], set(), None
        self.proc = self.thread = None
        self.stop_flag = False
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[gpu_index]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else gpu_index
            h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.smax = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"}

            def loop():
                while not self.stop_flag:
                    try:
                        self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                        for bit, nm in names.items():
                            if r & bit:
                                self.reasons.add(nm)
                        self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
                    except Exception:
                        pass
                    time.sleep(period_s)
            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
            self.how = "nvml, %d ms" % int(period_s * 1000)
        except Exception:
            self.thread = None
            self.how = "nvidia-smi -lms 100"
            self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
            try:
                self.f = open(self.path, "w")
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                              "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
            except OSError:
                self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "how": self.how}
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
        elif self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.f.close()
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            smax = []
            with open(self.path) as f:
                for line in f:
                    t = [x.strip() for x in line.split(",")]
                    if len(t) < 8:
                        continue
                    try:
                        self.sm.append(float(t[1]))
                        smax.append(float(t[2]))
                        self.power.append(float(t[3]))
                    except ValueError:
                        continue
                    for nm, v in zip(names, t[4:8]):
                        if v.lower().startswith("active"):
                            self.reasons.add(nm)
            os.unlink(self.path)
            self.smax = max(smax) if smax else None
        if self.sm:
            out.update(sm_mhz=float(np.median(self.sm)), sm_min_mhz=float(min(self.sm)), sm_max_mhz=self.smax, samples=len(self.sm))
        if self.power:
            out["power_w_max"] = float(max(self.power))
        out["reasons"] = sorted(self.reasons)
        return out


# ----------------------------------------------------------------------------- CPU legs

def ring_macs(n, d, col_start, masking=True):
    from linreg_mpc_b200.phase1 import Schedule
    return Schedule(d, col_start).ring_macs(n, masking=masking)


class ReferenceSample:
    """The reference's own phase1.c (oracle/_ref, compiled unmodified against shims) on a
    bounded sample of the workload: same n as the workload when it fits, few columns."""

    def __init__(self, wl, d_sample, n_sample=None, use_ot=False):
        from oracle import pyoracle as po
        self.po = po
        self.p = min(wl["p"], d_sample)
        self.n = int(n_sample or min(wl["n"], 1_000_000))
        self.d = d_sample
        self.wl = wl
        self.use_ot = use_ot
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
        self.macs = ring_macs(self.n, self.d, self.cs, masking=not use_ot) + (0 if not use_ot else self.n * self.cross_pairs())
        self.threads = (self.p + 1 if not use_ot else self.p * self.p) if kind == "reference" else (os.cpu_count() or 1)

    def cross_pairs(self):
        from linreg_mpc_b200.phase1 import Schedule
        return Schedule(self.d, self.cs).num_cross

    def run(self):
        """returns (seconds of phase-1 work, shares)"""
        wl = self.wl
        if self.kind == "reference":
            shares, stats, _ = self.ref.run(self.path, self.d, self.p, wl["precision"], wl["precision_p2"], self.use_ot)
            wall = float(np.max(stats[:, 0]))
            # every party's run_party starts with the text parse (src/linear.c:27-102), which is
            # outside the path: subtract one (concurrent) parse
            return max(wall - self.parse_s, 1e-9), shares
        t0 = time.perf_counter()
        shares = self.orc.run(self.X, self.y, self.cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"],
                              self.use_ot, threads=self.threads)
        return time.perf_counter() - t0, shares

    def describe(self):
        return ("n=%d d=%d providers=%d %s, %d-bit: all %d pairs of that instance, %d of them cross pairs (%.3g ring MACs); "
                "%s; text parse (%.2f s) excluded; the shims' protobuf pack/unpack and loop-back copies are included" %
                (self.n, self.d, self.p, "--use_ot" if self.use_ot else "TI masking", self.wl["bits_p1"],
                 self.d * (self.d + 1) // 2 + self.d, self.cross_pairs(), self.macs,
                 "reference phase1.c compiled unmodified + loopback shims, one thread per party (p*p threads with --use_ot)"
                 if self.kind == "reference" else "oracle port, %d threads" % self.threads, self.parse_s))

    def close(self):
        if self.path and os.path.exists(self.path):
            os.unlink(self.path)


def run_reference_arm(args, wl, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    use_ot = args.aggregation == "ot"
    total = max(1, args.steps + args.warmup)
    budget = 240.0 / total                      # seconds per step
    # ~0.2 s per cross pair at n = 1M (measured with this shim build; ~8 s with --use_ot); d=5 -> 14, 10 -> 48, 20 -> 176 cross pairs
    per_pair = (8.0 if use_ot else 0.2) * (min(wl["n"], 1_000_000) / 1e6)
    d_s, n_s = 5, None
    for cand, pairs in ((20, 176), (10, 48)):
        if per_pair * pairs <= budget * 1.05:      # a tolerance: 0.2 * 48 is 9.600000000000001, the driver's budget 9.6
            d_s = cand
            break
    if use_ot and per_pair * 14 > budget:
        n_s = max(4096, int(min(wl["n"], 1_000_000) * budget / (per_pair * 14)) // 256 * 256)
    smp = ReferenceSample(wl, d_s, n_sample=n_s, use_ot=use_ot)
    for _ in range(args.warmup):
        smp.run()
    secs = [smp.run()[0] for _ in range(args.steps)]
    smp.close()
    t = float(np.mean(secs))
    value = smp.macs / t / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int64" if wl["bits_p1"] == 64 else "int32", "data": "synthetic", "config": cfg,
            "step_is": "one pass of the reference over the bounded sample described in cpu_baseline.sample (a rate, not the full workload)",
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": smp.threads, "kind": smp.kind, "sample": smp.describe()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "host_cores": os.cpu_count(), "d_sample": smp.d, "n_sample": smp.n, "cross_pairs_sample": smp.cross_pairs()}
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
    ap.add_argument("--aggregation", default="masking", choices=["masking", "masking-split", "masking-seeded", "ot"],
                    help="masking: fused masking loop (one kernel per batch); masking-split: TI / party kernels one by one; "
                         "masking-seeded: providers expand their own masks; ot: --use_ot correlated-OT aggregation on a sample of the cross pairs")
    ap.add_argument("--ot-pairs", type=int, default=1000, help="cross pairs (inner products) of the --use_ot schedule per step")
    ap.add_argument("--gram", default="auto", choices=["auto", "imad", "tcgen05"])
    ap.add_argument("--batch-pairs", type=int, default=0)
    ap.add_argument("--pipes", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--no-dropin", action="store_true", help="skip the drop-in host file vs reference phase1.c timing (n=1M, d=10, 3 providers)")
    ap.add_argument("--opt", action="append", default=[], help="library tuning knob name=value (p1_ctx_set_option), repeatable")
    ap.add_argument("--wire-sample", type=int, default=3300, help="cross pairs to time with GPU (de)serialisation of every message (0 = skip)")
    ap.add_argument("--gram-p1", type=int, default=1, help="1: also time the one-provider tcgen05 Gram of the instance (gram_p1 block); 0: skip")
    ap.add_argument("--ot-sample", type=int, default=250, help="masking workloads: inner products PER GPU of the --use_ot schedule timed as the ot_workload block (0 = skip)")
    ap.add_argument("--min-sampled-s", type=float, default=2.0, help="short steps are repeated for at least this long under the clock sampler")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    use_ot = args.aggregation == "ot"
    agg = {"masking": "TI masking protocol", "masking-split": "TI masking protocol (split kernels)",
           "masking-seeded": "TI masking protocol (CSP-seeded masks)",
           "ot": "--use_ot correlated-OT aggregation, first %d cross pairs of the schedule" % args.ot_pairs}[args.aggregation]
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
    from linreg_mpc_b200.phase1 import Phase1Box, shard_rows
    lib = pkg.P1Lib()
    n, d, p = wl["n"], wl["d"], wl["p"]
    cs = even_col_start(d, p)
    w = wl["bits_p1"] // 8
    # row shards: boundaries on multiples of 256 rows (16-byte aligned mask slices, whole OT tiles)
    row0, n_rows = shard_rows(n, world, rank)
    per = shard_rows(n, world, 0)[1]
    assert n_rows > 0, "more ranks than 256-row blocks"

    # synthetic instance per experiments/generate_tests.py:159-169 (X ~ N(0,1), columns scaled to max |.| = 1,
    # beta ~ U(0,1), y = X beta + N(0, 0.1^2)) from the documented numpy seed, one stream per 4096-row chunk: every N and
    # every rank sees its rows of the same matrix, and the CPU side can regenerate any of them (synthetic_rows)
    cfg_index = CFG_INDEX[args.workload]
    slab = 64 * GEN_CHUNK
    Xd = torch.empty((n_rows, d), dtype=torch.float64, device=dev)
    noise = torch.empty(n_rows, dtype=torch.float64, device=dev)
    for r in range(0, n_rows, slab):
        rows = min(slab, n_rows - r)
        xs, es = synthetic_rows(cfg_index, row0 + r, rows, d, threads=max(1, min(16, (os.cpu_count() or 8) // max(1, min(world, 8)))))
        Xd[r:r + rows] = torch.from_numpy(xs).to(dev)
        noise[r:r + rows] = torch.from_numpy(es).to(dev)
    cmax = Xd.abs().amax(dim=0)
    if group is not None:
        dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
    Xd /= cmax
    beta = torch.from_numpy(np.random.default_rng(20170401 + cfg_index).random(d)).to(dev)
    yd = Xd @ beta + 0.1 * noise
    del noise
    torch.cuda.synchronize()

    gram = {"auto": pkg.ffi.P1_GRAM_AUTO, "imad": pkg.ffi.P1_GRAM_IMAD, "tcgen05": pkg.ffi.P1_GRAM_TCGEN05}[args.gram]
    # pairs per launch: the fused loop keeps two message buffers per pipe (<= 1 GB each), the split kernels four (<= 256 MB each).
    # Row shards get proportionally more pairs per launch (up to 2048): a launch of the persistent grid should hold
    # many work units per warp, or its last round of units leaves SMs idle (245 units per pair at 125k rows)
    fused_step = args.aggregation == "masking"
    batch = args.batch_pairs or max(1, min(2048 if fused_step else 256, ((1 << 30) if fused_step else (256 << 20)) // max(1, n_rows * w)))
    box = Phase1Box(lib, n, d, cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"], device=local,
                    row0=row0, n_rows=n_rows, batch_pairs=batch, pipes=args.pipes, gram_algo=gram, group=group)
    for o in args.opt:
        name, val = o.split("=")
        for c in box.ctxs:
            lib.check(lib.p1_ctx_set_option(c, name.encode(), int(val)))
    key = bytes(range(16))
    ot_key = bytes(range(16, 32))
    seeded = args.aggregation == "masking-seeded"
    fused = args.aggregation == "masking"
    C_all = box.sched.num_cross
    C = min(C_all, args.ot_pairs) if use_ot else C_all     # cross pairs one step executes
    L_pairs = int(np.count_nonzero(box.sched.kind == 1))
    # ring MACs of the reference's schedule: n (L + 3C) with masking; with --use_ot a cross pair is
    # one Gilboa product = n ring multiplications
    macs = n * (L_pairs + (C if use_ot else 3 * C))
    macs_gram_p1 = n * (d * (d + 1) // 2)   # the p = 1 full-Gram figure

    def barrier():
        torch.cuda.synchronize()
        if group is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def step(X, y):
        if not use_ot:
            return box.run_masking(X, y, key, seeded=seeded, fused=fused)
        box.reset()
        box.encode(X, y)
        box.local_blocks()
        box.cross_pairs_ot(ot_key, max_pairs=C)
        box.all_reduce()
        return box.finalize()

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

    # ---- measured dense int8 tensor peak of this GPU (no int8 entry in MEASURED_PEAKS.json): tcgen05 kind::i8 probe
    int8 = None
    if rank == 0:
        int8 = {}
        for ncols in (256, 192):
            tops, ms = ctypes.c_double(), ctypes.c_double()
            smp = ClockSampler(local)
            lib.check(lib.p1_probe_tensor_i8(box.ctxs[0], ncols, 400.0, ctypes.byref(tops), ctypes.byref(ms), None))
            ck = smp.stop()
            int8["n%d" % ncols] = {"tops": tops.value, "ms": ms.value, "sm_mhz": ck["sm_mhz"], "reasons": ck["reasons"], "samples": ck["samples"]}
        int8["how"] = ("p1_probe_tensor_i8: one CTA per SM, back-to-back tcgen05.mma cta_group::1 kind::i8 M=128 N=256|192 K=32 on "
                       "smem-resident operands for >= 400 ms (k_probe.cu); n192 is the Gram kernel's instruction shape for 3-digit words")
    barrier()

    # ---- warm-up, then the timed region (device-resident inputs)
    shares = None
    for _ in range(max(args.warmup, 3)):
        shares = step(Xd, yd)
    launches0 = box.launches()
    sampler = ClockSampler(local) if rank == 0 else None
    ms_step, ms_wall = timed(lambda: step(Xd, yd), args.steps)
    clocks = sampler.stop() if sampler else None
    launches = (box.launches() - launches0) // max(args.steps, 1)
    # ---- per-kernel table: ONE more step with every launch bracketed by CUDA events, on a single pipe.  (In the timed
    # steps above two pipes keep two persistent kernels in flight; an event bracket around one of them would include the
    # time it waits for the other's SMs, so the brackets are taken with the kernels back to back on one stream.)
    box.pipe_limit = 1
    step(Xd, yd)
    for c in box.ctxs:
        lib.check(lib.p1_ctx_profile(c, 1))
    ms_prof, _ = timed(lambda: step(Xd, yd), 1)
    prof = {}
    for k in range(pkg.ffi.P1_K_COUNT):
        tot, cnt = 0.0, 0
        for c in box.ctxs:
            ms, nl = ctypes.c_double(), ctypes.c_uint64()
            lib.check(lib.p1_ctx_profile_read(c, k, ctypes.byref(ms), ctypes.byref(nl)))
            tot += ms.value
            cnt += nl.value
        if cnt:
            prof[lib.p1_kernel_class_name(k).decode()] = {"ms_per_step": tot, "launches_per_step": cnt, "avg_ms": tot / cnt}
    for c in box.ctxs:
        lib.check(lib.p1_ctx_profile(c, 0))
    box.pipe_limit = None

    # ---- short steps: the same step repeated for >= min_sampled_s under the sampler, so that the clock record
    # has enough samples (a 3 ms step x 5 is shorter than any sampling period) and the rate is a sustained one
    sustained = None
    if ms_wall * args.steps < args.min_sampled_s * 1e3:
        reps = int(math.ceil(args.min_sampled_s * 1e3 / max(ms_wall, 1e-3)))
        smp2 = ClockSampler(local) if rank == 0 else None
        ms_sus, _ = timed(lambda: step(Xd, yd), reps)
        ck2 = smp2.stop() if smp2 else None
        sustained = {"steps": reps, "ms_per_step": ms_sus, "value": macs / (ms_sus * 1e-3) / 1e9, "clocks": ck2}
        if clocks is not None and ck2 is not None and clocks["samples"] < 10:
            clocks = dict(ck2, note="sampled over the %d back-to-back repetitions of the step that follow the timed region "
                                    "(the %d timed steps last %.1f ms in total)" % (reps, args.steps, ms_wall * args.steps))

    # ---- roofline of the dominant kernel (algorithmic bytes / flops per launch)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
    if group is not None:   # rank 0 measured the int8 peak
        t8 = torch.tensor([int8["n256"]["tops"] if int8 else 0.0], dtype=torch.float64, device=dev)
        dist.broadcast(t8, src=0)
        int8_peak = float(t8[0])
    else:
        int8_peak = int8["n256"]["tops"]
    int8_nominal = 4500.0
    nb = n_rows * w
    sm_clk = (clocks or {}).get("sm_mhz") or 1965.0
    lsu_lanes = 148 * 32 * sm_clk * 1e6            # shared-memory lookups per second at one 32-lane LDS per clock per SM
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
    cots = C * n_rows * wl["bits_p1"] if use_ot else 0
    alg_bytes = {  # per step on this rank
        "encode": (8 + w) * (n_rows * d + n_rows),
        "ti_generate": C * (2 * nb) * (1 if not seeded else 2),     # keystream words written
        "party_b_mask": C * 3 * nb,                                  # column + mask read, message written
        "party_a_step": C * 4 * nb,                                  # column, mask, peer message read; message written
        "party_b_share": C * 2 * nb,                                 # peer message + column read
        "masking_fused": C * 4 * nb,                                 # columns a_i, b_j read; messages b->a, a->b written
                                                                     # (a_i is shared by consecutive pairs and comes from L2: see hbm_bytes_expected)
        "gram_imad": (d + 1) * nb,                                   # every owned column read once
        "gram_limbs": limb_bytes,                                    # words read, digit planes written
        "cot_recv": cots * (16 + w),                                 # U written (16 B / OT), pads stashed
        "cot_send": cots * (16 + w),                                 # U read, C written
        "cot_finish": cots * 2 * w,                                  # pads + C read
    }
    # shared-memory lookups per unit for the kernels whose limiter is the LDS port (T-table AES):
    # TI / fused loop: 135 per 16-byte keystream block; OT tile kernels per OT: receiver 2 PRG blocks
    # (4 + 144 lookups each with the column key in registers) + 1 hash (160), sender 1 PRG + 2 hashes
    lds_per_unit = {"ti_generate": 135.0, "masking_fused": 135.0, "cot_recv": 2 * 148 + 160.0, "cot_send": 148 + 320.0}
    units = {"ti_generate": C * (2 * nb + 16) / 16.0 * (1 if not seeded else 2), "masking_fused": C * (2 * nb + 16) / 16.0,
             "cot_recv": float(cots), "cot_send": float(cots)}
    kernels = {}
    for name, pr in prof.items():
        ent = dict(pr)
        sec = pr["ms_per_step"] * 1e-3
        if alg_bytes.get(name):
            gbs = alg_bytes[name] / sec / 1e9
            ent.update(bound="hbm", achieved=gbs, peak=hbm_peak, unit="GB/s", frac=gbs / hbm_peak)
        if name == "gram_tc" and tc_ops:
            tops = tc_ops / sec / 1e12
            ent.update(bound="tensor", achieved=tops, peak=int8_peak, unit="TOP/s (int8)", frac=tops / int8_peak,
                       frac_of_nominal_dense_int8=tops / int8_nominal, ring_gmacs=local_macs / sec / 1e9,
                       peak_source="measured in this run by p1_probe_tensor_i8 (tcgen05 kind::i8, M=128 N=256 K=32, operands resident in "
                                   "shared memory); nominal dense int8 is 4500")
        if name in lds_per_unit and units.get(name):
            rate = units[name] / sec
            peak = lsu_lanes / lds_per_unit[name]
            ent.update(lsu_bound={"achieved_per_s": rate, "peak_per_s": peak, "frac": rate / peak,
                                  "unit": "AES blocks" if name in ("ti_generate", "masking_fused") else "correlated OTs",
                                  "lookups_per_unit": lds_per_unit[name],
                                  "note": "shared-memory port: one 32-lane LDS per clock per SM at the sampled SM clock (%.0f MHz)" % sm_clk})
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
        if dominant == "masking_fused":
            # a_i is re-used by the consecutive pairs of a launch and stays in L2, so HBM sees ~3 n w per pair
            exp_b = C * 3 * nb / max(kd["launches_per_step"], 1)
            roofline["hbm_bytes_expected_per_launch"] = exp_b
            roofline["achieved_on_expected_hbm_bytes"] = exp_b / (kd["avg_ms"] * 1e-3) / 1e9
            roofline["frac_on_expected_hbm_bytes"] = roofline["achieved_on_expected_hbm_bytes"] / kd["peak"]
        roofline["avg_launch_ms_note"] = ("from one extra profiled step on a single pipe with event brackets around every launch "
                                          "(p1_ctx_profile); the timed step overlaps two pipes and is about 1 % faster per launch")
        if "lsu_bound" in kd:
            roofline["limiter"] = "lsu"
            roofline["lsu_bound"] = kd["lsu_bound"]
            roofline["note"] = ("this kernel is bound by shared-memory lookups (T-table AES), not by HBM: lsu_bound is its roofline; "
                                "the hbm figures are the required keys")
        try:   # DRAM bytes per launch of this kernel from the committed ncu --set full capture, if the shape matches
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tr = json.load(f).get(dominant)
            if tr and "lsu_bound" in roofline and tr.get("l1tex_data_stage"):
                roofline["lsu_bound"] = dict(roofline["lsu_bound"], ncu_l1tex_data_stage=tr["l1tex_data_stage"])
            if tr and tr.get("pairs") is None:
                # a capture is evidence for the shape it was taken on only: rows, columns, providers and ring width must match
                if (tr.get("n_rows") == n_rows and tr.get("ring_bits") == wl["bits_p1"] and tr.get("d") == d
                        and tr.get("providers", p) == p):
                    roofline["traffic"] = tr["bytes_per_launch"]
                    roofline["traffic_source"] = tr["source"]
                else:
                    roofline["traffic_source"] = "no ncu --set full capture of %s at this shape (profiles/ncu_traffic.json holds: %s)" % (dominant, tr.get("launch"))
            elif tr and tr["n_rows"] == n_rows and tr["ring_bits"] == wl["bits_p1"]:
                # captured for tr["pairs"] pairs per launch; the traffic is linear in the pairs of a launch
                per_launch = C / max(kd["launches_per_step"], 1)
                roofline["traffic"] = tr["bytes_per_launch"] / tr["pairs"] * per_launch
                roofline["traffic_source"] = tr["source"] + "; scaled from %d to the %.1f pairs of an average launch of this step" % (tr["pairs"], per_launch)
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
        del Xh, yh

    # ---- parity of the timed configuration (outside the timed region, every N; collective)
    check = None
    if not args.no_check:
        from oracle import pyoracle as po
        from tools.parity_check import check_run
        shares = step(Xd, yd)
        only = None
        if use_ot:      # the sample covers the first C inner products of the schedule: those slots, plus every local pair
            done, only = 0, []
            for ps, pr, items in box.ot_schedule():
                for (_, _, slot) in items:
                    if done < C:
                        only.append(slot)
                        done += 1
            L1 = box.sched.kind == 1
            only = np.unique(np.concatenate([np.array(only, dtype=np.int64), box.sched.slot[L1]]))
        # masking with the single-key stream: both providers' share words of 32 cross pairs of the TIMED schedule replayed
        # on the CPU at their true stream positions (row shards: partial dots, ring sum over ranks)
        check = check_run(lib, po.Oracle(), box, shares, Xd, yd, samples=16, diag_samples=4, only_slots=only,
                          ti_key=None if (use_ot or seeded) else key, pairs_done=C, cross_samples=32)
        check["shares_nonzero"] = bool(np.any(shares))
        check["what"] = ("sum over providers of every pre-shift share word == one-provider tcgen05 Gram (all off-diagonal + share_b words), "
                         "all diagonal words and statistics, sampled words vs the CPU oracle's ring dot / sequential double diagonal; "
                         "sampled_cross_pair_*: both providers' masked share words of sampled cross pairs vs oracle.ti_pair + the "
                         "party algebra at the pair's true stream offset (src/phase1.c:186-222,270-278)")

    # ---- the same protocol with every message through the protobuf wire format on the GPU (what separate
    # boxes execute per pair, minus the network): a bounded sample of the cross pairs, extrapolated
    wire = None
    if args.wire_sample > 0 and C_all and not use_ot:
        npairs = min(C_all, args.wire_sample)

        def wire_step(wv):
            box.cross_pairs_masking(key, seeded=seeded, wire=wv, max_pairs=npairs, fused=(fused and not wv))
            box.sync()
        wire_step(True)
        ms_w, _ = timed(lambda: wire_step(True), 1)
        ms_p, _ = timed(lambda: wire_step(False), 1)
        ms_all_wire = ms_step + (ms_w - ms_p) / npairs * C_all
        wire = {"pairs": npairs, "ms_with_wire": ms_w, "ms_plain": ms_p, "us_per_pair_with_wire": ms_w * 1e3 / npairs,
                "us_per_pair_plain": ms_p * 1e3 / npairs, "all_cross_pairs_ms_extrapolated": ms_w / npairs * C_all,
                "step_ms_with_wire_extrapolated": ms_all_wire, "value_with_wire": macs / (ms_all_wire * 1e-3) / 1e9,
                "messages_per_pair": 4, "note": "split TI / party kernels with p1_msg_pack + p1_msg_unpack around each of the 4 messages of a pair "
                "(src/phase1.c:100-145): what parties on separate boxes execute, minus the network; the headline step keeps all "
                "parties on one device and hands messages over in HBM"}

    # ---- BASELINE config 3 as written (--use_ot) inside the default line, at every N: the reference's OT schedule
    # (src/phase1.c:386-449) over enough inner products that session set-up is a small, separately reported share; the
    # whole workload is the same loop 100 400 times (--aggregation ot runs it as the step itself)
    ot = None
    if args.ot_sample > 0 and C_all and not use_ot:
        P_ot = min(C_all, args.ot_sample * world)         # strong scaling: rows shrink with N, so the pair count grows with it
        box.cross_pairs_ot(ot_key, max_pairs=min(P_ot, 2 * args.pipes))    # warm-up: buffers, pads, both kernels
        box.reset()
        box.encode(Xd, yd)
        box.local_blocks()
        ms_ot, wall_ot = timed(lambda: box.cross_pairs_ot(ot_key, max_pairs=P_ot), 1)
        setup = torch.tensor([box.ot_setup_s * 1e3], dtype=torch.float64, device=dev)
        if group is not None:
            dist.all_reduce(setup, op=dist.ReduceOp.MAX)
        setup_ms = float(setup[0])
        box.all_reduce()
        ot_shares = box.finalize()
        c_s = P_ot * n * wl["bits_p1"]
        total_cots = C_all * n * wl["bits_p1"]
        rate = c_s / (max(ms_ot - setup_ms, 1e-3) * 1e-3)
        ot = {"inner_products": P_ot, "sessions": box.ot_sessions, "cots": c_s, "ms": ms_ot, "wall_ms": wall_ot,
              "session_setup_ms": setup_ms, "session_setup_frac": setup_ms / ms_ot,
              "cot_per_s": rate, "cot_per_s_incl_setup": c_s / (ms_ot * 1e-3), "s_per_1000_pairs": (ms_ot - setup_ms) / P_ot,
              "wire_bytes_per_cot": 16 + w, "workload_cots": total_cots,
              "workload_sessions": p * (p - 1) // 2,
              "workload_seconds_extrapolated": total_cots / rate + p * (p - 1) // 2 * setup_ms * 1e-3 / max(box.ot_sessions, 1),
              "published": PUBLISHED.get(args.workload, {}).get("use_ot_wall_s_extrapolated") if PUBLISHED.get(args.workload) else None,
              "note": "device-timed (CUDA events on the main stream, max over ranks) over the first %d inner products of the reference's OT "
                      "schedule; session_setup_ms = host time of derive_cot_session + p1_cot_*_new + p1_cot_free (max over ranks), reported apart "
                      "and excluded from cot_per_s; base OTs are inputs (host/base_ot.c in the drop-in)" % P_ot}
        if not args.no_check:
            from oracle import pyoracle as po
            from tools.parity_check import check_run
            done, only = 0, []
            for ps_, pr_, items in box.ot_schedule():
                for (_, _, slot) in items:
                    if done < P_ot:
                        only.append(slot)
                        done += 1
            only = np.unique(np.concatenate([np.array(only, dtype=np.int64), box.sched.slot[box.sched.kind == 1]]))
            ck = check_run(lib, po.Oracle(), box, ot_shares, Xd, yd, samples=4, diag_samples=1, only_slots=only)
            ot["check"] = {k: ck[k] for k in ("words_checked", "sum_invariant_all_words", "all_diagonals_match", "ok")}
            ot["check"]["what"] = ("the two providers' OT shares of every executed inner product sum to the plain ring Gram word (+ all local words); "
                                   "share words of whole --use_ot runs are compared with the reference word for word in tests/test_gpu_fullsize.py at n = 100 000")

    # ---- the one-provider Gram of this instance (the p = 1 full-Gram figure; north_star: >= 60 % of the dense
    # int8 tensor peak): one provider owning all d + 1 columns, tcgen05 kernel forced, the kernel's own launch time from event
    # brackets - alone (first launch after a sync) and back to back (5 launches in a row)
    gram_p1 = None
    if args.gram_p1 and not use_ot and d >= 16:
        one = Phase1Box(lib, n, d, [0], wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"], device=local,
                        row0=row0, n_rows=n_rows, pipes=1, gram_algo=pkg.ffi.P1_GRAM_TCGEN05, group=group)
        one.reset()
        one.encode(Xd, yd)
        one.local_blocks()
        one.sync()
        Lk = ctypes.c_int(0)
        lib.check(lib.p1_gram_digits(one.ctxs[0], one.data, 0, d, 1, ctypes.byref(Lk)))
        m1 = d + 1
        ops1 = 2.0 * Lk.value ** 2 * n_rows * (m1 * (m1 - 1) // 2)

        def gram_launch_ms(reps):
            lib.check(lib.p1_ctx_profile(one.ctxs[0], 1))
            barrier()
            for _ in range(reps):
                one.local_blocks()
            one.sync()
            ms, nl = ctypes.c_double(), ctypes.c_uint64()
            lib.check(lib.p1_ctx_profile_read(one.ctxs[0], pkg.ffi.P1_K_GRAM_TC, ctypes.byref(ms), ctypes.byref(nl)))
            lib.check(lib.p1_ctx_profile(one.ctxs[0], 0))
            return ms.value, nl.value
        ms_a, nl_a = gram_launch_ms(1)
        ms_b, nl_b = gram_launch_ms(5)
        v = torch.tensor([ms_a / max(nl_a, 1), (ms_b - ms_a) / max(nl_b - nl_a, 1)], dtype=torch.float64, device=dev)
        if group is not None:
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
        alone, b2b = float(v[0]), float(v[1])
        one.close()
        gram_p1 = {"columns": m1, "rows_per_rank": n_rows, "digits": Lk.value, "int8_ops_per_launch_per_rank": ops1,
                   "ms_alone": alone, "ms_back_to_back": b2b}
        for tag, ms_ in (("alone", alone), ("back_to_back", b2b)):
            tops = ops1 / (ms_ * 1e-3) / 1e12
            gram_p1[tag] = {"tops_int8_credited_per_gpu": tops, "frac_of_measured_int8_peak": tops / int8_peak,
                            "frac_of_nominal_dense_int8": tops / int8_nominal,
                            "ring_gmacs_all_gpus": world * n_rows * (m1 * (m1 - 1) // 2) / (ms_ * 1e-3) / 1e9}
        gram_p1["note"] = ("credited = 2 L^2 int8 ops per WANTED ring MAC (lower triangle only, no padding); the kernel also computes the upper "
                           "halves of the diagonal tiles and the 501 -> 512 padding, which is why a tensor pipe that ncu shows 92 %% active credits less; "
                           "measured peak %.0f TOP/s (p1_probe_tensor_i8 in this run), nominal 4500" % int8_peak)

    # ---- CPU baseline (rank 0, N = 1 only): the reference's own phase-1 code on a bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        smp = ReferenceSample(wl, 5 if use_ot else 20, n_sample=min(n, 131072) if use_ot else None, use_ot=use_ot)
        secs, ref_shares = smp.run()
        cpu = {"value": smp.macs / secs / 1e9, "unit": UNIT, "cores": smp.threads, "kind": smp.kind,
               "sample": smp.describe(), "seconds": secs, "host_cores": os.cpu_count()}
        # the same sample through the GPU path: every share word must match the reference's
        sbox = Phase1Box(lib, smp.n, smp.d, smp.cs, wl["precision"], wl["precision_p2"], wl["bits_p1"], wl["bits_p2"],
                         device=local, batch_pairs=8)
        got = sbox.run_ot(smp.X, smp.y, ot_key) if use_ot else sbox.run_masking(smp.X, smp.y, key)
        sbox.close()
        cpu["gpu_shares_match_sample"] = bool(np.array_equal(got, ref_shares))
        smp.close()

    box.close()

    # ---- the drop-in boundary itself (rank 0, N = 1): the reference's run_party / run_trusted_initializer entry points
    # served by linreg-mpc_b200/host/phase1_gpu.c vs the reference's own phase1.c, same loop-back driver, same instance file
    dropin = None
    if rank == 0 and world == 1 and not args.no_dropin and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libp1gpuhost_64_64.so")):
        from tools.dropin_timing import measure as dropin_measure
        torch.cuda.empty_cache()
        dropin = dropin_measure(1_000_000, 10, 3, use_ot=False, bits=64)

    if rank == 0:
        value = macs / (ms_step * 1e-3) / 1e9
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "int64" if w == 8 else "int32", "data": "synthetic",
                "config": dict(cfg, rows_per_rank=per, batch_pairs=batch, gram=args.gram, pipes=args.pipes,
                               cross_pairs_per_step=C, cross_pairs_of_workload=C_all,
                               l2="inputs (%.1f GB of columns + %.1f GB of messages per step) exceed the 126 MB L2"
                                  % ((d + 1) * nb / 1e9, C * (2 if not use_ot else 3 * 8) * nb / 1e9)),
                "ring_macs_per_step": macs, "share_time_ms": ms_step, "wall_ms_per_step": ms_wall,
                "gram_p1_equiv_macs": macs_gram_p1,
                "gpu_launches": int(launches) * world, "kernels": kernels,
                "kernels_note": "per-launch CUDA events of one extra step on a single pipe (%.1f ms: kernels back to back on one stream, "
                                "event overhead included); the timed steps run %d pipes" % (ms_prof, args.pipes),
                "roofline": roofline, "clocks": clocks,
                "sustained": sustained, "int8_peak_probe": int8,
                "e2e": e2e, "e2e_dropin": dropin, "wire_format_sample": wire, "ot_workload": ot, "gram_p1": gram_p1, "cpu_baseline": cpu, "check": check}
        line["published"] = PUBLISHED.get(args.workload)
        if wire:
            line["value_with_wire"] = wire["value_with_wire"]
        if use_ot:
            cot_total = C * n * wl["bits_p1"]
            line["use_ot"] = {"cot_per_s": cot_total / (ms_step * 1e-3), "cots_per_step": cot_total, "inner_products_per_step": C,
                              "wire_bytes_per_cot": 16 + w,
                              "workload_cots": C_all * n * wl["bits_p1"],
                              "workload_seconds_extrapolated": ms_step * 1e-3 / max(C, 1) * C_all,
                              "note": "the step includes encode, the local blocks and the finalize tail; the extrapolation to all %d "
                                      "cross pairs of the workload is linear in the pair count" % C_all}
        print(json.dumps(line))
    if group is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
