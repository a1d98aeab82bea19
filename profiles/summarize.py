#!/usr/bin/env python
"""Turn ncu output brought back in gpurun_out/ into the small text summaries kept under profiles/.

  python profiles/summarize.py full  gpurun_out/X.ncu-rep  profiles/X_full.txt     (ncu --set full capture)
  python profiles/summarize.py list  gpurun_out/X_launches.csv profiles/X_launches.txt  (launch list -> per-kernel shares)
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct",
    "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct",
    "smsp__warp_issue_stalled_wait_per_warp_active.pct", "smsp__warp_issue_stalled_no_instruction_per_warp_active.pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
]


def full(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write("# source: %s (ncu --set full --clock-control none), one block per captured launch\n" % rep)
        for r in rows[2:]:
            f.write("\n== %s\n" % r[idx["Kernel Name"]])
            for m in METRICS:
                if m in idx and r[idx[m]] != "":
                    f.write("%-75s %s %s\n" % (m, r[idx[m]], units[idx[m]]))
            # any other tensor-pipe activity counter (not the constant peak_sustained / attribute columns)
            tens = [h for h in hdr if "tensor" in h and h not in METRICS and "peak_sustained" not in h.split(".")[-1]
                    and not h.startswith("device__") and "ops_path" not in h and r[idx[h]] not in ("", "0")]
            for m in tens[:12]:
                f.write("%-75s %s %s\n" % (m, r[idx[m]], units[idx[m]]))


def launches(path, out):
    txt = open(path).read()
    start = txt.find('"ID"')
    rows = list(csv.DictReader(io.StringIO(txt[start:])))
    agg = OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# source: %s (ncu --metrics gpu__time_duration.sum --clock-control none); cold-cache, serialised: compare shares\n" % path)
        f.write("%-70s %8s %12s %10s %7s\n" % ("kernel", "launches", "total_us", "avg_us", "share"))
        for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%-70s %8d %12.1f %10.2f %6.1f%%\n" % (k, n, us, us / n, 100 * us / tot))
        f.write("%-70s %8d %12.1f\n" % ("TOTAL", sum(a[0] for a in agg.values()), tot))


if __name__ == "__main__":
    {"full": full, "list": launches}[sys.argv[1]](sys.argv[2], sys.argv[3])
