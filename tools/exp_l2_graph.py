#!/usr/bin/env python
"""Experiment (not a benchmark): does the masking loop gain from batches small enough for the four message
buffers to stay in the 126 MB L2?  Small batches are launch-bound from Python, so the loop over PAIRS cross
pairs is captured into a CUDA graph (both pipes) and replayed; prints us per pair for the real step."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import linreg_mpc_b200 as pkg  # noqa: E402
from linreg_mpc_b200.phase1 import Phase1Box  # noqa: E402
from tools.exp_overlap import loop  # noqa: E402


def main():
    lib = pkg.P1Lib()
    n, d, p = int(os.environ.get("N", "1000000")), 500, 5
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(1)
    X = torch.randn((n, d), dtype=torch.float64, device=dev, generator=g)
    X /= X.abs().amax(dim=0)
    y = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    pipes = int(os.environ.get("PIPES", "2"))
    batch = int(os.environ.get("BATCH", "2"))
    pairs = int(os.environ.get("PAIRS", "528"))
    mode = os.environ.get("MODE", "full")
    box = Phase1Box(lib, n, d, [k * (d // p) for k in range(p)], 56, batch_pairs=batch, pipes=pipes)
    box.reset(); box.encode(X, y); box.sync()
    key = bytes(range(16))
    frac = pairs / box.sched.num_cross
    loop(box, key, mode, frac)                       # eager once: buffers, function attributes
    s0 = torch.cuda.ExternalStream(int(lib.p1_ctx_stream(box.ctxs[0])), device=dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=s0, capture_error_mode="relaxed"):
        loop_nosync(box, key, mode, frac)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = int(os.environ.get("REPS", "20"))
    with torch.cuda.stream(s0):
        graph.replay()
        e0.record()
        for _ in range(reps):
            graph.replay()
        e1.record()
    torch.cuda.synchronize()
    C = int(box.sched.num_cross * frac)
    print(json.dumps({"pipes": pipes, "batch": batch, "mode": mode, "pairs": C, "us_per_pair": round(e0.elapsed_time(e1) * 1e3 / reps / C, 3)}))
    box.close()


def loop_nosync(box, key, mode, frac):
    sync = box.sync
    box.sync = lambda: None
    try:
        loop(box, key, mode, frac)
    finally:
        box.sync = sync


if __name__ == "__main__":
    main()
