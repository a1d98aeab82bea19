"""World-size-2 `gloo` tests (CPU) of the host-side row-shard logic of SURVEY.md 8e:
shard boundaries, the wrapping integer all-reduce that combines partial shares, the
rank-ordered chain of the sequential double diagonal, and bench.py's rank handling of
the reference arm.  The per-shard partial sums come from the CPU oracle (the checker);
on the GPU box the same helpers run over NCCL with the kernels' accumulators."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_rows_cover_and_align():
    from linreg_mpc_b200.phase1 import shard_rows
    for n in (1, 255, 256, 257, 1000, 4096, 1_000_000, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            shards = [shard_rows(n, world, r) for r in range(world)]
            assert sum(k for _, k in shards) == n
            pos = 0
            for row0, k in shards:
                assert row0 == pos or k == 0
                assert row0 % 256 == 0 or k == 0     # 16-byte aligned mask slices, whole OT tiles
                pos += k


def _worker(rank, world, port, tmp):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from linreg_mpc_b200.phase1 import shard_rows, ring_all_reduce, chain_in_rank_order, Schedule
    from oracle import pyoracle as po
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    orc = po.Oracle()
    n, d, prec = 2000, 6, 56
    X, y, _ = po.generate_lin_regression(n, d, seed=5)
    norm = orc.normalizer(prec, n)
    Xw = orc.encode(X, norm, prec).reshape(n, d)
    yw = orc.encode(y, norm, prec)
    row0, k = shard_rows(n, world, rank)
    S = Schedule(d, [0])                       # one provider: every pair is local
    # partial ring sums of this shard (what k_gram_* / k_party leave in the accumulator)
    part = np.zeros(S.nA + d, dtype=np.uint64)
    for i, j, slot, kind in zip(S.i, S.j, S.slot, S.kind):
        if kind == 1:
            a = Xw[row0:row0 + k, i] if i < d else yw[row0:row0 + k]
            part[slot] = np.uint64(orc.dot(np.ascontiguousarray(a), np.ascontiguousarray(Xw[row0:row0 + k, j])))
    # make the wrap-around visible: add a large constant on every rank
    part += np.uint64(0xC000000000000000)
    flat = torch.from_numpy(part.view(np.int64).copy())
    ring_all_reduce(flat, dist.group.WORLD)
    got = flat.numpy().view(np.uint64)

    # sequential double diagonal of column 0, chained in rank order (src/phase1.c:563-566)
    scale = float(1 << prec)

    def step(xy):
        acc = float(xy[0])
        for wv in Xw[row0:row0 + k, 0].view(np.int64):
            v = float(wv) / scale
            acc = acc + (v * v) * scale
        return np.array([acc])
    diag = chain_in_rank_order(step, 1, dist.group.WORLD)
    if rank == 0:
        with open(tmp, "w") as f:
            json.dump({"acc": [int(v) for v in got], "diag": float(diag[0])}, f)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_combine_exactly(tmp_path, oracle):
    import torch.multiprocessing as mp
    from oracle import pyoracle as po
    from linreg_mpc_b200.phase1 import Schedule
    out = str(tmp_path / "res.json")
    port = _free_port()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    res = json.load(open(out))
    n, d, prec = 2000, 6, 56
    X, y, _ = po.generate_lin_regression(n, d, seed=5)
    norm = oracle.normalizer(prec, n)
    Xw = oracle.encode(X, norm, prec).reshape(n, d)
    yw = oracle.encode(y, norm, prec)
    S = Schedule(d, [0])
    want = np.zeros(S.nA + d, dtype=np.uint64)
    for i, j, slot, kind in zip(S.i, S.j, S.slot, S.kind):
        if kind == 1:
            a = Xw[:, i] if i < d else yw
            want[slot] = np.uint64(oracle.dot(np.ascontiguousarray(a), np.ascontiguousarray(Xw[:, j])))
    with np.errstate(over="ignore"):
        want += np.uint64(2) * np.uint64(0xC000000000000000)    # wraps mod 2^64
    assert [int(v) for v in want] == res["acc"]
    # the chained diagonal equals the unsharded sequential loop bit for bit
    scale = float(1 << prec)
    acc = 0.0
    for wv in Xw[:, 0].view(np.int64):
        v = float(wv) / scale
        acc = acc + (v * v) * scale
    assert acc == res["diag"]


def test_reference_arm_other_ranks_exit_quietly():
    """under torchrun only rank 0 runs the CPU reference arm; the others print nothing and exit 0"""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
