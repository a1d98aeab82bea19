"""The drop-in itself: linreg-mpc_b200/host/phase1_gpu.c exports the reference's own entry points
(run_trusted_initializer / run_party, src/phase1.h:14-29).  Built against the reference's headers and
the loop-back shims, it runs every party of an instance as a thread - exactly like the build of the
reference's unmodified phase1.c next to it (oracle/_ref) - with all arithmetic, the text ingestion and
the protobuf (de)serialisation on the GPU.  Every share word of every provider must equal the
reference's, and the bytes on every socket must add up to the same totals."""
import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

CASES = [  # b1, b2, precision, precision_p2
    (64, 64, 56, 0), (32, 32, 24, 0), (64, 32, 56, 24)]


def both(path, d, num_dp, b1, b2, prec, prec2, use_ot):
    if not (po.RefLib.available(b1, b2) and po.RefLib.available(b1, b2, gpuhost=True)):
        pytest.skip("oracle/_ref was not prebuilt (needs the reference sources: build container only)")
    ref = po.RefLib(b1, b2).run(path, d, num_dp, prec, prec2, use_ot=use_ot)
    gpu = po.RefLib(b1, b2, gpuhost=True).run(path, d, num_dp, prec, prec2, use_ot=use_ot)
    return ref, gpu


@pytest.mark.parametrize("b1,b2,prec,prec2", CASES)
def test_run_party_masking_matches_reference(p1lib, tmp_path, b1, b2, prec, prec2, monkeypatch):
    monkeypatch.setenv("P1GPU_BATCH", "3")            # runs of 4 columns -> batches of 3 + 1
    n, d, cs = 1500, 12, [0, 4, 8]
    X, y, _ = po.generate_lin_regression(n, d, seed=21)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, rsent), (gs, _, gsent) = both(path, d, len(cs), b1, b2, prec, prec2, False)
    assert np.array_equal(gs, rs)
    assert np.array_equal(gsent, rsent), "bytes per socket differ from the reference's"


def test_run_party_matches_committed_reference_golden(p1lib, golden, tmp_path):
    """tests/golden/phase1_ref_shares.json: share words (and bytes per socket in masking mode) the
    reference's own phase1.c produced for examples/readme_example.in and the other committed
    instances, every width combination, masking and --use_ot"""
    ran = 0
    for k, c in enumerate(golden["cases"]):
        if not po.RefLib.available(c["bits_p1"], c["bits_p2"], gpuhost=True):
            pytest.skip("oracle/_ref/libp1gpuhost_* was not prebuilt (needs the reference headers: build container only)")
        path = str(tmp_path / ("g%d.in" % k))
        po.write_instance(path, c["X"], c["y"], c["col_start"])
        gs, _, sent = po.RefLib(c["bits_p1"], c["bits_p2"], gpuhost=True).run(
            path, c["d"], len(c["col_start"]), c["precision"], c["precision_p2"], use_ot=c["use_ot"],
            ti_key=golden["ti_key"], ot_key=golden["ot_key"])
        assert np.array_equal(gs, c["shares"]), "case %d (%s, use_ot=%s)" % (k, c["instance"], c["use_ot"])
        if not c["use_ot"]:
            assert np.array_equal(sent, np.array(c["bytes_sent"], dtype=np.uint64).reshape(sent.shape))
        ran += 1
    assert ran >= 6


@pytest.mark.parametrize("b1,b2,prec,prec2", CASES)
def test_run_party_use_ot_matches_reference(p1lib, tmp_path, b1, b2, prec, prec2):
    n, d, cs = 130, 5, [0, 2, 4]      # three providers: both role assignments of :392 occur
    X, y, _ = po.generate_lin_regression(n, d, seed=22)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, _), (gs, _, _) = both(path, d, len(cs), b1, b2, prec, prec2, True)
    assert np.array_equal(gs, rs)


def test_run_party_large_vectors_sequential_diagonal(p1lib, tmp_path):
    """words above 2^26 (tiny n): the exact diagonal shortcut does not apply and the literal loop runs"""
    n, d, cs = 40, 6, [0, 3]
    X, y, _ = po.generate_lin_regression(n, d, seed=23)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, _), (gs, _, _) = both(path, d, len(cs), 64, 64, 56, 0, False)
    assert np.array_equal(gs, rs)


def test_run_party_medium_instance_default_batch(p1lib, tmp_path):
    """four providers, runs longer than one message tile, default batching (whole runs per batch)"""
    n, d, cs = 20000, 16, [0, 4, 8, 12]
    X, y, _ = po.generate_lin_regression(n, d, seed=24)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, rsent), (gs, _, gsent) = both(path, d, len(cs), 64, 64, 56, 0, False)
    assert np.array_equal(gs, rs)
    assert np.array_equal(gsent, rsent)


def _devices(k):
    """k members for P1GPU_DEVICES: distinct GPUs when the box has them, else the same GPU k times (two
    contexts and two row slabs on one device: the sharding logic runs on the regular one-GPU box)"""
    import torch
    n = torch.cuda.device_count()
    return ",".join(str(g if n >= k else 0) for g in range(k))


@pytest.mark.parametrize("ndev", [2, 3])
@pytest.mark.parametrize("b1,b2,prec,prec2", CASES)
def test_run_party_row_sharded_over_devices(p1lib, tmp_path, monkeypatch, ndev, b1, b2, prec, prec2):
    """P1GPU_DEVICES lists several devices: run_party / run_trusted_initializer shard the rows over them behind
    the unchanged entry points (p1_group); shares and bytes per socket still equal the reference's"""
    monkeypatch.setenv("P1GPU_DEVICES", _devices(ndev))
    monkeypatch.setenv("P1GPU_BATCH", "3")
    n, d, cs = 2100, 12, [0, 4, 8]
    X, y, _ = po.generate_lin_regression(n, d, seed=25)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, rsent), (gs, _, gsent) = both(path, d, len(cs), b1, b2, prec, prec2, False)
    assert np.array_equal(gs, rs)
    assert np.array_equal(gsent, rsent), "bytes per socket differ from the reference's"


@pytest.mark.parametrize("b1,b2,prec,prec2", CASES)
def test_run_party_use_ot_row_sharded_over_devices(p1lib, tmp_path, monkeypatch, b1, b2, prec, prec2):
    monkeypatch.setenv("P1GPU_DEVICES", _devices(2))
    n, d, cs = 900, 5, [0, 2, 4]
    X, y, _ = po.generate_lin_regression(n, d, seed=26)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, _), (gs, _, _) = both(path, d, len(cs), b1, b2, prec, prec2, True)
    assert np.array_equal(gs, rs)


def test_run_party_sequential_diagonal_row_sharded(p1lib, tmp_path, monkeypatch):
    """large values: the sequential double diagonal chained over two row slabs in device order"""
    monkeypatch.setenv("P1GPU_DEVICES", _devices(2))
    n, d, cs = 700, 6, [0, 3]
    X, y, _ = po.generate_lin_regression(n, d, seed=27)
    X *= 300.0
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    (rs, _, _), (gs, _, _) = both(path, d, len(cs), 64, 64, 56, 0, False)
    assert np.array_equal(gs, rs)


def test_run_party_accepts_tokens_the_gpu_parser_defers(p1lib, tmp_path):
    """fscanf("%lf") accepts hex floats, more than 19 digits and infinities; the GPU parser defers them and the
    host file converts them with strtod (src/linear.c:46,82): same shares as the reference"""
    n, d, cs = 64, 4, [0, 2]
    X, y, _ = po.generate_lin_regression(n, d, seed=28)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    text = open(path).read().split("\n")
    row = text[6].split()            # first row of X (after "n d p", 2 + 2 endpoint lines, "n d")
    row[0] = float(X[0, 0]).hex()
    row[1] = "%.30f" % X[0, 1]
    X[0, 1] = float(row[1])
    row[2] = "inf"
    text[6] = " ".join(row)
    open(path, "w").write("\n".join(text))
    (rs, _, _), (gs, _, _) = both(path, d, len(cs), 64, 64, 56, 0, False)
    assert np.array_equal(gs, rs)


def _seeded_party_key(ti_key, node):
    """K_p of host/phase1_gpu.c (seeded masks): SHA-256(label || K || LE32(node index))[0..16)"""
    import hashlib
    return hashlib.sha256(b"linreg-mpc p1gpu seeded mask key" + bytes(ti_key) + int(node).to_bytes(4, "little")).digest()[:16]


@pytest.mark.parametrize("ndev", [1, 2])
@pytest.mark.parametrize("b1,b2,prec,prec2", CASES)
def test_run_party_seeded_masks(p1lib, oracle, tmp_path, monkeypatch, ndev, b1, b2, prec, prec2):
    """P1GPU_SEEDED_MASKS=1 (host half): the TI ships one 16-byte key per provider and then only the
    value of every pair; the providers expand their own mask on the GPU.  Checked word for word against the
    oracle's restatement of the seeded definition (x under the key of b, y under the key of a, r under the TI's
    key, same stream positions; share algebra of src/phase1.c:186-196,201-222), every other word against the
    reference's phase1.c; the TI's sockets carry O(1) bytes per pair."""
    if not (po.RefLib.available(b1, b2) and po.RefLib.available(b1, b2, gpuhost=True)):
        pytest.skip("oracle/_ref was not prebuilt (needs the reference sources: build container only)")
    monkeypatch.setenv("P1GPU_BATCH", "3")
    if ndev > 1:
        monkeypatch.setenv("P1GPU_DEVICES", _devices(ndev))
    n, d, cs = 1500, 12, [0, 4, 8]
    X, y, _ = po.generate_lin_regression(n, d, seed=29)
    path = str(tmp_path / "inst.in")
    po.write_instance(path, X, y, cs)
    ref = po.RefLib(b1, b2)
    rs, _, rsent = ref.run(path, d, len(cs), prec, prec2, use_ot=False)
    monkeypatch.setenv("P1GPU_SEEDED_MASKS", "1")
    gs, _, gsent = po.RefLib(b1, b2, gpuhost=True).run(path, d, len(cs), prec, prec2, use_ot=False)
    Xw, yw, _ = ref.read_inputs(path, n, d, prec)
    dt = po.word_dtype(b1)
    Xw, yw = Xw.astype(dt), yw.astype(dt)
    col = lambda c: yw if c == d else np.ascontiguousarray(Xw[:, c])
    sch = oracle.schedule(d, cs)
    nA = d * (d + 1) // 2
    mask = (1 << b1) - 1
    fin = (lambda v: oracle.downconvert(v, prec, prec2)) if b1 != b2 else (lambda v: v & mask)
    want = rs.copy()
    cross = 0
    with np.errstate(over="ignore"):
        for i, j, oi, oj, kind, q in zip(sch["i"], sch["j"], sch["owner_i"], sch["owner_j"], sch["kind"], sch["q"]):
            if kind != 2:
                continue
            cross += 1
            a, b = col(int(i)), col(int(j))
            xo = oracle.ti_pair(_seeded_party_key(po.TI_KEY, oj + 2), int(q), n, b1)[0]
            yo = oracle.ti_pair(_seeded_party_key(po.TI_KEY, oi + 2), int(q), n, b1)[1]
            r = oracle.ti_pair(po.TI_KEY, int(q), n, b1)[3]
            slot = int(i) * (int(i) + 1) // 2 + int(j) if i < d else nA + int(j)
            sa = (oracle.dot(b + xo, yo, b1) - (oracle.dot(xo, yo, b1) - r)) & mask
            sb = (oracle.dot(a - yo, b, b1) - r) & mask
            want[oi, slot] = fin(sa)
            want[oj, slot] = fin(sb)
    assert cross > 0
    assert np.array_equal(gs, want)
    if b1 == b2:   # the sum over the parties does not depend on the masks: the reference's, word for word
        assert np.array_equal(gs.sum(axis=0) & np.uint64(mask), rs.sum(axis=0) & np.uint64(mask))
    # the TI's sockets: a key per provider and <= 8 + 11 bytes per message, against ~9.5 n per message
    ti_bytes = int(gsent[0].sum())
    assert ti_bytes <= len(cs) * (8 + 16) + 2 * cross * (8 + 11)
    assert ti_bytes * 100 < int(rsent[0].sum())
