"""CPU suite, part 2: host-side logic of the package (schedule, index maps) against the
oracle, and the C-ABI library's exported surface.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import numpy as np
import pytest

import linreg_mpc_b200 as pkg
from linreg_mpc_b200 import ffi, phase1
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("d,cs", [(5, [0, 1, 2]), (12, [0, 3, 7]), (9, [0, 2, 4, 6]), (7, [0]), (6, [0, 3, 3]),
                                  (100, [0, 33, 66])])
def test_schedule_matches_oracle(oracle, d, cs):
    s = phase1.Schedule(d, cs)
    o = oracle.schedule(d, cs)
    assert np.array_equal(s.i, o["i"]) and np.array_equal(s.j, o["j"])
    assert np.array_equal(s.owner_i, o["owner_i"]) and np.array_equal(s.owner_j, o["owner_j"])
    assert np.array_equal(s.kind, o["kind"])
    assert np.array_equal(np.flatnonzero(s.kind == 2), s.cross)
    # slots: idx(i,j) for feature rows, nA + j for the target row (src/phase1.c:581-585)
    nA = d * (d + 1) // 2
    for t in range(len(s.i)):
        i, j = int(s.i[t]), int(s.j[t])
        want = oracle.lib.p1o_idx(i, j) if i < d else nA + j
        assert s.slot[t] == want


def test_idx_is_symmetric_and_packed():
    i = np.array([0, 3, 1, 499]); j = np.array([0, 1, 3, 498])
    assert list(phase1.idx(i, j)) == [0, 7, 7, 499 * 500 // 2 + 498]


def test_owner_rule_target_goes_to_last_provider():
    assert list(phase1.owners([0, 1, 2, 3, 4, 5], [0, 1, 2])) == [0, 1, 2, 2, 2, 2]
    with pytest.raises(ValueError):
        phase1.owners([0], [3, 1])


def test_ring_mac_count():
    s = phase1.Schedule(500, po.even_col_start(500, 5))
    assert s.num_cross == 100400
    assert s.ring_macs(10 ** 6) == 10 ** 6 * (24850 + 3 * 100400)


def test_library_loads_and_exports_every_declared_symbol():
    pkg.build()
    lib = pkg.P1Lib()
    header = open(os.path.join(ROOT, "include", "p1gpu.h")).read()
    declared = set(re.findall(r"\b(p1_[a-z0-9_]+)\s*\(", header))
    assert declared == set(ffi.SIGNATURES), declared ^ set(ffi.SIGNATURES)
    raw = ctypes.CDLL(lib.path)
    for name in sorted(declared):
        assert getattr(raw, name) is not None


def test_no_cpu_fallback_without_gpu():
    """the product path fails loudly when no CUDA device is visible"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = pkg.P1Lib()
    ctx = ctypes.c_void_p()
    assert lib.p1_ctx_create(0, 64, 64, ctypes.byref(ctx)) == 1
    assert b"no CUDA device" in lib.lib.p1_last_error()
    with pytest.raises(ffi.P1Error):
        lib.check(lib.p1_ctx_create(0, 64, 64, ctypes.byref(ctx)))


def test_missing_library_raises():
    with pytest.raises(ffi.P1Error):
        pkg.P1Lib("/nonexistent/libp1gpu.so")


def test_product_package_does_not_import_oracle():
    for root, _, files in os.walk(os.path.join(ROOT, "linreg-mpc_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".h", ".cpp")):
                text = open(os.path.join(root, f)).read()
                assert "pyoracle" not in text and "liboracle" not in text and "p1_oracle" not in text, f


def test_pow5_table_is_current_and_conversion_matches_python_float():
    """the Eisel-Lemire constants of k_text.cu (f4) are exactly what tools/gen_pow5_table.py
    derives, and the integer restatement of the device conversion agrees with Python's
    correctly rounded float() (the same result glibc's strtod gives fscanf("%lf"))."""
    import random
    import struct
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_pow5_table as g
    lines = [l for l in open(os.path.join(ROOT, "linreg-mpc_b200", "csrc", "pow5_table.inc")) if l.startswith("{")]
    assert len(lines) == len(g.TABLE) == 651
    for l, t in zip(lines, g.TABLE):
        hi, lo = (int(v, 16) for v in re.findall(r"0x([0-9a-f]{16})", l))
        assert (hi << 64) | lo == t
    rnd = random.Random(3)
    for _ in range(20000):
        w = rnd.randint(1, 10 ** rnd.randint(1, 19) - 1)
        q = rnd.randint(-345, 310)
        try:
            want = float("%de%d" % (w, q))
        except OverflowError:
            want = float("inf")
        assert struct.pack("<d", g.el_double(w, q)) == struct.pack("<d", want), (w, q)


def test_dropin_host_file_exports_the_reference_entry_points():
    """linreg-mpc_b200/host/phase1_gpu.c, built against the reference's headers (oracle/Makefile gpuhost),
    defines exactly the two symbols the reference's callers link against (src/phase1.h:14-29) plus the
    hooks it documents; its compute calls resolve into libp1gpu.so (no second implementation inside)."""
    import subprocess
    for widths in ("64_64", "32_32", "64_32"):
        path = os.path.join(ROOT, "oracle", "_ref", "libp1gpuhost_%s.so" % widths)
        if not os.path.exists(path):
            pytest.skip("oracle/_ref/libp1gpuhost_* is built only where the reference sources are present")
        syms = subprocess.run(["nm", "-D", path], capture_output=True, text=True, check=True).stdout
        defined = {l.split()[-1] for l in syms.splitlines() if " T " in l or " W " in l}
        undefined = {l.split()[-1] for l in syms.splitlines() if " U " in l}
        assert {"run_party", "run_trusted_initializer"} <= defined
        assert {"p1_group_create", "p1_ti_generate", "p1_group_ti_generate", "p1_msg_pack", "p1_msg_unpack", "p1_parse_doubles",
                "p1_group_encode", "p1_group_finalize", "p1_party_a_step", "p1_group_party_a_step", "p1_group_local_block",
                "p1_group_cot_sender_step", "p1_group_cot_receiver_extend"} <= undefined
        lib = ctypes.CDLL(path)     # resolves libp1gpu.so through its rpath; needs no GPU to load
        assert lib.run_party is not None and lib.run_trusted_initializer is not None


def test_default_base_ot_provider_of_the_dropin():
    """linreg-mpc_b200/host/base_ot.c, the hook behind honestOTExt{Sender,Recver}New (src/phase1.c:394,397) in the
    drop-in host file: SHA-256 against hashlib, the Edwards25519 ladder against big-integer arithmetic, and the
    Chou-Orlandi exchange itself over a loop-back connection - the extension sender ends up with exactly
    k_i^{Delta_i} of the receiver's pairs and with nothing that matches the other key"""
    import ctypes
    import hashlib
    path = os.path.join(ROOT, "oracle", "_ref", "libp1gpuhost_64_64.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libp1gpuhost_* is built only where the reference sources are present")
    L = ctypes.CDLL(path)
    rng = np.random.default_rng(3)
    for n in (0, 3, 55, 56, 63, 64, 65, 300):
        msg = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        out = (ctypes.c_uint8 * 32)()
        L.p1host_sha256(msg, len(msg), out)
        assert bytes(out) == hashlib.sha256(msg).digest()
    p = 2 ** 255 - 19
    dd = (-121665 * pow(121666, -1, p)) % p

    def add(P, Q):
        (x1, y1), (x2, y2) = P, Q
        t = dd * x1 * x2 * y1 * y2 % p
        return ((x1 * y2 + x2 * y1) * pow(1 + t, -1, p) % p, (y1 * y2 + x1 * x2) * pow(1 - t, -1, p) % p)

    def mul(s, P):
        R = (0, 1)
        while s:
            if s & 1:
                R = add(R, P)
            P = add(P, P)
            s >>= 1
        return R
    y = 4 * pow(5, -1, p) % p
    x2 = (y * y - 1) * pow(dd * y * y + 1, -1, p) % p
    x = pow(x2, (p + 3) // 8, p)
    if (x * x - x2) % p:
        x = x * pow(2, (p - 1) // 4, p) % p
    if x & 1:
        x = p - x
    for _ in range(3):
        s = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
        out = (ctypes.c_uint8 * 64)()
        L.p1host_ed25519_scalarmult_base(s, out)
        X, Y = mul(int.from_bytes(s, "little"), (x, y))
        assert bytes(out) == X.to_bytes(32, "little") + Y.to_bytes(32, "little")
    s0, s1, ss = (np.zeros((128, 16), np.uint8) for _ in range(3))
    dl = np.zeros(16, np.uint8)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    assert L.p1ref_base_ot_selftest(vp(s0), vp(s1), vp(ss), vp(dl)) == 0
    bits = np.unpackbits(dl, bitorder="little").astype(bool)
    assert 20 < bits.sum() < 108
    assert np.array_equal(np.where(bits[:, None], s1, s0), ss)
    assert not np.any(np.all(np.where(bits[:, None], s0, s1) == ss, axis=1))
    assert len({bytes(r) for r in np.concatenate([s0, s1])}) == 256


def test_seeded_mask_messages_of_the_dropin():
    """host/phase1_gpu.c with P1GPU_SEEDED_MASKS=1: the TI's value-only message is SecureMultiplication.Msg with an
    empty vector (field 2, varint) and round-trips for every width; malformed messages are refused; the per-provider
    key is SHA-256(label || K || LE32(p))[0..16)"""
    import ctypes
    import hashlib
    path = os.path.join(ROOT, "oracle", "_ref", "libp1gpuhost_64_64.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libp1gpuhost_* is built only where the reference sources are present")
    L = ctypes.CDLL(path)
    L.p1host_selftest_value_msg.restype = ctypes.c_size_t
    L.p1host_selftest_value_msg.argtypes = [ctypes.c_uint64, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_int)]
    L.p1host_selftest_value_msg_decode.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]

    def varint(v):
        out = bytearray()
        while True:
            b = v & 0x7f
            v >>= 7
            out.append(b | (0x80 if v else 0))
            if not v:
                return bytes(out)
    for v in (0, 1, 127, 128, 300, 2 ** 32 - 1, 2 ** 32, 2 ** 56 - 1, 2 ** 63, 2 ** 64 - 1):
        buf = (ctypes.c_uint8 * 16)()
        back, st = ctypes.c_uint64(), ctypes.c_int(-1)
        n = L.p1host_selftest_value_msg(v, buf, ctypes.byref(back), ctypes.byref(st))
        assert bytes(buf[:n]) == b"\x10" + varint(v) and st.value == 0 and back.value == v
    val = ctypes.c_uint64()
    for bad in (b"", b"\x10", b"\x08\x01", b"\x10\x80", b"\x10\x01\x01", b"\x10" + b"\xff" * 10 + b"\x01"):
        assert L.p1host_selftest_value_msg_decode(bad, len(bad), ctypes.byref(val)) != 0, bad
    key = bytes(range(16))
    for party in (2, 3, 21):
        out = (ctypes.c_uint8 * 16)()
        L.p1host_selftest_party_key(key, party, out)
        assert bytes(out) == hashlib.sha256(b"linreg-mpc p1gpu seeded mask key" + key + party.to_bytes(4, "little")).digest()[:16]


def test_bench_instance_is_the_same_matrix_for_every_row_shard():
    """bench.py's synthetic instance comes from one numpy stream per 4096-row chunk: a rank that generates
    only its shard gets exactly its rows of the matrix a single GPU generates, whatever N is"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    d, n = 7, 3 * b.GEN_CHUNK + 300
    X, e = b.synthetic_rows(3, 0, n, d, threads=2)
    for row0, rows in ((256, 5000), (b.GEN_CHUNK, b.GEN_CHUNK), (2 * b.GEN_CHUNK + 256, n - 2 * b.GEN_CHUNK - 256)):
        Xs, es = b.synthetic_rows(3, row0, rows, d, threads=3)
        assert np.array_equal(Xs, X[row0:row0 + rows]) and np.array_equal(es, e[row0:row0 + rows])
    assert not np.array_equal(b.synthetic_rows(4, 0, 256, d)[0], X[:256])      # another config, another matrix
