"""CPU suite, part 1: pins the oracle (oracle/p1_oracle.c) against
 - the committed golden vectors, which come from the reference's own phase1.c (tests/golden/make_golden.py),
 - the README known answer (examples/readme_example.in -> beta, README.md:83-87),
 - oracle/_ref itself when it is present (build container; also prebuilt on the GPU box),
 - published AES-128 known-answer vectors for the PRG."""
import math
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as po

WIDTHS = [(64, 64, 56, 0), (32, 32, 24, 0), (64, 32, 56, 24)]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_aes_known_answer(oracle):
    import ctypes
    # FIPS-197 Appendix C.1
    key = bytes(range(16))
    pt = bytes.fromhex("00112233445566778899aabbccddeeff")
    k = ctypes.create_string_buffer(256)
    out = ctypes.create_string_buffer(16)
    oracle.lib.aes128_set_key(k, key)
    oracle.lib.aes128_encrypt_block(k, pt, out)
    assert out.raw.hex() == "69c4e0d86a7b0430d8cdb78070b4c55a"
    oracle.lib.aes128_force_portable(k)
    oracle.lib.aes128_encrypt_block(k, pt, out)
    assert out.raw.hex() == "69c4e0d86a7b0430d8cdb78070b4c55a"


def test_ctr_stream_matches_library_aes(oracle):
    """zero IV, 128-bit big-endian counter: same bytes as a stock AES-CTR implementation"""
    crypto = pytest.importorskip("cryptography.hazmat.primitives.ciphers")
    from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
    key = po.TI_KEY
    enc = Cipher(algorithms.AES(key), modes.CTR(b"\0" * 16)).encryptor()
    ref = enc.update(b"\0" * 4096)
    got = oracle.prg(key, 0, 4096).tobytes()
    assert got == ref
    # random access at unaligned offsets and lengths (carry-over of partial blocks)
    for off, ln in [(1, 15), (8, 24), (17, 100), (1000, 1001), (4090, 6)]:
        assert oracle.prg(key, off, ln).tobytes() == ref[off:off + ln]


def test_portable_and_aesni_agree(oracle):
    import ctypes
    k = ctypes.create_string_buffer(256)
    oracle.lib.aes128_set_key(k, po.OT_KEY)
    a = np.empty(1000, np.uint8)
    b = np.empty(1000, np.uint8)
    oracle.lib.aes128_ctr_keystream(k, 7, a.ctypes.data, 1000)
    oracle.lib.aes128_force_portable(k)
    oracle.lib.aes128_ctr_keystream(k, 7, b.ctypes.data, 1000)
    assert np.array_equal(a, b)


def test_schedule_counts_match_survey_table(oracle):
    # (d, starts) -> T, D, L, C
    cases = [
        (5, [0, 1, 2], 20, 5, 6, 9),
        (100, [0, 33, 66], 5150, 100, 1651, 3399),
        (500, po.even_col_start(500, 5), 125750, 500, 24850, 100400),
        (200, [0], 20300, 200, 20100, 0),
    ]
    for d, cs, T, D, L, C in cases:
        s = oracle.schedule(d, cs)
        assert len(s["kind"]) == T
        assert int(np.count_nonzero(s["kind"] == 0)) == D
        assert int(np.count_nonzero(s["kind"] == 1)) == L
        assert int(np.count_nonzero(s["kind"] == 2)) == C
        q = s["q"][s["kind"] == 2]
        assert np.array_equal(q, np.arange(C))


def test_encode_readme_first_words(oracle, golden):
    c = next(c for c in golden["cases"] if c["instance"] == "readme_example" and c["bits_p1"] == 64 and c["bits_p2"] == 64)
    nz = oracle.normalizer(56, c["n"])
    assert nz == math.sqrt(2.0 ** 56 * 10)
    Xe = oracle.encode(c["X"], nz, 56).view(np.int64)
    ye = oracle.encode(c["y"], nz, 56).view(np.int64)
    assert Xe[0, 0] == -42010611 and ye[0] == 39355553  # SURVEY.md §8c, reproduced by _ref
    assert Xe[0, 0] == c["encoded_X_first"] and ye[0] == c["encoded_y_first"]


def test_oracle_matches_golden_reference_shares(oracle, golden):
    """every share word of every party, both protocols, three width builds"""
    assert len(golden["cases"]) == 24
    for c in golden["cases"]:
        got = oracle.run(c["X"], c["y"], c["col_start"], c["precision"], c["precision_p2"],
                         c["bits_p1"], c["bits_p2"], c["use_ot"], golden["ti_key"], golden["ot_key"], threads=2)
        assert np.array_equal(got, c["shares"]), (c["instance"], c["bits_p1"], c["bits_p2"], c["use_ot"])


def test_readme_known_answer(golden):
    """phase 2 restated in floating point on the golden shares reproduces README beta"""
    beta = np.array(golden["readme_beta"])
    for use_ot in (False, True):
        c = next(c for c in golden["cases"] if c["instance"] == "readme_example" and c["bits_p1"] == 64
                 and c["bits_p2"] == 64 and c["use_ot"] == use_ot)
        d, p = c["d"], c["precision"]
        tot = c["shares"].sum(axis=0, dtype=np.uint64)  # wrapping sum over parties (linear.oc:31-65)
        assert tot[1] == 0x000b6a9e8adf5744 and tot[15] == 0x0050db723de7540c
        A = np.zeros((d, d))
        for i in range(d):
            for j in range(i + 1):
                v = float(np.int64(tot[i * (i + 1) // 2 + j])) / 2.0 ** p
                A[i, j] = A[j, i] = v if i == j else v / d
        b = np.array([float(np.int64(v)) for v in tot[15:]]) / 2.0 ** p / d
        got = np.linalg.solve(A + 0.001 * np.eye(d), b)
        assert np.max(np.abs(got - beta)) < 1e-6


def test_share_sum_invariant(oracle):
    """sum over parties of every off-diagonal share == plain ring Gram"""
    X, y, _ = po.generate_lin_regression(50, 8, seed=5)
    cs = [0, 3, 5]
    for use_ot in (False, True):
        shares, Xe, ye = oracle.run(X, y, cs, 56, use_ot=use_ot, want_encoded=True)
        tot = shares.sum(axis=0, dtype=np.uint64)
        for i in range(9):
            for j in range(min(i + 1, 8)):
                if i == j:
                    continue
                a = Xe[:, i] if i < 8 else ye
                want = np.sum(a * Xe[:, j], dtype=np.uint64)
                got = tot[i * (i + 1) // 2 + j] if i < 8 else tot[36 + j]
                assert got == want


@pytest.mark.parametrize("b1,b2,prec,prec2", WIDTHS)
def test_oracle_matches_ref_live(oracle, b1, b2, prec, prec2):
    if not po.RefLib.available(b1, b2):
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    R = po.RefLib(b1, b2)
    X, y, _ = po.generate_lin_regression(150, 10, seed=77)
    y[3] = 1e30      # saturating / out-of-range encodes: x86 "integer indefinite"
    X[5, 2] = -1e300
    cs = [0, 4, 7]
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "t.in")
        po.write_instance(path, X, y, cs)
        Xr, yr, _ = R.read_inputs(path, 150, 10, prec)
        nz = oracle.normalizer(prec, 150)
        assert np.array_equal(oracle.encode(X, nz, prec, b1, b2).astype(np.int64 if b1 == 64 else np.int32).astype(np.int64), Xr)
        for use_ot in (False, True):
            sr, _, _ = R.run(path, 10, 3, prec, prec2, use_ot)
            so = oracle.run(X, y, cs, prec, prec2, b1, b2, use_ot)
            assert np.array_equal(sr, so)


@pytest.mark.parametrize("b1,b2", [(64, 64), (32, 32), (64, 32)])
def test_scalar_codec_matches_reference_cast(oracle, b1, b2):
    if not po.RefLib.available(b1, b2):
        pytest.skip("oracle/_ref not built")
    R = po.RefLib(b1, b2)
    vals = [0.0, -0.0, 0.5, -0.5, 1.0 - 2 ** -53, 123.456, -123.456, 2.0 ** 31, -2.0 ** 31, 2.0 ** 31 - 0.5,
            -2.0 ** 31 - 0.5, -2.0 ** 31 - 1.0, 2.0 ** 62, 2.0 ** 63, -2.0 ** 63, -2.0 ** 63 * 1.5, 1e300, -1e300,
            float("inf"), float("-inf"), float("nan"), 5e-324]
    for p in (0, 1, 24, 30, 31, 56, 62):
        for v in vals:
            assert oracle.lib.p1o_double_to_fixed(v, p, b2) == R.lib.p1ref_double_to_fixed(v, p), (v, p)
            assert oracle.lib.p1o_double_to_fixed(v, p, b1) == R.lib.p1ref_double_to_fixed_p1(v, p), (v, p)
    for i, j in [(0, 0), (3, 1), (1, 3), (499, 498), (500, 0)]:
        assert oracle.lib.p1o_idx(i, j) == R.lib.p1ref_idx(i, j)


def test_diag_matches_naive_python(oracle):
    rng = np.random.default_rng(3)
    col = rng.integers(-2 ** 27, 2 ** 27, 500).astype(np.int64)
    p = 56
    xy = 0.0
    for v in col:
        f = float(v) / float(1 << p)
        xy = xy + (f * f) * float(1 << p)
    want = int((xy / 7.0) * float(1 << p))
    assert oracle.diag(col.view(np.uint64), p, 7.0) == want % (1 << 64)


def test_downconvert(oracle):
    assert oracle.lib.p1o_downconvert(0xFFFFFFFF00000000, 56, 24) == 0xFFFFFFFF  # arithmetic shift of a negative share
    assert oracle.lib.p1o_downconvert(0x0000123400000000, 56, 24) == 0x00001234


def test_cot_transpose_and_correlation(oracle):
    """OT algebra (src/phase1.c:43-96): sender + receiver shares == <xs, xr>; wire sizes"""
    rng = np.random.default_rng(9)
    for bits in (64, 32):
        dt = po.word_dtype(bits)
        xs = rng.integers(0, 2 ** bits, 37, dtype=np.uint64).astype(dt)
        xr = rng.integers(0, 2 ** bits, 37, dtype=np.uint64).astype(dt)
        ss = oracle.lib.p1o_ot_session_new(po.OT_KEY, 2, 4)
        for _ in range(2):  # the session cursor advances between inner products
            s, r, u, c = oracle.ot_inner_product(ss, xs, xr, bits, want_wire=True)
            want = int(np.sum(xs.astype(np.uint64) * xr.astype(np.uint64), dtype=np.uint64)) % (1 << bits)
            assert (s + r) % (1 << bits) == want
            assert u.size == 2048 * ((37 * bits + 127) // 128)
        oracle.lib.p1o_ot_session_free(ss)


# --------------------------------------------------------------------------- f2 / f3
def _wire_cases():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "msg_wire.json")) as f:
        return json.load(f)["cases"]


def test_msg_wire_matches_google_protobuf_golden(oracle):
    """oracle pack/unpack against bytes serialised by the google.protobuf runtime
    (tests/golden/make_wire_golden.py): the format protobuf-c emits in send_pmsg."""
    cases = _wire_cases()
    assert len(cases) >= 20
    for c in cases:
        bits, vec, val = c["bits"], [int(v) for v in c["vector"]], int(c["value"])
        wire = oracle.msg_pack(np.array(vec, dtype=po.word_dtype(bits)), val, bits)
        assert wire.tobytes().hex() == c["wire_hex"]
        back = oracle.msg_unpack(np.frombuffer(bytes.fromhex(c["wire_hex"]), dtype=np.uint8), bits)
        assert back is not None and [int(v) for v in back[0]] == vec and back[1] == val


def test_msg_unpack_rejects_malformed(oracle):
    good = oracle.msg_pack(np.array([1, 300, 2 ** 63], dtype=np.uint64), 9, 64)
    assert oracle.msg_unpack(good, 64) is not None
    assert oracle.msg_unpack(good[:-1], 64) is None                       # trailer cut
    bad = good.copy(); bad[0] = 0x12
    assert oracle.msg_unpack(bad, 64) is None                             # wrong first tag
    bad = good.copy(); bad[1] += 1
    assert oracle.msg_unpack(bad, 64) is None                             # payload length past the trailer
    assert oracle.msg_unpack(np.concatenate([good, [0]]).astype(np.uint8), 64) is None  # trailing garbage


def test_share_bits_is_lsb_first(oracle):
    sel = oracle.share_bits(np.array([5, 1 << 63], dtype=np.uint64), 64)
    assert sel.shape == (128,) and list(np.flatnonzero(sel)) == [0, 2, 64 + 63]
    sel = oracle.share_bits(np.array([0x80000001], dtype=np.uint32), 32)
    assert list(np.flatnonzero(sel)) == [0, 31]
