"""TEST INFRASTRUCTURE: ctypes bindings for the two CPU checkers.

* ``Oracle``  -> oracle/liboracle.so, this repo's C restatement (oracle/p1_oracle.c).
* ``RefLib``  -> oracle/_ref/libp1ref_<P1>_<P2>.so, the reference's own phase1.c /
  linear.c / fixed.c / config.c compiled unmodified against oracle/ref_shims.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg may import this module.  The product path (linreg-mpc_b200/) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_SRC = "/root/reference/src"

TI_KEY = bytes(range(16))          # TI key 00 01 .. 0f
OT_KEY = bytes(range(16, 32))      # OT master key 10 11 .. 1f

c_u8p = ctypes.c_char_p
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_i = ctypes.c_int
c_u64 = ctypes.c_uint64
c_i64 = ctypes.c_int64
c_d = ctypes.c_double


def build(ref=True):
    """Compile liboracle.so and, when the reference sources are present, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir(REFERENCE_SRC):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
        if os.path.exists(os.path.join(HERE, "..", "linreg-mpc_b200", "libp1gpu.so")):
            # the product's host file needs the reference's headers too: built here, travels prebuilt
            subprocess.check_call(["make", "-s", "-C", HERE, "gpuhost"])


def word_dtype(bits):
    return np.uint64 if bits == 64 else np.uint32


def _ptr(a):
    return a.ctypes.data_as(c_vp) if a is not None else None


class Oracle:
    def __init__(self, path=None):
        path = path or os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = self.lib = ctypes.CDLL(path)
        L.p1o_double_to_fixed.argtypes = [c_d, c_i, c_i]
        L.p1o_double_to_fixed.restype = c_i64
        L.p1o_fixed_to_double.argtypes = [c_i64, c_i]
        L.p1o_fixed_to_double.restype = c_d
        L.p1o_normalizer.argtypes = [c_i, c_d]
        L.p1o_normalizer.restype = c_d
        L.p1o_encode.argtypes = [c_vp, c_sz, c_d, c_i, c_i, c_i, c_vp]
        L.p1o_idx.argtypes = [c_sz, c_sz]
        L.p1o_idx.restype = c_sz
        L.p1o_owner.argtypes = [c_sz, c_i, c_vp]
        L.p1o_owner.restype = c_i
        L.p1o_schedule.argtypes = [c_sz, c_i, c_vp] + [c_vp] * 6
        L.p1o_schedule.restype = c_sz
        L.p1o_dot.argtypes = [c_vp, c_sz, c_vp, c_sz, c_sz, c_i]
        L.p1o_dot.restype = c_u64
        L.p1o_diag.argtypes = [c_vp, c_sz, c_sz, c_i, c_d, c_i]
        L.p1o_diag.restype = c_u64
        L.p1o_diag_accumulate.argtypes = [c_vp, c_sz, c_sz, c_i, c_i, c_vp]
        L.p1o_diag_finish.argtypes = [c_d, c_d, c_i, c_i]
        L.p1o_diag_finish.restype = c_u64
        L.p1o_ti_offset.argtypes = [c_u64, c_sz, c_i]
        L.p1o_ti_offset.restype = c_u64
        L.p1o_prg.argtypes = [c_u8p, c_u64, c_vp, c_sz]
        L.p1o_ti_pair.argtypes = [c_u8p, c_u64, c_sz, c_i, c_vp, c_vp, c_vp, c_vp]
        L.p1o_party_a_mask.argtypes = [c_vp, c_sz, c_vp, c_sz, c_i, c_vp]
        L.p1o_party_b_mask.argtypes = [c_vp, c_sz, c_vp, c_sz, c_i, c_vp]
        L.p1o_party_a_share.argtypes = [c_vp, c_vp, c_u64, c_sz, c_i]
        L.p1o_party_a_share.restype = c_u64
        L.p1o_party_b_share.argtypes = [c_vp, c_vp, c_sz, c_u64, c_sz, c_i]
        L.p1o_party_b_share.restype = c_u64
        L.p1o_ot_session_new.argtypes = [c_u8p, c_i, c_i]
        L.p1o_ot_session_new.restype = c_vp
        L.p1o_ot_session_free.argtypes = [c_vp]
        L.p1o_ot_session_set_cursor.argtypes = [c_vp, c_u64, c_u64]
        L.p1o_ot_inner_product.argtypes = [c_vp, c_vp, c_sz, c_vp, c_sz, c_sz, c_i, c_vp, c_vp, c_vp, c_vp]
        L.p1o_downconvert.argtypes = [c_u64, c_i, c_i]
        L.p1o_downconvert.restype = c_u64
        L.p1o_run.argtypes = [c_vp, c_vp, c_sz, c_sz, c_i, c_vp, c_i, c_i, c_i, c_i, c_i,
                              c_u8p, c_u8p, c_i, c_vp, c_vp, c_vp]
        L.p1o_run.restype = c_i
        L.p1o_write_instance.argtypes = [c_u8p, c_vp, c_vp, c_sz, c_sz, c_i, c_vp]
        L.p1o_write_instance.restype = c_i
        L.p1o_msg_packed_size.argtypes = [c_vp, c_sz, c_u64, c_i]
        L.p1o_msg_packed_size.restype = c_sz
        L.p1o_msg_pack.argtypes = [c_vp, c_sz, c_u64, c_i, c_vp]
        L.p1o_msg_pack.restype = c_sz
        L.p1o_msg_unpack.argtypes = [c_vp, c_sz, c_i, c_vp, c_sz, c_vp, c_vp]
        L.p1o_msg_unpack.restype = c_i
        L.p1o_share_bits.argtypes = [c_vp, c_sz, c_i, c_vp]
        L.p1o_parse_doubles.argtypes = [c_vp, c_sz, c_vp, c_vp, c_sz]
        L.p1o_parse_doubles.restype = c_sz
        L.aes128_set_key.argtypes = [c_vp, c_u8p]
        L.aes128_force_portable.argtypes = [c_vp]
        L.aes128_ctr_keystream.argtypes = [c_vp, c_u64, c_vp, c_sz]
        L.aes128_encrypt_block.argtypes = [c_vp, c_u8p, c_vp]

    # ---- scalar / array helpers ----
    def encode(self, vals, normalizer, precision, bits_p1=64, bits_p2=64):
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        out = np.empty(vals.shape, dtype=word_dtype(bits_p1))
        self.lib.p1o_encode(_ptr(vals), vals.size, normalizer, precision, bits_p1, bits_p2, _ptr(out))
        return out

    def normalizer(self, precision, n):
        return self.lib.p1o_normalizer(precision, float(n))

    def schedule(self, d, col_start):
        cs = np.ascontiguousarray(col_start, dtype=np.int64)
        T = self.lib.p1o_schedule(d, len(cs), _ptr(cs), None, None, None, None, None, None)
        pi = np.empty(T, np.int64); pj = np.empty(T, np.int64); q = np.empty(T, np.int64)
        oi = np.empty(T, np.int32); oj = np.empty(T, np.int32); kind = np.empty(T, np.int32)
        self.lib.p1o_schedule(d, len(cs), _ptr(cs), _ptr(pi), _ptr(pj), _ptr(oi), _ptr(oj), _ptr(kind), _ptr(q))
        return dict(i=pi, j=pj, owner_i=oi, owner_j=oj, kind=kind, q=q)

    def dot(self, x, y, bits=64):
        x = np.ascontiguousarray(x); y = np.ascontiguousarray(y)
        return int(self.lib.p1o_dot(_ptr(x), 1, _ptr(y), 1, x.size, bits))

    def diag(self, col, precision, normalizer2, bits=64):
        col = np.ascontiguousarray(col)
        return int(self.lib.p1o_diag(_ptr(col), 1, col.size, precision, float(normalizer2), bits))

    def downconvert(self, share, precision, precision_p2):
        """64 -> 32 hand-off of a share word (src/phase1.c:609-620,629-635)"""
        self.lib.p1o_downconvert.argtypes = [c_u64, c_i, c_i]
        self.lib.p1o_downconvert.restype = c_u64
        return int(self.lib.p1o_downconvert(int(share) & (2 ** 64 - 1), precision, precision_p2))

    def prg(self, key, byte_offset, nbytes):
        out = np.empty(nbytes, dtype=np.uint8)
        self.lib.p1o_prg(key, byte_offset, _ptr(out), nbytes)
        return out

    def ti_pair(self, key, q, n, bits=64):
        x = np.empty(n, word_dtype(bits)); y = np.empty(n, word_dtype(bits))
        va = c_u64(); vb = c_u64()
        self.lib.p1o_ti_pair(key, q, n, bits, _ptr(x), _ptr(y), ctypes.byref(va), ctypes.byref(vb))
        return x, y, va.value, vb.value

    def ot_inner_product(self, session, xs, xr, bits=64, want_wire=False):
        xs = np.ascontiguousarray(xs); xr = np.ascontiguousarray(xr)
        n = xs.size
        ss = c_u64(); sr = c_u64()
        u = c = None
        if want_wire:
            tiles = (n * bits + 127) // 128
            u = np.empty(2048 * tiles, np.uint8); c = np.empty(n * bits * (bits // 8), np.uint8)
        self.lib.p1o_ot_inner_product(session, _ptr(xs), 1, _ptr(xr), 1, n, bits,
                                      ctypes.byref(ss), ctypes.byref(sr), _ptr(u), _ptr(c))
        return (ss.value, sr.value, u, c) if want_wire else (ss.value, sr.value)

    def msg_pack(self, vec, value, bits=64):
        """SecureMultiplication__Msg{vector, value} -> proto2 wire bytes (src/phase1.c:137-139)"""
        vec = np.ascontiguousarray(vec, dtype=word_dtype(bits))
        size = self.lib.p1o_msg_packed_size(_ptr(vec), len(vec), int(value), bits)
        out = np.zeros(size, dtype=np.uint8)
        got = self.lib.p1o_msg_pack(_ptr(vec), len(vec), int(value), bits, _ptr(out))
        assert got == size
        return out

    def msg_unpack(self, wire, bits=64, cap=None):
        """wire bytes -> (vector, value), or None when malformed (src/phase1.c:111-112)"""
        wire = np.ascontiguousarray(wire, dtype=np.uint8)
        cap = len(wire) if cap is None else cap
        vec = np.zeros(max(cap, 1), dtype=word_dtype(bits))
        n, v = c_sz(), c_u64()
        if self.lib.p1o_msg_unpack(_ptr(wire), len(wire), bits, _ptr(vec), cap, ctypes.byref(n), ctypes.byref(v)):
            return None
        return vec[:n.value].copy(), int(v.value)

    def share_bits(self, shares, bits_p2=64):
        """dcsSendIntArray's choice bits, src/input.c:41"""
        shares = np.ascontiguousarray(shares, dtype=word_dtype(bits_p2))
        sel = np.zeros(len(shares) * bits_p2, dtype=np.uint8)
        self.lib.p1o_share_bits(_ptr(shares), len(shares), bits_p2, _ptr(sel))
        return sel

    def parse_doubles(self, text):
        """fscanf("%lf") token by token (src/linear.c:46,82): (values, whole-token flags)"""
        buf = np.frombuffer(bytes(text), dtype=np.uint8)
        cap = len(buf) // 2 + 1
        out = np.zeros(cap, dtype=np.float64)
        whole = np.zeros(cap, dtype=np.uint8)
        n = self.lib.p1o_parse_doubles(_ptr(buf), len(buf), _ptr(out), _ptr(whole), cap)
        return out[:n].copy(), whole[:n].astype(bool)

    def write_instance(self, path, X, y, col_start):
        """fast C writer of the reference's input-file format (same text as write_instance below)"""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        cs = np.ascontiguousarray(col_start, dtype=np.int64)
        assert self.lib.p1o_write_instance(path.encode(), _ptr(X), _ptr(y), X.shape[0], X.shape[1], len(cs), _ptr(cs)) == 0

    def run(self, X, y, col_start, precision, precision_p2=0, bits_p1=64, bits_p2=64, use_ot=False,
            ti_key=TI_KEY, ot_key=OT_KEY, threads=1, want_encoded=False):
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        cs = np.ascontiguousarray(col_start, dtype=np.int64)
        shares = np.zeros((len(cs), d * (d + 1) // 2 + d), dtype=np.uint64)
        Xe = np.empty((n, d), word_dtype(bits_p1)) if want_encoded else None
        ye = np.empty(n, word_dtype(bits_p1)) if want_encoded else None
        st = self.lib.p1o_run(_ptr(X), _ptr(y), n, d, len(cs), _ptr(cs), precision, precision_p2,
                              bits_p1, bits_p2, int(use_ot), ti_key, ot_key, threads,
                              _ptr(shares), _ptr(Xe), _ptr(ye))
        assert st == 0
        return (shares, Xe, ye) if want_encoded else shares


class RefLib:
    """The reference's own phase 1 (see oracle/ref_shims/ref_driver.c)."""

    def __init__(self, bits_p1=64, bits_p2=64, gpuhost=False):
        """gpuhost=True: the same driver and shims around linreg-mpc_b200/host/phase1_gpu.c (the
        product's run_party / run_trusted_initializer on top of libp1gpu.so) instead of the
        reference's phase1.c - the drop-in under test, not the checker."""
        path = os.path.join(HERE, "_ref", "libp1%s_%d_%d.so" % ("gpuhost" if gpuhost else "ref", bits_p1, bits_p2))
        if not os.path.exists(path):
            build(ref=True)
        L = self.lib = ctypes.CDLL(path)
        self.bits_p1, self.bits_p2 = bits_p1, bits_p2
        L.p1ref_run.argtypes = [c_u8p, c_i, c_i, c_i, c_u8p, c_u8p, c_vp, c_vp, c_vp]
        L.p1ref_run.restype = c_i
        L.p1ref_read_inputs.argtypes = [c_u8p, c_i, c_vp, c_vp, c_vp]
        L.p1ref_read_inputs.restype = c_i
        L.p1ref_double_to_fixed.argtypes = [c_d, c_i]
        L.p1ref_double_to_fixed.restype = c_i64
        L.p1ref_double_to_fixed_p1.argtypes = [c_d, c_i]
        L.p1ref_double_to_fixed_p1.restype = c_i64
        L.p1ref_fixed_to_double_p1.argtypes = [c_i64, c_i]
        L.p1ref_fixed_to_double_p1.restype = c_d
        L.p1ref_idx.argtypes = [c_sz, c_sz]
        L.p1ref_idx.restype = c_sz
        assert L.p1ref_bits_p1() == bits_p1 and L.p1ref_bits_p2() == bits_p2

    @staticmethod
    def available(bits_p1=64, bits_p2=64, gpuhost=False):
        return os.path.exists(os.path.join(HERE, "_ref", "libp1%s_%d_%d.so" % ("gpuhost" if gpuhost else "ref", bits_p1, bits_p2)))

    def run(self, path, d, num_dp, precision, precision_p2=0, use_ot=False, ti_key=TI_KEY, ot_key=OT_KEY):
        P = num_dp + 2
        shares = np.zeros((num_dp, d * (d + 1) // 2 + d), dtype=np.uint64)
        stats = np.zeros((P, 3)); sent = np.zeros((P, P), dtype=np.uint64)
        # the reference prints a debug line per local pair in OT mode (phase1.c:373,380)
        saved = None
        if use_ot:
            import sys
            sys.stdout.flush()
            saved = os.dup(1)
            devnull = os.open(os.devnull, os.O_WRONLY)
            os.dup2(devnull, 1)
            os.close(devnull)
        try:
            st = self.lib.p1ref_run(path.encode(), precision, precision_p2, int(use_ot), ti_key, ot_key,
                                    _ptr(shares), _ptr(stats), _ptr(sent))
        finally:
            if saved is not None:
                ctypes.CDLL(None).fflush(None)
                os.dup2(saved, 1)
                os.close(saved)
        assert st == 0, "reference phase 1 failed"
        return shares, stats, sent

    def read_inputs(self, path, n, d, precision):
        X = np.empty((n, d), np.int64); y = np.empty(n, np.int64); sec = c_d()
        st = self.lib.p1ref_read_inputs(path.encode(), precision, _ptr(X), _ptr(y), ctypes.byref(sec))
        assert st == 0
        return X, y, sec.value


# ---------------------------------------------------------------------------
# instance files and synthetic inputs
# ---------------------------------------------------------------------------

def even_col_start(d, num_dp):
    """start_k = k * (d // p)  (experiments/generate_tests.py:54-61)"""
    return [k * (d // num_dp) for k in range(num_dp)]


def generate_lin_regression(n, d, sigma=0.1, seed=20170401):
    """Synthetic instance per experiments/generate_tests.py:159-169 (sigma as in
    experiments/test_phase1_aws.py:197): X ~ N(0,1), columns scaled to max|.| = 1,
    beta ~ U(0,1), y = X beta + N(0, sigma^2).  Draw order X (row-major), beta, noise."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    X /= np.max(np.abs(X), axis=0, keepdims=True)
    beta = rng.uniform(0.0, 1.0, d)
    e = rng.normal(0.0, sigma, n)
    y = X @ beta + e
    return X, y, beta


def write_instance(path, X, y, col_start):
    """Input-file format parsed by src/config.c:24-44 and src/linear.c:27-102
    (layout of experiments/generate_tests.py:41-74).  repr() round-trips doubles."""
    n, d = X.shape
    with open(path, "w") as f:
        f.write("%d %d %d\n" % (n, d, len(col_start)))
        f.write("localhost:1234\nlocalhost:1235\n")
        for k, s in enumerate(col_start):
            f.write("localhost:%d %d\n" % (1236 + k, s))
        f.write("%d %d\n" % (n, d))
        for i in range(n):
            f.write(" ".join(repr(float(v)) for v in X[i]))
            f.write("\n")
        f.write("%d\n" % n)
        f.write(" ".join(repr(float(v)) for v in y))
        f.write("\n")


def read_instance(path):
    """Parse an instance file (same fields as src/config.c + src/linear.c read)."""
    with open(path) as f:
        tok = f.read().split()
    n, d, p = int(tok[0]), int(tok[1]), int(tok[2])
    pos = 3 + 2  # two endpoints without an index
    col_start = []
    for _ in range(p):
        col_start.append(int(tok[pos + 1]))
        pos += 2
    assert int(tok[pos]) == n and int(tok[pos + 1]) == d
    pos += 2
    X = np.array([float(t) for t in tok[pos:pos + n * d]]).reshape(n, d)
    pos += n * d
    assert int(tok[pos]) == n
    pos += 1
    y = np.array([float(t) for t in tok[pos:pos + n]])
    return X, y, col_start
