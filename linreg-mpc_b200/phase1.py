"""Host-side orchestration of one phase-1 run on top of the C-ABI (include/p1gpu.h).

Mirrors the reference's control flow — the pair schedule of run_trusted_initializer
(src/phase1.c:257-267) and run_party (:534-587), the OT thread schedule (:353-450) and
the tail of run_party (:605-641) — for ALL parties of one instance that share one GPU
(or, with row sharding, one GPU per rank: every rank holds rows [row0, row0+n_rows) of
every provider's columns and the ranks combine their partial shares with one wrapping
uint64 sum all-reduce).  Arithmetic happens only inside libp1gpu.so; this file does
bookkeeping with numpy index arrays and torch device buffers (plumbing).
"""
import ctypes
import math

import time

import numpy as np

from . import ffi
from .ffi import c_vp, c_sz, ptr


def idx(i, j):
    """packed lower-triangular index, src/linear.c:11-16 (vectorised)"""
    i, j = np.maximum(i, j), np.minimum(i, j)
    return i * (i + 1) // 2 + j


def owners(rows, col_start):
    """get_owner, src/phase1.c:25-33, as provider index 0..p-1 (row d -> last provider)"""
    cs = np.asarray(col_start, dtype=np.int64)
    if np.any(np.diff(cs) < 0):
        raise ValueError("provider column starts must be non-decreasing")
    return (np.searchsorted(cs, np.asarray(rows, dtype=np.int64), side="right") - 1).astype(np.int32)


class Schedule:
    """Pair enumeration i = 0..d, j = 0..min(i, d-1) (src/phase1.c:257-258,534,544) with
    the class of every pair.  Cross pairs are numbered q = 0,1,.. in this order, which is
    the order of the TI's random stream and of every message on the wire."""

    def __init__(self, d, col_start):
        self.d = d
        self.col_start = [int(c) for c in col_start]
        self.num_dp = len(self.col_start)
        self.nA = d * (d + 1) // 2
        ii = np.repeat(np.arange(d + 1, dtype=np.int64), np.minimum(np.arange(d + 1) + 1, d))
        jj = np.concatenate([np.arange(min(i + 1, d), dtype=np.int64) for i in range(d + 1)])
        oi, oj = owners(ii, col_start), owners(jj, col_start)
        cross = oi != oj
        self.i, self.j, self.owner_i, self.owner_j = ii, jj, oi, oj
        self.kind = np.where(cross, 2, np.where(ii == jj, 0, 1)).astype(np.int32)
        # accumulator slot of a pair (src/phase1.c:581-585)
        self.slot = np.where(ii < d, idx(ii, jj), self.nA + jj).astype(np.int64)
        self.cross = np.flatnonzero(cross)
        self.num_cross = len(self.cross)

    def col_range(self, k):
        lo = self.col_start[k]
        hi = self.col_start[k + 1] if k + 1 < self.num_dp else self.d
        return lo, hi

    def ring_macs(self, n, masking=True):
        """ring MACs the reference's schedule executes: n * (L + 3C)"""
        L = int(np.count_nonzero(self.kind == 1))
        return n * (L + (3 * self.num_cross if masking else 0))


def shard_rows(n_total, world, rank, align=256):
    """Row shard of rank `rank`: equal slices whose boundaries fall on multiples
    of `align` rows, so that every shard's slice of a mask stream is 16-byte aligned and
    holds whole OT tiles (128/W rows); the last shards may be short or empty.
    Returns (row0, n_rows)."""
    per = (int(n_total) + world - 1) // world
    per = (per + align - 1) // align * align
    row0 = min(rank * per, int(n_total))
    return row0, max(0, min(int(n_total), row0 + per) - row0)


def ring_all_reduce(flat, group):
    """Exact combination of the row shards' partial shares: every entry is a sum over rows in
    Z_2^64 (a Z_2^32 share is the low half), so one wrapping integer sum all-reduce (NCCL over
    NVLink on the GPU box, gloo in the CPU tests) is bit-exact in any reduction order."""
    import torch.distributed as dist
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def chain_in_rank_order(step, width, group, device=None):
    """The one quantity that is NOT order independent, the sequential double diagonal
    (src/phase1.c:563-566): rank g continues the running sums of rank g-1 over its own rows
    (`step(xy) -> xy`, numpy float64[width]) and hands them on; the last rank's result is
    broadcast.  Returns the final float64[width] on every rank."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    gr = lambda r: dist.get_global_rank(group, r) if group is not None else r
    buf = torch.zeros(width, dtype=torch.float64, device=device)
    xy = np.zeros(width, dtype=np.float64)
    if rank > 0:
        dist.recv(buf, src=gr(rank - 1), group=group)
        xy = buf.cpu().numpy().copy()
    xy = np.ascontiguousarray(step(xy), dtype=np.float64)
    buf.copy_(torch.from_numpy(xy))
    if rank + 1 < world:
        dist.send(buf, dst=gr(rank + 1), group=group)
    dist.broadcast(buf, src=gr(world - 1), group=group)
    return buf.cpu().numpy().copy()


def _aes_ecb(key):
    from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
    return Cipher(algorithms.AES(bytes(key)), modes.ECB()).encryptor()


def _bits128(block):
    return np.unpackbits(np.frombuffer(bytes(block), dtype=np.uint8), bitorder="little").astype(bool)


def derive_cot_session(ot_key, sender_party, receiver_party):
    """Deterministic stand-in for the kappa = 128 base OTs of one sender/receiver session
    (host-side setup; in the reference this is Obliv-C's honestOTExt*New, src/phase1.c:394,397).
    ksess = AES_otkey(BE128(sender*65536 + receiver)); k_i^b = AES_ksess(BE128(2i+b));
    Delta = AES_ksess(BE128(256)).  Returns (seeds0[128,16], seeds1[128,16], delta[16])."""
    be = lambda v: int(v).to_bytes(16, "big")
    ksess = _aes_ecb(ot_key).update(be(sender_party * 65536 + receiver_party))
    out = _aes_ecb(ksess).update(b"".join(be(v) for v in range(257)))
    blocks = np.frombuffer(out, dtype=np.uint8).reshape(257, 16)
    seeds0 = np.ascontiguousarray(blocks[0:256:2])
    seeds1 = np.ascontiguousarray(blocks[1:256:2])
    return seeds0, seeds1, np.ascontiguousarray(blocks[256])


class Phase1Box:
    def __init__(self, lib, n_total, d, col_start, precision, precision_p2=0, bits_p1=64, bits_p2=64,
                 device=0, row0=0, n_rows=None, batch_pairs=32, pipes=2, gram_algo=ffi.P1_GRAM_AUTO,
                 group=None):
        import torch
        self.torch = torch
        self.lib = lib
        self.n_total, self.d = int(n_total), int(d)
        self.row0 = int(row0)
        self.n_rows = int(n_rows if n_rows is not None else n_total - row0)
        self.precision, self.precision_p2 = precision, precision_p2
        self.bits_p1, self.bits_p2 = bits_p1, bits_p2
        self.w = bits_p1 // 8
        self.device = device
        self.group = group
        self.gram_algo = gram_algo
        self.sched = Schedule(d, col_start)
        self.acc_len = int(lib.p1_acc_len(d))
        self.batch_pairs = int(batch_pairs)
        self.pipe_limit = None          # use only the first k pipes (bench.py's per-kernel pass: one stream, kernels back to back)
        self.dev = torch.device("cuda", device)
        self.ctxs = []
        for _ in range(max(1, pipes)):
            c = c_vp()
            lib.check(lib.p1_ctx_create(device, bits_p1, bits_p2, ctypes.byref(c)))
            self.ctxs.append(c)
        # every provider's block is computed from this one column store, so the digit planes of all
        # columns are used: let the encode pass write them (a single party would use only its own columns)
        lib.check(lib.p1_ctx_set_option(self.ctxs[0], b"encode_planes", 1))
        self.data = None
        S, C = self.sched, self.sched.num_cross
        # one flat u64 buffer so that a single all-reduce combines row shards:
        #   [num_dp x acc_len accumulators | C partial <x,y>]
        self.flat = torch.zeros(S.num_dp * self.acc_len + C, dtype=torch.int64, device=self.dev)
        self.acc = self.flat[:S.num_dp * self.acc_len]
        self.xy = self.flat[S.num_dp * self.acc_len:]
        self.r = torch.zeros(max(C, 1), dtype=torch.int64, device=self.dev)
        cr = S.cross
        to_dev = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dt).to(self.dev)
        self.cols_a = to_dev(S.i[cr], torch.int32)            # row i of the pair (d = target)
        self.cols_b = to_dev(S.j[cr], torch.int32)
        self.slot_a = to_dev(S.owner_i[cr].astype(np.int64) * self.acc_len + S.slot[cr], torch.int64)
        self.slot_b = to_dev(S.owner_j[cr].astype(np.int64) * self.acc_len + S.slot[cr], torch.int64)
        self.bufs = None
        self.normalizer = math.sqrt(math.pow(2, precision) * float(self.n_total))  # src/phase1.c:473
        # one step is stream-ordered from reset() to the D2H of the shares: torch work (zeroing, the all-reduce over
        # row shards) is issued on pipe 0's stream, so nothing in between needs the host
        self.s0 = torch.cuda.ExternalStream(int(lib.p1_ctx_stream(self.ctxs[0])), device=self.dev)
        self.col_start_dev = to_dev(np.asarray(list(col_start) + [self.d]), torch.int32)
        self.out_dev = torch.zeros(S.num_dp * (S.nA + self.d), dtype=torch.int64, device=self.dev)
        self.need_dev = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.out_host = torch.zeros(S.num_dp * (S.nA + self.d), dtype=torch.int64).pin_memory()
        self.need_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._has_y = None
        torch.cuda.current_stream(self.dev).synchronize()

    # ------------------------------------------------------------------
    def close(self):
        if self.data is not None:
            self.lib.p1_data_free(self.data)
            self.data = None
        for c in self.ctxs:
            self.lib.p1_ctx_destroy(c)
        self.ctxs = []

    def sync(self):
        for c in self.ctxs:
            self.lib.check(self.lib.p1_ctx_sync(c))

    def launches(self):
        return sum(int(self.lib.p1_ctx_launches(c)) for c in self.ctxs)

    def _fan_out(self):
        """later pipes wait for what pipe 0 has enqueued (encode, zeroing)"""
        for c in self.ctxs[1:]:
            self.lib.check(self.lib.p1_ctx_wait(c, self.ctxs[0]))

    def _fan_in(self):
        for c in self.ctxs[1:]:
            self.lib.check(self.lib.p1_ctx_wait(self.ctxs[0], c))

    # ------------------------------------------------------------------
    def encode(self, X, y):
        """X: this shard's rows (n_rows x d doubles; numpy = host, torch cuda tensor = device)."""
        if self.data is not None and self._has_y == (y is not None):
            # same shape as the last run: re-encode in place (no allocation, no host synchronisation)
            self.lib.check(self.lib.p1_encode_again(self.ctxs[0], self.data, ptr(X), ptr(y), self.normalizer, self.precision))
            return
        if self.data is not None:
            self.lib.p1_data_free(self.data)
            self.data = None
        self._has_y = y is not None
        out = c_vp()
        self.lib.check(self.lib.p1_encode(self.ctxs[0], ptr(X), ptr(y), self.n_rows, self.d,
                                          self.normalizer, self.precision, ctypes.byref(out)))
        self.data = out
        ld = c_sz()
        self.lib.p1_data_colmajor_dev(out, ctypes.byref(ld))
        self.ld = int(ld.value)

    def reset(self):
        with self.torch.cuda.stream(self.s0):
            self.flat.zero_()
            self.r.zero_()

    def local_blocks(self):
        S = self.sched
        for k in range(S.num_dp):
            lo, hi = S.col_range(k)
            acc_k = self.acc.data_ptr() + 8 * k * self.acc_len
            self.lib.check(self.lib.p1_local_block(self.ctxs[0], self.data, lo, hi, int(k == S.num_dp - 1),
                                                   c_vp(acc_k), self.gram_algo))

    def _ensure_bufs(self):
        if self.bufs is not None:
            return
        t = self.torch
        words = self.batch_pairs * self.ld
        dt = t.int64 if self.w == 8 else t.int32
        self.bufs = [[t.empty(words, dtype=dt, device=self.dev) for _ in range(4)] for _ in self.ctxs]
        t.cuda.current_stream(self.dev).synchronize()

    def _ensure_wire(self):
        if getattr(self, "wire", None) is not None:
            return
        t = self.torch
        self.wire_stride = int(self.lib.p1_msg_max_bytes(self.ctxs[0], self.n_rows))
        mk = lambda dt, k: t.zeros(k, dtype=dt, device=self.dev)
        self.wire = [dict(buf=t.empty(self.batch_pairs * self.wire_stride, dtype=t.uint8, device=self.dev),
                          sizes=mk(t.int64, self.batch_pairs), vals=mk(t.int64, self.batch_pairs),
                          zero=mk(t.int64, self.batch_pairs), status=mk(t.int32, self.batch_pairs)) for _ in self.ctxs]
        t.cuda.current_stream(self.dev).synchronize()

    def _over_the_wire(self, ctx, pipe, vec, B):
        """what a message costs between two boxes, minus the network: serialised by the sender
        (send_pmsg, src/phase1.c:131-141) and parsed by the receiver (recv_pmsg, :111) in place"""
        L, wr = self.lib, self.wire[pipe]
        L.check(L.p1_msg_pack(ctx, vec, self.ld, self.n_rows, B, ptr(wr["zero"]), ptr(wr["buf"]), self.wire_stride, ptr(wr["sizes"])))
        L.check(L.p1_msg_unpack(ctx, ptr(wr["buf"]), self.wire_stride, ptr(wr["sizes"]), B, self.n_rows, vec, self.ld,
                                ptr(wr["vals"]), ptr(wr["status"])))

    def cross_pairs_masking(self, ti_key, seeded=False, wire=False, max_pairs=None, fused=None, keep_masks=False):
        """TI masking protocol for every cross pair (src/phase1.c:257-285 and :148-236).
        fused (default: whenever neither seeded nor wire is asked for): the whole batch in ONE
                      persistent kernel (p1_masking_fused) - mask expansion, b's masked column, a's step and
                      b's share on the same rows, the masks never leave the SM; the two party messages are
                      written as before.  keep_masks=True also materialises x and y (tests).  fused=False:
                      the four kernels a deployment with parties on different devices runs.
        seeded=False: the TI materialises x and y once (what the reference ships);
        seeded=True : the TI computes only the values; each provider expands its own
                      mask from (key, q) - same share words.
        wire=True   : every one of the four messages of a pair additionally goes through the
                      protobuf wire format on the GPU (pack by the sender, unpack by the receiver),
                      as between separate boxes; the share words do not change."""
        C = self.sched.num_cross if max_pairs is None else min(self.sched.num_cross, int(max_pairs))
        if not C:
            return
        if fused is None:
            fused = not seeded and not wire
        if fused and (seeded or wire):
            raise ValueError("the fused masking loop hands messages over in HBM: not with seeded / wire")
        self._ensure_bufs()
        if wire:
            self._ensure_wire()
        L = self.lib
        self._fan_out()
        for b, q0 in enumerate(range(0, C, self.batch_pairs)):
            B = min(self.batch_pairs, C - q0)
            pipe = b % (self.pipe_limit or len(self.ctxs))
            ctx = self.ctxs[pipe]
            xb, yb, mb, ma = (c_vp(tn.data_ptr()) for tn in self.bufs[pipe])
            xy = c_vp(self.xy.data_ptr() + 8 * q0)
            r = c_vp(self.r.data_ptr() + 8 * q0)
            ca = c_vp(self.cols_a.data_ptr() + 4 * q0)
            cb = c_vp(self.cols_b.data_ptr() + 4 * q0)
            sa = c_vp(self.slot_a.data_ptr() + 8 * q0)
            sb = c_vp(self.slot_b.data_ptr() + 8 * q0)
            acc = c_vp(self.acc.data_ptr())
            args = (ti_key, q0, B, self.n_total, self.row0, self.n_rows)
            if fused:
                L.check(L.p1_masking_fused(ctx, ti_key, q0, B, self.n_total, self.row0, self.data, ca, cb, sa, sb,
                                           mb, ma, self.ld, xb if keep_masks else None, yb if keep_masks else None,
                                           xy, r, acc))
                continue
            if not seeded:
                L.check(L.p1_ti_generate(ctx, *args, xb, yb, self.ld, xy, r))
            else:
                L.check(L.p1_ti_generate(ctx, *args, None, None, self.ld, xy, r))      # TI: values only
                L.check(L.p1_ti_generate(ctx, *args, xb, None, self.ld, None, None))   # party b expands x
                L.check(L.p1_ti_generate(ctx, *args, None, yb, self.ld, None, None))   # party a expands y
            if wire:
                self._over_the_wire(ctx, pipe, xb, B)      # TI -> b
                self._over_the_wire(ctx, pipe, yb, B)      # TI -> a
            L.check(L.p1_party_b_mask(ctx, self.data, cb, B, xb, self.ld, mb, self.ld))
            if wire:
                self._over_the_wire(ctx, pipe, mb, B)      # b -> a
            L.check(L.p1_party_a_step(ctx, self.data, ca, sa, B, yb, self.ld, mb, self.ld, ma, self.ld, acc))
            if wire:
                self._over_the_wire(ctx, pipe, ma, B)      # a -> b
            L.check(L.p1_party_b_share(ctx, self.data, cb, sb, B, ma, self.ld, acc))
        self._fan_in()

    def all_reduce(self):
        """exact ring combination of row shards: wrapping integer sum (NCCL over NVLink)"""
        if self.group is None:
            return
        with self.torch.cuda.stream(self.s0):      # ordered behind the kernels, no host synchronisation
            ring_all_reduce(self.flat, self.group)

    def apply_values(self):
        """share -= value, once, after the row shards were combined (src/phase1.c:196,222)"""
        C = self.sched.num_cross
        if not C:
            return
        L, ctx, t = self.lib, self.ctxs[0], self.torch
        if getattr(self, "_va", None) is None:
            self._va = t.empty(C, dtype=t.int64, device=self.dev)
            self._vb = t.empty(C, dtype=t.int64, device=self.dev)
            t.cuda.current_stream(self.dev).synchronize()
        L.check(L.p1_ti_values(ctx, ptr(self.xy), ptr(self.r), C, ptr(self._va), ptr(self._vb)))
        L.check(L.p1_acc_sub_values(ctx, ptr(self.acc), ptr(self.slot_a), ptr(self._va), C))
        L.check(L.p1_acc_sub_values(ctx, ptr(self.acc), ptr(self.slot_b), ptr(self._vb), C))

    def _sequential_diag(self, lo, hi):
        """literal double loop, chained over row shards in rank order"""
        def step(xy):
            xy = np.ascontiguousarray(xy, dtype=np.float64)
            self.lib.check(self.lib.p1_diag_sequential(self.ctxs[0], self.data, lo, hi, self.precision, ptr(xy)))
            return xy
        if self.group is None:
            return step(np.zeros(hi - lo, dtype=np.float64))
        return chain_in_rank_order(step, hi - lo, self.group, self.dev)

    def accumulators(self):
        """host copy of every provider's uint64 accumulator, [num_dp, p1_acc_len(d)] (the share words before
        the diagonal / 64->32 tail of p1_finalize; combined over ranks once all_reduce() has run)"""
        self.sync()
        return self.acc.cpu().numpy().view(np.uint64).reshape(self.sched.num_dp, self.acc_len).copy()

    def finalize(self):
        """tail of run_party for every provider (src/phase1.c:605-641): returns
        shares[num_dp, d(d+1)/2 + d] as ufixed_t widened to uint64.  One kernel for all providers behind the
        all-reduce and one D2H; only a diagonal whose exact-integer shortcut does not apply (values far outside
        [-1, 1]) sends the step through the host path with the sequential double sum."""
        S, L, ctx = self.sched, self.lib, self.ctxs[0]
        L.check(L.p1_finalize_async(ctx, ptr(self.acc), S.num_dp, self.d, ptr(self.col_start_dev), self.precision,
                                    self.precision_p2, float(self.d), None, ptr(self.need_dev), ptr(self.out_dev)))
        L.check(L.p1_copy_d2h_async(ctx, c_vp(self.out_host.data_ptr()), ptr(self.out_dev), self.out_host.numel() * 8))
        L.check(L.p1_copy_d2h_async(ctx, c_vp(self.need_host.data_ptr()), ptr(self.need_dev), 4))
        self.sync()
        if int(self.need_host[0]) == 0:
            return self.out_host.numpy().view(np.uint64).reshape(S.num_dp, S.nA + self.d).copy()
        return self._finalize_host()

    def _finalize_host(self):
        S, L = self.sched, self.lib
        self.sync()
        acc_host = self.acc.cpu().numpy().view(np.uint64).reshape(S.num_dp, self.acc_len)
        out = np.zeros((S.num_dp, S.nA + self.d), dtype=np.uint64)
        wdt = np.uint64 if self.bits_p2 == 64 else np.uint32
        for k in range(S.num_dp):
            lo, hi = S.col_range(k)
            a = acc_host[k]
            sq_lo = a[S.nA + self.d + lo:S.nA + self.d + hi]
            sq_hi = a[S.nA + 2 * self.d + lo:S.nA + 2 * self.d + hi]
            viol = a[S.nA + 3 * self.d + lo:S.nA + 3 * self.d + hi]
            exact = [(int(v) == 0 and (int(h) << 32) + int(l) <= (1 << 53)) for l, h, v in zip(sq_lo, sq_hi, viol)]
            diag_xy = None
            if not all(exact):
                diag_xy = self._sequential_diag(lo, hi)
            A, B = c_vp(), c_vp()
            L.check(L.p1_finalize(self.ctxs[0], ptr(np.ascontiguousarray(a)), self.d, lo, hi, self.precision,
                                  self.precision_p2, float(self.d), ptr(diag_xy), None,
                                  ctypes.byref(A), ctypes.byref(B)))
            out[k, :S.nA] = np.ctypeslib.as_array(ctypes.cast(A, ctypes.POINTER(np.ctypeslib.as_ctypes_type(wdt))), (S.nA,))
            out[k, S.nA:] = np.ctypeslib.as_array(ctypes.cast(B, ctypes.POINTER(np.ctypeslib.as_ctypes_type(wdt))), (self.d,))
            L.free(A)
            L.free(B)
        return out

    # ------------------------------------------------------------------
    def ot_schedule(self):
        """The inner products of the --use_ot aggregation in the order of src/phase1.c:386-438: for every
        pair of providers (me < peer) one session, sender by the parity rule of :392; inside a session
        sender columns x receiver columns (:404-414), then (i, target) when the receiver is the last
        provider (:415-424) and (target, j) when the sender is (:426-438).
        Yields (sender provider, receiver provider, [(col_s, col_r, slot), ...])."""
        S, P, d = self.sched, self.sched.num_dp, self.d
        for me in range(P):
            for peer in range(me + 1, P):
                me_sends = ((me % 2 == peer % 2) == (me < peer))
                ps, pr = (me, peer) if me_sends else (peer, me)
                i_lo, i_hi = S.col_range(ps)
                j_lo, j_hi = S.col_range(pr)
                items = []
                for i in range(i_lo, i_hi):
                    items += [(i, j, int(idx(np.int64(i), np.int64(j)))) for j in range(j_lo, j_hi)]
                    if pr == P - 1:
                        items.append((i, d, S.nA + i))          # (i, target) -> share_b[i]
                if ps == P - 1:
                    items += [(d, j, S.nA + j) for j in range(j_lo, j_hi)]   # (target, j) -> share_b[j]
                yield ps, pr, items

    def cross_pairs_ot(self, ot_key, max_pairs=None):
        """--use_ot aggregation (src/phase1.c:386-449).  The base-OT outputs (seeds, Delta) are inputs.
        Inner products of a session alternate over the box's pipes (streams): every pipe holds its own pair
        of session objects built from the same seeds and skips the inner products of the others
        (p1_cot_advance), so each one runs at the cursor the reference's single loop gives it while the
        receiver's finish of one overlaps the next one's extension.  max_pairs: stop after that many inner
        products (bench samples).  Returns the number of inner products executed."""
        L, t = self.lib, self.torch
        ctx0 = self.ctxs[0]
        npipe = self.pipe_limit or len(self.ctxs)
        u_bytes = int(L.p1_cot_u_bytes(ctx0, self.n_rows))
        c_bytes = int(L.p1_cot_c_bytes(ctx0, self.n_rows))
        if getattr(self, "_ot_bufs", None) is None:
            self._ot_bufs = [(t.empty(u_bytes, dtype=t.uint8, device=self.dev), t.empty(c_bytes, dtype=t.uint8, device=self.dev))
                             for _ in self.ctxs]
            t.cuda.current_stream(self.dev).synchronize()
        acc = self.acc.data_ptr()
        done = 0
        self.ot_setup_s = 0.0           # host seconds of this call spent on session set-up and tear-down (bench.py reports it apart)
        self.ot_sessions = 0
        self._fan_out()
        for ps, pr, items in self.ot_schedule():
            if max_pairs is not None and done >= max_pairs:
                break
            t_setup = time.perf_counter()
            self.ot_sessions += 1
            seeds0, seeds1, delta = derive_cot_session(ot_key, ps + 2, pr + 2)
            chosen = np.ascontiguousarray(np.where(_bits128(delta)[:, None], seeds1, seeds0))
            sess = []
            for ctx in self.ctxs[:npipe]:
                snd, rcv = c_vp(), c_vp()
                L.check(L.p1_cot_sender_new(ctx, delta.tobytes(), ptr(chosen), ctypes.byref(snd)))
                L.check(L.p1_cot_receiver_new(ctx, ptr(seeds0), ptr(seeds1), ctypes.byref(rcv)))
                sess.append((snd, rcv))
            self.ot_setup_s += time.perf_counter() - t_setup
            for k, (col_s, col_r, slot) in enumerate(items):
                if max_pairs is not None and done >= max_pairs:
                    break
                pipe = k % npipe
                ctx, (snd, rcv), (ubuf, cbuf) = self.ctxs[pipe], sess[pipe], self._ot_bufs[pipe]
                L.check(L.p1_cot_receiver_extend(ctx, rcv, self.data, col_r, self.n_total, self.row0, ptr(ubuf)))
                L.check(L.p1_cot_sender_step(ctx, snd, self.data, col_s, self.n_total, self.row0, ptr(ubuf), ptr(cbuf),
                                             c_vp(acc), ps * self.acc_len + slot))
                L.check(L.p1_cot_receiver_finish(ctx, rcv, ptr(cbuf), c_vp(acc), pr * self.acc_len + slot))
                for q, (s2, r2) in enumerate(sess):
                    if q != pipe:
                        L.check(L.p1_cot_advance(s2, self.n_total, 1))
                        L.check(L.p1_cot_advance(r2, self.n_total, 1))
                done += 1
            for ctx, (snd, rcv) in zip(self.ctxs[:npipe], sess):
                L.check(L.p1_ctx_sync(ctx))
            t_setup = time.perf_counter()
            for snd, rcv in sess:
                L.p1_cot_free(snd)
                L.p1_cot_free(rcv)
            self.ot_setup_s += time.perf_counter() - t_setup
        self._fan_in()
        return done

    def run_ot(self, X, y, ot_key):
        """whole phase 1 with the OT-based aggregation; returns every provider's shares"""
        self.reset()
        self.encode(X, y)
        self.local_blocks()
        self.cross_pairs_ot(ot_key)
        self.all_reduce()
        return self.finalize()

    def run_masking(self, X, y, ti_key, seeded=False, wire=False, fused=None):
        """whole phase 1 with the TI masking protocol; returns every provider's shares"""
        self.reset()
        self.encode(X, y)
        self.local_blocks()
        self.cross_pairs_masking(ti_key, seeded=seeded, wire=wire, fused=fused)
        self.all_reduce()
        self.apply_values()
        return self.finalize()
