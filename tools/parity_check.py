"""TEST INFRASTRUCTURE: full-size parity properties of one phase-1 run,
used by tests/test_gpu_fullsize.py and by bench.py's `check` block (outside the timed region).

At BASELINE sizes the CPU oracle cannot replay every word (hours), so a run is checked through
  1. the share algebra on ALL words: for every off-diagonal index and every share_b entry the
     providers' (pre-shift) share words sum, mod 2^W, to the plain ring Gram of the encoded data,
     computed by an independent path - ONE provider owning every column, tcgen05 Gram over the
     rank's rows, the ranks combined by the same wrapping-u64 all-reduce;
  2. all diagonal words: the owner's word equals the one-provider run's, every other provider
     holds 0 (calloc, src/phase1.c:482-483,559-560), and the three statistics rows agree;
  3. sampled words of that Gram against the CPU oracle (`oracle.dot`, `oracle.diag`, i.e.
     src/phase1.c:14-20 and :562-567) on columns read back from the device, rank slices combined
     with the ring sum (dot) or gathered in rank order (the sequential double diagonal);
  3b. sampled CROSS pairs, both providers' masked share words against the CPU oracle at the pair's true
     position q of the TI stream (`oracle.ti_pair`, src/phase1.c:270-278): share_a = <b_j + x, y> - (<x,y> - r)
     (:186-196) in the accumulator of i's owner, share_b = <a_i - y, b_j> - r (:201-222) in j's; the sum invariant
     of 1. cannot see a wrong mask (it cancels between the two parties), these words can.  The sample holds
     q = 0, the last pair and the pairs around the first and the last multiple of 2^32 keystream blocks;
  4. the 64 -> 32 hand-off (src/phase1.c:609-620,629-635) on every word against
     `oracle.downconvert` when the configuration has one.
The product path never imports this file.
"""
import ctypes

import numpy as np

c_vp = ctypes.c_void_p


def column(lib, ctx, data, c, n_rows, bits):
    """ring words of column c (d = the target) of a p1_data, read back from the device"""
    from oracle import pyoracle as po
    ld = ctypes.c_size_t()
    base = lib.p1_data_colmajor_dev(data, ctypes.byref(ld))
    w = bits // 8
    out = np.empty(n_rows, dtype=po.word_dtype(bits))
    lib.check(lib.p1_copy_d2h(ctx, out.ctypes.data_as(c_vp), c_vp(base + c * ld.value * w), n_rows * w))
    return out


def _ring_sum_over_ranks(vals, group, dev):
    """wrapping u64 sum over the ranks of a small host array"""
    if group is None:
        return vals
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(vals.view(np.int64).copy()).to(dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy().view(np.uint64)


def _gather_rows(col, per, group, dev):
    """the whole column on every rank, rank slices in order (short slices zero padded: a zero row
    adds +0.0 to the sequential double sum and 0 to a ring dot product)"""
    if group is None:
        return col
    import torch
    import torch.distributed as dist
    pad = np.zeros(per, dtype=col.dtype)
    pad[:len(col)] = col
    t = torch.from_numpy(pad.view(np.int64 if col.dtype == np.uint64 else np.int32).copy()).to(dev)
    out = torch.empty(per * dist.get_world_size(group), dtype=t.dtype, device=dev)
    dist.all_gather_into_tensor(out, t, group=group)
    return out.cpu().numpy().view(col.dtype)


def sample_cross_pairs(C, n_total, bits, count, rng):
    """pair numbers q to replay: both ends, the pairs around the first and the last crossing of a multiple of 2^32
    AES blocks of the TI stream (pair q starts at word q * (2 n + 1)), the rest drawn uniformly"""
    if C <= count:
        return list(range(C))
    words_per_block = 128 // bits
    picks = {0, 1, C - 2, C - 1}
    per_pair = 2 * n_total + 1
    last_block = (C * per_pair) // words_per_block
    for m in sorted({1, last_block >> 32} - {0}):
        qx = (m << 32) * words_per_block // per_pair             # the pair that holds block m * 2^32
        picks |= {q for q in (qx - 1, qx, qx + 1) if 0 <= q < C}
    while len(picks) < count:
        picks.add(int(rng.integers(0, C)))
    return sorted(picks)


def check_cross_pairs(lib, oracle, box, acc, ti_key, pairs_done=None, count=16, seed=12):
    """3b of the module text.  acc: box.accumulators() (combined over ranks, values applied).  Collective over box.group."""
    S, d, bits = box.sched, box.d, box.bits_p1
    mask = np.uint64((1 << bits) - 1)
    C = S.num_cross if pairs_done is None else min(S.num_cross, int(pairs_done))
    if not C:
        return {"sampled_cross_pairs": 0}
    qs = sample_cross_pairs(C, box.n_total, bits, count, np.random.default_rng(seed))
    ctx = box.ctxs[0]
    cols = {}
    col = lambda c: cols.setdefault(int(c), column(lib, ctx, box.data, int(c), box.n_rows, bits))
    rows = slice(box.row0, box.row0 + box.n_rows)
    partial, vals, got = [], [], []
    with np.errstate(over="ignore"):
        for q in qs:
            t = S.cross[q]
            i, j = int(S.i[t]), int(S.j[t])
            x, y, va, vb = oracle.ti_pair(ti_key, q, box.n_total, bits)        # the whole vectors at the pair's offset
            x, y = x[rows], y[rows]
            a_i, b_j = col(i), col(j)
            partial.append(oracle.dot(b_j + x, y, bits))                         # party a: <b_j + x, y>
            partial.append(oracle.dot(a_i - y, b_j, bits))                       # party b: <a_i - y, b_j>
            vals += [va, vb]
            slot = int(S.slot[t])
            got += [acc[int(S.owner_i[t]), slot], acc[int(S.owner_j[t]), slot]]
        dots = _ring_sum_over_ranks(np.array(partial, dtype=np.uint64), box.group, box.dev)
        want = (dots - np.array(vals, dtype=np.uint64)) & mask
    got = np.array(got, dtype=np.uint64) & mask
    blocks = [q * (2 * box.n_total + 1) * bits // 128 >> 32 for q in qs]
    return {"sampled_cross_pairs": len(qs), "sampled_cross_pair_words": 2 * len(qs),
            "sampled_cross_pair_words_match": bool(np.array_equal(got, want)),
            "sampled_cross_pair_q": [int(q) for q in qs],
            "sampled_cross_pair_2p32_block_epochs": sorted(set(int(b) for b in blocks))}


def check_run(lib, oracle, box, shares, X, y, samples=16, diag_samples=4, seed=11, only_slots=None,
              ti_key=None, pairs_done=None, cross_samples=16):
    """`box`: the Phase1Box that produced `shares` (its accumulators and column store are still
    alive); X, y: the rank's rows as given to box.encode.  Collective over box.group: every rank
    must call it.  only_slots: restrict the sum invariant to these accumulator slots (a run that executed
    only part of the cross-pair schedule).  ti_key: the TI key of a masking run with the single-key stream (not seeded
    masks): both providers' words of `cross_samples` cross pairs (among the first pairs_done, default all) are replayed on
    the CPU.  Returns a dict of booleans / counts (meaningful on every rank)."""
    from linreg_mpc_b200 import phase1
    S, d, nA = box.sched, box.d, box.sched.nA
    bits, bits2 = box.bits_p1, box.bits_p2
    mask = np.uint64((1 << bits) - 1)
    group, dev = box.group, box.dev
    acc = box.accumulators()                         # [num_dp, acc_len] u64, combined over ranks, values applied
    cross = check_cross_pairs(lib, oracle, box, acc, ti_key, pairs_done, cross_samples, seed + 1) if ti_key is not None else {}
    # ---- the independent path: one provider, tcgen05 Gram over all d (+ target) columns
    one = phase1.Phase1Box(lib, box.n_total, d, [0], box.precision, box.precision_p2, bits, bits2, device=box.device,
                           row0=box.row0, n_rows=box.n_rows, pipes=1, gram_algo=phase1.ffi.P1_GRAM_TCGEN05, group=group)
    plain_shares = one.run_masking(X, y, bytes(16))[0]
    plain_acc = one.accumulators()[0]
    out = {}
    with np.errstate(over="ignore"):
        total = np.zeros(nA + d, dtype=np.uint64)
        for k in range(S.num_dp):
            total += acc[k, :nA + d]
    diag_idx = np.array([i * (i + 1) // 2 + i for i in range(d)])
    off = np.ones(nA + d, dtype=bool)
    off[diag_idx] = False
    if only_slots is not None:
        sel = np.zeros(nA + d, dtype=bool)
        sel[np.asarray(only_slots, dtype=np.int64)] = True
        off &= sel
    out["words_checked"] = int(off.sum())
    out["sum_invariant_all_words"] = bool(np.array_equal(total[off] & mask, plain_acc[:nA + d][off] & mask))
    # ---- diagonals: owner's word == one-provider run's, zero elsewhere; statistics rows equal
    own = phase1.owners(np.arange(d), S.col_start)
    ok = True
    for k in range(S.num_dp):
        mine = own == k
        ok &= bool(np.array_equal(shares[k][diag_idx[mine]], plain_shares[diag_idx[mine]]))
        ok &= not shares[k][diag_idx[~mine]].any()
        for row in range(1, 4):
            sl = slice(nA + row * d, nA + (row + 1) * d)
            ok &= bool(np.array_equal(acc[k, sl][mine], plain_acc[sl][mine]))
    out["all_diagonals_match"] = bool(ok)
    # ---- the hand-off shift on every word
    if bits == 64 and bits2 == 32:
        sh = box.precision - box.precision_p2
        want = (acc[:, :nA + d].view(np.int64) >> sh).astype(np.uint32).astype(np.uint64)
        offs = np.flatnonzero(off)
        if not len(offs):
            offs = np.arange(1)
        okc = bool(np.array_equal(shares[:, offs], want[:, offs]))
        rng = np.random.default_rng(seed + 1)
        for k, j in zip(rng.integers(0, S.num_dp, 256), rng.choice(offs, 256)):       # the C oracle on a sample
            okc &= int(shares[k, j]) == oracle.lib.p1o_downconvert(int(acc[k, j]), box.precision, box.precision_p2)
        out["downconvert_all_words"] = okc
    # ---- sampled words against the CPU oracle on read-back columns
    ctx = one.ctxs[0]
    rng = np.random.default_rng(seed)
    cols = {}
    col = lambda c: cols.setdefault(int(c), column(lib, ctx, one.data, int(c), box.n_rows, bits))
    picks, partial = [], []
    for s in range(samples):
        if s % 4 == 3:                                  # share_b words
            j = int(rng.integers(0, d))
            picks.append(nA + j)
            partial.append(oracle.dot(col(d), col(j), bits))
        else:
            i, j = sorted(rng.choice(d, size=2, replace=False))[::-1] if d > 1 else (0, 0)
            if i == j:
                continue
            picks.append(int(i) * (int(i) + 1) // 2 + int(j))
            partial.append(oracle.dot(col(i), col(j), bits))
    got = _ring_sum_over_ranks(np.array(partial, dtype=np.uint64), group, dev) & mask
    out["sampled_words"] = len(picks)
    out["sampled_words_match"] = bool(np.array_equal(got, plain_acc[np.array(picks, dtype=np.int64)] & mask)) if picks else True
    per = phase1.shard_rows(box.n_total, 1 if group is None else __import__("torch").distributed.get_world_size(group), 0)[1]
    okd = True
    for i in rng.choice(d, size=min(diag_samples, d), replace=False):
        whole = _gather_rows(col(i), per, group, dev)
        word = oracle.diag(whole, box.precision, float(d), bits)
        if bits == 64 and bits2 == 32:
            word = oracle.lib.p1o_downconvert(word, box.precision, box.precision_p2)
        okd &= int(plain_shares[diag_idx[int(i)]]) == word
    out["sampled_diagonals"] = int(min(diag_samples, d))
    out["sampled_diagonals_match"] = bool(okd)
    one.close()
    out.update(cross)
    out["ok"] = all(v for k, v in out.items() if isinstance(v, bool))
    return out
