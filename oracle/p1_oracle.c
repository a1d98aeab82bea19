/* TEST INFRASTRUCTURE (oracle). Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference leg may call this; the product path never does.
 *
 * CPU restatement of linreg-mpc's phase-1 secure aggregation.  Every function cites
 * the reference lines it follows (paths are under /root/reference/).  The restatement
 * is pinned word-for-word against oracle/_ref (the reference's own phase1.c compiled
 * unmodified against shims) by tests/test_oracle.py, and through it against the
 * README golden vector (examples/readme_example.in, README.md:83-87).
 *
 * Ring words: bits_p1 in {64,32} selects Z_2^64 / Z_2^32 (src/fixed.h:4-16); buffers of
 * ring words are uint64_t[] or uint32_t[] accordingly.  bits_p2 is the phase-2 width
 * (src/fixed.h:18-30) that the encoder casts through (src/linear.c:51,87).
 */
#define _POSIX_C_SOURCE 201605L
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/types.h>

#include "aes128.h"
#include "cot_ext.h"

/* ---------------- word accessors ---------------- */

static inline uint64_t ld(const void *p, size_t i, int bits) {
	return bits == 64 ? ((const uint64_t *) p)[i] : (uint64_t) ((const uint32_t *) p)[i];
}
static inline void st(void *p, size_t i, int bits, uint64_t v) {
	if(bits == 64) ((uint64_t *) p)[i] = v;
	else ((uint32_t *) p)[i] = (uint32_t) v;
}
static inline uint64_t ring(uint64_t v, int bits) {
	return bits == 64 ? v : (v & 0xffffffffull);
}
/* signed value of a ring word (fixed_p1_t) */
static inline int64_t sgn(uint64_t v, int bits) {
	return bits == 64 ? (int64_t) v : (int64_t) (int32_t) (uint32_t) v;
}

/* ---------------- a2: scalar codec (src/fixed.c:3-17) ---------------- */

/* C cast double -> intN as x86-64 gcc compiles it (cvttsd2si): truncate toward zero,
 * "integer indefinite" (INT_MIN of the target width) when out of range or NaN. */
static inline int64_t cast_x86(double v, int bits) {
	if(bits == 64) {
		if(!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) return INT64_MIN;
		return (int64_t) v;
	}
	if(!(v > -2147483649.0 && v < 2147483648.0)) return (int64_t) INT32_MIN;
	return (int64_t) (int32_t) v;
}

/* double_to_fixed / double_to_fixed_p1 (src/fixed.c:3-5,11-13): (T)(d * (1ll << p)) */
int64_t p1o_double_to_fixed(double d, int p, int bits) {
	return cast_x86(d * (double) (1ll << p), bits);
}

/* fixed_to_double[_p1] (src/fixed.c:7-9,15-17): ((double) f) / (1ll << p) */
double p1o_fixed_to_double(int64_t f, int p) {
	return ((double) f) / (double) (1ll << p);
}

/* a4: normalizer = sqrt(pow(2,precision) * normalizer1), normalizer1 = n (src/phase1.c:473, src/config.c:47) */
double p1o_normalizer(int precision, double n) {
	return sqrt(pow(2, precision) * n);
}

/* a3: arithmetic of read_matrix / read_vector (src/linear.c:46-51,83-87): val /= normalizer;
 * word = (fixed_p1_t) double_to_fixed(val, precision) — cast through the PHASE-2 type. */
void p1o_encode(const double *vals, size_t count, double normalizer, int precision,
	int bits_p1, int bits_p2, void *out) {
	for(size_t i = 0; i < count; i++) {
		double val = vals[i];
		val /= normalizer;
		int64_t f = p1o_double_to_fixed(val, precision, bits_p2);
		st(out, i, bits_p1, (uint64_t) f); /* sign-extends 32->64 when P1=64, P2=32 */
	}
}

/* ---------------- a5/a6: index map and ownership ---------------- */

/* idx (src/linear.c:11-16) */
size_t p1o_idx(size_t i, size_t j) {
	if(j > i) {
		size_t t = i;
		i = j;
		j = t;
	}
	return (i * (i + 1)) / 2 + j;
}

/* get_owner (src/phase1.c:25-33), expressed over data providers only: col_start[k] is
 * index_owned[k+2]; returns k in [0, num_dp), the last provider whose start <= row.
 * Row d (the target) therefore belongs to the last provider. */
int p1o_owner(size_t row, int num_dp, const int64_t *col_start) {
	int k = 0;
	for(; k + 1 < num_dp && col_start[k + 1] <= (int64_t) row; k++);
	return k;
}

/* pair schedule (TI loop src/phase1.c:257-267; DP loop :534-579).  kind: 0 = diagonal,
 * 1 = local, 2 = cross.  q = running index of the cross pair in TI order (else -1).
 * Returns the number of pairs T = d(d+1)/2 + d; arrays may be NULL to just count. */
size_t p1o_schedule(size_t d, int num_dp, const int64_t *col_start,
	int64_t *pi, int64_t *pj, int32_t *owner_i, int32_t *owner_j, int32_t *kind, int64_t *q) {
	size_t t = 0;
	int64_t nq = 0;
	for(size_t i = 0; i <= d; i++) {
		for(size_t j = 0; j <= i && j < d; j++) {
			int oi = p1o_owner(i, num_dp, col_start), oj = p1o_owner(j, num_dp, col_start);
			int k = (oi != oj) ? 2 : (i == j ? 0 : 1);
			if(pi) {
				pi[t] = (int64_t) i;
				pj[t] = (int64_t) j;
				owner_i[t] = oi;
				owner_j[t] = oj;
				kind[t] = k;
				q[t] = k == 2 ? nq : -1;
			}
			if(k == 2) nq++;
			t++;
		}
	}
	return t;
}

/* ---------------- a7: inner_product_local (src/phase1.c:14-20) ---------------- */

uint64_t p1o_dot(const void *x, size_t sx, const void *y, size_t sy, size_t n, int bits) {
	uint64_t xy = 0;
	for(size_t i = 0; i < n; i++) {
		xy += ld(x, i * sx, bits) * ld(y, i * sy, bits);
	}
	return ring(xy, bits);
}

/* ---------------- a8: diagonal (src/phase1.c:562-567, OT copy :364-369) ---------------- */

/* continues the sequential double sum from *acc over rows [0,n) of one column */
void p1o_diag_accumulate(const void *col, size_t stride, size_t n, int precision, int bits, double *acc) {
	double xy = *acc;
	double scale = (double) (1ll << precision); /* pow(2, precision), exact */
	for(size_t k = 0; k < n; k++) {
		double v = p1o_fixed_to_double(sgn(ld(col, k * stride, bits), bits), precision);
		double sq = v * v;      /* pow(v, 2) */
		double term = sq * scale;
		xy = xy + term;          /* three separately rounded operations, no FMA */
	}
	*acc = xy;
}

/* share = double_to_fixed_p1(xy / normalizer2, precision) (src/phase1.c:567) */
uint64_t p1o_diag_finish(double xy, double normalizer2, int precision, int bits) {
	return ring((uint64_t) p1o_double_to_fixed(xy / normalizer2, precision, bits), bits);
}

uint64_t p1o_diag(const void *col, size_t stride, size_t n, int precision, double normalizer2, int bits) {
	double xy = 0;
	p1o_diag_accumulate(col, stride, n, precision, bits, &xy);
	return p1o_diag_finish(xy, normalizer2, precision, bits);
}

/* ---------------- a9: trusted initializer (src/phase1.c:270-278) ---------------- */

/* byte offset in the TI stream of cross pair q: per pair the TI draws x (n words),
 * y (n words), r (1 word), in this order (src/phase1.c:271-273) */
uint64_t p1o_ti_offset(uint64_t q, size_t n, int bits) {
	return q * (2 * (uint64_t) n + 1) * (uint64_t) (bits / 8);
}

void p1o_prg(const uint8_t key[16], uint64_t byte_offset, void *out, size_t len) {
	aes128_key k;
	aes128_set_key(&k, key);
	aes128_ctr_keystream(&k, byte_offset, (uint8_t *) out, len);
}

/* masks and values of cross pair q: value_a = <x,y> - r goes with y to owner(i),
 * value_b = r goes with x to owner(j) (src/phase1.c:253-254,274-278) */
void p1o_ti_pair(const uint8_t key[16], uint64_t q, size_t n, int bits,
	void *x, void *y, uint64_t *value_a, uint64_t *value_b) {
	aes128_key k;
	size_t w = (size_t) bits / 8;
	uint64_t off = p1o_ti_offset(q, n, bits), r = 0;
	aes128_set_key(&k, key);
	aes128_ctr_keystream(&k, off, (uint8_t *) x, n * w);
	aes128_ctr_keystream(&k, off + n * w, (uint8_t *) y, n * w);
	aes128_ctr_keystream(&k, off + 2 * n * w, (uint8_t *) &r, w);
	uint64_t xy = p1o_dot(x, 1, y, 1, n, bits);
	*value_a = ring(xy - r, bits);
	*value_b = ring(r, bits);
}

/* ---------------- a10: inner_product_ti party algebra (src/phase1.c:186-196,201-222) ---------------- */

/* party a (owns row i): out = a_i - T (T = y) */
void p1o_party_a_mask(const void *a_i, size_t stride, const void *T, size_t n, int bits, void *out) {
	for(size_t k = 0; k < n; k++) st(out, k, bits, ld(a_i, k * stride, bits) - ld(T, k, bits));
}
/* party b (owns row j): out = b_j + T (T = x) */
void p1o_party_b_mask(const void *b_j, size_t stride, const void *T, size_t n, int bits, void *out) {
	for(size_t k = 0; k < n; k++) st(out, k, bits, ld(b_j, k * stride, bits) + ld(T, k, bits));
}
/* party a: share = <in, T> - value, in = b_j + x */
uint64_t p1o_party_a_share(const void *in, const void *T, uint64_t value, size_t n, int bits) {
	return ring(p1o_dot(in, 1, T, 1, n, bits) - value, bits);
}
/* party b: share = <in, b_j> - value, in = a_i - y */
uint64_t p1o_party_b_share(const void *in, const void *b_j, size_t stride, uint64_t value, size_t n, int bits) {
	return ring(p1o_dot(in, 1, b_j, stride, n, bits) - value, bits);
}

/* ---------------- a12: OT-based inner product (src/phase1.c:43-96) ---------------- */

typedef struct {
	cot_sender s;
	cot_receiver r;
} p1o_ot_session;

p1o_ot_session *p1o_ot_session_new(const uint8_t ot_key[16], int sender_party, int receiver_party) {
	p1o_ot_session *ss = malloc(sizeof(*ss));
	cot_sender_init(&ss->s, ot_key, sender_party, receiver_party);
	cot_receiver_init(&ss->r, ot_key, sender_party, receiver_party);
	return ss;
}
void p1o_ot_session_free(p1o_ot_session *ss) {
	free(ss);
}
/* resume a session at a given PRG block / OT counter (tests of long sessions) */
void p1o_ot_session_set_cursor(p1o_ot_session *ss, uint64_t pos, uint64_t ot_base) {
	ss->s.pos = ss->r.pos = pos;
	ss->s.ot_base = ss->r.ot_base = ot_base;
}

/* One inner product <xs, xr>: sender holds xs (inner_product_ot_sender :53-72), receiver
 * holds xr (inner_product_ot_recver :74-96).  OT index ni = k*W + bit (:45-46); choice
 * bit = bit `bit` of xr[k] (:78-83); correlator opt1 = (1<<bit)*xs[k] + opt0 (:50);
 * sender share = -sum opt0 (:66-68); receiver share = +sum dest (:90-92).
 * u_out / c_out (optional) receive the two wire messages of cot_ext.h. */
void p1o_ot_inner_product(p1o_ot_session *ss, const void *xs, size_t stride_s,
	const void *xr, size_t stride_r, size_t n, int bits,
	uint64_t *share_sender, uint64_t *share_receiver, uint8_t *u_out, uint8_t *c_out) {
	size_t W = (size_t) bits, w = W / 8, m = n * W, tiles = (m + 127) / 128;
	uint64_t acc_s = 0, acc_r = 0;
	uint8_t u[2048], trows[2048], qrows[2048], h0[16], h1[16], ht[16], v[16];
	for(size_t t = 0; t < tiles; t++) {
		uint8_t choice[16] = {0};
		for(size_t b = 0; b < 128; b++) {
			size_t ni = 128 * t + b;
			if(ni >= m) break;
			uint64_t a = ld(xr, (ni / W) * stride_r, bits);
			if((a >> (ni % W)) & 1) choice[b / 8] |= (uint8_t) (1u << (b % 8));
		}
		cot_receiver_tile(&ss->r, ss->r.pos + t, choice, u, trows);
		cot_sender_tile(&ss->s, ss->s.pos + t, u, qrows);
		if(u_out) memcpy(u_out + 2048 * t, u, 2048);
		for(size_t b = 0; b < 128; b++) {
			size_t ni = 128 * t + b;
			if(ni >= m) break;
			uint64_t J = ss->s.ot_base + ni;
			size_t k = ni / W;
			int bit = (int) (ni % W);
			/* sender */
			cot_hash(&ss->s.hash, J, qrows + 16 * b, h0);
			uint64_t opt0 = 0, pad1 = 0, padt = 0;
			memcpy(&opt0, h0, w);
			uint64_t opt1 = ring((((uint64_t) 1) << bit) * ld(xs, k * stride_s, bits) + opt0, bits);
			for(int z = 0; z < 16; z++) v[z] = qrows[16 * b + z] ^ ss->s.delta[z];
			cot_hash(&ss->s.hash, J, v, h1);
			memcpy(&pad1, h1, w);
			uint64_t c = opt1 ^ pad1;
			if(c_out) memcpy(c_out + ni * w, &c, w);
			acc_s -= opt0;
			/* receiver */
			cot_hash(&ss->r.hash, J, trows + 16 * b, ht);
			memcpy(&padt, ht, w);
			int sel = (choice[b / 8] >> (b % 8)) & 1;
			acc_r += sel ? (padt ^ c) : padt;
		}
	}
	ss->s.pos += tiles;
	ss->r.pos += tiles;
	ss->s.ot_base += 128 * tiles;
	ss->r.ot_base += 128 * tiles;
	*share_sender = ring(acc_s, bits);
	*share_receiver = ring(acc_r, bits);
}

/* ---------------- a14: 64 -> 32 hand-off (src/phase1.c:609-620,629-635) ---------------- */

uint64_t p1o_downconvert(uint64_t share, int precision, int precision_p2) {
	return (uint64_t) (uint32_t) (((int64_t) share) >> (precision - precision_p2));
}

/* ---------------- whole phase 1, all parties (src/phase1.c:241-287,453-656) ---------------- */

typedef struct {
	size_t n, d;
	int num_dp, bits, precision, use_ot;
	const int64_t *col_start;
	const void *X; /* n x d row-major ring words */
	const void *y; /* n ring words */
	const uint8_t *ti_key;
	uint64_t *shares; /* num_dp x (nA + d), phase-1 ring words widened to u64 */
	/* work list */
	size_t npairs;
	int64_t *pi, *pj, *q;
	int32_t *oi, *oj, *kind;
	size_t next;
	pthread_mutex_t mu;
} run_ctx;

static const void *row_ptr(const run_ctx *rc, size_t i, size_t *stride) {
	/* src/phase1.c:537-551: rows < d are columns of X (stride d), row d is the target */
	size_t w = (size_t) rc->bits / 8;
	if(i < rc->d) {
		*stride = rc->d;
		return (const uint8_t *) rc->X + i * w;
	}
	*stride = 1;
	return rc->y;
}

static void store_share(run_ctx *rc, int party, size_t i, size_t j, uint64_t share) {
	/* src/phase1.c:581-585 */
	size_t nA = rc->d * (rc->d + 1) / 2;
	uint64_t *row = rc->shares + (size_t) party * (nA + rc->d);
	if(i < rc->d) row[p1o_idx(i, j)] = share;
	else row[nA + j] = share;
}

static void *masking_worker(void *v) {
	run_ctx *rc = v;
	size_t n = rc->n, w = (size_t) rc->bits / 8;
	void *x = malloc(n * w), *y = malloc(n * w), *m_a = malloc(n * w), *m_b = malloc(n * w);
	for(;;) {
		pthread_mutex_lock(&rc->mu);
		size_t t = rc->next++;
		pthread_mutex_unlock(&rc->mu);
		if(t >= rc->npairs) break;
		size_t i = (size_t) rc->pi[t], j = (size_t) rc->pj[t], si, sj;
		const void *ri = row_ptr(rc, i, &si), *rj = row_ptr(rc, j, &sj);
		if(rc->kind[t] == 0) {
			store_share(rc, rc->oi[t], i, j, p1o_diag(ri, si, n, rc->precision, (double) rc->d, rc->bits));
		} else if(rc->kind[t] == 1) {
			store_share(rc, rc->oi[t], i, j, p1o_dot(ri, si, rj, sj, n, rc->bits));
		} else {
			uint64_t va, vb;
			p1o_ti_pair(rc->ti_key, (uint64_t) rc->q[t], n, rc->bits, x, y, &va, &vb);
			p1o_party_b_mask(rj, sj, x, n, rc->bits, m_b); /* b -> a : b_j + x */
			p1o_party_a_mask(ri, si, y, n, rc->bits, m_a); /* a -> b : a_i - y */
			store_share(rc, rc->oi[t], i, j, p1o_party_a_share(m_b, y, va, n, rc->bits));
			store_share(rc, rc->oj[t], i, j, p1o_party_b_share(m_a, rj, sj, vb, n, rc->bits));
		}
	}
	free(x);
	free(y);
	free(m_a);
	free(m_b);
	return NULL;
}

static void run_ot(run_ctx *rc, const uint8_t *ot_key) {
	size_t n = rc->n, d = rc->d;
	int P = rc->num_dp;
	/* local blocks (src/phase1.c:359-384) */
	for(size_t t = 0; t < rc->npairs; t++) {
		size_t i = (size_t) rc->pi[t], j = (size_t) rc->pj[t], si, sj;
		const void *ri = row_ptr(rc, i, &si), *rj = row_ptr(rc, j, &sj);
		if(rc->kind[t] == 0) store_share(rc, rc->oi[t], i, j, p1o_diag(ri, si, n, rc->precision, (double) d, rc->bits));
		else if(rc->kind[t] == 1) store_share(rc, rc->oi[t], i, j, p1o_dot(ri, si, rj, sj, n, rc->bits));
	}
	/* one OT session per unordered provider pair; role rule src/phase1.c:392 on the
	 * 0-based party index (party-1 = provider index + 2; parity is unchanged by +2) */
	for(int me = 0; me < P; me++) {
		for(int peer = me + 1; peer < P; peer++) {
			int me_sends = ((me % 2 == peer % 2) == (me < peer));
			int ps = me_sends ? me : peer, pr = me_sends ? peer : me; /* party_i, party_j */
			size_t i_lo = (size_t) rc->col_start[ps], i_hi = ps + 1 < P ? (size_t) rc->col_start[ps + 1] : d;
			size_t j_lo = (size_t) rc->col_start[pr], j_hi = pr + 1 < P ? (size_t) rc->col_start[pr + 1] : d;
			p1o_ot_session *ss = p1o_ot_session_new(ot_key, ps + 2, pr + 2);
			uint64_t sh_s, sh_r;
			size_t si, sj;
			for(size_t i = i_lo; i < i_hi; i++) { /* src/phase1.c:404-425 */
				const void *ri = row_ptr(rc, i, &si);
				for(size_t j = j_lo; j < j_hi; j++) {
					const void *rj = row_ptr(rc, j, &sj);
					p1o_ot_inner_product(ss, ri, si, rj, sj, n, rc->bits, &sh_s, &sh_r, NULL, NULL);
					store_share(rc, ps, i, j, sh_s);
					store_share(rc, pr, i, j, sh_r);
				}
				if(pr == P - 1) {
					const void *rt = row_ptr(rc, d, &sj);
					p1o_ot_inner_product(ss, ri, si, rt, sj, n, rc->bits, &sh_s, &sh_r, NULL, NULL);
					store_share(rc, ps, d, i, sh_s);
					store_share(rc, pr, d, i, sh_r);
				}
			}
			if(ps == P - 1) { /* src/phase1.c:426-438 */
				const void *rt = row_ptr(rc, d, &si);
				for(size_t j = j_lo; j < j_hi; j++) {
					const void *rj = row_ptr(rc, j, &sj);
					p1o_ot_inner_product(ss, rt, si, rj, sj, n, rc->bits, &sh_s, &sh_r, NULL, NULL);
					store_share(rc, ps, d, j, sh_s);
					store_share(rc, pr, d, j, sh_r);
				}
			}
			p1o_ot_session_free(ss);
		}
	}
}

/* X, y: raw doubles as they appear in the instance file (row-major n x d, n).
 * shares_out: num_dp x (d(d+1)/2 + d) words of the PHASE-2 type widened to u64
 * (i.e. after the 64->32 conversion when bits_p1=64, bits_p2=32).
 * encoded_X / encoded_y (optional): the phase-1 ring words, bits_p1 wide. */
int p1o_run(const double *X, const double *y, size_t n, size_t d, int num_dp, const int64_t *col_start,
	int precision, int precision_p2, int bits_p1, int bits_p2, int use_ot,
	const uint8_t ti_key[16], const uint8_t ot_key[16], int threads,
	uint64_t *shares_out, void *encoded_X, void *encoded_y) {
	size_t w = (size_t) bits_p1 / 8, nA = d * (d + 1) / 2;
	run_ctx rc = {.n = n, .d = d, .num_dp = num_dp, .bits = bits_p1, .precision = precision, .use_ot = use_ot,
		.col_start = col_start, .ti_key = ti_key, .shares = shares_out};
	void *Xe = malloc(n * d * w), *ye = malloc(n * w);
	double normalizer = p1o_normalizer(precision, (double) n);
	p1o_encode(X, n * d, normalizer, precision, bits_p1, bits_p2, Xe);
	p1o_encode(y, n, normalizer, precision, bits_p1, bits_p2, ye);
	if(encoded_X) memcpy(encoded_X, Xe, n * d * w);
	if(encoded_y) memcpy(encoded_y, ye, n * w);
	rc.X = Xe;
	rc.y = ye;
	memset(shares_out, 0, (size_t) num_dp * (nA + d) * sizeof(uint64_t)); /* calloc, src/phase1.c:482-483 */

	rc.npairs = p1o_schedule(d, num_dp, col_start, NULL, NULL, NULL, NULL, NULL, NULL);
	rc.pi = malloc(rc.npairs * sizeof(int64_t));
	rc.pj = malloc(rc.npairs * sizeof(int64_t));
	rc.q = malloc(rc.npairs * sizeof(int64_t));
	rc.oi = malloc(rc.npairs * sizeof(int32_t));
	rc.oj = malloc(rc.npairs * sizeof(int32_t));
	rc.kind = malloc(rc.npairs * sizeof(int32_t));
	p1o_schedule(d, num_dp, col_start, rc.pi, rc.pj, rc.oi, rc.oj, rc.kind, rc.q);

	if(use_ot) {
		run_ot(&rc, ot_key);
	} else {
		if(threads < 1) threads = 1;
		pthread_t *th = malloc((size_t) threads * sizeof(pthread_t));
		pthread_mutex_init(&rc.mu, NULL);
		rc.next = 0;
		for(int t = 0; t < threads; t++) pthread_create(&th[t], NULL, masking_worker, &rc);
		for(int t = 0; t < threads; t++) pthread_join(th[t], NULL);
		pthread_mutex_destroy(&rc.mu);
		free(th);
	}
	if(bits_p1 == 64 && bits_p2 == 32) {
		for(size_t k = 0; k < (size_t) num_dp * (nA + d); k++) shares_out[k] = p1o_downconvert(shares_out[k], precision, precision_p2);
	}
	free(Xe);
	free(ye);
	free(rc.pi);
	free(rc.pj);
	free(rc.q);
	free(rc.oi);
	free(rc.oj);
	free(rc.kind);
	return 0;
}

/* ---------------- instance files (bench.py cpu_baseline leg) ----------------
 * Writes X, y in the input-file format parsed by src/config.c:24-44 and
 * src/linear.c:27-102 (layout of experiments/generate_tests.py:41-74); %.17g
 * round-trips every double through the reference's fscanf("%lf"). */
#include <stdio.h>
int p1o_write_instance(const char *path, const double *X, const double *y, size_t n, size_t d,
	int num_dp, const int64_t *col_start) {
	FILE *f = fopen(path, "w");
	if(!f) return 1;
	static char buf[1 << 20];
	setvbuf(f, buf, _IOFBF, sizeof(buf));
	fprintf(f, "%zu %zu %d\nlocalhost:1234\nlocalhost:1235\n", n, d, num_dp);
	for(int k = 0; k < num_dp; k++) fprintf(f, "localhost:%d %lld\n", 1236 + k, (long long) col_start[k]);
	fprintf(f, "%zu %zu\n", n, d);
	for(size_t i = 0; i < n; i++) {
		for(size_t j = 0; j < d; j++) fprintf(f, j + 1 < d ? "%.17g " : "%.17g\n", X[i * d + j]);
	}
	fprintf(f, "%zu\n", n);
	for(size_t i = 0; i < n; i++) fprintf(f, i + 1 < n ? "%.17g " : "%.17g\n", y[i]);
	return fclose(f) ? 1 : 0;
}

/* ---------------- f2: SecureMultiplication__Msg on the wire ----------------
 * src/protobuf/secure_multiplication{,_32}.proto:
 *     message msg { repeated uintW vector = 1 [packed=true]; required uintW value = 2; }
 * serialised by protobuf-c's generated pack_to_buffer (src/phase1.c:137-139) and parsed by
 * secure_multiplication__msg__unpack (:111).  protobuf-c is a system package, absent from
 * /root/reference; this restates the published proto2 wire format:
 *     [0x0A][varint payload_bytes][varint(vector[0]) .. varint(vector[n-1])]   (omitted when n == 0)
 *     [0x10][varint(value)]
 * base-128 varints, least significant group first, MSB = continuation.  Pinned against the
 * google.protobuf runtime by tests/golden/msg_wire.json (tests/golden/make_wire_golden.py). */

static size_t vi_size(uint64_t v) {
	size_t s = 1;
	for(; v >= 0x80; v >>= 7) s++;
	return s;
}
static size_t vi_put(uint8_t *out, uint64_t v) {
	size_t s = 0;
	for(; v >= 0x80; v >>= 7) out[s++] = (uint8_t) (v | 0x80);
	out[s++] = (uint8_t) v;
	return s;
}

size_t p1o_msg_packed_size(const void *vec, size_t n, uint64_t value, int bits) {
	size_t payload = 0;
	for(size_t i = 0; i < n; i++) payload += vi_size(ld(vec, i, bits));
	return (n ? 1 + vi_size(payload) + payload : 0) + 1 + vi_size(ring(value, bits));
}

size_t p1o_msg_pack(const void *vec, size_t n, uint64_t value, int bits, uint8_t *out) {
	size_t o = 0;
	if(n) {
		size_t payload = 0;
		for(size_t i = 0; i < n; i++) payload += vi_size(ld(vec, i, bits));
		out[o++] = 0x0A;
		o += vi_put(out + o, payload);
		for(size_t i = 0; i < n; i++) o += vi_put(out + o, ld(vec, i, bits));
	}
	out[o++] = 0x10;
	o += vi_put(out + o, ring(value, bits));
	return o;
}

/* varint of at most 10 bytes; bits above 64 are dropped as protobuf parsers do */
static int vi_get(const uint8_t **p, const uint8_t *end, uint64_t *v) {
	uint64_t r = 0;
	for(int k = 0; k < 10; k++) {
		if(*p >= end) return 0;
		uint8_t b = *(*p)++;
		if(7 * k < 64) r |= (uint64_t) (b & 0x7f) << (7 * k);
		if(!(b & 0x80)) {
			*v = r;
			return 1;
		}
	}
	return 0;
}

/* canonical form only (field 1 packed, then field 2).  Returns 0 on success; *n_out words
 * are written to vec (capacity cap), the value to *value. */
int p1o_msg_unpack(const uint8_t *in, size_t len, int bits, void *vec, size_t cap, size_t *n_out, uint64_t *value) {
	const uint8_t *p = in, *end = in + len;
	size_t n = 0;
	uint64_t v;
	if(p < end && *p == 0x0A) {
		p++;
		if(!vi_get(&p, end, &v) || v > (uint64_t) (end - p)) return 1;
		const uint8_t *pe = p + v;
		while(p < pe) {
			if(!vi_get(&p, pe, &v)) return 1;
			if(n >= cap) return 1;
			st(vec, n++, bits, v);
		}
	}
	if(p >= end || *p != 0x10) return 1;
	p++;
	if(!vi_get(&p, end, &v) || p != end) return 1;
	*value = ring(v, bits);
	*n_out = n;
	return 0;
}

/* ---------------- f3: share hand-off to phase 2 (src/input.c:37-42) ----------------
 * dcsSendIntArray: sel[i*intsize + j] = (input[i] >> j) & 1, intsize = FIXED_BIT_SIZE_P2,
 * one bool (byte) per bit: the OT choice bits of the garbled-circuit input. */
void p1o_share_bits(const void *shares, size_t n, int bits_p2, uint8_t *sel) {
	for(size_t i = 0; i < n; i++)
		for(int j = 0; j < bits_p2; j++) sel[i * (size_t) bits_p2 + j] = (uint8_t) ((ld(shares, i, bits_p2) >> j) & 1);
}

/* ---------------- f4: text ingestion (src/linear.c:27-102) ----------------
 * read_matrix / read_vector consume the file with fscanf("%lf"): skip white space, then the
 * longest prefix strtod accepts.  The restatement walks the white-space separated tokens and
 * converts each with glibc's strtod (what fscanf uses); whole[k] tells whether strtod consumed
 * the entire token with the plain decimal grammar (no inf / nan / hex float), which is the
 * set the GPU path converts itself.  Returns the number of tokens. */
#include <ctype.h>
size_t p1o_parse_doubles(const char *text, size_t bytes, double *out, uint8_t *whole, size_t max_tokens) {
	size_t k = 0, i = 0;
	char tmp[512];
	while(i < bytes) {
		while(i < bytes && isspace((unsigned char) text[i])) i++;
		if(i >= bytes) break;
		size_t j = i;
		while(j < bytes && !isspace((unsigned char) text[j])) j++;
		if(k < max_tokens) {
			size_t len = j - i;
			int ok = len < sizeof(tmp);
			double v = NAN;
			if(ok) {
				memcpy(tmp, text + i, len);
				tmp[len] = 0;
				char *end = NULL;
				v = strtod(tmp, &end);
				ok = end == tmp + len;
				for(size_t c = 0; c < len && ok; c++) {
					char ch = tmp[c];
					if(!(isdigit((unsigned char) ch) || ch == '+' || ch == '-' || ch == '.' || ch == 'e' || ch == 'E')) ok = 0;
				}
			}
			out[k] = v;
			whole[k] = (uint8_t) ok;
		}
		k++;
		i = j;
	}
	return k;
}
