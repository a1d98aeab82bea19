/* base_ot.c — the kappa = 128 base OTs of one OT-extension session, over the peer socket.
 *
 * The reference gets them inside Obliv-C: honestOTExtSenderNew / honestOTExtRecverNew
 * (src/phase1.c:394,397) run Naor-Pinkas base OTs (dhRandomInit, :391) that Obliv-C implements on
 * libgcrypt.  Obliv-C is not part of the reference tree, so the drop-in host file ships its own
 * provider behind the hook p1host_base_ot (phase1_gpu.c): the "simplest OT" of Chou and Orlandi
 * (LATINCRYPT 2015) on the twisted Edwards curve of Ed25519, semi-honest like the reference's
 * "honest" OT extension, with only osend / orecv for transport.  A build against Obliv-C may override
 * the hook with the npot glue of INTEGRATION.md instead.
 *
 * Roles are swapped as in every IKNP-style extension: the extension SENDER holds the choice string
 * Delta and learns k_i^{Delta_i}; the extension RECEIVER learns both k_i^0, k_i^1.
 *
 *   base-OT sender S (extension receiver)            base-OT receiver R (extension sender)
 *   a <- Z_q, A = aG                   -- A -->      Delta <- {0,1}^128
 *                                                    b_i <- Z_q, B_i = b_i G + Delta_i A
 *                                      <-- B_0..127
 *   k_i^0 = H(i, a B_i)                              k_i^{Delta_i} = H(i, b_i A)
 *   k_i^1 = H(i, a B_i - a A)
 *
 * H = SHA-256 over a domain tag, the two party indices, i, A and B_i as sent, and the affine shared point, cut to
 * 16 bytes.  Points travel as affine (x, y), 64 bytes, and are checked to lie on the curve; an A of small order
 * (8 A = neutral) is refused; scalars are multiples of the cofactor 8, which removes small-order components.
 * Secrets are wiped with explicit_bzero.  SEMI-HONEST ONLY, like the reference's "honest" OT extension: nothing here
 * defends against a peer that deviates from the protocol beyond those two point checks.
 * Field and curve arithmetic: radix-2^16 limbs in int64 (constant-time selects), the textbook extended-coordinates
 * addition law; the limb carry chain follows TweetNaCl's car25519 (public domain), as does the 16-limb layout.
 */
#define _POSIX_C_SOURCE 201605L
#define _DEFAULT_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/random.h>

typedef struct ProtocolDesc ProtocolDesc;
int osend(ProtocolDesc *pd, int dest, const void *data, size_t len);
int orecv(ProtocolDesc *pd, int src, void *data, size_t len);

typedef int64_t gf[16];

static const gf GF0 = {0}, GF1 = {1};
/* d = -121665/121666, 2d, and the base point (x, 4/5) of Ed25519, little-endian 16-bit limbs */
static const gf ED_D = {0x78a3, 0x1359, 0x4dca, 0x75eb, 0xd8ab, 0x4141, 0x0a4d, 0x0070, 0xe898, 0x7779, 0x4079, 0x8cc7, 0xfe73, 0x2b6f, 0x6cee, 0x5203};
static const gf ED_D2 = {0xf159, 0x26b2, 0x9b94, 0xebd6, 0xb156, 0x8283, 0x149a, 0x00e0, 0xd130, 0xeef3, 0x80f2, 0x198e, 0xfce7, 0x56df, 0xd9dc, 0x2406};
static const gf ED_X = {0xd51a, 0x8f25, 0x2d60, 0xc956, 0xa7b2, 0x9525, 0xc760, 0x692c, 0xdc5c, 0xfdd6, 0xe231, 0xc0a4, 0x53fe, 0xcd6e, 0x36d3, 0x2169};
static const gf ED_Y = {0x6658, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666, 0x6666};

static void gf_set(gf r, const gf a) { for(int i = 0; i < 16; i++) r[i] = a[i]; }

static void gf_carry(gf o) {
	for(int i = 0; i < 16; i++) {
		o[i] += (1LL << 16);
		const int64_t c = o[i] >> 16;
		o[(i + 1) * (i < 15)] += c - 1 + 37 * (c - 1) * (i == 15);
		o[i] -= c << 16;
	}
}

/* swap p and q iff b == 1, without a branch */
static void gf_cswap(gf p, gf q, int b) {
	const int64_t c = ~((int64_t) b - 1);
	for(int i = 0; i < 16; i++) {
		const int64_t t = c & (p[i] ^ q[i]);
		p[i] ^= t;
		q[i] ^= t;
	}
}

/* canonical little-endian bytes of n mod 2^255 - 19 */
static void gf_pack(uint8_t o[32], const gf n) {
	gf m, t;
	gf_set(t, n);
	gf_carry(t);
	gf_carry(t);
	gf_carry(t);
	for(int j = 0; j < 2; j++) {
		m[0] = t[0] - 0xffed;
		for(int i = 1; i < 15; i++) {
			m[i] = t[i] - 0xffff - ((m[i - 1] >> 16) & 1);
			m[i - 1] &= 0xffff;
		}
		m[15] = t[15] - 0x7fff - ((m[14] >> 16) & 1);
		const int b = (int) ((m[15] >> 16) & 1);
		m[14] &= 0xffff;
		gf_cswap(t, m, 1 - b);
	}
	for(int i = 0; i < 16; i++) {
		o[2 * i] = (uint8_t) (t[i] & 0xff);
		o[2 * i + 1] = (uint8_t) (t[i] >> 8);
	}
}

static void gf_unpack(gf o, const uint8_t n[32]) {
	for(int i = 0; i < 16; i++) o[i] = n[2 * i] + ((int64_t) n[2 * i + 1] << 8);
	o[15] &= 0x7fff;
}

static void gf_add(gf o, const gf a, const gf b) { for(int i = 0; i < 16; i++) o[i] = a[i] + b[i]; }
static void gf_sub(gf o, const gf a, const gf b) { for(int i = 0; i < 16; i++) o[i] = a[i] - b[i]; }

static void gf_mul(gf o, const gf a, const gf b) {
	int64_t t[31];
	for(int i = 0; i < 31; i++) t[i] = 0;
	for(int i = 0; i < 16; i++) for(int j = 0; j < 16; j++) t[i + j] += a[i] * b[j];
	for(int i = 0; i < 15; i++) t[i] += 38 * t[i + 16];
	for(int i = 0; i < 16; i++) o[i] = t[i];
	gf_carry(o);
	gf_carry(o);
}

static void gf_inv(gf o, const gf in) {      /* in^(p-2) */
	gf c;
	gf_set(c, in);
	for(int a = 253; a >= 0; a--) {
		gf_mul(c, c, c);
		if(a != 2 && a != 4) gf_mul(c, c, in);
	}
	gf_set(o, c);
}

static int gf_equal(const gf a, const gf b) {
	uint8_t x[32], y[32];
	gf_pack(x, a);
	gf_pack(y, b);
	uint8_t d = 0;
	for(int i = 0; i < 32; i++) d |= x[i] ^ y[i];
	return d == 0;
}

/* points in extended coordinates (X : Y : Z : T), x = X/Z, y = Y/Z, xy = T/Z */
typedef struct { gf x, y, z, t; } ge;

static void ge_add(ge *p, const ge *q) {      /* p += q, unified (also doubles) */
	gf a, b, c, d, t, e, f, g, h;
	gf_sub(a, p->y, p->x);
	gf_sub(t, q->y, q->x);
	gf_mul(a, a, t);
	gf_add(b, p->x, p->y);
	gf_add(t, q->x, q->y);
	gf_mul(b, b, t);
	gf_mul(c, p->t, q->t);
	gf_mul(c, c, ED_D2);
	gf_mul(d, p->z, q->z);
	gf_add(d, d, d);
	gf_sub(e, b, a);
	gf_sub(f, d, c);
	gf_add(g, d, c);
	gf_add(h, b, a);
	gf_mul(p->x, e, f);
	gf_mul(p->y, h, g);
	gf_mul(p->z, g, f);
	gf_mul(p->t, e, h);
}

static void ge_cswap(ge *p, ge *q, int b) {
	gf_cswap(p->x, q->x, b);
	gf_cswap(p->y, q->y, b);
	gf_cswap(p->z, q->z, b);
	gf_cswap(p->t, q->t, b);
}

static void ge_neg(ge *p) {
	gf_sub(p->x, GF0, p->x);
	gf_sub(p->t, GF0, p->t);
}

/* r = s * q, Montgomery ladder over the 256 bits of s */
static void ge_scalarmult(ge *r, const ge *q_in, const uint8_t s[32]) {
	ge p, q = *q_in;
	gf_set(p.x, GF0);
	gf_set(p.y, GF1);
	gf_set(p.z, GF1);
	gf_set(p.t, GF0);
	for(int i = 255; i >= 0; i--) {
		const int b = (s[i / 8] >> (i & 7)) & 1;
		ge_cswap(&p, &q, b);
		ge_add(&q, &p);
		ge_add(&p, &p);
		ge_cswap(&p, &q, b);
	}
	*r = p;
}

static void ge_base(ge *g) {
	gf_set(g->x, ED_X);
	gf_set(g->y, ED_Y);
	gf_set(g->z, GF1);
	gf_mul(g->t, ED_X, ED_Y);
}

static void ge_to_bytes(uint8_t out[64], const ge *p) {      /* affine x || y */
	gf zi, x, y;
	gf_inv(zi, p->z);
	gf_mul(x, p->x, zi);
	gf_mul(y, p->y, zi);
	gf_pack(out, x);
	gf_pack(out + 32, y);
}

/* 0 on success; rejects anything that is not a canonical point of the curve -x^2 + y^2 = 1 + d x^2 y^2 */
static int ge_from_bytes(ge *p, const uint8_t in[64]) {
	gf x2, y2, l, r;
	uint8_t chk[64];
	gf_unpack(p->x, in);
	gf_unpack(p->y, in + 32);
	gf_set(p->z, GF1);
	gf_mul(p->t, p->x, p->y);
	gf_pack(chk, p->x);
	gf_pack(chk + 32, p->y);
	if(memcmp(chk, in, 64)) return 1;               /* non-canonical encoding */
	gf_mul(x2, p->x, p->x);
	gf_mul(y2, p->y, p->y);
	gf_sub(l, y2, x2);
	gf_mul(r, x2, y2);
	gf_mul(r, r, ED_D);
	gf_add(r, r, GF1);
	return gf_equal(l, r) ? 0 : 1;
}

/* ------------------------------------------------------------------ SHA-256 (FIPS 180-4) */

static const uint32_t SHA_K[64] = {
	0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3,
	0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
	0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
	0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
	0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
	0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static uint32_t rotr(uint32_t v, int k) { return (v >> k) | (v << (32 - k)); }

static void sha256_block(uint32_t h[8], const uint8_t blk[64]) {
	uint32_t w[64];
	for(int i = 0; i < 16; i++) w[i] = ((uint32_t) blk[4 * i] << 24) | ((uint32_t) blk[4 * i + 1] << 16) | ((uint32_t) blk[4 * i + 2] << 8) | blk[4 * i + 3];
	for(int i = 16; i < 64; i++) {
		const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
		const uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
		w[i] = w[i - 16] + s0 + w[i - 7] + s1;
	}
	uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
	for(int i = 0; i < 64; i++) {
		const uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + SHA_K[i] + w[i];
		const uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
		hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
	}
	h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

void p1host_sha256(const uint8_t *msg, size_t len, uint8_t out[32]) {
	uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
	uint8_t blk[64];
	size_t off = 0;
	for(; off + 64 <= len; off += 64) sha256_block(h, msg + off);
	size_t rem = len - off;
	memset(blk, 0, 64);
	memcpy(blk, msg + off, rem);
	blk[rem] = 0x80;
	if(rem >= 56) {
		sha256_block(h, blk);
		memset(blk, 0, 64);
	}
	const uint64_t bits = (uint64_t) len * 8;
	for(int i = 0; i < 8; i++) blk[63 - i] = (uint8_t) (bits >> (8 * i));
	sha256_block(h, blk);
	for(int i = 0; i < 8; i++) {
		out[4 * i] = (uint8_t) (h[i] >> 24);
		out[4 * i + 1] = (uint8_t) (h[i] >> 16);
		out[4 * i + 2] = (uint8_t) (h[i] >> 8);
		out[4 * i + 3] = (uint8_t) h[i];
	}
}

/* s * G as affine x || y (tests: compared with an independent big-integer implementation) */
void p1host_ed25519_scalarmult_base(const uint8_t s[32], uint8_t out[64]) {
	ge G, R;
	ge_base(&G);
	ge_scalarmult(&R, &G, s);
	ge_to_bytes(out, &R);
}

/* ------------------------------------------------------------------ the protocol */

static int random_scalar(uint8_t s[32]) {
	if(getrandom(s, 32, 0) != 32) return 1;
	s[0] &= 248;            /* multiple of the cofactor */
	s[31] &= 63;            /* below 2^254 < 8q: no reduction needed for the ladder */
	s[31] |= 32;
	return 0;
}

/* the key of OT i: bound to the session (tag, parties, i) and to the transcript (A and B_i as they travelled), so
 * that a key is tied to the messages that produced it and not to the shared point alone */
static void seed_of(uint8_t seed[16], int sender_party, int receiver_party, int i, const uint8_t A_wire[64],
	const uint8_t B_wire[64], const ge *p) {
	uint8_t msg[16 + 12 + 3 * 64], dig[32];
	memcpy(msg, "p1gpu-base-ot-v2", 16);
	const uint32_t v[3] = {(uint32_t) sender_party, (uint32_t) receiver_party, (uint32_t) i};
	for(int k = 0; k < 3; k++) for(int b = 0; b < 4; b++) msg[16 + 4 * k + b] = (uint8_t) (v[k] >> (8 * b));
	memcpy(msg + 28, A_wire, 64);
	memcpy(msg + 92, B_wire, 64);
	ge_to_bytes(msg + 156, p);
	p1host_sha256(msg, sizeof(msg), dig);
	memcpy(seed, dig, 16);
	explicit_bzero(msg, sizeof(msg));
	explicit_bzero(dig, sizeof(dig));
}

/* 1 if 8 P is the neutral element (0, 1): P lies in the small-order subgroup, and a scalar that is a multiple of
 * the cofactor would send it to the neutral element - every key derived from it would be public */
static int ge_small_order(const ge *p) {
	static const uint8_t eight[32] = {8};
	uint8_t enc[64];
	ge r;
	ge_scalarmult(&r, p, eight);
	ge_to_bytes(enc, &r);
	int nz = enc[32] ^ 1;
	for(int k = 0; k < 64; k++) if(k != 32) nz |= enc[k];
	return nz == 0;
}

/* role, seeds0, seeds1, delta as for p1host_base_ot (phase1_gpu.c): role 1 = extension sender */
int p1host_base_ot_co15(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]) {
	ge G, A;
	uint8_t wire[64], all[128 * 64];
	ge_base(&G);
	if(role == 0) {
		/* extension receiver = base-OT sender: learns both keys of every OT */
		uint8_t a[32];
		ge aA, P0, P1, B;
		if(random_scalar(a)) { fprintf(stderr, "p1host_base_ot: getrandom failed\n"); return 1; }
		ge_scalarmult(&A, &G, a);
		ge_to_bytes(wire, &A);
		if(osend(pd, 0, wire, 64) < 0) return 1;
		orecv(pd, 0, NULL, 0);
		if(orecv(pd, 0, all, sizeof(all)) != (int) sizeof(all)) return 1;
		ge_scalarmult(&aA, &A, a);
		ge_neg(&aA);
		for(int i = 0; i < 128; i++) {
			if(ge_from_bytes(&B, all + 64 * i)) { fprintf(stderr, "p1host_base_ot: peer sent a point that is not on the curve\n"); return 1; }
			ge_scalarmult(&P0, &B, a);
			P1 = P0;
			ge_add(&P1, &aA);                     /* a B - a A */
			seed_of(seeds0[i], sender_party, receiver_party, i, wire, all + 64 * i, &P0);
			seed_of(seeds1[i], sender_party, receiver_party, i, wire, all + 64 * i, &P1);
		}
		memset(delta, 0, 16);
		explicit_bzero(a, sizeof(a));
		explicit_bzero(&aA, sizeof(aA));
		explicit_bzero(&P0, sizeof(P0));
		explicit_bzero(&P1, sizeof(P1));
	} else {
		/* extension sender = base-OT receiver with the choice string Delta */
		uint8_t b[32];
		ge B, K;
		if(getrandom(delta, 16, 0) != 16) { fprintf(stderr, "p1host_base_ot: getrandom failed\n"); return 1; }
		if(orecv(pd, 0, wire, 64) != 64) return 1;
		if(ge_from_bytes(&A, wire)) { fprintf(stderr, "p1host_base_ot: peer sent a point that is not on the curve\n"); return 1; }
		if(ge_small_order(&A)) { fprintf(stderr, "p1host_base_ot: peer sent a point of small order\n"); return 1; }
		for(int i = 0; i < 128; i++) {
			const int bit = (delta[i / 8] >> (i % 8)) & 1;
			if(random_scalar(b)) { fprintf(stderr, "p1host_base_ot: getrandom failed\n"); return 1; }
			ge_scalarmult(&B, &G, b);
			ge sum = B;
			ge_add(&sum, &A);
			/* B_i = b_i G + Delta_i A, selected without a secret-dependent branch */
			ge_cswap(&B, &sum, bit);
			ge_to_bytes(all + 64 * i, &B);
			ge_scalarmult(&K, &A, b);
			seed_of(seeds0[i], sender_party, receiver_party, i, wire, all + 64 * i, &K);
			explicit_bzero(&sum, sizeof(sum));
		}
		if(osend(pd, 0, all, sizeof(all)) < 0) return 1;
		orecv(pd, 0, NULL, 0);
		memset(seeds1, 0, 128 * 16);
		explicit_bzero(b, sizeof(b));
		explicit_bzero(&K, sizeof(K));
		explicit_bzero(&B, sizeof(B));
	}
	return 0;
}

/* the default provider behind the hook of phase1_gpu.c (a strong definition elsewhere wins) */
__attribute__((weak)) int p1host_base_ot(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]) {
	return p1host_base_ot_co15(pd, sender_party, receiver_party, role, seeds0, seeds1, delta);
}
