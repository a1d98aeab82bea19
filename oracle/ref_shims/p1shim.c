/* TEST INFRASTRUCTURE (oracle/_ref build only). Not part of the product path.
 * Implementations behind p1shim.h / protobuf/secure_multiplication.pb-c.h: the
 * externals of /root/reference/src/phase1.c that are not vendored (Obliv-C, protobuf-c).
 */
#define _POSIX_C_SOURCE 201605L
#include "p1shim.h"
#include "p1shim_internal.h"
#include "protobuf/secure_multiplication.pb-c.h"
#include "../aes128.h"
#include "../cot_ext.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ */
/* transport: blocking bounded byte stream per direction (TCP-like)    */
/* ------------------------------------------------------------------ */

struct p1shim_pipe {
	pthread_mutex_t mu;
	pthread_cond_t cv;
	uint8_t *buf;
	size_t cap, head, len; /* ring buffer */
	uint64_t total;        /* bytes ever written (network volume accounting) */
};

p1shim_pipe *p1shim_pipe_new(size_t cap) {
	p1shim_pipe *p = calloc(1, sizeof(*p));
	pthread_mutex_init(&p->mu, NULL);
	pthread_cond_init(&p->cv, NULL);
	p->buf = malloc(cap);
	p->cap = cap;
	return p;
}

void p1shim_pipe_free(p1shim_pipe *p) {
	if(!p) return;
	pthread_mutex_destroy(&p->mu);
	pthread_cond_destroy(&p->cv);
	free(p->buf);
	free(p);
}

uint64_t p1shim_pipe_total(p1shim_pipe *p) { return p->total; }

static void pipe_write(p1shim_pipe *p, const uint8_t *data, size_t len) {
	pthread_mutex_lock(&p->mu);
	while(len) {
		while(p->len == p->cap) pthread_cond_wait(&p->cv, &p->mu);
		size_t tail = (p->head + p->len) % p->cap;
		size_t room = p->cap - p->len;
		size_t run = p->cap - tail < room ? p->cap - tail : room;
		if(run > len) run = len;
		memcpy(p->buf + tail, data, run);
		p->len += run;
		p->total += run;
		data += run;
		len -= run;
		pthread_cond_broadcast(&p->cv);
	}
	pthread_mutex_unlock(&p->mu);
}

static void pipe_read(p1shim_pipe *p, uint8_t *data, size_t len) {
	pthread_mutex_lock(&p->mu);
	while(len) {
		while(p->len == 0) pthread_cond_wait(&p->cv, &p->mu);
		size_t run = p->cap - p->head < p->len ? p->cap - p->head : p->len;
		if(run > len) run = len;
		memcpy(data, p->buf + p->head, run);
		p->head = (p->head + run) % p->cap;
		p->len -= run;
		data += run;
		len -= run;
		pthread_cond_broadcast(&p->cv);
	}
	pthread_mutex_unlock(&p->mu);
}

ProtocolDesc *p1shim_pd_new(p1shim_pipe *tx, p1shim_pipe *rx, int self, int peer) {
	ProtocolDesc *pd = calloc(1, sizeof(*pd));
	pd->tx = tx;
	pd->rx = rx;
	pd->self = self;
	pd->peer = peer;
	return pd;
}

int osend(ProtocolDesc *pd, int dest, const void *data, size_t len) {
	(void) dest;
	if(len) pipe_write(pd->tx, data, len);
	return (int) len;
}

int orecv(ProtocolDesc *pd, int src, void *data, size_t len) {
	(void) src;
	if(len) pipe_read(pd->rx, data, len); /* zero-length call is Obliv-C's flush idiom */
	return (int) len;
}

/* ------------------------------------------------------------------ */
/* PRG                                                                  */
/* ------------------------------------------------------------------ */

static uint8_t g_ti_key[16];
static uint8_t g_ot_key[16];

void p1shim_set_keys(const uint8_t ti_key[16], const uint8_t ot_key[16]) {
	memcpy(g_ti_key, ti_key, 16);
	memcpy(g_ot_key, ot_key, 16);
}

struct BCipherRandomGen {
	aes128_ctr_gen g;
};

BCipherRandomGen *newBCipherRandomGen(void) {
	BCipherRandomGen *gen = malloc(sizeof(*gen));
	aes128_ctr_gen_init(&gen->g, g_ti_key); /* reference: random key; here injected */
	return gen;
}

void randomizeBuffer(BCipherRandomGen *gen, char *dest, size_t len) {
	aes128_ctr_gen_read(&gen->g, dest, len);
}

void releaseBCipherRandomGen(BCipherRandomGen *gen) {
	free(gen);
}

/* ------------------------------------------------------------------ */
/* correlated OT extension (oracle/cot_ext.h)                           */
/* ------------------------------------------------------------------ */

struct HonestOTExtSender {
	ProtocolDesc *pd;
	cot_sender s;
};
struct HonestOTExtRecver {
	ProtocolDesc *pd;
	cot_receiver r;
};

void dhRandomInit(void) {}

struct HonestOTExtSender *honestOTExtSenderNew(ProtocolDesc *pd, int destparty) {
	(void) destparty;
	struct HonestOTExtSender *s = malloc(sizeof(*s));
	s->pd = pd;
	cot_sender_init(&s->s, g_ot_key, pd->self, pd->peer);
	return s;
}

struct HonestOTExtRecver *honestOTExtRecverNew(ProtocolDesc *pd, int srcparty) {
	(void) srcparty;
	struct HonestOTExtRecver *r = malloc(sizeof(*r));
	r->pd = pd;
	cot_receiver_init(&r->r, g_ot_key, pd->peer, pd->self);
	return r;
}

void honestOTExtSenderRelease(struct HonestOTExtSender *s) { free(s); }
void honestOTExtRecverRelease(struct HonestOTExtRecver *r) { free(r); }

#define TILES_PER_CHUNK 64

void honestCorrelatedOTExtSend1Of2(struct HonestOTExtSender *s, char *opt0, char *opt1,
	int n, int len, OcOtCorrelator corr, void *corrArg) {
	size_t m = (size_t) n, tiles = (m + 127) / 128;
	uint8_t *u = malloc(2048 * TILES_PER_CHUNK);
	uint8_t *c = malloc((size_t) len * 128 * TILES_PER_CHUNK);
	uint8_t q[2048], h[16], v[16];
	for(size_t t0 = 0; t0 < tiles; t0 += TILES_PER_CHUNK) {
		size_t nt = tiles - t0 < TILES_PER_CHUNK ? tiles - t0 : TILES_PER_CHUNK;
		orecv(s->pd, 0, u, 2048 * nt);
		size_t cbytes = 0;
		for(size_t t = 0; t < nt; t++) {
			cot_sender_tile(&s->s, s->s.pos + t0 + t, u + 2048 * t, q);
			for(int r = 0; r < 128; r++) {
				size_t j = 128 * (t0 + t) + r;
				if(j >= m) break;
				uint64_t J = s->s.ot_base + j;
				cot_hash(&s->s.hash, J, q + 16 * r, h);
				memcpy(opt0 + j * len, h, len);
				corr(opt1 + j * len, opt0 + j * len, (int) j, corrArg);
				for(int b = 0; b < 16; b++) v[b] = q[16 * r + b] ^ s->s.delta[b];
				cot_hash(&s->s.hash, J, v, h);
				for(int b = 0; b < len; b++) c[cbytes + b] = (uint8_t) opt1[j * len + b] ^ h[b];
				cbytes += len;
			}
		}
		osend(s->pd, 0, c, cbytes);
	}
	s->s.pos += tiles;
	s->s.ot_base += 128 * tiles;
	free(u);
	free(c);
}

void honestCorrelatedOTExtRecv1Of2(struct HonestOTExtRecver *r, char *dest, const bool *sel,
	int n, int len) {
	size_t m = (size_t) n, tiles = (m + 127) / 128;
	uint8_t *u = malloc(2048 * TILES_PER_CHUNK);
	uint8_t *trows = malloc(2048 * TILES_PER_CHUNK);
	uint8_t *c = malloc((size_t) len * 128 * TILES_PER_CHUNK);
	uint8_t h[16];
	for(size_t t0 = 0; t0 < tiles; t0 += TILES_PER_CHUNK) {
		size_t nt = tiles - t0 < TILES_PER_CHUNK ? tiles - t0 : TILES_PER_CHUNK;
		size_t cbytes = 0;
		for(size_t t = 0; t < nt; t++) {
			uint8_t choice[16] = {0};
			for(int b = 0; b < 128; b++) {
				size_t j = 128 * (t0 + t) + b;
				if(j < m && sel[j]) choice[b / 8] |= (uint8_t) (1u << (b % 8));
			}
			cot_receiver_tile(&r->r, r->r.pos + t0 + t, choice, u + 2048 * t, trows + 2048 * t);
			size_t lim = m - 128 * (t0 + t);
			cbytes += (size_t) len * (lim < 128 ? lim : 128);
		}
		osend(r->pd, 0, u, 2048 * nt);
		orecv(r->pd, 0, c, cbytes);
		size_t off = 0;
		for(size_t t = 0; t < nt; t++) {
			for(int b = 0; b < 128; b++) {
				size_t j = 128 * (t0 + t) + b;
				if(j >= m) break;
				cot_hash(&r->r.hash, r->r.ot_base + j, trows + 2048 * t + 16 * b, h);
				for(int k = 0; k < len; k++) {
					dest[j * len + k] = (char) (h[k] ^ (sel[j] ? c[off + k] : 0));
				}
				off += len;
			}
		}
	}
	r->r.pos += tiles;
	r->r.ot_base += 128 * tiles;
	free(u);
	free(trows);
	free(c);
}

/* ------------------------------------------------------------------ */
/* protobuf wire format for SecureMultiplication__Msg                   */
/* ------------------------------------------------------------------ */

static inline size_t varint_size(uint64_t v) {
	size_t s = 1;
	while(v >= 0x80) {
		v >>= 7;
		s++;
	}
	return s;
}

static inline size_t varint_put(uint8_t *out, uint64_t v) {
	size_t s = 0;
	while(v >= 0x80) {
		out[s++] = (uint8_t) (v | 0x80);
		v >>= 7;
	}
	out[s++] = (uint8_t) v;
	return s;
}

void secure_multiplication__msg__init(SecureMultiplication__Msg *m) {
	memset(m, 0, sizeof(*m));
}

static size_t vector_payload_size(const SecureMultiplication__Msg *m) {
	size_t s = 0;
	for(size_t i = 0; i < m->n_vector; i++) s += varint_size(m->vector[i]);
	return s;
}

size_t secure_multiplication__msg__get_packed_size(const SecureMultiplication__Msg *m) {
	size_t s = 0;
	if(m->n_vector) {
		size_t payload = vector_payload_size(m);
		s += 1 + varint_size(payload) + payload;
	}
	s += 1 + varint_size(m->value);
	return s;
}

size_t secure_multiplication__msg__pack_to_buffer(const SecureMultiplication__Msg *m, ProtobufCBuffer *buffer) {
	uint8_t chunk[8192 + 16];
	size_t fill = 0, total = 0;
	if(m->n_vector) {
		chunk[fill++] = (1 << 3) | 2;
		fill += varint_put(chunk + fill, vector_payload_size(m));
		for(size_t i = 0; i < m->n_vector; i++) {
			fill += varint_put(chunk + fill, m->vector[i]);
			if(fill >= 8192) {
				buffer->append(buffer, fill, chunk);
				total += fill;
				fill = 0;
			}
		}
	}
	chunk[fill++] = (2 << 3) | 0;
	fill += varint_put(chunk + fill, m->value);
	buffer->append(buffer, fill, chunk);
	return total + fill;
}

static int varint_get(const uint8_t **p, const uint8_t *end, uint64_t *v) {
	uint64_t r = 0;
	int shift = 0;
	while(*p < end) {
		uint8_t b = *(*p)++;
		r |= (uint64_t) (b & 0x7f) << shift;
		if(!(b & 0x80)) {
			*v = r;
			return 1;
		}
		shift += 7;
		if(shift > 63) return 0;
	}
	return 0;
}

SecureMultiplication__Msg *secure_multiplication__msg__unpack(ProtobufCAllocator *a, size_t len, const uint8_t *data) {
	(void) a;
	SecureMultiplication__Msg *m = calloc(1, sizeof(*m));
	const uint8_t *p = data, *end = data + len;
	int have_value = 0;
	while(p < end) {
		uint64_t tag, v;
		if(!varint_get(&p, end, &tag)) goto fail;
		if(tag == ((1 << 3) | 2)) {
			if(!varint_get(&p, end, &v) || v > (uint64_t) (end - p)) goto fail;
			const uint8_t *pe = p + v;
			size_t count = 0;
			for(const uint8_t *q = p; q < pe; q++) count += !(*q & 0x80);
			m->vector = realloc(m->vector, (m->n_vector + count) * sizeof(p1shim_word_t));
			while(p < pe) {
				if(!varint_get(&p, pe, &v)) goto fail;
				m->vector[m->n_vector++] = (p1shim_word_t) v;
			}
		} else if(tag == ((2 << 3) | 0)) {
			if(!varint_get(&p, end, &v)) goto fail;
			m->value = (p1shim_word_t) v;
			have_value = 1;
		} else {
			goto fail;
		}
	}
	if(!have_value) goto fail;
	return m;
fail:
	free(m->vector);
	free(m);
	return NULL;
}

void secure_multiplication__msg__free_unpacked(SecureMultiplication__Msg *m, ProtobufCAllocator *a) {
	(void) a;
	if(!m) return;
	free(m->vector);
	free(m);
}

/* ------------------------------------------------------------------ */
/* hooks of linreg-mpc_b200/host/phase1_gpu.c (strong definitions override its weak ones):  */
/* the same injected keys as the stand-ins above, so that the GPU host file and the           */
/* reference's phase1.c produce the same share words in the loop-back driver.                 */
/* ------------------------------------------------------------------ */

int p1host_ti_key(uint8_t key[16]) {
	memcpy(key, g_ti_key, 16);
	return 0;
}

int p1host_base_ot(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]) {
	(void) pd;
	uint8_t s0[COT_KAPPA][16], s1[COT_KAPPA][16], dl[16];
	cot_session_derive(g_ot_key, sender_party, receiver_party, s0, s1, dl);
	if(role) {   /* sender: Delta and k_i^{Delta_i} */
		memcpy(delta, dl, 16);
		for(int i = 0; i < COT_KAPPA; i++) memcpy(seeds0[i], ((dl[i / 8] >> (i % 8)) & 1) ? s1[i] : s0[i], 16);
		memset(seeds1, 0, COT_KAPPA * 16);
	} else {     /* receiver: both seeds of every base OT */
		memcpy(seeds0, s0, sizeof(s0));
		memcpy(seeds1, s1, sizeof(s1));
		memset(delta, 0, 16);
	}
	return 0;
}
