/* TEST INFRASTRUCTURE (oracle). See cot_ext.h for the protocol definition and for
 * which reference call sites (/root/reference/src/phase1.c:58-65,84-89,391-397) it backs. */
#include "cot_ext.h"
#include <string.h>

const uint8_t COT_HASH_KEY[16] = {
	0x6c, 0x69, 0x6e, 0x72, 0x65, 0x67, 0x2d, 0x6d, 0x70, 0x63, 0x2d, 0x63, 0x6f, 0x74, 0x2d, 0x48
}; /* "linreg-mpc-cot-H" */

static void be128(uint64_t v, uint8_t out[16]) {
	memset(out, 0, 8);
	for(int i = 0; i < 8; i++) out[15 - i] = (uint8_t) (v >> (8 * i));
}

void cot_session_derive(const uint8_t ot_key[16], int sender_party, int receiver_party,
	uint8_t seed0[COT_KAPPA][16], uint8_t seed1[COT_KAPPA][16], uint8_t delta[16]) {
	aes128_key master, sess;
	uint8_t blk[16], ksess[16];
	aes128_set_key(&master, ot_key);
	be128((uint64_t) sender_party * 65536u + (uint64_t) receiver_party, blk);
	aes128_encrypt_block(&master, blk, ksess);
	aes128_set_key(&sess, ksess);
	for(int i = 0; i < COT_KAPPA; i++) {
		be128(2 * (uint64_t) i, blk);
		aes128_encrypt_block(&sess, blk, seed0[i]);
		be128(2 * (uint64_t) i + 1, blk);
		aes128_encrypt_block(&sess, blk, seed1[i]);
	}
	be128(256, blk);
	aes128_encrypt_block(&sess, blk, delta);
}

void cot_sender_init(cot_sender *s, const uint8_t ot_key[16], int sender_party, int receiver_party) {
	uint8_t s0[COT_KAPPA][16], s1[COT_KAPPA][16];
	cot_session_derive(ot_key, sender_party, receiver_party, s0, s1, s->delta);
	for(int i = 0; i < COT_KAPPA; i++) {
		int bit = (s->delta[i / 8] >> (i % 8)) & 1;
		memcpy(s->seed[i], bit ? s1[i] : s0[i], 16);
		aes128_set_key(&s->col[i], s->seed[i]);
	}
	aes128_set_key(&s->hash, COT_HASH_KEY);
	s->pos = 0;
	s->ot_base = 0;
}

void cot_receiver_init(cot_receiver *r, const uint8_t ot_key[16], int sender_party, int receiver_party) {
	uint8_t delta_unused[16];
	cot_session_derive(ot_key, sender_party, receiver_party, r->seed0, r->seed1, delta_unused);
	for(int i = 0; i < COT_KAPPA; i++) {
		aes128_set_key(&r->col0[i], r->seed0[i]);
		aes128_set_key(&r->col1[i], r->seed1[i]);
	}
	aes128_set_key(&r->hash, COT_HASH_KEY);
	r->pos = 0;
	r->ot_base = 0;
}

/* 64x64 bit transpose, LSB-first: element (r,c) = bit c of a[r] */
static void transpose64(uint64_t a[64]) {
	uint64_t m = 0x00000000FFFFFFFFull;
	for(int j = 32; j; j >>= 1, m ^= m << j) {
		for(int k = 0; k < 64; k = ((k | j) + 1) & ~j) {
			uint64_t t = ((a[k] >> j) ^ a[k | j]) & m;
			a[k] ^= t << j;
			a[k | j] ^= t;
		}
	}
}

void cot_transpose128(const uint8_t in[2048], uint8_t out[2048]) {
	uint64_t b00[64], b01[64], b10[64], b11[64];
	for(int i = 0; i < 64; i++) {
		memcpy(&b00[i], in + 16 * i, 8);
		memcpy(&b01[i], in + 16 * i + 8, 8);
		memcpy(&b10[i], in + 16 * (i + 64), 8);
		memcpy(&b11[i], in + 16 * (i + 64) + 8, 8);
	}
	transpose64(b00);
	transpose64(b01);
	transpose64(b10);
	transpose64(b11);
	for(int j = 0; j < 64; j++) {
		memcpy(out + 16 * j, &b00[j], 8);
		memcpy(out + 16 * j + 8, &b10[j], 8);
		memcpy(out + 16 * (j + 64), &b01[j], 8);
		memcpy(out + 16 * (j + 64) + 8, &b11[j], 8);
	}
}

void cot_receiver_tile(const cot_receiver *r, uint64_t block, const uint8_t choice[16],
	uint8_t u[2048], uint8_t trows[2048]) {
	uint8_t tcols[2048];
	for(int i = 0; i < COT_KAPPA; i++) {
		uint8_t g1[16];
		aes128_ctr_keystream(&r->col0[i], 16 * block, tcols + 16 * i, 16);
		aes128_ctr_keystream(&r->col1[i], 16 * block, g1, 16);
		for(int b = 0; b < 16; b++) u[16 * i + b] = tcols[16 * i + b] ^ g1[b] ^ choice[b];
	}
	cot_transpose128(tcols, trows);
}

void cot_sender_tile(const cot_sender *s, uint64_t block, const uint8_t u[2048], uint8_t qrows[2048]) {
	uint8_t qcols[2048];
	for(int i = 0; i < COT_KAPPA; i++) {
		aes128_ctr_keystream(&s->col[i], 16 * block, qcols + 16 * i, 16);
		if((s->delta[i / 8] >> (i % 8)) & 1) {
			for(int b = 0; b < 16; b++) qcols[16 * i + b] ^= u[16 * i + b];
		}
	}
	cot_transpose128(qcols, qrows);
}

void cot_hash(const aes128_key *hk, uint64_t J, const uint8_t v[16], uint8_t out[16]) {
	uint8_t x[16], e[16];
	memcpy(x, v, 16);
	for(int i = 0; i < 8; i++) x[i] ^= (uint8_t) (J >> (8 * i));
	aes128_encrypt_block(hk, x, e);
	for(int i = 0; i < 16; i++) out[i] = e[i] ^ x[i];
}
