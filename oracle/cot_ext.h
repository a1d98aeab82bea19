/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Correlated-OT extension used by the `--use_ot` aggregation, CPU restatement.
 *
 * The reference calls Obliv-C's honestOTExt{Sender,Recver}New and
 * honestCorrelatedOTExt{Send,Recv}1Of2 (/root/reference/src/phase1.c:58-65,84-89,
 * 391-397).  Obliv-C is not vendored and is unpinned (Dockerfile:20-23), so the
 * construction below is this repo's DEFINITION of that box (semi-honest IKNP'03
 * extension in its correlated form, the family Obliv-C's "honest" OT belongs to).
 * What the reference itself pins — the correlator (phase1.c:43-51), the choice bits
 * (:78-83), share = -sum(opt0) (:66-68) and share = +sum(dest) (:90-92) — is kept
 * verbatim.  PARITY UNPINNED against real Obliv-C output; pinned against this file.
 *
 * Protocol, kappa = 128, for one call of m OTs (m padded up to M = 128*ceil(m/128)):
 *   session keys (stand-in for the kappa base OTs, see cot_session_derive):
 *       receiver R holds seeds k_i^0, k_i^1 (i < 128); sender S holds Delta in
 *       {0,1}^128 and k_i^{Delta_i}.
 *   G(k)   = AES-128-CTR keystream of oracle/aes128.h under key k; a session keeps a
 *            block cursor `pos`, each call consumes blocks [pos, pos + M/128).
 *   R: t^i = G(k_i^0), u^i = t^i ^ G(k_i^1) ^ r      (r = choice bits, bit j of the
 *            call = bit (j%8) of byte j/8, zero padded); sends U.
 *   S: q^i = G(k_i^{Delta_i}) ^ (Delta_i ? u^i : 0).
 *   rows : t_j, q_j in {0,1}^128 with bit i = t^i[j], q^i[j]  (q_j = t_j ^ r_j*Delta).
 *   H(J,v) = AES128_{COT_HASH_KEY}(v ^ tw(J)) ^ v ^ tw(J), tw(J) = LE64(J) || 0^64,
 *            J = session OT counter (ot_base + j); output truncated to the first
 *            `len` bytes.
 *   S: opt0_j = H(J, q_j); opt1_j = corr(opt0_j, j); sends c_j = opt1_j ^ H(J, q_j ^ Delta).
 *   R: dest_j = H(J, t_j) ^ (r_j ? c_j : 0).
 * Wire layout of U: tiles of 128 OTs; tile b holds u^i[128b .. 128b+127] as 16 bytes,
 * for i = 0..127 consecutively (2048 bytes per tile).  c is `len` bytes per OT.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>
#include "aes128.h"

#ifdef __cplusplus
extern "C" {
#endif

#define COT_KAPPA 128

extern const uint8_t COT_HASH_KEY[16];

typedef struct {
	uint8_t delta[16];
	uint8_t seed[COT_KAPPA][16];  /* k_i^{Delta_i} */
	aes128_key col[COT_KAPPA];
	aes128_key hash;
	uint64_t pos;                 /* PRG block cursor */
	uint64_t ot_base;             /* OT counter */
} cot_sender;

typedef struct {
	uint8_t seed0[COT_KAPPA][16], seed1[COT_KAPPA][16];
	aes128_key col0[COT_KAPPA], col1[COT_KAPPA];
	aes128_key hash;
	uint64_t pos;
	uint64_t ot_base;
} cot_receiver;

/* Deterministic stand-in for the base-OT phase: both roles' outputs for the session
 * (sender party index, receiver party index) derived from a 16-byte master key:
 *   ksess   = AES_{ot_key}(BE128(sender*65536 + receiver))
 *   k_i^b   = AES_{ksess}(BE128(2i+b)),  Delta = AES_{ksess}(BE128(256)). */
void cot_session_derive(const uint8_t ot_key[16], int sender_party, int receiver_party,
	uint8_t seed0[COT_KAPPA][16], uint8_t seed1[COT_KAPPA][16], uint8_t delta[16]);

void cot_sender_init(cot_sender *s, const uint8_t ot_key[16], int sender_party, int receiver_party);
void cot_receiver_init(cot_receiver *r, const uint8_t ot_key[16], int sender_party, int receiver_party);

/* One tile = 128 consecutive OTs of the current call.
 * receiver: choice = 16 bytes of choice bits; writes u (2048 B) and the 128 rows t_j (2048 B). */
void cot_receiver_tile(const cot_receiver *r, uint64_t block, const uint8_t choice[16],
	uint8_t u[2048], uint8_t trows[2048]);
/* sender: reads u, writes the 128 rows q_j. */
void cot_sender_tile(const cot_sender *s, uint64_t block, const uint8_t u[2048], uint8_t qrows[2048]);

void cot_hash(const aes128_key *hk, uint64_t J, const uint8_t v[16], uint8_t out[16]);

/* 128x128 bit-matrix transpose: in[i] = 16-byte row i; out[j] bit i = in[i] bit j */
void cot_transpose128(const uint8_t in[2048], uint8_t out[2048]);

#ifdef __cplusplus
}
#endif
