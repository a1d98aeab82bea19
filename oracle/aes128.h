/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * AES-128 in CTR mode, written from FIPS-197 / SP 800-38A.  It stands in for the
 * PRG that the reference borrows from Obliv-C (`bcrandom.h`: newBCipherRandomGen /
 * randomizeBuffer, used at /root/reference/src/phase1.c:243,271-273), which is an
 * AES-128-CTR generator over libgcrypt and is NOT vendored under /root/reference
 * (Dockerfile:20-23 clones samee/obliv-c HEAD, unpinned).
 *
 * Stream definition used by every component of this repo (oracle, _ref shims, GPU):
 *   S(key)[16*b .. 16*b+15] = AES128_key( BE128(b) ),  b = 0,1,2,...
 * i.e. CTR with an all-zero IV and a 128-bit big-endian block counter.  A generator
 * is a byte cursor into S: every randomizeBuffer(len) call returns the next `len`
 * bytes (left-over bytes of a partially used block are carried to the next call,
 * as libgcrypt's CTR mode and OpenSSL's EVP CTR both do).
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
	uint8_t rk[11][16]; /* expanded round keys, byte order as in FIPS-197 */
	int use_aesni;
} aes128_key;

void aes128_set_key(aes128_key *k, const uint8_t key[16]);
void aes128_encrypt_block(const aes128_key *k, const uint8_t in[16], uint8_t out[16]);

/* writes S(key)[byte_offset .. byte_offset+len) into out */
void aes128_ctr_keystream(const aes128_key *k, uint64_t byte_offset, uint8_t *out, size_t len);

/* sequential generator (byte cursor) */
typedef struct {
	aes128_key key;
	uint64_t pos; /* next unread byte of S */
} aes128_ctr_gen;

void aes128_ctr_gen_init(aes128_ctr_gen *g, const uint8_t key[16]);
void aes128_ctr_gen_read(aes128_ctr_gen *g, void *out, size_t len);

#ifdef __cplusplus
}
#endif
