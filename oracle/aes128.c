/* TEST INFRASTRUCTURE (oracle). See aes128.h for what this replaces and why. */
#include "aes128.h"
#include <string.h>

#if defined(__x86_64__)
#include <immintrin.h>
#define HAVE_X86 1
#else
#define HAVE_X86 0
#endif

/* ---- portable AES-128 (FIPS-197) ---- */

static uint8_t SBOX[256];
static int sbox_ready = 0;

static uint8_t gf_mul(uint8_t a, uint8_t b) {
	uint8_t p = 0;
	for(int i = 0; i < 8; i++) {
		if(b & 1) p ^= a;
		uint8_t hi = a & 0x80;
		a <<= 1;
		if(hi) a ^= 0x1b;
		b >>= 1;
	}
	return p;
}

static void sbox_init(void) {
	/* S(x) = affine(x^-1) computed from the field definition, no tables copied */
	uint8_t inv[256];
	inv[0] = 0;
	for(int x = 1; x < 256; x++) {
		/* x^254 = x^-1 in GF(2^8) */
		uint8_t r = 1, b = (uint8_t) x;
		int e = 254;
		while(e) {
			if(e & 1) r = gf_mul(r, b);
			b = gf_mul(b, b);
			e >>= 1;
		}
		inv[x] = r;
	}
	for(int x = 0; x < 256; x++) {
		uint8_t v = inv[x], s = v;
		for(int i = 1; i <= 4; i++) {
			s ^= (uint8_t) ((v << i) | (v >> (8 - i)));
		}
		SBOX[x] = s ^ 0x63;
	}
	sbox_ready = 1;
}

void aes128_set_key(aes128_key *k, const uint8_t key[16]) {
	if(!sbox_ready) sbox_init();
	memcpy(k->rk[0], key, 16);
	uint8_t rcon = 1;
	for(int r = 1; r <= 10; r++) {
		const uint8_t *p = k->rk[r - 1];
		uint8_t *q = k->rk[r];
		uint8_t t[4] = { SBOX[p[13]], SBOX[p[14]], SBOX[p[15]], SBOX[p[12]] };
		t[0] ^= rcon;
		rcon = gf_mul(rcon, 2);
		for(int i = 0; i < 4; i++) q[i] = p[i] ^ t[i];
		for(int i = 4; i < 16; i++) q[i] = p[i] ^ q[i - 4];
	}
#if HAVE_X86
	k->use_aesni = __builtin_cpu_supports("aes") ? 1 : 0;
#else
	k->use_aesni = 0;
#endif
}

static void encrypt_portable(const aes128_key *k, const uint8_t in[16], uint8_t out[16]) {
	uint8_t s[16], t[16];
	for(int i = 0; i < 16; i++) s[i] = in[i] ^ k->rk[0][i];
	for(int r = 1; r <= 10; r++) {
		/* SubBytes + ShiftRows: state is column-major, byte i = row i%4, col i/4 */
		for(int c = 0; c < 4; c++) {
			for(int row = 0; row < 4; row++) {
				t[4 * c + row] = SBOX[s[4 * ((c + row) & 3) + row]];
			}
		}
		if(r < 10) {
			for(int c = 0; c < 4; c++) {
				uint8_t a0 = t[4*c], a1 = t[4*c+1], a2 = t[4*c+2], a3 = t[4*c+3];
				uint8_t x = a0 ^ a1 ^ a2 ^ a3;
				s[4*c]   = a0 ^ x ^ gf_mul(a0 ^ a1, 2);
				s[4*c+1] = a1 ^ x ^ gf_mul(a1 ^ a2, 2);
				s[4*c+2] = a2 ^ x ^ gf_mul(a2 ^ a3, 2);
				s[4*c+3] = a3 ^ x ^ gf_mul(a3 ^ a0, 2);
			}
		} else {
			memcpy(s, t, 16);
		}
		for(int i = 0; i < 16; i++) s[i] ^= k->rk[r][i];
	}
	memcpy(out, s, 16);
}

/* ---- AES-NI path (same function, hardware rounds) ---- */
#if HAVE_X86
__attribute__((target("aes,sse4.1")))
static void encrypt_aesni(const aes128_key *k, const uint8_t in[16], uint8_t out[16]) {
	__m128i b = _mm_loadu_si128((const __m128i *) in);
	b = _mm_xor_si128(b, _mm_loadu_si128((const __m128i *) k->rk[0]));
	for(int r = 1; r < 10; r++) {
		b = _mm_aesenc_si128(b, _mm_loadu_si128((const __m128i *) k->rk[r]));
	}
	b = _mm_aesenclast_si128(b, _mm_loadu_si128((const __m128i *) k->rk[10]));
	_mm_storeu_si128((__m128i *) out, b);
}

/* whole blocks [b0, b0+nblk) of the keystream, 8 blocks in flight */
__attribute__((target("aes,sse4.1,ssse3")))
static void ctr_blocks_aesni(const aes128_key *k, uint64_t b0, uint8_t *out, size_t nblk) {
	__m128i rk[11];
	for(int r = 0; r < 11; r++) rk[r] = _mm_loadu_si128((const __m128i *) k->rk[r]);
	/* counter block = 64 zero bits || BE64(b): high half never carries for b < 2^64 */
	const __m128i bswap = _mm_set_epi8(8, 9, 10, 11, 12, 13, 14, 15, 0, 1, 2, 3, 4, 5, 6, 7);
	size_t i = 0;
	for(; i + 8 <= nblk; i += 8) {
		__m128i s[8];
		for(int j = 0; j < 8; j++) {
			__m128i c = _mm_set_epi64x((long long) (b0 + i + j), 0);
			s[j] = _mm_xor_si128(_mm_shuffle_epi8(c, bswap), rk[0]);
		}
		for(int r = 1; r < 10; r++) {
			for(int j = 0; j < 8; j++) s[j] = _mm_aesenc_si128(s[j], rk[r]);
		}
		for(int j = 0; j < 8; j++) {
			s[j] = _mm_aesenclast_si128(s[j], rk[10]);
			_mm_storeu_si128((__m128i *) (out + 16 * (i + j)), s[j]);
		}
	}
	for(; i < nblk; i++) {
		__m128i c = _mm_set_epi64x((long long) (b0 + i), 0);
		__m128i s = _mm_xor_si128(_mm_shuffle_epi8(c, bswap), rk[0]);
		for(int r = 1; r < 10; r++) s = _mm_aesenc_si128(s, rk[r]);
		s = _mm_aesenclast_si128(s, rk[10]);
		_mm_storeu_si128((__m128i *) (out + 16 * i), s);
	}
}
#endif

void aes128_encrypt_block(const aes128_key *k, const uint8_t in[16], uint8_t out[16]) {
#if HAVE_X86
	if(k->use_aesni) {
		encrypt_aesni(k, in, out);
		return;
	}
#endif
	encrypt_portable(k, in, out);
}

static void counter_block(uint64_t b, uint8_t ctr[16]) {
	memset(ctr, 0, 8);
	for(int i = 0; i < 8; i++) ctr[15 - i] = (uint8_t) (b >> (8 * i));
}

void aes128_ctr_keystream(const aes128_key *k, uint64_t byte_offset, uint8_t *out, size_t len) {
	uint8_t ctr[16], ks[16];
	uint64_t b = byte_offset / 16;
	size_t skip = (size_t) (byte_offset % 16);
	if(skip && len) {
		counter_block(b, ctr);
		aes128_encrypt_block(k, ctr, ks);
		size_t take = 16 - skip < len ? 16 - skip : len;
		memcpy(out, ks + skip, take);
		out += take;
		len -= take;
		b++;
	}
	size_t nblk = len / 16;
#if HAVE_X86
	if(k->use_aesni && nblk) {
		ctr_blocks_aesni(k, b, out, nblk);
		out += 16 * nblk;
		len -= 16 * nblk;
		b += nblk;
		nblk = 0;
	}
#endif
	for(size_t i = 0; i < nblk; i++, b++) {
		counter_block(b, ctr);
		aes128_encrypt_block(k, ctr, out);
		out += 16;
		len -= 16;
	}
	if(len) {
		counter_block(b, ctr);
		aes128_encrypt_block(k, ctr, ks);
		memcpy(out, ks, len);
	}
}

void aes128_ctr_gen_init(aes128_ctr_gen *g, const uint8_t key[16]) {
	aes128_set_key(&g->key, key);
	g->pos = 0;
}

void aes128_ctr_gen_read(aes128_ctr_gen *g, void *out, size_t len) {
	aes128_ctr_keystream(&g->key, g->pos, (uint8_t *) out, len);
	g->pos += len;
}

/* exported for tests: force the portable path regardless of CPU features */
void aes128_force_portable(aes128_key *k) {
	k->use_aesni = 0;
}
