// AES-128 on the GPU: T-table rounds out of shared memory.
// The table T0[x] = (2S, S, S, 3S) (little-endian bytes, FIPS-197 MixColumns of the
// S-box output) is replicated REPL times so that with REPL = 32 lane l only ever
// touches bank l: every lookup is conflict-free regardless of the data.  T1..T3 are
// byte rotations of T0, either recomputed with PRMT (NTAB = 1) or stored (NTAB = 4).
#pragma once
#include <stdint.h>

struct AesKey {
	uint32_t rk[44]; // round keys as little-endian words, rk[4*r + c]
};

// host: expand a 16-byte key; build T0
void aes_host_expand(const uint8_t key[16], AesKey *out);
const uint32_t *aes_host_T0(void);
struct p1_ctx;
// device copy of T0 (1 KB in global memory, uploaded once per device)
int aes_device_T0(p1_ctx *ctx, const uint32_t **dT0);

#ifdef __CUDACC__

template <int REPL, int NTAB>
__device__ __forceinline__ void aes_load_tables(uint32_t *sT, const uint32_t *__restrict__ gT0) {
	for(int i = threadIdx.x; i < 256 * REPL; i += blockDim.x) {
		uint32_t v = gT0[i / REPL];
		sT[i] = v;
		if(NTAB == 4) {
			sT[256 * REPL + i] = __byte_perm(v, v, 0x2103);
			sT[2 * 256 * REPL + i] = __byte_perm(v, v, 0x1032);
			sT[3 * 256 * REPL + i] = __byte_perm(v, v, 0x0321);
		}
	}
}

template <int REPL, int NTAB, int T>
__device__ __forceinline__ uint32_t aes_T(const uint32_t *sT, uint32_t x, uint32_t lane_off) {
	if(NTAB == 4) return sT[T * 256 * REPL + x * REPL + lane_off];
	uint32_t v = sT[x * REPL + lane_off];
	if(T == 0) return v;
	if(T == 1) return __byte_perm(v, v, 0x2103);
	if(T == 2) return __byte_perm(v, v, 0x1032);
	return __byte_perm(v, v, 0x0321);
}

// key accessors: round-key word i = 4*round + column
struct AesConstKey {           // one key for the whole grid, in the kernel parameter bank
	const AesKey &K;
	__device__ __forceinline__ AesConstKey(const AesKey &k) : K(k) {}
	__device__ __forceinline__ uint32_t operator()(int i) const { return K.rk[i]; }
};
struct AesSmemKey {            // one key per thread: sK[i * stride + t]
	const uint32_t *p;
	int stride;
	__device__ __forceinline__ AesSmemKey(const uint32_t *base, int stride_, int t) : p(base + t), stride(stride_) {}
	__device__ __forceinline__ uint32_t operator()(int i) const { return p[i * stride]; }
};

// AES-128 encryption of the block (i0,i1,i2,i3) (little-endian words)
template <int REPL, int NTAB, typename KA>
__device__ __forceinline__ void aes_encrypt(const KA &rk, const uint32_t *sT, uint32_t lane_off,
	uint32_t i0, uint32_t i1, uint32_t i2, uint32_t i3, uint32_t s[4]) {
	uint32_t s0 = i0 ^ rk(0), s1 = i1 ^ rk(1), s2 = i2 ^ rk(2), s3 = i3 ^ rk(3);
#pragma unroll
	for(int r = 1; r < 10; r++) {
		uint32_t t0 = aes_T<REPL, NTAB, 0>(sT, s0 & 0xff, lane_off) ^ aes_T<REPL, NTAB, 1>(sT, (s1 >> 8) & 0xff, lane_off) ^
		              aes_T<REPL, NTAB, 2>(sT, (s2 >> 16) & 0xff, lane_off) ^ aes_T<REPL, NTAB, 3>(sT, s3 >> 24, lane_off) ^ rk(4 * r);
		uint32_t t1 = aes_T<REPL, NTAB, 0>(sT, s1 & 0xff, lane_off) ^ aes_T<REPL, NTAB, 1>(sT, (s2 >> 8) & 0xff, lane_off) ^
		              aes_T<REPL, NTAB, 2>(sT, (s3 >> 16) & 0xff, lane_off) ^ aes_T<REPL, NTAB, 3>(sT, s0 >> 24, lane_off) ^ rk(4 * r + 1);
		uint32_t t2 = aes_T<REPL, NTAB, 0>(sT, s2 & 0xff, lane_off) ^ aes_T<REPL, NTAB, 1>(sT, (s3 >> 8) & 0xff, lane_off) ^
		              aes_T<REPL, NTAB, 2>(sT, (s0 >> 16) & 0xff, lane_off) ^ aes_T<REPL, NTAB, 3>(sT, s1 >> 24, lane_off) ^ rk(4 * r + 2);
		uint32_t t3 = aes_T<REPL, NTAB, 0>(sT, s3 & 0xff, lane_off) ^ aes_T<REPL, NTAB, 1>(sT, (s0 >> 8) & 0xff, lane_off) ^
		              aes_T<REPL, NTAB, 2>(sT, (s1 >> 16) & 0xff, lane_off) ^ aes_T<REPL, NTAB, 3>(sT, s2 >> 24, lane_off) ^ rk(4 * r + 3);
		s0 = t0; s1 = t1; s2 = t2; s3 = t3;
	}
	// last round: SubBytes + ShiftRows + AddRoundKey; S[x] is byte 1 of T0[x]
#define P1_SB(a, b, c, d) __byte_perm(__byte_perm(sT[((a) & 0xff) * REPL + lane_off], sT[(((b) >> 8) & 0xff) * REPL + lane_off], 0x0051), \
	__byte_perm(sT[(((c) >> 16) & 0xff) * REPL + lane_off], sT[((d) >> 24) * REPL + lane_off], 0x0051), 0x5410)
	s[0] = P1_SB(s0, s1, s2, s3) ^ rk(40);
	s[1] = P1_SB(s1, s2, s3, s0) ^ rk(41);
	s[2] = P1_SB(s2, s3, s0, s1) ^ rk(42);
	s[3] = P1_SB(s3, s0, s1, s2) ^ rk(43);
#undef P1_SB
}

// keystream block: encrypts the counter block 0^64 || BE64(block); the 16 keystream
// bytes come back as four little-endian words s[0..3]
template <int REPL, int NTAB, typename KA>
__device__ __forceinline__ void aes_ctr_block_k(const KA &rk, const uint32_t *sT, uint32_t lane_off, uint64_t block, uint32_t s[4]) {
	aes_encrypt<REPL, NTAB>(rk, sT, lane_off, 0u, 0u, __byte_perm((uint32_t) (block >> 32), 0, 0x0123),
		__byte_perm((uint32_t) block, 0, 0x0123), s);
}
template <int REPL, int NTAB>
__device__ __forceinline__ void aes_ctr_block(const AesKey &K, const uint32_t *sT, uint32_t lane_off, uint64_t block, uint32_t s[4]) {
	aes_ctr_block_k<REPL, NTAB>(AesConstKey(K), sT, lane_off, block, s);
}

#endif
