// AES-128 on the GPU: T-table rounds out of shared memory, laid out for one PRMT per
// lookup address and conflict-free LDS.
//
// Tables.  T0[x] = (2S, S, S, 3S) (little-endian bytes: MixColumns of the S-box
// output, FIPS-197 5.1), T1..T3 = T0 rotated left by 8, 16, 24 bits.  Shared memory
// holds two arrays of 256 rows x 256 bytes:
//     array 0, row x : bytes [0,128) = T0[x] replicated for lanes 0..31, [128,256) = T1[x]
//     array 1, row x : T2[x], T3[x]
// so lane l reads word l of a half row: every lookup hits bank l (no conflicts for any
// data).  The arrays start at a shared-memory address that is a multiple of 64 KB, so
// the address of a lookup is  base | (x << 8) | (4*l): ONE PRMT drops the state byte
// into byte 1 of the lane's base register; array/half select is the LDS immediate.
// Per round and column: 4 PRMT + 4 LDS + 2 LOP3 (5-input xor).  Budget per 32 blocks
// (one warp instruction each): ~236 ALU-pipe + ~150 LDS; the LSU (128 B/clk/SM = one
// 32-lane LDS per clock) is the floor at ~150 clk.  Dynamic shared memory starts 1 KB
// into the window on sm_100, so the kernels allocate AES_SMEM_BYTES and use the gap
// below the tables for their own data (aes_table_offset).
#pragma once
#include <stdint.h>

struct AesKey {
	uint32_t rk[44]; // round keys as little-endian words, rk[4*r + c]
};

// CTR key: round keys + round 1 of the three counter-block columns that do not depend
// on the low 32 bits of the block counter (valid while block >> 32 == hi)
struct AesCtrKey {
	AesKey k;
	uint32_t hi;     // high half of the block counter the partial round was computed for
	uint32_t c1[4];  // round-1 output columns without the s3 lookups (includes rk[4..7])
};

#define AES_TAB_BYTES 131072
// dynamic shared memory a kernel must request to hold the 64 KB-aligned tables when its
// window starts at 0x400 (checked on the device, see aes_table_offset)
#define AES_SMEM_GAP 64512
#define AES_SMEM_BYTES (AES_SMEM_GAP + AES_TAB_BYTES)

// host: expand a 16-byte key; build T0; CTR key for the stream blocks with block >> 32 == hi
void aes_host_expand(const uint8_t key[16], AesKey *out);
void aes_host_ctr_key(const uint8_t key[16], uint32_t hi, AesCtrKey *out);
const uint32_t *aes_host_T0(void);
struct p1_ctx;
// device copy of T0 (1 KB in global memory, uploaded once per device)
int aes_device_T0(p1_ctx *ctx, const uint32_t **dT0);

#ifdef __CUDACC__

// byte offset into the kernel's dynamic shared memory at which the tables start (the
// next multiple of 64 KB in the shared window that leaves `reserve` bytes below it).
// Traps if the tables (+ `above` bytes after them) do not fit the launch's allocation.
__device__ __forceinline__ uint32_t aes_table_offset(const void *dyn_smem, uint32_t reserve, uint32_t above) {
	const uint32_t base = (uint32_t) __cvta_generic_to_shared(dyn_smem);
	const uint32_t off = ((base + reserve + 0xffffu) & ~0xffffu) - base;
	uint32_t dyn;
	asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
	if(off + AES_TAB_BYTES + above > dyn) {
		if(threadIdx.x == 0) printf("p1gpu: AES tables do not fit: window base 0x%x, offset %u, allocation %u\n", base, off, dyn);
		__trap();
	}
	return off;
}

__device__ __forceinline__ void aes_load_tables(uint8_t *sT, const uint32_t *__restrict__ gT0) {
	for(int i = threadIdx.x; i < 256 * 32; i += blockDim.x) {
		const uint32_t x = (uint32_t) i >> 5, l = (uint32_t) i & 31u;
		const uint32_t v = gT0[x];
		uint8_t *row = sT + x * 256 + l * 4;
		*reinterpret_cast<uint32_t *>(row) = v;
		*reinterpret_cast<uint32_t *>(row + 128) = __byte_perm(v, v, 0x2103);
		*reinterpret_cast<uint32_t *>(row + 65536) = __byte_perm(v, v, 0x1032);
		*reinterpret_cast<uint32_t *>(row + 65536 + 128) = __byte_perm(v, v, 0x0321);
	}
}

// the lane's lookup base: shared address of table row 0 + 4*lane (byte 1 is zero).
// Call after the barrier that follows aes_load_tables.
__device__ __forceinline__ uint32_t aes_lane_base(const uint8_t *sT, uint32_t lane) {
	uint32_t b = (uint32_t) __cvta_generic_to_shared(sT) + 4 * lane;
	asm volatile("" : "+r"(b) :: "memory");   // lookups may not move above the barrier
	return b;
}

// T_T[byte K of s]
template <int T, int K>
__device__ __forceinline__ uint32_t aes_lk(uint32_t s, uint32_t lb) {
	const uint32_t a = __byte_perm(lb, s, 0x3240 + (K << 4));
	uint32_t v;
	asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"((T >> 1) * 65536 + (T & 1) * 128));
	return v;
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

#define AES_ROUND(rk, r, s0, s1, s2, s3, t0, t1, t2, t3) \
	t0 = aes_lk<0, 0>(s0, lb) ^ aes_lk<1, 1>(s1, lb) ^ aes_lk<2, 2>(s2, lb) ^ aes_lk<3, 3>(s3, lb) ^ rk(4 * (r)); \
	t1 = aes_lk<0, 0>(s1, lb) ^ aes_lk<1, 1>(s2, lb) ^ aes_lk<2, 2>(s3, lb) ^ aes_lk<3, 3>(s0, lb) ^ rk(4 * (r) + 1); \
	t2 = aes_lk<0, 0>(s2, lb) ^ aes_lk<1, 1>(s3, lb) ^ aes_lk<2, 2>(s0, lb) ^ aes_lk<3, 3>(s1, lb) ^ rk(4 * (r) + 2); \
	t3 = aes_lk<0, 0>(s3, lb) ^ aes_lk<1, 1>(s0, lb) ^ aes_lk<2, 2>(s1, lb) ^ aes_lk<3, 3>(s2, lb) ^ rk(4 * (r) + 3);

// last round: SubBytes + ShiftRows + AddRoundKey.  S[x] sits in byte 0 of T2[x], bytes
// 1 and 2 of T0[x] and byte 3 of T1[x], so the four lookups already have the S-box output
// in its destination byte and three PRMTs merge them.
#define AES_LAST(a, b, c, d) \
	__byte_perm(__byte_perm(aes_lk<2, 0>(a, lb), aes_lk<0, 1>(b, lb), 0x0050), \
	            __byte_perm(aes_lk<0, 2>(c, lb), aes_lk<1, 3>(d, lb), 0x7200), 0x7610)

// rounds R0..10 on the state after round R0-1
template <int R0, typename KA>
__device__ __forceinline__ void aes_rounds_to_10(const KA &rk, uint32_t lb,
	uint32_t s0, uint32_t s1, uint32_t s2, uint32_t s3, uint32_t s[4]) {
#pragma unroll
	for(int r = R0; r < 10; r++) {
		uint32_t t0, t1, t2, t3;
		AES_ROUND(rk, r, s0, s1, s2, s3, t0, t1, t2, t3)
		s0 = t0; s1 = t1; s2 = t2; s3 = t3;
	}
	s[0] = AES_LAST(s0, s1, s2, s3) ^ rk(40);
	s[1] = AES_LAST(s1, s2, s3, s0) ^ rk(41);
	s[2] = AES_LAST(s2, s3, s0, s1) ^ rk(42);
	s[3] = AES_LAST(s3, s0, s1, s2) ^ rk(43);
}
template <typename KA>
__device__ __forceinline__ void aes_rounds_2_10(const KA &rk, uint32_t lb,
	uint32_t s0, uint32_t s1, uint32_t s2, uint32_t s3, uint32_t s[4]) {
	aes_rounds_to_10<2>(rk, lb, s0, s1, s2, s3, s);
}

// AES-128 encryption of the block (i0,i1,i2,i3) (little-endian words)
template <typename KA>
__device__ __forceinline__ void aes_encrypt(const KA &rk, uint32_t lb,
	uint32_t i0, uint32_t i1, uint32_t i2, uint32_t i3, uint32_t s[4]) {
	uint32_t s0 = i0 ^ rk(0), s1 = i1 ^ rk(1), s2 = i2 ^ rk(2), s3 = i3 ^ rk(3);
	uint32_t t0, t1, t2, t3;
	AES_ROUND(rk, 1, s0, s1, s2, s3, t0, t1, t2, t3)
	aes_rounds_2_10(rk, lb, t0, t1, t2, t3, s);
}

// keystream block: encrypts the counter block 0^64 || BE64(block); the 16 keystream
// bytes come back as four little-endian words s[0..3]
template <typename KA>
__device__ __forceinline__ void aes_ctr_block_k(const KA &rk, uint32_t lb, uint64_t block, uint32_t s[4]) {
	aes_encrypt(rk, lb, 0u, 0u, __byte_perm((uint32_t) (block >> 32), 0, 0x0123),
		__byte_perm((uint32_t) block, 0, 0x0123), s);
}

// same under a grid-wide CTR key: while block >> 32 == K.hi only the four round-1
// lookups that depend on the low counter word are done (the rest is folded into K.c1)
__device__ __forceinline__ void aes_ctr_block(const AesCtrKey &K, uint32_t lb, uint64_t block, uint32_t s[4]) {
	const AesConstKey rk(K.k);
	const uint32_t s3 = __byte_perm((uint32_t) block, 0, 0x0123) ^ K.k.rk[3];
	uint32_t t0, t1, t2, t3;
	if((uint32_t) (block >> 32) == K.hi) {
		t0 = K.c1[0] ^ aes_lk<3, 3>(s3, lb);
		t1 = K.c1[1] ^ aes_lk<2, 2>(s3, lb);
		t2 = K.c1[2] ^ aes_lk<1, 1>(s3, lb);
		t3 = K.c1[3] ^ aes_lk<0, 0>(s3, lb);
	} else {
		const uint32_t s0 = K.k.rk[0], s1 = K.k.rk[1], s2 = __byte_perm((uint32_t) (block >> 32), 0, 0x0123) ^ K.k.rk[2];
		AES_ROUND(rk, 1, s0, s1, s2, s3, t0, t1, t2, t3)
	}
	aes_rounds_2_10(rk, lb, t0, t1, t2, t3, s);
}

// CTR blocks that agree in all counter bits above the low byte (a 256-block group) share
// three of the four round-1 columns, hence 12 of the 16 round-2 lookups: they are kept in
// registers (g) and a block of the group costs 1 + 4 lookups for rounds 1-2 instead of 20.
struct AesCtrGroup {
	uint32_t id;     // (uint32_t) block >> 8 of the group g belongs to; 0xffffffff = none
	uint32_t g[4];   // round-2 output columns without the lookups of round-1 column 0 (includes rk[8..11])
};

__device__ __forceinline__ void aes_ctr_group(const AesCtrKey &K, uint32_t lb, uint32_t lo, AesCtrGroup &G) {
	const uint32_t s3 = __byte_perm(lo, 0, 0x0123) ^ K.k.rk[3];
	const uint32_t t1 = K.c1[1] ^ aes_lk<2, 2>(s3, lb);
	const uint32_t t2 = K.c1[2] ^ aes_lk<1, 1>(s3, lb);
	const uint32_t t3 = K.c1[3] ^ aes_lk<0, 0>(s3, lb);
	G.g[0] = aes_lk<1, 1>(t1, lb) ^ aes_lk<2, 2>(t2, lb) ^ aes_lk<3, 3>(t3, lb) ^ K.k.rk[8];
	G.g[1] = aes_lk<0, 0>(t1, lb) ^ aes_lk<1, 1>(t2, lb) ^ aes_lk<2, 2>(t3, lb) ^ K.k.rk[9];
	G.g[2] = aes_lk<0, 0>(t2, lb) ^ aes_lk<1, 1>(t3, lb) ^ aes_lk<3, 3>(t1, lb) ^ K.k.rk[10];
	G.g[3] = aes_lk<0, 0>(t3, lb) ^ aes_lk<2, 2>(t1, lb) ^ aes_lk<3, 3>(t2, lb) ^ K.k.rk[11];
	G.id = lo >> 8;
}

// keystream block of the group G caches (block >> 32 == K.hi and (uint32_t) block >> 8 == G.id): 1 + 4 lookups for
// rounds 1-2, then rounds 3-10
__device__ __forceinline__ void aes_ctr_block_in_group(const AesCtrKey &K, uint32_t lb, uint32_t lo, const AesCtrGroup &G, uint32_t s[4]) {
	const AesConstKey rk(K.k);
	const uint32_t s3 = __byte_perm(lo, 0, 0x0123) ^ K.k.rk[3];
	const uint32_t t0 = K.c1[0] ^ aes_lk<3, 3>(s3, lb);
	const uint32_t u0 = G.g[0] ^ aes_lk<0, 0>(t0, lb);
	const uint32_t u1 = G.g[1] ^ aes_lk<3, 3>(t0, lb);
	const uint32_t u2 = G.g[2] ^ aes_lk<2, 2>(t0, lb);
	const uint32_t u3 = G.g[3] ^ aes_lk<1, 1>(t0, lb);
	aes_rounds_to_10<3>(rk, lb, u0, u1, u2, u3, s);
}

// keystream block through the group cache G (one cache per stream a lane walks)
__device__ __forceinline__ void aes_ctr_block_g(const AesCtrKey &K, uint32_t lb, uint64_t block, AesCtrGroup &G, uint32_t s[4]) {
	const uint32_t lo = (uint32_t) block;
	if((uint32_t) (block >> 32) == K.hi) {
		if((lo >> 8) != G.id) aes_ctr_group(K, lb, lo, G);
		aes_ctr_block_in_group(K, lb, lo, G, s);
	} else {
		aes_ctr_block_k(AesConstKey(K.k), lb, block, s);
	}
}

#endif
