// Correlated-OT extension for the --use_ot aggregation (src/phase1.c:43-96): the
// construction of oracle/cot_ext.h on the GPU.  Tile = 128 consecutive OTs.
//   per tile and column i (thread i of a 128-thread group): one AES-CTR block per seed
//   (per-thread key schedules in shared memory), U = t ^ G(k1) ^ choice (receiver) or
//   q = G(k) ^ Delta_i*U (sender); 128x128 bit transpose with warp ballots; per OT row
//   (thread j): fixed-key AES hash, correlation (1<<bit)*x + opt0, wire word C.
#include "p1_common.cuh"
#include "aes_dev.cuh"
#include <string.h>

static const uint8_t COT_HASH_KEY[16] = { // "linreg-mpc-cot-H", as in oracle/cot_ext.c
	0x6c, 0x69, 0x6e, 0x72, 0x65, 0x67, 0x2d, 0x6d, 0x70, 0x63, 0x2d, 0x63, 0x6f, 0x74, 0x2d, 0x48};

struct p1_cot {
	int device, bits;
	int role;            // 0 receiver, 1 sender
	uint32_t *keys_dev;  // [nk][44][128] round-key words, nk = 2 (receiver: k^0, k^1) or 1 (sender)
	size_t keys_bytes;
	uint32_t delta[4];
	uint64_t pos, ot_base;
	// receiver state between _extend and _finish
	void *pads;
	size_t pads_bytes;
	const void *col_ptr;
	size_t n_rows;
};

#define COT_GROUPS 4
#define COT_THREADS (128 * COT_GROUPS)

struct CotParams {
	AesKey hash;
	const uint32_t *T0;
	const uint32_t *keys;   // device key schedules
	uint32_t delta[4];
	uint64_t ctr0;          // PRG block of local tile 0
	uint64_t J0;            // OT counter of local OT 0
	uint64_t tiles;         // local tiles
	uint64_t n_rows;
	const void *col;        // this party's column (choice bits / correlator input)
	void *u;                // wire U
	void *c;                // wire C (sender)
	void *pads;             // receiver stash
	uint64_t *acc;          // sender: acc[slot] -= sum opt0
};

__device__ __forceinline__ void group_sync(int g) {
	asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(128) : "memory");
}

// A thread keeps the CTR key of ITS column (thread t <-> column t for the whole launch) in
// registers: round keys 8..43, rk[3], and round 1 of the counter block 0^64 || BE32(hi) || BE32(lo)
// without the four lookups that depend on lo (as AesCtrKey does for the grid-wide TI key).  A PRG
// block then costs 4 + 144 shared-memory lookups and no round-key loads (was 160 + 44).
struct AesRegCtrKey {
	uint32_t hi, rk3, c1[4], rk[36];
};
struct AesRegAcc {
	const AesRegCtrKey &R;
	__device__ __forceinline__ AesRegAcc(const AesRegCtrKey &r) : R(r) {}
	__device__ __forceinline__ uint32_t operator()(int i) const { return R.rk[i - 8]; }
};
__device__ __forceinline__ void aes_regkey_init(const AesSmemKey &k, uint32_t lb, uint32_t hi, AesRegCtrKey &R) {
	R.hi = hi;
	R.rk3 = k(3);
#pragma unroll
	for(int i = 0; i < 36; i++) R.rk[i] = k(8 + i);
	const uint32_t s0 = k(0), s1 = k(1), s2 = __byte_perm(hi, 0, 0x0123) ^ k(2);
	R.c1[0] = aes_lk<0, 0>(s0, lb) ^ aes_lk<1, 1>(s1, lb) ^ aes_lk<2, 2>(s2, lb) ^ k(4);
	R.c1[1] = aes_lk<0, 0>(s1, lb) ^ aes_lk<1, 1>(s2, lb) ^ aes_lk<3, 3>(s0, lb) ^ k(5);
	R.c1[2] = aes_lk<0, 0>(s2, lb) ^ aes_lk<2, 2>(s0, lb) ^ aes_lk<3, 3>(s1, lb) ^ k(6);
	R.c1[3] = aes_lk<1, 1>(s0, lb) ^ aes_lk<2, 2>(s1, lb) ^ aes_lk<3, 3>(s2, lb) ^ k(7);
}
// keystream block `block` (block >> 32 == R.hi)
__device__ __forceinline__ void aes_ctr_block_r(const AesRegCtrKey &R, uint32_t lb, uint32_t lo, uint32_t s[4]) {
	const uint32_t s3 = __byte_perm(lo, 0, 0x0123) ^ R.rk3;
	const uint32_t t0 = R.c1[0] ^ aes_lk<3, 3>(s3, lb);
	const uint32_t t1 = R.c1[1] ^ aes_lk<2, 2>(s3, lb);
	const uint32_t t2 = R.c1[2] ^ aes_lk<1, 1>(s3, lb);
	const uint32_t t3 = R.c1[3] ^ aes_lk<0, 0>(s3, lb);
	aes_rounds_to_10<2>(AesRegAcc(R), lb, t0, t1, t2, t3, s);
}

// 32x32 bit-matrix transpose across a warp: lane r holds row r, returns row r of the transpose
// (five butterfly steps: swap the off-diagonal j x j blocks of every 2j x 2j block).  Straight-line code:
// the byte-granular steps (j = 16, 8) are one PRMT each; a bit-granular step takes the partner's word
// rotated by j (lower lane of the pair) or 32 - j (upper lane) - the bits that wrap around land outside
// the half that is kept - and one bit-select.
template <int J>
__device__ __forceinline__ uint32_t cot_bfly_bits(uint32_t x, bool up) {
	constexpr uint32_t m = J == 4 ? 0x0f0f0f0fu : J == 2 ? 0x33333333u : 0x55555555u;
	const uint32_t y = __shfl_xor_sync(0xffffffffu, x, J);
	const uint32_t r = __funnelshift_l(y, y, up ? 32 - J : J);
	const uint32_t keep = up ? ~m : m;
	return (x & keep) | (r & ~keep);
}
__device__ __forceinline__ uint32_t cot_transpose32(uint32_t x, uint32_t lane) {
	uint32_t y = __shfl_xor_sync(0xffffffffu, x, 16);
	x = __byte_perm(x, y, (lane & 16) ? 0x3276 : 0x5410);
	y = __shfl_xor_sync(0xffffffffu, x, 8);
	x = __byte_perm(x, y, (lane & 8) ? 0x3715 : 0x6240);
	x = cot_bfly_bits<4>(x, lane & 4);
	x = cot_bfly_bits<2>(x, lane & 2);
	return cot_bfly_bits<1>(x, lane & 1);
}

// ROLE 0: receiver extend, ROLE 1: sender step.  TPI = tiles a 128-thread group works on per iteration
// (independent AES chains per thread: 2 for both roles - the receiver's two PRG blocks of one tile, the
// sender's PRG blocks of two tiles; then 1 / 4 independent hashes).
// Shared memory: [transpose rows: COT_GROUPS x TPI x 2 buffers x 128 x 5 words | gap | AES tables].  The rows
// are double buffered, so one named barrier per iteration suffices: a group writes buffer i&1 in
// iteration i while stragglers may still read buffer (i-1)&1.
template <typename W, int ROLE>
__global__ void __launch_bounds__(COT_THREADS, 1) k_cot_tiles(const CotParams P) {
	extern __shared__ __align__(256) uint8_t smem[];
	constexpr int TPI = ROLE == 1 ? 2 : 1;
	constexpr int WB = sizeof(W) * 8;
	constexpr int ROWS = 128 * 5;                     // words of one transposed tile (row stride 5: conflict free)
	uint32_t *sR = reinterpret_cast<uint32_t *>(smem);
	uint8_t *sT = smem + aes_table_offset(smem, COT_GROUPS * TPI * 2 * ROWS * 4, 0);
	aes_load_tables(sT, P.T0);
	__syncthreads();
	const int g = threadIdx.x >> 7, t = threadIdx.x & 127, wq = t >> 5, lane = t & 31;
	const uint32_t lb = aes_lane_base(sT, (uint32_t) lane);
	uint32_t *myR = sR + g * (TPI * 2 * ROWS);
	const bool dbit = ROLE == 1 && ((P.delta[t >> 5] >> (t & 31)) & 1);
	// the column keys of thread t: expanded schedules in global memory (read once; again only when the PRG
	// block counter crosses a multiple of 2^32)
	const AesSmemKey k0s(P.keys, 128, t), k1s(P.keys + 44 * 128, 128, t);
	AesRegCtrKey K0, K1;
	aes_regkey_init(k0s, lb, (uint32_t) (P.ctr0 >> 32), K0);
	if(ROLE == 0) aes_regkey_init(k1s, lb, (uint32_t) (P.ctr0 >> 32), K1);
	uint64_t dot = 0;
	const uint64_t stride = (uint64_t) gridDim.x * COT_GROUPS * TPI;
	uint32_t it = 0;
	for(uint64_t tile0 = ((uint64_t) blockIdx.x * COT_GROUPS + g) * TPI; tile0 < P.tiles; tile0 += stride, it++) {
		uint32_t *buf = myR + (it & 1) * (TPI * ROWS);
		// ---- global operands of this iteration, requested before the AES work that hides their latency
		uint4 uv[TPI];
		W xk[TPI];
#pragma unroll
		for(int z = 0; z < TPI; z++) {
			const uint64_t tile = tile0 + z;
			uv[z] = make_uint4(0, 0, 0, 0);
			xk[z] = 0;
			if(tile >= P.tiles) continue;
			if(ROLE == 0) uv[z] = __ldg(reinterpret_cast<const uint4 *>(P.col) + tile);   // 128 choice bits = 16 bytes of the column
			else {
				uv[z] = __ldcs(reinterpret_cast<const uint4 *>((const char *) P.u + tile * 2048) + t);
				const uint64_t k = (tile * 128 + t) / WB;
				if(k < P.n_rows) xk[z] = __ldg(reinterpret_cast<const W *>(P.col) + k);
			}
		}
		// ---- column phase: thread t = column t of the kappa x m matrix.  Straight-line code for all TPI tiles
		// (a tile past the end computes on zeros and stores nothing), so the independent AES chains interleave.
		const uint64_t blk0 = P.ctr0 + tile0;
		const bool fast = (uint32_t) (blk0 >> 32) == K0.hi && (uint32_t) ((blk0 + TPI - 1) >> 32) == K0.hi;   // uniform over the group
		uint32_t a[TPI][4];
		if(fast) {
#pragma unroll
			for(int z = 0; z < TPI; z++) aes_ctr_block_r(K0, lb, (uint32_t) (blk0 + z), a[z]);
			if(ROLE == 0) {
				uint32_t b[4];
				aes_ctr_block_r(K1, lb, (uint32_t) blk0, b);
#pragma unroll
				for(int c = 0; c < 4; c++) b[c] ^= a[0][c];
				if(tile0 < P.tiles) __stcs(reinterpret_cast<uint4 *>((char *) P.u + tile0 * 2048) + t,
					make_uint4(b[0] ^ uv[0].x, b[1] ^ uv[0].y, b[2] ^ uv[0].z, b[3] ^ uv[0].w));
			}
		} else {
			// the PRG block counter crosses a multiple of 2^32 here: schedules from global memory, then re-key
#pragma unroll
			for(int z = 0; z < TPI; z++) aes_ctr_block_k(k0s, lb, blk0 + z, a[z]);
			if(ROLE == 0) {
				uint32_t b[4];
				aes_ctr_block_k(k1s, lb, blk0, b);
				if(tile0 < P.tiles) __stcs(reinterpret_cast<uint4 *>((char *) P.u + tile0 * 2048) + t,
					make_uint4(a[0][0] ^ b[0] ^ uv[0].x, a[0][1] ^ b[1] ^ uv[0].y, a[0][2] ^ b[2] ^ uv[0].z, a[0][3] ^ b[3] ^ uv[0].w));
			}
			const uint32_t hi_next = (uint32_t) ((blk0 + stride) >> 32);
			aes_regkey_init(k0s, lb, hi_next, K0);
			if(ROLE == 0) aes_regkey_init(k1s, lb, hi_next, K1);
		}
#pragma unroll
		for(int z = 0; z < TPI; z++) {
			if(ROLE == 1 && dbit) { a[z][0] ^= uv[z].x; a[z][1] ^= uv[z].y; a[z][2] ^= uv[z].z; a[z][3] ^= uv[z].w; }
			// 128x128 bit transpose: four 32x32 butterflies per warp (word b of the warp's 32 columns ->
			// word wq of rows 32b .. 32b+31), then the rows are gathered through shared memory
#pragma unroll
			for(int b = 0; b < 4; b++) buf[z * ROWS + (32 * b + lane) * 5 + wq] = cot_transpose32(a[z][b], (uint32_t) lane);
		}
		group_sync(g);
		// ---- row phase: thread t = OT t of each tile; all hashes first (independent chains), then the stores
		uint32_t x[TPI][4], h0[TPI][4], h1[TPI][4];
#pragma unroll
		for(int z = 0; z < TPI; z++) {
			const uint64_t J = P.J0 + (tile0 + z) * 128 + t;
			x[z][0] = buf[z * ROWS + t * 5 + 0] ^ (uint32_t) J;
			x[z][1] = buf[z * ROWS + t * 5 + 1] ^ (uint32_t) (J >> 32);
			x[z][2] = buf[z * ROWS + t * 5 + 2];
			x[z][3] = buf[z * ROWS + t * 5 + 3];
		}
#pragma unroll
		for(int z = 0; z < TPI; z++) {
			aes_encrypt(AesConstKey(P.hash), lb, x[z][0], x[z][1], x[z][2], x[z][3], h0[z]);
			if(ROLE == 1) aes_encrypt(AesConstKey(P.hash), lb, x[z][0] ^ P.delta[0], x[z][1] ^ P.delta[1], x[z][2] ^ P.delta[2], x[z][3] ^ P.delta[3], h1[z]);
		}
#pragma unroll
		for(int z = 0; z < TPI; z++) {
			const uint64_t tile = tile0 + z;
			if(tile >= P.tiles) continue;
			const W pad0 = sizeof(W) == 8 ? (W) ((((uint64_t) (h0[z][1] ^ x[z][1])) << 32) | (h0[z][0] ^ x[z][0])) : (W) (h0[z][0] ^ x[z][0]);
			if(ROLE == 0) {
				__stcs(reinterpret_cast<W *>(P.pads) + tile * 128 + t, pad0);
			} else {
				const uint32_t y0 = x[z][0] ^ P.delta[0], y1 = x[z][1] ^ P.delta[1];
				const W pad1 = sizeof(W) == 8 ? (W) ((((uint64_t) (h1[z][1] ^ y1)) << 32) | (h1[z][0] ^ y0)) : (W) (h1[z][0] ^ y0);
				const uint64_t ni = tile * 128 + t, k = ni / WB;
				const int bit = (int) (ni % WB);
				if(k < P.n_rows) {
					const W opt1 = (W) (xk[z] << bit) + pad0; // (1 << bit) * x + opt0, src/phase1.c:50
					__stcs(reinterpret_cast<W *>(P.c) + ni, (W) (opt1 ^ pad1));
					dot -= (uint64_t) pad0;             // share = -sum opt0, src/phase1.c:66-68
				}
			}
		}
	}
	if(ROLE == 1) {
		__shared__ uint64_t part[COT_THREADS / 32];
		dot = p1_warp_sum(dot);
		if((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = dot;
		__syncthreads();
		if(threadIdx.x < 32) {
			uint64_t s = threadIdx.x < COT_THREADS / 32 ? part[threadIdx.x] : 0;
			s = p1_warp_sum(s);
			if(threadIdx.x == 0 && s) p1_red_add(P.acc, s);
		}
	}
}

// receiver second half: dest = pad ^ (choice ? C : 0); share = +sum dest (src/phase1.c:90-92)
template <typename W>
__global__ void __launch_bounds__(256) k_cot_finish(const W *__restrict__ pads, const W *__restrict__ c, const W *__restrict__ col,
	uint64_t m, uint64_t *__restrict__ acc) {
	constexpr int WB = sizeof(W) * 8;
	uint64_t dot = 0;
	for(uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; j < m; j += (uint64_t) gridDim.x * blockDim.x) {
		W a = col[j / WB];
		W v = pads[j];
		if((a >> (j % WB)) & 1) v ^= c[j];
		dot += (uint64_t) v;
	}
	__shared__ uint64_t part[8];
	dot = p1_warp_sum(dot);
	if((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = dot;
	__syncthreads();
	if(threadIdx.x < 32) {
		uint64_t s = threadIdx.x < 8 ? part[threadIdx.x] : 0;
		s = p1_warp_sum(s);
		if(threadIdx.x == 0 && s) p1_red_add(acc, s);
	}
}

// ------------------------------------------------------------------ host


static int cot_new(p1_ctx *ctx, int role, const uint8_t (*seeds_a)[16], const uint8_t (*seeds_b)[16], const uint8_t *delta, p1_cot **out) {
	P1_CHECK(out, "null output");
	*out = NULL;
	int nk = role == 0 ? 2 : 1;
	uint32_t *h = (uint32_t *) malloc((size_t) nk * 44 * 128 * sizeof(uint32_t));
	P1_CHECK(h, "out of memory");
	for(int kk = 0; kk < nk; kk++) {
		for(int t = 0; t < 128; t++) {
			AesKey K;
			aes_host_expand(kk == 0 ? seeds_a[t] : seeds_b[t], &K);
			for(int i = 0; i < 44; i++) h[((size_t) kk * 44 + i) * 128 + t] = K.rk[i];
		}
	}
	p1_cot *s = new p1_cot();
	s->device = ctx->device;
	s->bits = ctx->bits_p1;
	s->role = role;
	s->pos = s->ot_base = 0;
	s->pads = NULL;
	s->pads_bytes = 0;
	s->col_ptr = NULL;
	s->n_rows = 0;
	memset(s->delta, 0, sizeof(s->delta));
	if(delta) memcpy(s->delta, delta, 16); // little-endian words: bit i of Delta = bit (i%8) of byte i/8
	s->keys_bytes = (size_t) nk * 44 * 128 * sizeof(uint32_t);
	s->keys_dev = (uint32_t *) p1_pool_dev_take(ctx->device, s->keys_bytes);
	cudaError_t e = s->keys_dev ? cudaMemcpy(s->keys_dev, h, s->keys_bytes, cudaMemcpyHostToDevice) : cudaErrorMemoryAllocation;
	free(h);
	if(e != cudaSuccess) {
		delete s;
		return p1_fail("p1_cot_new: %s", cudaGetErrorString(e));
	}
	*out = s;
	return 0;
}

extern "C" int p1_cot_sender_new(p1_ctx *ctx, const uint8_t delta[16], const uint8_t seeds[128][16], p1_cot **s) {
	P1_ENTER(ctx);
	P1_CHECK(delta && seeds, "p1_cot_sender_new: arguments may not be null");
	return cot_new(ctx, 1, seeds, NULL, delta, s);
}

extern "C" int p1_cot_receiver_new(p1_ctx *ctx, const uint8_t seeds0[128][16], const uint8_t seeds1[128][16], p1_cot **r) {
	P1_ENTER(ctx);
	P1_CHECK(seeds0 && seeds1, "p1_cot_receiver_new: arguments may not be null");
	return cot_new(ctx, 0, seeds0, seeds1, NULL, r);
}

extern "C" void p1_cot_free(p1_cot *s) {
	if(!s) return;
	cudaSetDevice(s->device);
	cudaDeviceSynchronize();
	p1_pool_dev_give_secret(s->device, s->keys_dev, s->keys_bytes);    // expanded round keys of the base-OT seeds
	p1_pool_dev_give_secret(s->device, s->pads, s->pads_bytes);        // hashed pads H(t) of the last inner product
	delete s;
}

static size_t cot_tiles(int bits, size_t n_rows) { return (n_rows * (size_t) bits + 127) / 128; }

extern "C" size_t p1_cot_u_bytes(const p1_ctx *ctx, size_t n_rows) { return ctx ? 2048 * cot_tiles(ctx->bits_p1, n_rows) : 0; }
extern "C" size_t p1_cot_c_bytes(const p1_ctx *ctx, size_t n_rows) { return ctx ? 128 * (size_t) (ctx->bits_p1 / 8) * cot_tiles(ctx->bits_p1, n_rows) : 0; }

template <int ROLE>
static int cot_launch(p1_ctx *ctx, p1_cot *s, const p1_data *data, int32_t col, size_t n_total, size_t row0,
	void *u_dev, void *c_dev, uint64_t *acc_slot) {
	P1_CHECK(s && data, "null session or data");
	P1_CHECK(s->role == (ROLE == 0 ? 0 : 1), "session has the other role");
	P1_CHECK(s->bits == ctx->bits_p1 && data->bits == ctx->bits_p1, "ring width mismatch");
	P1_CHECK(col >= 0 && (size_t) col <= data->d && ((size_t) col < data->d || data->has_y), "invalid column %d", col);
	size_t per_tile = 128 / (size_t) s->bits; // rows per tile
	P1_CHECK(row0 % per_tile == 0 && row0 + data->n <= n_total, "row shard [%zu,%zu) must start on a multiple of %zu rows inside [0,%zu)",
		row0, row0 + data->n, per_tile, n_total);
	P1_CHECK(row0 + data->n == n_total || data->n % per_tile == 0, "only the last row shard may hold a partial OT tile");
	CotParams P;
	aes_host_expand(COT_HASH_KEY, &P.hash);
	if(aes_device_T0(ctx, &P.T0)) return 1;
	P.keys = s->keys_dev;
	memcpy(P.delta, s->delta, 16);
	size_t tile0 = row0 * (size_t) s->bits / 128;
	P.ctr0 = s->pos + tile0;
	P.J0 = s->ot_base + 128 * (uint64_t) tile0;
	P.tiles = cot_tiles(s->bits, data->n);
	P.n_rows = data->n;
	P.col = (const char *) data->xt + (size_t) col * data->ld * (size_t) (s->bits / 8);
	P.u = u_dev;
	P.c = c_dev;
	P.acc = acc_slot;
	P.pads = NULL;
	if(ROLE == 0) {
		size_t need = P.tiles * 128 * (size_t) (s->bits / 8);
		if(s->pads_bytes < need) {
			if(s->pads) {
				P1_CUDA(cudaDeviceSynchronize());
				p1_pool_dev_give_secret(s->device, s->pads, s->pads_bytes);
				s->pads_bytes = 0;
			}
			s->pads = p1_pool_dev_take(s->device, need);
			P1_CHECK(s->pads, "p1_cot_receiver_extend: cannot allocate %zu bytes", need);
			s->pads_bytes = need;
		}
		P.pads = s->pads;
		s->col_ptr = P.col;
		s->n_rows = data->n;
	}
	size_t smem = AES_SMEM_BYTES;                         // the transposed rows live in the gap below the tables
	static_assert(COT_GROUPS * 2 * 2 * 128 * 5 * 4 <= AES_SMEM_GAP, "transpose rows must fit below the AES tables");
	constexpr int TPI = ROLE == 1 ? 2 : 1;
	uint64_t grid = (P.tiles + COT_GROUPS * TPI - 1) / (COT_GROUPS * TPI);
	int per_sm = 0;
	if(s->bits == 64) {
		P1_CUDA(cudaFuncSetAttribute(k_cot_tiles<uint64_t, ROLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		P1_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_cot_tiles<uint64_t, ROLE>, COT_THREADS, smem));
	} else {
		P1_CUDA(cudaFuncSetAttribute(k_cot_tiles<uint32_t, ROLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		P1_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_cot_tiles<uint32_t, ROLE>, COT_THREADS, smem));
	}
	P1_CHECK(per_sm > 0, "k_cot_tiles does not fit on an SM");
	uint64_t cap = (uint64_t) ctx->sm_count * per_sm;
	if(grid > cap) grid = cap;
	p1_prof_begin(ctx, ROLE == 0 ? P1_K_COT_RECV : P1_K_COT_SEND);
	if(s->bits == 64) k_cot_tiles<uint64_t, ROLE><<<(unsigned) grid, COT_THREADS, smem, ctx->stream>>>(P);
	else k_cot_tiles<uint32_t, ROLE><<<(unsigned) grid, COT_THREADS, smem, ctx->stream>>>(P);
	P1_LAUNCHED(ctx);
	// the cursor moves by the WHOLE inner product, so every row shard stays in step
	size_t tiles_total = cot_tiles(s->bits, n_total);
	s->pos += tiles_total;
	s->ot_base += 128 * (uint64_t) tiles_total;
	return 0;
}

extern "C" int p1_cot_advance(p1_cot *s, size_t n_total, size_t count) {
	P1_CHECK(s, "p1_cot_advance: null session");
	const size_t tiles_total = cot_tiles(s->bits, n_total);
	s->pos += (uint64_t) count * tiles_total;
	s->ot_base += 128 * (uint64_t) count * tiles_total;
	return 0;
}

extern "C" int p1_cot_set_cursor(p1_cot *s, uint64_t prg_block, uint64_t ot_counter) {
	P1_CHECK(s, "p1_cot_set_cursor: null session");
	s->pos = prg_block;
	s->ot_base = ot_counter;
	return 0;
}

extern "C" int p1_cot_receiver_extend(p1_ctx *ctx, p1_cot *r, const p1_data *data, int32_t col,
	size_t n_total, size_t row0, void *u_dev) {
	P1_ENTER(ctx);
	P1_CHECK(u_dev, "p1_cot_receiver_extend: null U buffer");
	return cot_launch<0>(ctx, r, data, col, n_total, row0, u_dev, NULL, NULL);
}

extern "C" int p1_cot_sender_step(p1_ctx *ctx, p1_cot *s, const p1_data *data, int32_t col,
	size_t n_total, size_t row0, const void *u_dev, void *c_dev, uint64_t *acc_dev, int64_t slot) {
	P1_ENTER(ctx);
	P1_CHECK(u_dev && c_dev && acc_dev, "p1_cot_sender_step: arguments may not be null");
	return cot_launch<1>(ctx, s, data, col, n_total, row0, (void *) u_dev, c_dev, acc_dev + slot);
}

extern "C" int p1_cot_receiver_finish(p1_ctx *ctx, p1_cot *r, const void *c_dev, uint64_t *acc_dev, int64_t slot) {
	P1_ENTER(ctx);
	P1_CHECK(r && c_dev && acc_dev, "p1_cot_receiver_finish: arguments may not be null");
	P1_CHECK(r->role == 0 && r->col_ptr, "p1_cot_receiver_finish: no extend pending on this session");
	uint64_t m = (uint64_t) r->n_rows * (uint64_t) r->bits;
	unsigned grid = (unsigned) ((m + 256 * 8 - 1) / (256 * 8));
	if(grid > (unsigned) ctx->sm_count * 8) grid = ctx->sm_count * 8;
	p1_prof_begin(ctx, P1_K_COT_FINISH);
	if(r->bits == 64) k_cot_finish<uint64_t><<<grid, 256, 0, ctx->stream>>>((const uint64_t *) r->pads, (const uint64_t *) c_dev, (const uint64_t *) r->col_ptr, m, acc_dev + slot);
	else k_cot_finish<uint32_t><<<grid, 256, 0, ctx->stream>>>((const uint32_t *) r->pads, (const uint32_t *) c_dev, (const uint32_t *) r->col_ptr, m, acc_dev + slot);
	P1_LAUNCHED(ctx);
	r->col_ptr = NULL;
	return 0;
}
