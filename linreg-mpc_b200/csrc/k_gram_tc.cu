// Ring Gram of a column block on the 5th-gen tensor cores (tcgen05 kind::i8).
//
// Replaces the "own both" pair loop of src/phase1.c:569-571 (inner_product_local,
// :14-20) for a whole column block at once:  acc[idx(i,j)] += sum_k X[k][i]*X[k][j]
// mod 2^64 (mod 2^32 lives in the low half).
//
// Arithmetic.  Every ring word is read as a signed integer and split into L balanced
// base-256 digits  x = sum_l x_l 256^l,  x_l in [-128,127]  (k_limb_split; L is chosen
// from the per-column max |word| gathered by the encode pass; for L = W/8 the dropped
// carry is a multiple of 2^W, so full-range words are exact mod 2^W as well).  Then
//     x*y = sum_{l,m} x_l*y_m * 256^(l+m)                       (mod 2^W: l+m < W/8)
// is a set of s8 x s8 -> s32 GEMMs.  Products are grouped by shift s = l+m into one
// s32 TMEM accumulator each; a group receives at most L products per row, each
// |.| <= 2^14, so after at most floor((2^17-1)/L) rows the accumulators are drained
// into 64-bit registers (sum_s acc_s << 8s), which keeps every s32 sum exact.
//
// Data path.  k_limb_split writes digit planes P[kblock][l][col][128 rows] (K-major: the
// contraction index k = row is contiguous inside a 128-row block, and everything one
// pipeline stage needs is one contiguous region per limb - a stage touches one or two
// 2 MB pages instead of one page per column, which is what the per-SM TMA rate needs).
// One CTA owns a 128 x 64 block of the (column x column) output and a slice of the rows:
//   warp 0    TMA producer: box (128 rows, 128|64 cols, L limbs, 1 block) -> smem, SWIZZLE_128B
//   warp 1    MMA issuer:   for each K-step of 32 rows and each A-limb l one
//             tcgen05.mma  D[:, 64l .. 64(l+c)) += A_l (128 x 32) * [B_0;..;B_{c-1}]^T
//             (the B limb tiles are stacked in smem, so N = 64c covers c shift groups)
//   warps 2-5 epilogue: tcgen05.ld the groups, recombine into u64 registers, zero TMEM
//             for the next chunk; at the end red.global.add.u64 into the packed share.
#include "p1_common.cuh"
#include <cuda.h>

#define TC_THREADS 192
#define TC_BM 128          // output rows per CTA   (A operand columns of X)
#define TC_BN 64           // output columns per CTA (B operand columns of X)
#define TC_BK 128          // rows of X per pipeline stage (= one 128-byte swizzle row of s8)
#define TC_A_BYTES (TC_BM * TC_BK)
#define TC_B_BYTES (TC_BN * TC_BK)

struct TcSet {
	size_t col_lo;
	int mc, m;            // feature columns; + target
	size_t d, nA;
};

// accumulator slot of the pair (a > b) of the set (src/phase1.c:581-585)
__device__ __forceinline__ size_t tc_slot(const TcSet &gs, int a, int b) {
	size_t j = gs.col_lo + b;
	if(a >= gs.mc) return gs.nA + j;
	size_t i = gs.col_lo + a;
	return i * (i + 1) / 2 + j;
}

// ------------------------------------------------------------------ digit planes

// grid (ceil(ldp/1024), m), block 256: thread = 4 consecutive rows of one column; a warp
// writes one whole 128-byte row of P[kblock][l][col][.] per limb
template <typename W>
__global__ void __launch_bounds__(256) k_limb_split(const W *__restrict__ xt, size_t ld, TcSet gs, int L,
	uint8_t *__restrict__ planes, size_t ldp) {
	typedef typename p1_word<W>::s S;
	size_t r = ((size_t) blockIdx.x * 256 + threadIdx.x) * 4;
	if(r >= ldp) return;
	int a = blockIdx.y;
	size_t c = a < gs.mc ? gs.col_lo + a : gs.d;
	const W *src = xt + c * ld + r;
	int64_t v[4];
	if(r < ld) {
		if(sizeof(W) == 8) {
			ulonglong2 p0 = *reinterpret_cast<const ulonglong2 *>(src), p1 = *reinterpret_cast<const ulonglong2 *>(src + 2);
			v[0] = (int64_t) p0.x; v[1] = (int64_t) p0.y; v[2] = (int64_t) p1.x; v[3] = (int64_t) p1.y;
		} else {
			uint4 p = *reinterpret_cast<const uint4 *>(src);
			v[0] = (S) p.x; v[1] = (S) p.y; v[2] = (S) p.z; v[3] = (S) p.w;
		}
	} else {
		v[0] = v[1] = v[2] = v[3] = 0;
	}
	const size_t kb = r / TC_BK, rr = r % TC_BK;
	uint8_t *dst = planes + ((kb * L) * (size_t) gs.m + a) * TC_BK + rr;
	for(int l = 0; l < L; l++) {
		uint32_t packed = 0;
#pragma unroll
		for(int u = 0; u < 4; u++) {
			int64_t dg = (int64_t) (int8_t) (v[u] & 0xff);
			v[u] = (v[u] - dg) >> 8;
			packed |= ((uint32_t) dg & 0xffu) << (8 * u);
		}
		*reinterpret_cast<uint32_t *>(dst + (size_t) l * gs.m * TC_BK) = packed;
	}
}

#include "tc_ptx.cuh"

// ------------------------------------------------------------------ the GEMM kernel

struct TcParams {
	TcSet gs;
	int stages;
	uint32_t tmem_cols;    // power of two >= 64*G
	int nblk;              // K-blocks (of TC_BK rows) in the shard
	int blk_per_cta;       // K-blocks per CTA (grid.y slices)
	int chunk_blks;        // K-blocks per exact s32 accumulation chunk
	uint64_t *acc;
};


// L = digits per word, WD = ring bytes (8 or 4): pairs with l+m >= WD vanish mod 2^W.
// Everything the single MMA-issuing thread needs is a compile-time constant: the issue
// loop is two integer adds per tcgen05.mma (measured: with run-time L the issue thread
// itself was the bottleneck at ~200 clk per MMA).
template <int L, int WD>
__global__ void __launch_bounds__(TC_THREADS, 1) k_gram_tc(const __grid_constant__ CUtensorMap mapA,
	const __grid_constant__ CUtensorMap mapB, const TcParams P) {
	constexpr int G = 2 * L - 1 < WD ? 2 * L - 1 : WD;          // shift groups
	constexpr uint32_t STAGE_BYTES = (uint32_t) L * (TC_A_BYTES + TC_B_BYTES);
	extern __shared__ __align__(1024) uint8_t smem_raw[];
	__shared__ __align__(8) uint64_t bar_full[8], bar_empty[8], bar_tfull, bar_tempty;
	__shared__ uint32_t tmem_base_s;

	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int S = P.stages;
	uint8_t *smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);

	// tile decode: ta = 128-wide row tile, tb = 64-wide column tile with 64*tb < min(128*(ta+1), mc)
	int ta = 0, tb = (int) blockIdx.x;
	for(;;) {
		int lim = TC_BM * (ta + 1) < P.gs.mc ? TC_BM * (ta + 1) : P.gs.mc;
		int cnt = (lim + TC_BN - 1) / TC_BN;
		if(tb < cnt) break;
		tb -= cnt;
		ta++;
	}
	const int blk_lo = (int) blockIdx.y * P.blk_per_cta;
	const int blk_hi = blk_lo + P.blk_per_cta < P.nblk ? blk_lo + P.blk_per_cta : P.nblk;
	const int nb = blk_hi - blk_lo;          // > 0 by construction of the grid

	if(threadIdx.x == 0) {
		for(int s = 0; s < S; s++) {
			mbar_init(&bar_full[s], 1);
			mbar_init(&bar_empty[s], 1);
		}
		mbar_init(&bar_tfull, 1);
		mbar_init(&bar_tempty, 128);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapA)) : "memory");
		asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapB)) : "memory");
	}
	if(warp == 1) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(P.tmem_cols) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = tmem_base_s;

	if(warp == 0) {
		// ===== TMA producer
		if(lane == 0) {
			int s = 0;
			uint32_t ph = 0;
			for(int b = 0; b < nb; b++) {
				mbar_wait(&bar_empty[s], ph ^ 1u);
				mbar_expect_tx(&bar_full[s], STAGE_BYTES);
				uint8_t *sa = smem + (size_t) s * STAGE_BYTES, *sb = sa + (size_t) L * TC_A_BYTES;
				tma_load_4d(sa, &mapA, &bar_full[s], 0, ta * TC_BM, 0, blk_lo + b);
				tma_load_4d(sb, &mapB, &bar_full[s], 0, tb * TC_BN, 0, blk_lo + b);
				if(++s == S) { s = 0; ph ^= 1u; }
			}
		}
	} else if(warp == 1) {
		// ===== MMA issuer (one thread)
		if(lane == 0) {
			const uint32_t lo0 = ((smem_u32(smem) >> 4) & 0x3fffu) | (1u << 16);
			int s = 0, in_chunk = 0;
			uint32_t ph = 0, chunk = 0;
			for(int b = 0; b < nb; b++) {
				if(in_chunk == 0) {
					// the epilogue has drained and zeroed the accumulators of the previous chunk
					mbar_wait(&bar_tempty, chunk & 1u);
					tc_fence_after();
				}
				mbar_wait(&bar_full[s], ph);
				tc_fence_after();
				const uint32_t a_lo = lo0 + (uint32_t) s * (STAGE_BYTES >> 4), b_lo = a_lo + (uint32_t) L * (TC_A_BYTES >> 4);
				// limb-major order: the four K-steps of one A limb accumulate into the same TMEM columns
#pragma unroll
				for(int l = 0; l < L; l++) {
					const int cnt = L < WD - l ? L : WD - l;      // B limbs m with l+m < WD (compile-time after unrolling)
					if(cnt > 0) {
						const int c0 = cnt < 4 ? cnt : 4;
#pragma unroll
						for(int k4 = 0; k4 < TC_BK / 32; k4++) {
							const uint64_t da = tc_desc(a_lo + (uint32_t) (l * (TC_A_BYTES >> 4) + k4 * 2));
							tc_mma_i8(tmem + (uint32_t) (TC_BN * l), da, tc_desc(b_lo + (uint32_t) (k4 * 2)), tc_idesc(TC_BN * c0), 1u);
							if(cnt > 4)
								tc_mma_i8(tmem + (uint32_t) (TC_BN * (l + 4)), da, tc_desc(b_lo + (uint32_t) (4 * (TC_B_BYTES >> 4) + k4 * 2)),
									tc_idesc(TC_BN * (cnt - 4)), 1u);
						}
					}
				}
				tc_commit(&bar_empty[s]);          // frees the smem stage once these MMAs have read it
				if(++s == S) { s = 0; ph ^= 1u; }
				if(++in_chunk == P.chunk_blks || b + 1 == nb) {
					tc_commit(&bar_tfull);         // accumulators of this chunk are complete
					chunk++;
					in_chunk = 0;
				}
			}
		}
	} else {
		// ===== epilogue: 4 warps, warp q = warp % 4 owns TMEM lanes [32q, 32q+32)
		const int q = warp & 3;
		const uint32_t tlane = tmem + ((uint32_t) (32 * q) << 16);
		uint64_t out[TC_BN];
#pragma unroll
		for(int c = 0; c < TC_BN; c++) out[c] = 0;
#pragma unroll
		for(int g = 0; g < 2 * G; g++) tmem_zero32(tlane + (uint32_t) (32 * g));
		tmem_wait_st();
		tc_fence_before();
		mbar_arrive(&bar_tempty);
		const int nchunks = (nb + P.chunk_blks - 1) / P.chunk_blks;
		for(int ch = 0; ch < nchunks; ch++) {
			mbar_wait(&bar_tfull, (uint32_t) ch & 1u);
			tc_fence_after();
#pragma unroll
			for(int s = 0; s < G; s++) {
#pragma unroll
				for(int h = 0; h < 2; h++) {
					uint32_t v[32];
					tmem_ld32(tlane + (uint32_t) (TC_BN * s + 32 * h), v);
#pragma unroll
					for(int c = 0; c < 32; c++) out[32 * h + c] += (uint64_t) (int64_t) (int32_t) v[c] << (8 * s);
				}
			}
			if(ch + 1 < nchunks) {
#pragma unroll
				for(int g = 0; g < 2 * G; g++) tmem_zero32(tlane + (uint32_t) (32 * g));
				tmem_wait_st();
				tc_fence_before();
				mbar_arrive(&bar_tempty);
			}
		}
		// row of the output tile = TMEM lane
		const int a = ta * TC_BM + 32 * q + lane;
		if(a < P.gs.m) {
#pragma unroll
			for(int c = 0; c < TC_BN; c++) {
				int b = tb * TC_BN + c;
				if(b < a && b < P.gs.mc) p1_red_add(P.acc + tc_slot(P.gs, a, b), out[c]);
			}
		}
	}
	tc_fence_before();
	__syncthreads();
	if(warp == 1) {
		tc_fence_after();
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(P.tmem_cols) : "memory");
	}
}

// ------------------------------------------------------------------ one-digit blocks of up to 256 columns
//
// A block whose words fit ONE digit (|word| <= 127: the 32-bit configurations, where a value is ~10 units of the
// scale) and that has at most 256 columns is multiplied whole by every CTA: a pipeline stage is the block's
// 128 rows x cols digit tile, loaded ONCE, and serves as both operands - B = all columns (N = cols), A = its
// first / second 128 columns (two MMAs per K-step into TMEM columns [0,128) and [128,384)).  The tiled kernel
// above moves every digit through L2 once per 128 x 64 tile that touches it (6 tiles at 201 columns: 5.7x the
// bytes of the planes, which capped it at 25 % tensor-pipe activity); here every digit crosses L2 -> SM once and
// the kernel runs at the rate HBM delivers the planes.  One s32 chunk holds 131071 rows, so a CTA's row slice is
// normally a single chunk: the epilogue adds TMEM straight into the packed share (red.global.add.u64).
struct TsParams {
	TcSet gs;
	int stages;
	uint32_t blk_bytes;    // cols * 128: one K-block of the block's digit tile
	int kpb;               // K-blocks per pipeline stage (narrow blocks: several, so that a stage is worth a TMA)
	int n0, n1;            // N of the two MMAs (0: no second row tile)
	int nblk, blk_per_cta, chunk_blks;   // in stages (groups of kpb K-blocks)
	uint64_t *acc;
};

__global__ void __launch_bounds__(TC_THREADS, 1) k_gram_whole(const __grid_constant__ CUtensorMap map, const TsParams P) {
	extern __shared__ __align__(1024) uint8_t smem_raw[];
	__shared__ __align__(8) uint64_t bar_full[8], bar_empty[8], bar_tfull, bar_tempty;
	__shared__ uint32_t tmem_base_s;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int S = P.stages;
	uint8_t *smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
	const int blk_lo = (int) blockIdx.x * P.blk_per_cta;
	const int blk_hi = blk_lo + P.blk_per_cta < P.nblk ? blk_lo + P.blk_per_cta : P.nblk;
	const int nb = blk_hi - blk_lo;          // > 0 by construction of the grid
	const uint32_t stage_bytes = P.blk_bytes * (uint32_t) P.kpb;
	if(threadIdx.x == 0) {
		for(int s = 0; s < S; s++) {
			mbar_init(&bar_full[s], 1);
			mbar_init(&bar_empty[s], 1);
		}
		mbar_init(&bar_tfull, 1);
		mbar_init(&bar_tempty, 128);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map)) : "memory");
	}
	if(warp == 1) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = tmem_base_s;

	if(warp == 0) {
		if(lane == 0) {
			int s = 0;
			uint32_t ph = 0;
			for(int b = 0; b < nb; b++) {
				mbar_wait(&bar_empty[s], ph ^ 1u);
				mbar_expect_tx(&bar_full[s], stage_bytes);
				tma_load_4d(smem + (size_t) s * stage_bytes, &map, &bar_full[s], 0, 0, 0, (blk_lo + b) * P.kpb);
				if(++s == S) { s = 0; ph ^= 1u; }
			}
		}
	} else if(warp == 1) {
		if(lane == 0) {
			const uint32_t lo0 = ((smem_u32(smem) >> 4) & 0x3fffu) | (1u << 16);
			const uint32_t id0 = tc_idesc(P.n0), id1 = tc_idesc(P.n1 ? P.n1 : 16);
			int s = 0, in_chunk = 0;
			uint32_t ph = 0, chunk = 0;
			for(int b = 0; b < nb; b++) {
				if(in_chunk == 0) {
					mbar_wait(&bar_tempty, chunk & 1u);     // the epilogue has read the previous chunk out of TMEM
					tc_fence_after();
				}
				mbar_wait(&bar_full[s], ph);
				tc_fence_after();
				uint32_t t_lo = lo0 + (uint32_t) s * (stage_bytes >> 4);
				for(int kb = 0; kb < P.kpb; kb++, t_lo += P.blk_bytes >> 4) {
#pragma unroll
					for(int k4 = 0; k4 < TC_BK / 32; k4++) {
						const uint32_t keep = (in_chunk | kb | k4) ? 1u : 0u;  // the first K-step of a chunk overwrites the accumulators
						const uint64_t db = tc_desc(t_lo + (uint32_t) (k4 * 2));
						tc_mma_i8(tmem, db, db, id0, keep);
						if(P.n1) tc_mma_i8(tmem + 128u, tc_desc(t_lo + (uint32_t) ((TC_BM * TC_BK >> 4) + k4 * 2)), db, id1, keep);
					}
				}
				tc_commit(&bar_empty[s]);
				if(++s == S) { s = 0; ph ^= 1u; }
				if(++in_chunk == P.chunk_blks || b + 1 == nb) {
					tc_commit(&bar_tfull);
					chunk++;
					in_chunk = 0;
				}
			}
		}
	} else {
		// epilogue: warp q owns TMEM lanes [32q, 32q+32) = output rows 32q+lane (and 128+32q+lane)
		const int q = warp & 3;
		const uint32_t tlane = tmem + ((uint32_t) (32 * q) << 16);
		mbar_arrive(&bar_tempty);
		const int nchunks = (nb + P.chunk_blks - 1) / P.chunk_blks;
		for(int ch = 0; ch < nchunks; ch++) {
			mbar_wait(&bar_tfull, (uint32_t) ch & 1u);
			tc_fence_after();
			for(int ta = 0; ta < (P.n1 ? 2 : 1); ta++) {
				const int a = TC_BM * ta + 32 * q + lane;
				const int ncol = ta ? P.n1 : P.n0;
				// columns b < a only: groups past the warp's last row are all upper triangle
				const int gmax = (TC_BM * ta + 32 * q + 31) / 32;
				for(int g = 0; g <= gmax && 32 * g < ncol; g++) {
					uint32_t v[32];
					tmem_ld32(tlane + (uint32_t) (128 * ta + 32 * g), v);
					if(a < P.gs.m) {
#pragma unroll
						for(int c = 0; c < 32; c++) {
							const int b = 32 * g + c;
							if(b < a && b < P.gs.mc && v[c]) p1_red_add(P.acc + tc_slot(P.gs, a, b), (uint64_t) (int64_t) (int32_t) v[c]);
						}
					}
				}
			}
			if(ch + 1 < nchunks) {
				tc_fence_before();
				mbar_arrive(&bar_tempty);
			}
		}
	}
	tc_fence_before();
	__syncthreads();
	if(warp == 1) {
		tc_fence_after();
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
	}
}

typedef void (*tc_kernel_t)(const CUtensorMap, const CUtensorMap, const TcParams);
template <int WD> static tc_kernel_t tc_kernel_for(int L) {
	switch(L) {
	case 1: return k_gram_tc<1, WD>;
	case 2: return k_gram_tc<2, WD>;
	case 3: return k_gram_tc<3, WD>;
	case 4: return k_gram_tc<4, WD>;
	case 5: return WD == 8 ? k_gram_tc<5, 8> : NULL;
	case 6: return WD == 8 ? k_gram_tc<6, 8> : NULL;
	case 7: return WD == 8 ? k_gram_tc<7, 8> : NULL;
	case 8: return WD == 8 ? k_gram_tc<8, 8> : NULL;
	}
	return NULL;
}

// ------------------------------------------------------------------ host side

typedef CUresult (*tmap_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
	const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static tmap_encode_fn get_tmap_encode() {
	static tmap_encode_fn fn = NULL;
	if(!fn) {
		void *p = NULL;
		cudaDriverEntryPointQueryResult q;
		if(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
			fn = (tmap_encode_fn) p;
	}
	return fn;
}

// digits needed so that every |word| <= maxabs has a balanced base-256 expansion:
// L digits cover [-128*(256^L-1)/255, 127*(256^L-1)/255]; L = ring bytes always works (mod 2^W)
int p1_tc_digits(uint64_t maxabs, int ring_bytes) {
	unsigned __int128 span = 0;
	for(int L = 1; L < ring_bytes; L++) {
		span = span * 256 + 1;                       // (256^L - 1) / 255
		if((unsigned __int128) maxabs <= 127 * span) return L;
	}
	return ring_bytes;
}

// digits for the block [col_lo, col_hi) (+ target): from the per-column max |word| the encode pass left in data->stats
static int tc_block_digits(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, int *L) {
	std::vector<uint64_t> st((data->d + 1) * 4);
	P1_CUDA(cudaMemcpyAsync(st.data(), data->stats, st.size() * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	uint64_t maxabs = 0;
	for(size_t c = col_lo; c < col_hi; c++) if(st[4 * c + 3] > maxabs) maxabs = st[4 * c + 3];
	if(owns_target && st[4 * data->d + 3] > maxabs) maxabs = st[4 * data->d + 3];
	*L = p1_tc_digits(maxabs, data->bits / 8);
	return 0;
}

extern "C" int p1_gram_digits(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, int *digits) {
	P1_CHECK(ctx && data && digits, "p1_gram_digits: arguments may not be null");
	std::lock_guard<std::mutex> lock_(ctx->mu);
	P1_CUDA(cudaSetDevice(ctx->device));
	P1_CHECK(col_lo <= col_hi && col_hi <= data->d, "p1_gram_digits: invalid column range");
	return tc_block_digits(ctx, data, col_lo, col_hi, owns_target, digits);
}

int p1_launch_gram_tc(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, uint64_t *acc_dev) {
	tmap_encode_fn encode = get_tmap_encode();
	P1_CHECK(encode, "p1_local_block: cuTensorMapEncodeTiled is not available from this driver");
	TcSet gs;
	gs.col_lo = col_lo;
	gs.mc = (int) (col_hi - col_lo);
	gs.m = gs.mc + (owns_target ? 1 : 0);
	gs.d = data->d;
	gs.nA = p1_nA(data->d);
	const int WD = data->bits / 8;

	int L = 0;
	if(tc_block_digits(ctx, data, col_lo, col_hi, owns_target, &L)) return 1;
	const int G = 2 * L - 1 < WD ? 2 * L - 1 : WD;

	// digit planes P[kblock][l][a][128 rows], rows padded to a whole K-block
	const size_t ldp = p1_round_up(data->ld, TC_BK);
	// The encode pass may have left the planes of every column already (p1_api.cu, option "encode_planes"):
	// usable when they hold enough digits and the set is contiguous in the store's column order.
	const bool pre = data->planes && L <= data->Lg && ldp == data->ld && (!owns_target || col_hi == data->d);
	uint8_t *planes;
	size_t col_stride, blk_stride;              // bytes between digits of a K-block, between K-blocks
	if(pre) {
		planes = data->planes + col_lo * TC_BK;
		col_stride = (data->d + 1) * TC_BK;
		blk_stride = col_stride * data->Lg;
	} else {
		const size_t need = (size_t) gs.m * ldp * L;
		if(ctx->limb_ws_bytes < need) {
			if(ctx->limb_ws) {
				P1_CUDA(cudaDeviceSynchronize());
				p1_pool_dev_give(ctx->device, ctx->limb_ws, ctx->limb_ws_bytes);
				ctx->limb_ws_bytes = 0;
			}
			ctx->limb_ws = p1_pool_dev_take(ctx->device, need);
			P1_CHECK(ctx->limb_ws, "p1_local_block: cannot allocate %zu bytes of digit planes", need);
			ctx->limb_ws_bytes = need;
		}
		planes = (uint8_t *) ctx->limb_ws;
		col_stride = (size_t) gs.m * TC_BK;
		blk_stride = col_stride * L;
		dim3 grid((unsigned) ((ldp + 1023) / 1024), (unsigned) gs.m);
		p1_prof_begin(ctx, P1_K_GRAM_LIMBS);
		if(data->bits == 64) k_limb_split<uint64_t><<<grid, 256, 0, ctx->stream>>>((const uint64_t *) data->xt, data->ld, gs, L, planes, ldp);
		else k_limb_split<uint32_t><<<grid, 256, 0, ctx->stream>>>((const uint32_t *) data->xt, data->ld, gs, L, planes, ldp);
		P1_LAUNCHED(ctx);
	}

	// tensor maps over the planes: dims (row in block, column, limb, block), u8, 128-byte swizzle
	CUtensorMap mapA, mapB;
	// (declared as 32-bit elements: the swizzle and the bytes moved are the same as for u8)
	const int esz = 4;
	const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_UINT32;
	cuuint64_t gdim[4] = {(cuuint64_t) (TC_BK / esz), (cuuint64_t) gs.m, (cuuint64_t) L, (cuuint64_t) (ldp / TC_BK)};
	cuuint64_t gstr[3] = {TC_BK, (cuuint64_t) col_stride, (cuuint64_t) blk_stride};
	cuuint32_t estr[4] = {1, 1, 1, 1};
	cuuint32_t boxA[4] = {(cuuint32_t) (TC_BK / esz), TC_BM, (cuuint32_t) L, 1}, boxB[4] = {(cuuint32_t) (TC_BK / esz), TC_BN, (cuuint32_t) L, 1};
	CUresult r1 = encode(&mapA, dt, 4, planes, gdim, gstr, boxA, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
		CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	CUresult r2 = encode(&mapB, dt, 4, planes, gdim, gstr, boxB, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
		CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	P1_CHECK(r1 == CUDA_SUCCESS && r2 == CUDA_SUCCESS, "p1_local_block: cuTensorMapEncodeTiled failed (%d, %d)", (int) r1, (int) r2);

	if(L == 1 && gs.m <= 256 && ctx->opt_tc_whole) {
		// one-digit block of at most 256 columns: every CTA multiplies the whole block over its slice of the rows
		const int cols = (gs.m + 15) / 16 * 16;
		CUtensorMap map;
		cuuint32_t box[4] = {(cuuint32_t) (TC_BK / esz), (cuuint32_t) cols, 1, 1};
		TsParams Q;
		Q.gs = gs;
		Q.blk_bytes = (uint32_t) cols * TC_BK;
		// narrow blocks: several K-blocks per stage (one TMA box of up to 24 KB)
		Q.kpb = (int) ((size_t) 24576 / Q.blk_bytes);
		if(Q.kpb < 1) Q.kpb = 1;
		if(Q.kpb > 8) Q.kpb = 8;
		const int nblk = (int) ((data->n + TC_BK - 1) / TC_BK);
		while(Q.kpb > 1 && (nblk + Q.kpb - 1) / Q.kpb < ctx->sm_count) Q.kpb--;     // few rows: rather one stage per SM
		box[3] = (cuuint32_t) Q.kpb;
		CUresult r = encode(&map, dt, 4, planes, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
			CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
		P1_CHECK(r == CUDA_SUCCESS, "p1_local_block: cuTensorMapEncodeTiled failed (%d)", (int) r);
		const size_t stage_bytes = (size_t) Q.blk_bytes * Q.kpb;
		Q.n0 = cols < TC_BM ? cols : TC_BM;
		Q.n1 = cols > TC_BM ? cols : 0;
		// an A operand is always 128 rows: when the block has fewer than 128 (256) columns the last rows of the
		// (second) row tile lie past the K-block, for the last K-block of the last stage past the stage (their
		// products land in output rows >= m, which nobody reads) - the allocation covers them
		const size_t a_end = (size_t) (Q.n1 ? 2 : 1) * TC_BM * TC_BK;
		const size_t overrun = a_end > Q.blk_bytes ? a_end - Q.blk_bytes : 0;
		int stages = (int) (((size_t) 216 * 1024 - overrun) / stage_bytes);
		Q.stages = stages > 8 ? 8 : stages;
		Q.nblk = (nblk + Q.kpb - 1) / Q.kpb;
		Q.chunk_blks = 131071 / (TC_BK * Q.kpb);
		if(ctx->opt_tc_chunk_blocks > 0 && ctx->opt_tc_chunk_blocks < Q.chunk_blks) Q.chunk_blks = (int) ctx->opt_tc_chunk_blocks;
		int ctas = ctx->sm_count;
		if(ctx->opt_tc_max_slices > 0 && ctas > ctx->opt_tc_max_slices) ctas = (int) ctx->opt_tc_max_slices;
		if(ctas > Q.nblk) ctas = Q.nblk;
		Q.blk_per_cta = (Q.nblk + ctas - 1) / ctas;
		ctas = (Q.nblk + Q.blk_per_cta - 1) / Q.blk_per_cta;
		Q.acc = acc_dev;
		const size_t smem = (size_t) Q.stages * stage_bytes + overrun + 1024;
		P1_CUDA(cudaFuncSetAttribute(k_gram_whole, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		p1_prof_begin(ctx, P1_K_GRAM_TC);
		k_gram_whole<<<(unsigned) ctas, TC_THREADS, smem, ctx->stream>>>(map, Q);
		P1_LAUNCHED(ctx);
		return 0;
	}

	TcParams P;
	P.gs = gs;
	const size_t stage_bytes = (size_t) L * (TC_A_BYTES + TC_B_BYTES);
	const size_t smem_budget = 220 * 1024;
	int stages = (int) (smem_budget / stage_bytes);
	if(stages > 6) stages = 6;
	P1_CHECK(stages >= 1, "p1_local_block: a pipeline stage of %zu bytes does not fit shared memory", stage_bytes);
	P.stages = stages;
	P.tmem_cols = 64;
	while(P.tmem_cols < (uint32_t) (TC_BN * G)) P.tmem_cols *= 2;
	P.nblk = (int) ((data->n + TC_BK - 1) / TC_BK);
	P.chunk_blks = (131071 / L) / TC_BK;             // rows per chunk * L * 2^14 < 2^31
	if(ctx->opt_tc_chunk_blocks > 0 && ctx->opt_tc_chunk_blocks < P.chunk_blks) P.chunk_blks = (int) ctx->opt_tc_chunk_blocks;
	// tiles of the lower triangle (incl. the target row as the last row of the set)
	int tiles = 0;
	for(int ta = 0; ta * TC_BM < gs.m; ta++) {
		int lim = TC_BM * (ta + 1) < gs.mc ? TC_BM * (ta + 1) : gs.mc;
		tiles += (lim + TC_BN - 1) / TC_BN;
	}
	int slices = ctx->sm_count / tiles;
	if(ctx->opt_tc_max_slices > 0 && slices > ctx->opt_tc_max_slices) slices = (int) ctx->opt_tc_max_slices;
	if(slices < 1) slices = 1;
	if(slices > P.nblk) slices = P.nblk;
	P.blk_per_cta = (P.nblk + slices - 1) / slices;
	slices = (P.nblk + P.blk_per_cta - 1) / P.blk_per_cta;
	P.acc = acc_dev;
	const size_t smem = (size_t) stages * stage_bytes + 1024;
	tc_kernel_t kern = WD == 8 ? tc_kernel_for<8>(L) : tc_kernel_for<4>(L);
	P1_CHECK(kern, "p1_local_block: no tcgen05 kernel for %d digits in a %d-byte ring", L, WD);
	P1_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	dim3 grid((unsigned) tiles, (unsigned) slices);
	p1_prof_begin(ctx, P1_K_GRAM_TC);
	kern<<<grid, TC_THREADS, smem, ctx->stream>>>(mapA, mapB, P);
	P1_LAUNCHED(ctx);
	return 0;
}
