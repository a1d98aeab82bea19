// Encode + layout kernels (HBM-bound streaming).
//   k_rows_to_cols : row-major doubles (or ring words) -> column-major ring words,
//                    fixed-point encode fused (src/linear.c:46-51, src/fixed.c:3-5) and
//                    per-column statistics gathered in the same pass.
//   k_vec_encode   : same for the target vector (src/linear.c:83-87).
//   k_rows_to_cols_planes / k_vec_encode_planes : the same passes, also writing the digit planes
//                    of the tcgen05 Gram (option encode_planes).
//   k_cols_to_rows : column-major -> the reference's row-major layout (tests / download).
#include "p1_common.cuh"
#include <type_traits>

// (fixed_t) cast with x86 cvttsd2si semantics: truncate toward zero, INT_MIN of the
// target width when out of range or NaN ("integer indefinite").  CUDA's cvt.rzi
// saturates and maps NaN to 0, so the range test is explicit.
template <int BITS>
__device__ __forceinline__ int64_t cast_x86(double v) {
	if(BITS == 64) {
		if(!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) return INT64_MIN;
		return __double2ll_rz(v);
	} else {
		if(!(v > -2147483649.0 && v < 2147483648.0)) return (int64_t) INT32_MIN;
		return (int64_t) __double2int_rz(v);
	}
}

struct EncodeParams {
	double normalizer;
	double scale;   // (double)(1ll << precision)
	int bits_p2;
};

template <typename W, typename SRC>
__device__ __forceinline__ W load_word(SRC v, const EncodeParams &ep);

template <> __device__ __forceinline__ uint64_t load_word<uint64_t, double>(double v, const EncodeParams &ep) {
	double t = __dmul_rn(__ddiv_rn(v, ep.normalizer), ep.scale); // val /= normalizer; val * (1ll<<p)
	return (uint64_t) (ep.bits_p2 == 64 ? cast_x86<64>(t) : cast_x86<32>(t)); // sign-extends the 32-bit fixed_t
}
template <> __device__ __forceinline__ uint32_t load_word<uint32_t, double>(double v, const EncodeParams &ep) {
	double t = __dmul_rn(__ddiv_rn(v, ep.normalizer), ep.scale);
	return (uint32_t) cast_x86<32>(t);
}
template <> __device__ __forceinline__ uint64_t load_word<uint64_t, uint64_t>(uint64_t v, const EncodeParams &) { return v; }
template <> __device__ __forceinline__ uint32_t load_word<uint32_t, uint32_t>(uint32_t v, const EncodeParams &) { return v; }

// Per-column statistics of one thread.  sq = sum of (low 32 bits of |word|)^2: the exact sum of
// squares whenever no |word| reaches 2^26 (viol == 0), which is the only case its consumers
// (p1_finalize's exact diagonal) use it; it is flushed as two 32-bit halves so that the
// 64-bit accumulator entries cannot wrap when row shards and CTAs are summed.
struct ColStats {
	uint64_t sq, mx;
	uint32_t viol;
	__device__ __forceinline__ void init() { sq = mx = 0; viol = 0; }
	template <typename W> __device__ __forceinline__ void add(W word) {
		typedef typename p1_word<W>::s S;
		const int64_t s = (int64_t) (S) word;
		const uint64_t a = s < 0 ? (uint64_t) 0 - (uint64_t) s : (uint64_t) s;
		const uint32_t a32 = (uint32_t) a;
		sq += (uint64_t) a32 * a32;                     // one IMAD.WIDE.U32 with accumulate
		viol += (a >> P1_SQ_LIMIT_BITS) ? 1u : 0u;
		mx = a > mx ? a : mx;
	}
	__device__ __forceinline__ void warp_reduce() {
		sq = p1_warp_sum(sq);
		viol = __reduce_add_sync(0xffffffffu, viol);
#pragma unroll
		for(int o = 16; o; o >>= 1) {
			uint64_t t = __shfl_xor_sync(0xffffffffu, mx, o);
			mx = t > mx ? t : mx;
		}
	}
	__device__ __forceinline__ void flush(uint64_t *st) {
		p1_red_add(st + 0, sq & 0xffffffffull);
		p1_red_add(st + 1, sq >> 32);
		if(viol) p1_red_add(st + 2, (uint64_t) viol);
		atomicMax(reinterpret_cast<unsigned long long *>(st + 3), (unsigned long long) mx);
	}
};

#define ENC_ROWS_PER_CTA 512
#define ENC_TILE_ROWS 128     // rows per pass: 16 loads per thread in flight (32 KB per CTA)

// One CTA: ENC_ROWS_PER_CTA rows x 32 columns.  A pass moves a 128-row x 32-column tile: every
// thread issues its 16 source loads before the first use, and the loads of the next pass are
// issued before the transposed stores of the current one (software pipelining through
// registers), so ~32 KB per CTA stay in flight.  FULL = the CTA's rows and columns all exist:
// no bounds tests, addresses by pointer increments (the integer pipe, not HBM, was the limit
// with per-element tests and 64-bit index products: profiles/r1g_new_kernels_full.txt).
template <typename W, typename SRC, bool FULL>
__device__ __forceinline__ void enc_cta(const SRC *__restrict__ src, size_t rows, size_t d, size_t row_off,
	W *__restrict__ xt, size_t ld, uint64_t *__restrict__ stats, const EncodeParams &ep, W (*tile)[33]) {
	constexpr int NL = ENC_TILE_ROWS / 8;        // loads per thread per pass
	constexpr int NSUB = ENC_ROWS_PER_CTA / ENC_TILE_ROWS;
	const int tx = threadIdx.x, ty = threadIdx.y;
	const size_t c0 = (size_t) blockIdx.x * 32;
	const size_t rbase = (size_t) blockIdx.y * ENC_ROWS_PER_CTA;
	const size_t c_in = c0 + tx;
	ColStats cs[4];
#pragma unroll
	for(int u = 0; u < 4; u++) cs[u].init();
	SRC v[NL];
	const size_t step = 8 * d;
	const SRC *ip = src + (rbase + ty) * d + c_in;       // next element this thread loads
	auto fetch = [&](size_t r0) {
#pragma unroll
		for(int u = 0; u < NL; u++) {
			if(FULL) v[u] = __ldcs(ip);
			else v[u] = (r0 + ty + 8 * u < rows && c_in < d) ? __ldcs(ip) : (SRC) 0;
			ip += step;
		}
	};
	fetch(rbase);
	W *op[4];
#pragma unroll
	for(int u = 0; u < 4; u++) op[u] = xt + (c0 + ty + 8 * u) * ld + row_off + rbase + tx;
#pragma unroll 1
	for(int sub = 0; sub < NSUB; sub++) {
		const size_t r0 = rbase + (size_t) sub * ENC_TILE_ROWS;
		if(!FULL && r0 >= rows) break;
#pragma unroll
		for(int u = 0; u < NL; u++) {
			if(FULL) tile[ty + 8 * u][tx] = load_word<W, SRC>(v[u], ep);
			else tile[ty + 8 * u][tx] = (r0 + ty + 8 * u < rows && c_in < d) ? load_word<W, SRC>(v[u], ep) : (W) 0;
		}
		__syncthreads();
		if(sub + 1 < NSUB && (FULL || r0 + ENC_TILE_ROWS < rows)) fetch(r0 + ENC_TILE_ROWS);
#pragma unroll
		for(int u = 0; u < 4; u++) {
			const int cc = ty + 8 * u;
			if(FULL || c0 + cc < d) {
#pragma unroll
				for(int j = 0; j < ENC_TILE_ROWS / 32; j++) {
					const int rr = tx + 32 * j;
					if(FULL || r0 + rr < rows) {
						const W wv = tile[rr][cc];
						op[u][32 * j] = wv;
						cs[u].add<W>(wv);
					}
				}
			}
			op[u] += ENC_TILE_ROWS;
		}
		__syncthreads();
	}
#pragma unroll
	for(int u = 0; u < 4; u++) {
		cs[u].warp_reduce();
		const size_t c = c0 + ty + 8 * u;
		if(tx == 0 && c < d) cs[u].flush(stats + 4 * c);
	}
}

// grid (ceil(d/32), ceil(rows/ENC_ROWS_PER_CTA)), block (32,8)
template <typename W, typename SRC>
// resident CTAs per SM: 3 with 4-byte words (80 registers, a few spills, +10 % measured), 2 with 8-byte words
// (3 gained nothing there: 1.59 vs 1.58 ms)
__global__ void __launch_bounds__(256, (sizeof(W) == 4 ? 3 : 2)) k_rows_to_cols(const SRC *__restrict__ src, size_t rows, size_t d, size_t row_off,
	W *__restrict__ xt, size_t ld, uint64_t *__restrict__ stats, EncodeParams ep) {
	__shared__ W tile[ENC_TILE_ROWS][33];
	const bool full = (size_t) blockIdx.x * 32 + 32 <= d && (size_t) blockIdx.y * ENC_ROWS_PER_CTA + ENC_ROWS_PER_CTA <= rows;
	if(full) enc_cta<W, SRC, true>(src, rows, d, row_off, xt, ld, stats, ep, tile);
	else enc_cta<W, SRC, false>(src, rows, d, row_off, xt, ld, stats, ep, tile);
}

// ------------------------------------------------------------------ encode + digit planes
//
// Same pass as k_rows_to_cols, also writing the balanced base-256 digits of every word in the layout the
// tcgen05 Gram reads through TMA (k_gram_tc.cu): planes[row/128][digit][column 0..d][row%128], Lg digits per
// word.  That removes k_limb_split's re-read of the whole column store before every Gram.  A thread owns
// 4 consecutive rows of a column here (one 16/32-byte word store, one packed 32-bit store per digit), so the
// tile is read back along rows: the XOR swizzle keeps both the row-wise writes and these reads conflict-free.

__device__ __forceinline__ int enc_swz(int r, int c) { return r * 32 + (c ^ ((r >> 2) & 31)); }

template <typename W>
__device__ __forceinline__ void enc_store4(W *dst, const W (&w)[4]) {
	if(sizeof(W) == 8) {
		*reinterpret_cast<ulonglong2 *>(dst) = make_ulonglong2((uint64_t) w[0], (uint64_t) w[1]);
		*reinterpret_cast<ulonglong2 *>(dst + 2) = make_ulonglong2((uint64_t) w[2], (uint64_t) w[3]);
	} else {
		*reinterpret_cast<uint4 *>(dst) = make_uint4((uint32_t) w[0], (uint32_t) w[1], (uint32_t) w[2], (uint32_t) w[3]);
	}
}

// Digits of 4 words, digit l of all four packed into one 32-bit store at dst + l*stride (Lg <= 4).
// The balanced base-256 digits d_l in [-128,127] of v mod 2^32 are the bytes of (v + 0x80808080) with bit 7
// flipped back: v + sum 128*256^l = sum b_l 256^l  =>  d_l = b_l - 128 = (s8)(b_l ^ 0x80).  Only the low 32 bits
// of a word matter for the first four digits; a word that does not fit Lg digits makes the Gram split again
// (its digit count comes from the max |word| statistic), so the planes are never used for it.
template <typename W>
__device__ __forceinline__ void enc_digits4(const W (&w)[4], uint8_t *dst, size_t stride, int Lg) {
	uint32_t t[4];
#pragma unroll
	for(int i = 0; i < 4; i++) t[i] = ((uint32_t) w[i] + 0x80808080u) ^ 0x80808080u;
	// 4x4 byte transpose: word i byte l -> digit l byte i
	const uint32_t a = __byte_perm(t[0], t[1], 0x5140), b = __byte_perm(t[2], t[3], 0x5140);
	const uint32_t c = __byte_perm(t[0], t[1], 0x7362), d = __byte_perm(t[2], t[3], 0x7362);
	*reinterpret_cast<uint32_t *>(dst) = __byte_perm(a, b, 0x5410);
	if(Lg > 1) *reinterpret_cast<uint32_t *>(dst + stride) = __byte_perm(a, b, 0x7632);
	if(Lg > 2) *reinterpret_cast<uint32_t *>(dst + 2 * stride) = __byte_perm(c, d, 0x5410);
	if(Lg > 3) *reinterpret_cast<uint32_t *>(dst + 3 * stride) = __byte_perm(c, d, 0x7632);
}

template <typename W, bool FULL>
__device__ __forceinline__ void enc_cta_planes(const double *__restrict__ src, size_t rows, size_t d, size_t row_off,
	W *__restrict__ xt, size_t ld, uint64_t *__restrict__ stats, const EncodeParams &ep, W *tile, uint8_t *__restrict__ planes, int Lg) {
	constexpr int NL = ENC_TILE_ROWS / 8;
	constexpr int NSUB = ENC_ROWS_PER_CTA / ENC_TILE_ROWS;
	const int tx = threadIdx.x, ty = threadIdx.y;
	const size_t c0 = (size_t) blockIdx.x * 32;
	const size_t rbase = (size_t) blockIdx.y * ENC_ROWS_PER_CTA;
	const size_t c_in = c0 + tx;
	ColStats cs[4];
#pragma unroll
	for(int u = 0; u < 4; u++) cs[u].init();
	double v[NL];
	const size_t step = 8 * d;
	const double *ip = src + (rbase + ty) * d + c_in;
	auto fetch = [&](size_t r0) {
#pragma unroll
		for(int u = 0; u < NL; u++) {
			if(FULL) v[u] = __ldcs(ip);
			else v[u] = (r0 + ty + 8 * u < rows && c_in < d) ? __ldcs(ip) : 0.0;
			ip += step;
		}
	};
	fetch(rbase);
	const size_t pcol = (d + 1) * 128;                    // bytes between digits of a K-block
	W *op = xt + (c0 + ty) * ld + row_off + rbase + 4 * tx;
	uint8_t *pp = planes + ((row_off + rbase) / 128) * (size_t) Lg * pcol + (c0 + ty) * 128 + 4 * tx;
#pragma unroll 1
	for(int sub = 0; sub < NSUB; sub++) {
		const size_t r0 = rbase + (size_t) sub * ENC_TILE_ROWS;
		if(!FULL && r0 >= rows) break;
#pragma unroll
		for(int u = 0; u < NL; u++) {
			const int rr = ty + 8 * u;
			if(FULL) tile[enc_swz(rr, tx)] = load_word<W, double>(v[u], ep);
			else tile[enc_swz(rr, tx)] = (r0 + rr < rows && c_in < d) ? load_word<W, double>(v[u], ep) : (W) 0;
		}
		__syncthreads();
		if(sub + 1 < NSUB && (FULL || r0 + ENC_TILE_ROWS < rows)) fetch(r0 + ENC_TILE_ROWS);
#pragma unroll
		for(int u = 0; u < 4; u++) {
			const int cc = ty + 8 * u;
			if(FULL || (c0 + cc < d && r0 + 4 * tx < rows)) {
				W w[4];
#pragma unroll
				for(int i = 0; i < 4; i++) w[i] = tile[(4 * tx + i) * 32 + (cc ^ tx)];       // enc_swz(4tx+i, cc)
				W *dst = op + (size_t) (8 * u) * ld;
				if(FULL || r0 + 4 * tx + 4 <= rows) {
					enc_store4<W>(dst, w);
#pragma unroll
					for(int i = 0; i < 4; i++) cs[u].add<W>(w[i]);
				} else {
#pragma unroll
					for(int i = 0; i < 4; i++) if(r0 + 4 * tx + i < rows) { dst[i] = w[i]; cs[u].add<W>(w[i]); }
				}
				enc_digits4<W>(w, pp + (size_t) (8 * u) * 128, pcol, Lg);
			}
		}
		op += ENC_TILE_ROWS;
		pp += (size_t) Lg * pcol;
		__syncthreads();
	}
#pragma unroll
	for(int u = 0; u < 4; u++) {
		cs[u].warp_reduce();
		const size_t c = c0 + ty + 8 * u;
		if(tx == 0 && c < d) cs[u].flush(stats + 4 * c);
	}
}

// grid (ceil(d/32), ceil(rows/ENC_ROWS_PER_CTA)), block (32,8); row_off must be a multiple of 128
template <typename W>
__global__ void __launch_bounds__(256, 2) k_rows_to_cols_planes(const double *__restrict__ src, size_t rows, size_t d, size_t row_off,
	W *__restrict__ xt, size_t ld, uint64_t *__restrict__ stats, EncodeParams ep, uint8_t *__restrict__ planes, int Lg) {
	__shared__ __align__(16) W tile[ENC_TILE_ROWS * 32];
	const bool full = (size_t) blockIdx.x * 32 + 32 <= d && (size_t) blockIdx.y * ENC_ROWS_PER_CTA + ENC_ROWS_PER_CTA <= rows;
	if(full) enc_cta_planes<W, true>(src, rows, d, row_off, xt, ld, stats, ep, tile, planes, Lg);
	else enc_cta_planes<W, false>(src, rows, d, row_off, xt, ld, stats, ep, tile, planes, Lg);
}

// target vector: a thread owns 4 consecutive rows; pl = the planes of column d
template <typename W>
__global__ void __launch_bounds__(256) k_vec_encode_planes(const double *__restrict__ src, size_t n, W *__restrict__ col,
	uint64_t *__restrict__ stats, EncodeParams ep, uint8_t *__restrict__ pl, size_t pcol, int Lg) {
	ColStats cs;
	cs.init();
	for(size_t r = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * 4; r < n; r += (size_t) gridDim.x * blockDim.x * 4) {
		W w[4];
#pragma unroll
		for(int i = 0; i < 4; i++) w[i] = r + i < n ? load_word<W, double>(src[r + i], ep) : (W) 0;
		if(r + 4 <= n) {
			enc_store4<W>(col + r, w);
#pragma unroll
			for(int i = 0; i < 4; i++) cs.add<W>(w[i]);
		} else {
#pragma unroll
			for(int i = 0; i < 4; i++) if(r + i < n) { col[r + i] = w[i]; cs.add<W>(w[i]); }
		}
		enc_digits4<W>(w, pl + (r / 128) * (size_t) Lg * pcol + (r % 128), pcol, Lg);
	}
	cs.warp_reduce();
	if((threadIdx.x & 31) == 0) cs.flush(stats);
}

// ------------------------------------------------------------------ encode through an async-copy pipeline
//
// k_encode_async: the source tile goes global -> shared with 16-byte cp.async (no registers held by loads in
// flight), ENC_A_STAGES tiles of 64 rows x 32 columns per CTA in a ring, one __syncthreads per tile.  The
// 16-byte chunks of a tile row are XOR-swizzled with (row/4)%16, so that the transposed read - a half-warp
// per column, a lane per group of 4 consecutive rows, odd/even columns in the two halves - touches every bank
// pair exactly twice per 64-bit warp load (the minimum).  Each thread then owns 4 consecutive rows of a column:
// one 32/16-byte word store and, with PLANES, one packed 32-bit store per digit (see enc_digits4).
// Needs an even d and a 16-byte aligned source (every chunk is then 16-byte aligned and never straddles the
// last column); the other cases take the kernels above.
#define ENC_A_ROWS 64
#define ENC_A_STAGES 4
#define ENC_A_ROWS_PER_CTA 1024
#define ENC_A_STAGE_BYTES (ENC_A_ROWS * 32 * 8)

__device__ __forceinline__ void enc_cp_async16(uint32_t dst_smem, const void *src, bool valid) {
	const int sz = valid ? 16 : 0;                       // src-size 0: the 16 bytes are zero-filled
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(sz) : "memory");
}

template <typename W, bool PLANES>
__global__ void __launch_bounds__(256, 3) k_encode_async(const double *__restrict__ src, size_t rows, size_t d, size_t row_off,
	W *__restrict__ xt, size_t ld, uint64_t *__restrict__ stats, EncodeParams ep, uint8_t *__restrict__ planes, int Lg) {
	extern __shared__ __align__(128) unsigned char enc_smem[];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const size_t c0 = (size_t) blockIdx.x * 32;
	const size_t rbase = (size_t) blockIdx.y * ENC_A_ROWS_PER_CTA;
	const size_t rend = rbase + ENC_A_ROWS_PER_CTA < rows ? rbase + ENC_A_ROWS_PER_CTA : rows;
	const int ntiles = (int) ((rend - rbase + ENC_A_ROWS - 1) / ENC_A_ROWS);
	const uint32_t smem0 = (uint32_t) __cvta_generic_to_shared(enc_smem);

	// producer side: chunk id = tid + 256k -> row id/16, chunk id%16
	auto issue = [&](int t) {
		if(t < ntiles) {
			const size_t r0 = rbase + (size_t) t * ENC_A_ROWS;
			const uint32_t stage = smem0 + (uint32_t) (t % ENC_A_STAGES) * ENC_A_STAGE_BYTES;
#pragma unroll
			for(int k = 0; k < 4; k++) {
				const int id = tid + 256 * k, row = id >> 4, ch = id & 15;
				const size_t r = r0 + row, c = c0 + 2 * ch;
				const bool valid = r < rend && c < d;
				const double *g = valid ? src + r * d + c : src;
				enc_cp_async16(stage + row * 256 + ((ch ^ ((row >> 2) & 15)) << 4), g, valid);
			}
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
	};
#pragma unroll
	for(int t = 0; t < ENC_A_STAGES - 1; t++) issue(t);

	// consumer side: half h of the warp takes the odd/even column of a pair, lane g the rows 4g..4g+3
	const int h = lane >> 4, g = lane & 15;
	ColStats cs[2];
	cs[0].init();
	cs[1].init();
	const size_t pcol = (d + 1) * 128;
#pragma unroll 1
	for(int t = 0; t < ntiles; t++) {
		asm volatile("cp.async.wait_group %0;" ::"n"(ENC_A_STAGES - 2) : "memory");
		__syncthreads();
		issue(t + ENC_A_STAGES - 1);                     // refills the stage consumed in iteration t-1
		const size_t r0 = rbase + (size_t) t * ENC_A_ROWS;
		const unsigned char *stage = enc_smem + (size_t) (t % ENC_A_STAGES) * ENC_A_STAGE_BYTES;
		const size_t r4 = r0 + 4 * g;
		if(r4 < rend) {
#pragma unroll
			for(int u = 0; u < 2; u++) {
				const int cc = 2 * (warp + 8 * u) + h;
				const size_t c = c0 + cc;
				if(c < d) {
					W w[4];
#pragma unroll
					for(int i = 0; i < 4; i++) {
						const double v = *reinterpret_cast<const double *>(stage + (4 * g + i) * 256 + ((((cc >> 1) ^ g) << 4) | ((cc & 1) << 3)));
						w[i] = load_word<W, double>(v, ep);          // rows >= rend were zero-filled: word 0
					}
					W *dst = xt + c * ld + row_off + r4;
					if(r4 + 4 <= rend) {
						enc_store4<W>(dst, w);
#pragma unroll
						for(int i = 0; i < 4; i++) cs[u].add<W>(w[i]);
					} else {
#pragma unroll
						for(int i = 0; i < 4; i++) if(r4 + i < rend) { dst[i] = w[i]; cs[u].add<W>(w[i]); }
					}
					if(PLANES) {
						const size_t rg = row_off + r4;
						enc_digits4<W>(w, planes + (rg >> 7) * (size_t) Lg * pcol + c * 128 + (rg & 127), pcol, Lg);
					}
				}
			}
		}
	}
	asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
	for(int u = 0; u < 2; u++) {
		// reduce over the 16 lanes of the half (xor 8,4,2,1 stays inside it)
		cs[u].sq = cs[u].sq + __shfl_xor_sync(0xffffffffu, cs[u].sq, 8);
		cs[u].sq = cs[u].sq + __shfl_xor_sync(0xffffffffu, cs[u].sq, 4);
		cs[u].sq = cs[u].sq + __shfl_xor_sync(0xffffffffu, cs[u].sq, 2);
		cs[u].sq = cs[u].sq + __shfl_xor_sync(0xffffffffu, cs[u].sq, 1);
#pragma unroll
		for(int o = 8; o; o >>= 1) {
			cs[u].viol += __shfl_xor_sync(0xffffffffu, cs[u].viol, o);
			const uint64_t m = __shfl_xor_sync(0xffffffffu, cs[u].mx, o);
			cs[u].mx = m > cs[u].mx ? m : cs[u].mx;
		}
		const size_t c = c0 + 2 * (warp + 8 * u) + h;
		if(g == 0 && c < d) cs[u].flush(stats + 4 * c);
	}
}

template <typename W, typename SRC>
__global__ void __launch_bounds__(256) k_vec_encode(const SRC *__restrict__ src, size_t n, W *__restrict__ col,
	uint64_t *__restrict__ stats, EncodeParams ep) {
	ColStats cs;
	cs.init();
	for(size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
		W wv = load_word<W, SRC>(src[i], ep);
		col[i] = wv;
		cs.add<W>(wv);
	}
	cs.warp_reduce();
	if((threadIdx.x & 31) == 0) cs.flush(stats);
}

// grid (ceil(n/32), ceil(d/32)), block (32,8): column-major -> row-major (rows on grid.x: no 65535 limit)
template <typename W>
__global__ void __launch_bounds__(256) k_cols_to_rows(const W *__restrict__ xt, size_t ld, size_t n, size_t d, W *__restrict__ dst) {
	__shared__ W tile[32][33];
	const int tx = threadIdx.x, ty = threadIdx.y;
	const size_t c0 = (size_t) blockIdx.y * 32, r0 = (size_t) blockIdx.x * 32;
#pragma unroll
	for(int u = 0; u < 4; u++) {
		int cc = ty + 8 * u;
		size_t c = c0 + cc, r = r0 + tx;
		tile[cc][tx] = (c < d && r < n) ? xt[c * ld + r] : (W) 0;
	}
	__syncthreads();
#pragma unroll
	for(int u = 0; u < 4; u++) {
		int rr = ty + 8 * u;
		size_t r = r0 + rr, c = c0 + tx;
		if(r < n && c < d) dst[r * d + c] = tile[tx][rr];
	}
}

template <typename W, typename SRC>
static int launch_rows_to_cols(p1_ctx *ctx, const SRC *src, size_t rows, size_t d, size_t row_off, p1_data *out, EncodeParams ep) {
	dim3 grid((unsigned) ((d + 31) / 32), (unsigned) ((rows + ENC_ROWS_PER_CTA - 1) / ENC_ROWS_PER_CTA)), block(32, 8);
	p1_prof_begin(ctx, std::is_same<SRC, double>::value ? P1_K_ENCODE : P1_K_LAYOUT);
	k_rows_to_cols<W, SRC><<<grid, block, 0, ctx->stream>>>(src, rows, d, row_off, (W *) out->xt, out->ld, out->stats, ep);
	P1_LAUNCHED(ctx);
	return 0;
}

template <typename W, typename SRC>
static int launch_vec(p1_ctx *ctx, const SRC *src, p1_data *out, EncodeParams ep) {
	unsigned blocks = (unsigned) ((out->n + 256 * 8 - 1) / (256 * 8));
	if(blocks > (unsigned) ctx->sm_count * 8) blocks = ctx->sm_count * 8;
	p1_prof_begin(ctx, std::is_same<SRC, double>::value ? P1_K_ENCODE : P1_K_LAYOUT);
	k_vec_encode<W, SRC><<<blocks, 256, 0, ctx->stream>>>(src, out->n, (W *) out->xt + out->d * out->ld, out->stats + 4 * out->d, ep);
	P1_LAUNCHED(ctx);
	return 0;
}

int p1_launch_encode(p1_ctx *ctx, const double *X_dev, size_t rows, size_t d, size_t row_off,
	const double *y_dev, double normalizer, int precision, p1_data *out) {
	EncodeParams ep = {normalizer, (double) (1ll << precision), ctx->bits_p2};
	// grid.y carries the row tiles (<= 65535): longer matrices go in slabs of 2^24 rows
	const size_t slab = (size_t) 1 << 24;
	if(X_dev && rows > slab) {
		for(size_t r = 0; r < rows; r += slab) {
			const size_t part = rows - r < slab ? rows - r : slab;
			if(p1_launch_encode(ctx, X_dev + r * d, part, d, row_off + r, NULL, normalizer, precision, out)) return 1;
		}
		X_dev = NULL;
	}
	// async-copy pipeline: needs 16-byte aligned chunks (even d, aligned source) and 128-row aligned chunk starts
	if(X_dev && ctx->opt_encode_async && d % 2 == 0 && ((uintptr_t) X_dev & 15) == 0 && row_off % 128 == 0) {
		dim3 grid((unsigned) ((d + 31) / 32), (unsigned) ((rows + ENC_A_ROWS_PER_CTA - 1) / ENC_A_ROWS_PER_CTA));
		const int smem = ENC_A_STAGES * ENC_A_STAGE_BYTES;
		p1_prof_begin(ctx, P1_K_ENCODE);
#define ENC_A_LAUNCH(W, PL) do { \
			P1_CUDA(cudaFuncSetAttribute(k_encode_async<W, PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
			k_encode_async<W, PL><<<grid, 256, smem, ctx->stream>>>(X_dev, rows, d, row_off, (W *) out->xt, out->ld, out->stats, ep, out->planes, out->Lg); \
		} while(0)
		if(out->bits == 64) { if(out->planes) ENC_A_LAUNCH(uint64_t, true); else ENC_A_LAUNCH(uint64_t, false); }
		else { if(out->planes) ENC_A_LAUNCH(uint32_t, true); else ENC_A_LAUNCH(uint32_t, false); }
#undef ENC_A_LAUNCH
		P1_LAUNCHED(ctx);
		X_dev = NULL;                                    // the matrix is done; the target vector follows below
	}
	if(out->planes) {
		// one pass that also writes the digit planes of the tensor-core Gram
		P1_CHECK(row_off % 128 == 0, "p1_encode: staged chunks must start on a 128-row boundary");
		if(X_dev) {
			dim3 grid((unsigned) ((d + 31) / 32), (unsigned) ((rows + ENC_ROWS_PER_CTA - 1) / ENC_ROWS_PER_CTA)), block(32, 8);
			p1_prof_begin(ctx, P1_K_ENCODE);
			if(out->bits == 64) k_rows_to_cols_planes<uint64_t><<<grid, block, 0, ctx->stream>>>(X_dev, rows, d, row_off, (uint64_t *) out->xt, out->ld, out->stats, ep, out->planes, out->Lg);
			else k_rows_to_cols_planes<uint32_t><<<grid, block, 0, ctx->stream>>>(X_dev, rows, d, row_off, (uint32_t *) out->xt, out->ld, out->stats, ep, out->planes, out->Lg);
			P1_LAUNCHED(ctx);
		}
		if(y_dev) {
			unsigned blocks = (unsigned) ((out->n + 256 * 16 - 1) / (256 * 16));
			if(blocks > (unsigned) ctx->sm_count * 8) blocks = ctx->sm_count * 8;
			const size_t pcol = (out->d + 1) * 128;
			p1_prof_begin(ctx, P1_K_ENCODE);
			if(out->bits == 64) k_vec_encode_planes<uint64_t><<<blocks, 256, 0, ctx->stream>>>(y_dev, out->n, (uint64_t *) out->xt + out->d * out->ld, out->stats + 4 * out->d, ep, out->planes + out->d * 128, pcol, out->Lg);
			else k_vec_encode_planes<uint32_t><<<blocks, 256, 0, ctx->stream>>>(y_dev, out->n, (uint32_t *) out->xt + out->d * out->ld, out->stats + 4 * out->d, ep, out->planes + out->d * 128, pcol, out->Lg);
			P1_LAUNCHED(ctx);
		}
		return 0;
	}
	if(X_dev) {
		if(out->bits == 64) { if(launch_rows_to_cols<uint64_t, double>(ctx, X_dev, rows, d, row_off, out, ep)) return 1; }
		else { if(launch_rows_to_cols<uint32_t, double>(ctx, X_dev, rows, d, row_off, out, ep)) return 1; }
	}
	if(y_dev) {
		if(out->bits == 64) return launch_vec<uint64_t, double>(ctx, y_dev, out, ep);
		return launch_vec<uint32_t, double>(ctx, y_dev, out, ep);
	}
	return 0;
}

int p1_launch_words_in(p1_ctx *ctx, const void *Xw, const void *yw, p1_data *out) {
	EncodeParams ep = {1.0, 1.0, ctx->bits_p2};
	if(out->bits == 64) {
		if(launch_rows_to_cols<uint64_t, uint64_t>(ctx, (const uint64_t *) Xw, out->n, out->d, 0, out, ep)) return 1;
		if(yw) return launch_vec<uint64_t, uint64_t>(ctx, (const uint64_t *) yw, out, ep);
	} else {
		if(launch_rows_to_cols<uint32_t, uint32_t>(ctx, (const uint32_t *) Xw, out->n, out->d, 0, out, ep)) return 1;
		if(yw) return launch_vec<uint32_t, uint32_t>(ctx, (const uint32_t *) yw, out, ep);
	}
	return 0;
}

int p1_launch_words_out(p1_ctx *ctx, const p1_data *dt, void *Xw, void *yw) {
	size_t w = (size_t) dt->bits / 8;
	if(Xw) {
		dim3 grid((unsigned) ((dt->n + 31) / 32), (unsigned) ((dt->d + 31) / 32)), block(32, 8);
		p1_prof_begin(ctx, P1_K_LAYOUT);
		if(dt->bits == 64) k_cols_to_rows<uint64_t><<<grid, block, 0, ctx->stream>>>((const uint64_t *) dt->xt, dt->ld, dt->n, dt->d, (uint64_t *) Xw);
		else k_cols_to_rows<uint32_t><<<grid, block, 0, ctx->stream>>>((const uint32_t *) dt->xt, dt->ld, dt->n, dt->d, (uint32_t *) Xw);
		P1_LAUNCHED(ctx);
	}
	if(yw) {
		P1_CUDA(cudaMemcpyAsync(yw, (const char *) dt->xt + dt->d * dt->ld * w, dt->n * w, cudaMemcpyDeviceToDevice, ctx->stream));
	}
	return 0;
}
