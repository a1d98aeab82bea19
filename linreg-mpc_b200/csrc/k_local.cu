// Local-block kernels on CUDA cores.
//   k_diag_stats : copies the owned columns' exact sum-of-squares statistics into the
//                  share accumulator (diagonal shortcut, see p1_finalize in p1gpu.h).
//   k_diag_seq   : the literal sequential double loop of src/phase1.c:563-566.
//   k_gram_imad  : ring Gram of a column block, 64-bit IMAD (src/phase1.c:14-20,569-571),
//                  columns staged through shared memory, 4x4 register micro-tiles.
#include "p1_common.cuh"

__global__ void k_diag_stats(const uint64_t *__restrict__ stats, size_t col_lo, size_t m, size_t d, size_t nA, uint64_t *__restrict__ acc) {
	size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t >= m) return;
	size_t c = col_lo + t;
	p1_red_add(acc + nA + d + c, stats[4 * c + 0]);
	p1_red_add(acc + nA + 2 * d + c, stats[4 * c + 1]);
	p1_red_add(acc + nA + 3 * d + c, stats[4 * c + 2]);
}

int p1_launch_diag_stats(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, uint64_t *acc_dev) {
	size_t m = col_hi - col_lo;
	p1_prof_begin(ctx, P1_K_DIAG);
	k_diag_stats<<<(unsigned) ((m + 127) / 128), 128, 0, ctx->stream>>>(data->stats, col_lo, m, data->d, p1_nA(data->d), acc_dev);
	P1_LAUNCHED(ctx);
	return 0;
}

// One thread per column; k strictly increasing; mul, mul, add each rounded to binary64
// (no FMA contraction), exactly as gcc compiles the reference loop.
template <typename W>
__global__ void __launch_bounds__(32) k_diag_seq(const W *__restrict__ xt, size_t ld, size_t n, size_t col_lo, size_t m,
	double inv_scale, double scale, double *__restrict__ xy) {
	typedef typename p1_word<W>::s S;
	size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t >= m) return;
	const W *col = xt + (col_lo + t) * ld;
	double acc = xy[t];
	size_t k = 0;
	for(; k + 4 <= n; k += 4) {
		W w0 = col[k], w1 = col[k + 1], w2 = col[k + 2], w3 = col[k + 3];
		double v0 = __dmul_rn((double) (S) w0, inv_scale), v1 = __dmul_rn((double) (S) w1, inv_scale);
		double v2 = __dmul_rn((double) (S) w2, inv_scale), v3 = __dmul_rn((double) (S) w3, inv_scale);
		double t0 = __dmul_rn(__dmul_rn(v0, v0), scale), t1 = __dmul_rn(__dmul_rn(v1, v1), scale);
		double t2 = __dmul_rn(__dmul_rn(v2, v2), scale), t3 = __dmul_rn(__dmul_rn(v3, v3), scale);
		acc = __dadd_rn(acc, t0);
		acc = __dadd_rn(acc, t1);
		acc = __dadd_rn(acc, t2);
		acc = __dadd_rn(acc, t3);
	}
	for(; k < n; k++) {
		double v = __dmul_rn((double) (S) col[k], inv_scale);
		acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(v, v), scale));
	}
	xy[t] = acc;
}

int p1_launch_diag_seq(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int precision, double *xy_dev) {
	size_t m = col_hi - col_lo;
	double scale = (double) (1ll << precision), inv = 1.0 / scale; // both exact powers of two
	unsigned blocks = (unsigned) ((m + 31) / 32);
	p1_prof_begin(ctx, P1_K_DIAG);
	if(data->bits == 64) k_diag_seq<uint64_t><<<blocks, 32, 0, ctx->stream>>>((const uint64_t *) data->xt, data->ld, data->n, col_lo, m, inv, scale, xy_dev);
	else k_diag_seq<uint32_t><<<blocks, 32, 0, ctx->stream>>>((const uint32_t *) data->xt, data->ld, data->n, col_lo, m, inv, scale, xy_dev);
	P1_LAUNCHED(ctx);
	return 0;
}

// ------------------------------------------------------------------ IMAD Gram

#define GI_T 64        // columns per tile side
#define GI_THREADS 256 // 16 x 16 threads, 4x4 pairs each

struct GramSet {
	size_t col_lo;   // first owned feature column
	int mc;          // owned feature columns
	int m;           // mc (+1 if the target is appended as the last column of the set)
	size_t d, nA, ld, n;
};

__device__ __forceinline__ size_t gs_col(const GramSet &gs, int a) { return a < gs.mc ? gs.col_lo + a : gs.d; }
// accumulator slot of the pair (a > b) of the set (src/phase1.c:581-585)
__device__ __forceinline__ size_t gs_slot(const GramSet &gs, int a, int b) {
	size_t j = gs.col_lo + b;
	if(a >= gs.mc) return gs.nA + j;
	size_t i = gs.col_lo + a;
	return i * (i + 1) / 2 + j;
}

template <typename W, typename E, int KS>
__device__ __forceinline__ void gram_tile(const W *__restrict__ xt, const GramSet &gs, int ta, int tb, size_t row_lo, size_t row_hi,
	uint64_t *__restrict__ acc, E *sA, E *sB) {
	constexpr int KSP = KS + 1;
	const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, warp = tid >> 5, lane = tid & 31;
	uint64_t c[4][4];
#pragma unroll
	for(int u = 0; u < 4; u++)
#pragma unroll
		for(int v = 0; v < 4; v++) c[u][v] = 0;

	for(size_t k0 = row_lo; k0 < row_hi; k0 += KS) {
		// stage KS rows of the 64 + 64 columns: one warp per column, lanes along rows
		for(int cc = warp; cc < 2 * GI_T; cc += GI_THREADS / 32) {
			int a = (cc < GI_T ? ta : tb) * GI_T + (cc & (GI_T - 1));
			E *dst = (cc < GI_T ? sA : sB) + (cc & (GI_T - 1)) * KSP;
			const W *src = xt + gs_col(gs, a < gs.m ? a : 0) * gs.ld;
			bool live = a < gs.m;
#pragma unroll
			for(int r = lane; r < KS; r += 32) {
				size_t k = k0 + r;
				dst[r] = (live && k < row_hi) ? (E) src[k] : (E) 0;
			}
		}
		__syncthreads();
#pragma unroll 4
		for(int k = 0; k < KS; k++) {
			E a[4], b[4];
#pragma unroll
			for(int u = 0; u < 4; u++) {
				a[u] = sA[(ty + 16 * u) * KSP + k];
				b[u] = sB[(tx + 16 * u) * KSP + k];
			}
#pragma unroll
			for(int u = 0; u < 4; u++)
#pragma unroll
				for(int v = 0; v < 4; v++) {
					if(sizeof(E) == 4) c[u][v] += (uint64_t) ((int64_t) (int32_t) a[u] * (int64_t) (int32_t) b[v]); // IMAD.WIDE
					else c[u][v] += (uint64_t) a[u] * (uint64_t) b[v];
				}
		}
		__syncthreads();
	}
#pragma unroll
	for(int u = 0; u < 4; u++)
#pragma unroll
		for(int v = 0; v < 4; v++) {
			int a = ta * GI_T + ty + 16 * u, b = tb * GI_T + tx + 16 * v;
			if(a < gs.m && b < a && b < gs.mc) p1_red_add(acc + gs_slot(gs, a, b), c[u][v]);
		}
}

// grid (tile pairs, k slices)
template <typename W>
__global__ void __launch_bounds__(GI_THREADS) k_gram_imad(const W *__restrict__ xt, const uint64_t *__restrict__ stats, GramSet gs,
	size_t rows_per_cta, uint64_t *__restrict__ acc) {
	extern __shared__ __align__(16) unsigned char smem[];
	// tile pair index -> (ta >= tb)
	int p = blockIdx.x, ta = 0;
	while((ta + 1) * (ta + 2) / 2 <= p) ta++;
	int tb = p - ta * (ta + 1) / 2;
	size_t row_lo = (size_t) blockIdx.y * rows_per_cta, row_hi = row_lo + rows_per_cta;
	if(row_hi > gs.n) row_hi = gs.n;
	if(row_lo >= row_hi) return;
	// narrow path when every involved column fits int32: one IMAD.WIDE per ring MAC
	__shared__ int wide;
	if(threadIdx.x == 0) wide = 0;
	__syncthreads();
	if(sizeof(W) == 8) {
		for(int cc = threadIdx.x; cc < 2 * GI_T; cc += GI_THREADS) {
			int a = (cc < GI_T ? ta : tb) * GI_T + (cc & (GI_T - 1));
			if(a < gs.m && stats[4 * gs_col(gs, a) + 3] >= (1ull << 31)) wide = 1;
		}
		__syncthreads();
	}
	if(sizeof(W) == 8 && wide) {
		uint64_t *sA = (uint64_t *) smem, *sB = sA + GI_T * (64 + 1);
		gram_tile<W, uint64_t, 64>(xt, gs, ta, tb, row_lo, row_hi, acc, sA, sB);
	} else {
		uint32_t *sA = (uint32_t *) smem, *sB = sA + GI_T * (128 + 1);
		gram_tile<W, uint32_t, 128>(xt, gs, ta, tb, row_lo, row_hi, acc, sA, sB);
	}
}

int p1_launch_gram_imad(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, uint64_t *acc_dev) {
	GramSet gs;
	gs.col_lo = col_lo;
	gs.mc = (int) (col_hi - col_lo);
	gs.m = gs.mc + (owns_target ? 1 : 0);
	gs.d = data->d;
	gs.nA = p1_nA(data->d);
	gs.ld = data->ld;
	gs.n = data->n;
	int nt = (gs.m + GI_T - 1) / GI_T, pairs = nt * (nt + 1) / 2;
	size_t want = ((size_t) ctx->sm_count * 6 + pairs - 1) / pairs;
	size_t max_slices = (gs.n + 255) / 256;
	if(want > max_slices) want = max_slices;
	if(want < 1) want = 1;
	if(want > 65535) want = 65535;
	size_t rows_per_cta = p1_round_up((gs.n + want - 1) / want, 128);
	unsigned slices = (unsigned) ((gs.n + rows_per_cta - 1) / rows_per_cta);
	size_t smem = 2 * GI_T * (64 + 1) * sizeof(uint64_t); // wide path: 66 560 B; the narrow path needs 2*64*129*4 = 66 048 B
	dim3 grid((unsigned) pairs, slices);
	p1_prof_begin(ctx, P1_K_GRAM_IMAD);
	if(data->bits == 64) {
		P1_CUDA(cudaFuncSetAttribute(k_gram_imad<uint64_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		k_gram_imad<uint64_t><<<grid, GI_THREADS, smem, ctx->stream>>>((const uint64_t *) data->xt, data->stats, gs, rows_per_cta, acc_dev);
	} else {
		P1_CUDA(cudaFuncSetAttribute(k_gram_imad<uint32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
		k_gram_imad<uint32_t><<<grid, GI_THREADS, smem, ctx->stream>>>((const uint32_t *) data->xt, data->stats, gs, rows_per_cta, acc_dev);
	}
	P1_LAUNCHED(ctx);
	return 0;
}

// ------------------------------------------------------------------ finalize on the device

// (fixed_t) / (fixed_p1_t) cast of a double as x86-64 cvttsd2si does it (src/fixed.c:3-5,11-13)
__device__ __forceinline__ int64_t p1_cast_x86(double v, int bits) {
	if(bits == 64) {
		if(!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) return (int64_t) 0x8000000000000000ull;
		return (int64_t) __double2ll_rz(v);
	}
	if(!(v > -2147483649.0 && v < 2147483648.0)) return (int64_t) (int32_t) 0x80000000u;
	return (int64_t) __double2int_rz(v);
}

// One thread per share word of one provider: the tail of run_party (src/phase1.c:605-641) with the diagonal of the
// owned columns taken from the exact statistics (or diag_xy) - the same arithmetic as p1_finalize, word for word.
__global__ void __launch_bounds__(256) k_finalize(const uint64_t *__restrict__ acc, size_t num, size_t d, size_t nA,
	const int32_t *__restrict__ col_start, int b1, int b2, int shift, double inv_scale, double scale, double normalizer2,
	const double *__restrict__ diag_xy, int32_t *__restrict__ need, uint64_t *__restrict__ out) {
	size_t per = nA + d, t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t >= num * per) return;
	size_t p = t / per, k = t - p * per;
	const uint64_t *a = acc + p * (nA + 4 * d);
	uint64_t v = a[k];
	if(k < nA) {
		// k = i (i + 1) / 2 + j, j <= i
		size_t i = (size_t) ((sqrt(8.0 * (double) k + 1.0) - 1.0) * 0.5);
		while(i * (i + 1) / 2 > k) i--;
		while((i + 1) * (i + 2) / 2 <= k) i++;
		size_t j = k - i * (i + 1) / 2;
		if(j == i && (int32_t) i >= col_start[p] && (int32_t) i < col_start[p + 1]) {
			uint64_t lo = a[nA + d + i], hi = a[nA + 2 * d + i], viol = a[nA + 3 * d + i];
			// S = (hi << 32) + lo <= 2^53 ?
			uint64_t S = (hi << 32) + lo;
			bool exact = viol == 0 && (hi >> 32) == 0 && S >= (hi << 32) && S <= (1ull << 53);
			double xy;
			if(exact) xy = __dmul_rn((double) S, inv_scale);
			else if(diag_xy && !isnan(diag_xy[i])) xy = diag_xy[i];
			else {
				atomicAdd(need, 1);
				out[t] = 0;
				return;
			}
			v = (uint64_t) p1_cast_x86(__dmul_rn(__ddiv_rn(xy, normalizer2), scale), b1);
		}
	}
	if(b1 == 64 && b2 == 64) out[t] = v;
	else if(b1 == 32) out[t] = (uint32_t) v;
	else out[t] = (uint32_t) (((int64_t) v) >> shift); // src/phase1.c:617,633
}

int p1_launch_finalize(p1_ctx *ctx, const uint64_t *acc_dev, size_t num, size_t d, const int32_t *col_start_dev, int precision,
	int precision_p2, double normalizer2, const double *diag_xy_dev, int32_t *need_dev, uint64_t *out_dev) {
	size_t nA = p1_nA(d), total = num * (nA + d);
	if(cudaMemsetAsync(need_dev, 0, sizeof(int32_t), ctx->stream) != cudaSuccess) return p1_fail("p1_finalize_async: memset failed");
	if(!total) return 0;
	double scale = (double) (1ll << precision);
	p1_prof_begin(ctx, P1_K_DIAG);
	k_finalize<<<(unsigned) ((total + 255) / 256), 256, 0, ctx->stream>>>(acc_dev, num, d, nA, col_start_dev, ctx->bits_p1, ctx->bits_p2,
		precision - precision_p2, 1.0 / scale, scale, normalizer2, diag_xy_dev, need_dev, out_dev);
	P1_LAUNCHED(ctx);
	return 0;
}
