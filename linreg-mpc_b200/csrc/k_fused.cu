// The masking loop of one box as ONE persistent kernel: for a batch of cross pairs the
// trusted initializer's mask expansion (src/phase1.c:270-274), party b's masked column
// (:202-205), party a's step (:187-195) and party b's share (:221) run on the same rows
// in the same warp, so the masks x, y never touch HBM.  Per row k of pair p:
//     x = S[W0 + k], y = S[W0 + n_total + k]                   (AES-128-CTR, as k_ti_generate)
//     msg_b[p][k] = b_j[k] + x          (message b -> a, materialised: the protocol's data plane)
//     msg_a[p][k] = a_i[k] - y          (message a -> b, materialised)
//     xy[p]          += x * y           (TI: value_a = xy - r, value_b = r)
//     acc[slot_a[p]] += msg_b[p][k] * y (party a: <in, T>)
//     acc[slot_b[p]] += msg_a[p][k] * b_j[k]                   (party b: <in, col>)
// The share words are those of the four separate kernels (k_ti_generate, k_party<0|1|2>),
// which stay in the library for parties that live on different devices; with x_dev / y_dev
// given the masks are materialised too (used by the parity tests to prove identical words).
// HBM per pair: b_j read, two messages written (3 n w; a_i is shared by the consecutive
// pairs of a row i and comes from L2) instead of 11 n w.  Bound: the shared-memory lookups
// of the T-table AES (aes_dev.cuh), as for k_ti_generate.
#include "p1_common.cuh"
#include "aes_dev.cuh"

#define FU_U 8          // stream blocks per lane per work unit (one 256-block CTR group of the x stream)
#define FU_THREADS 512

struct FusedParams {
	AesCtrKey key, keyb;   // keyb: the same key with round 1 folded for the NEXT 2^32-block span of the counter (hi + 1)
	uint64_t q0;
	uint32_t count;
	uint64_t n_total, row0, n_rows;
	const void *xt;        // column store of the box
	uint64_t ld_x;
	const int32_t *cols_a, *cols_b;
	const int64_t *slots_a, *slots_b;
	void *msg_b, *msg_a;   // [count][ld] ring words
	uint64_t ld;
	void *x, *y;           // optional: the masks themselves
	uint64_t *xy, *r, *acc;
	uint32_t units_per_pair;
	uint32_t chunk;        // consecutive work units a warp takes at a time (ctx option "fused_chunk")
	uint32_t opt_fast;     // 0: every unit through the general loop (ctx option "fused_fast", for comparisons)
	const uint32_t *T0;
};

template <typename W> struct FuLine;   // a 16-byte line of ring words
template <> struct FuLine<uint64_t> {
	static constexpr int N = 2;
	static __device__ __forceinline__ void words(const uint32_t s[4], uint64_t o[2]) {
		o[0] = ((uint64_t) s[1] << 32) | s[0];
		o[1] = ((uint64_t) s[3] << 32) | s[2];
	}
	static __device__ __forceinline__ void load(const uint64_t *p, uint64_t o[2]) {
		const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(p));
		o[0] = v.x; o[1] = v.y;
	}
	static __device__ __forceinline__ void store(uint64_t *p, const uint64_t o[2]) {
		__stcs(reinterpret_cast<ulonglong2 *>(p), make_ulonglong2(o[0], o[1]));
	}
};
template <> struct FuLine<uint32_t> {
	static constexpr int N = 4;
	static __device__ __forceinline__ void words(const uint32_t s[4], uint32_t o[4]) {
#pragma unroll
		for(int e = 0; e < 4; e++) o[e] = s[e];
	}
	static __device__ __forceinline__ void load(const uint32_t *p, uint32_t o[4]) {
		const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
		o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
	}
	static __device__ __forceinline__ void store(uint32_t *p, const uint32_t o[4]) {
		__stcs(reinterpret_cast<uint4 *>(p), make_uint4(o[0], o[1], o[2], o[3]));
	}
};

template <typename W> struct FuPair {   // pointers and running sums of the pair a warp works on
	const W *a, *b;
	W *mb, *ma, *xo, *yo;
	uint64_t dxy, da, db;
};

// one row, scalar accesses (shard edges, and the half lines at the ends of a misaligned run)
template <typename W>
__device__ __forceinline__ void fu_elem(FuPair<W> &S, uint64_t k, W xv, W yv) {
	const W av = S.a[k], bv = S.b[k];
	const W mb = bv + xv, ma = av - yv;
	S.mb[k] = mb;
	S.ma[k] = ma;
	if(S.xo) S.xo[k] = xv;
	if(S.yo) S.yo[k] = yv;
	S.dxy += (uint64_t) xv * (uint64_t) yv;
	S.da += (uint64_t) mb * (uint64_t) yv;
	S.db += (uint64_t) ma * (uint64_t) bv;
}

// one 16-byte line of rows starting at row k (k * sizeof(W) is a multiple of 16)
template <typename W>
__device__ __forceinline__ void fu_line(FuPair<W> &S, uint64_t k, const uint32_t xs[4], const uint32_t ys[4]) {
	constexpr int N = FuLine<W>::N;
	W xv[N], yv[N], av[N], bv[N], mb[N], ma[N];
	FuLine<W>::load(S.a + k, av);
	FuLine<W>::load(S.b + k, bv);
	FuLine<W>::words(xs, xv);
	FuLine<W>::words(ys, yv);
#pragma unroll
	for(int e = 0; e < N; e++) {
		mb[e] = bv[e] + xv[e];
		ma[e] = av[e] - yv[e];
		S.dxy += (uint64_t) xv[e] * (uint64_t) yv[e];
		S.da += (uint64_t) mb[e] * (uint64_t) yv[e];
		S.db += (uint64_t) ma[e] * (uint64_t) bv[e];
	}
	FuLine<W>::store(S.mb + k, mb);
	FuLine<W>::store(S.ma + k, ma);
	if(S.xo) FuLine<W>::store(S.xo + k, xv);
	if(S.yo) FuLine<W>::store(S.yo + k, yv);
}

// words [lo, hi) of a keystream block as rows starting at byte offset ob of the shard's mask
template <typename W>
__device__ __forceinline__ void fu_words(FuPair<W> &S, int64_t ob, const uint32_t xs[4], const uint32_t ys[4], int lo, int hi) {
	constexpr int w = sizeof(W);
#pragma unroll
	for(int e = 0; e < 16 / w; e++) {
		const int w0 = e * (w / 4);
		if(w0 < lo || w0 >= hi) continue;
		W xv, yv;
		if(w == 8) {
			xv = (W) (((uint64_t) xs[2 * e + 1] << 32) | xs[2 * e]);
			yv = (W) (((uint64_t) ys[2 * e + 1] << 32) | ys[2 * e]);
		} else {
			xv = (W) xs[e];
			yv = (W) ys[e];
		}
		fu_elem<W>(S, (uint64_t) (ob + e * w) / w, xv, yv);
	}
}

// HW = 32-bit words of a keystream block that belong to the 16-byte line ending inside it
// (0: blocks and lines coincide).  The line of lane l is the tail of lane l-1's block and the
// head of its own; lane 0 takes the tail lane 31 had in the previous iteration (carry).
template <typename W, int HW>
__device__ __forceinline__ void fu_full(FuPair<W> &S, int64_t o, const uint32_t xs[4], const uint32_t ys[4],
	uint32_t cx[4], uint32_t cy[4], bool &carry, bool last, uint32_t lane) {
	if constexpr(HW == 0) {
		fu_line<W>(S, (uint64_t) o / sizeof(W), xs, ys);
	} else {
	uint32_t lx[4], ly[4];
#pragma unroll
	for(int i = 0; i < 4; i++) {
		if(i < 4 - HW) {
			const uint32_t rx = __shfl_sync(0xffffffffu, xs[HW + i], (lane + 31) & 31);
			const uint32_t ry = __shfl_sync(0xffffffffu, ys[HW + i], (lane + 31) & 31);
			lx[i] = lane ? rx : cx[i];
			ly[i] = lane ? ry : cy[i];
			cx[i] = rx;          // lane 0: the tail of lane 31, for the next iteration
			cy[i] = ry;
		} else {
			lx[i] = xs[i - (4 - HW)];
			ly[i] = ys[i - (4 - HW)];
		}
	}
	if(lane || carry) fu_line<W>(S, (uint64_t) (o + 4 * HW - 16) / sizeof(W), lx, ly);
	else fu_words<W>(S, o, xs, ys, 0, HW);                      // head of the run: no tail to join
	if(last && lane == 31) fu_words<W>(S, o, xs, ys, HW, 4);   // end of the run: nobody takes this tail
	carry = true;
	}
}

// The common case as a loop of its own: 64-bit ring, x and y streams 16 bytes apart modulo 16 (n_total even), a
// work unit (one 256-block group of the x stream) that lies wholly inside the shard's mask and inside one 2^32-block
// span of the counter for both streams, masks not materialised.  No bounds tests, no cross-lane traffic: for odd
// pairs (HW = 2: a block starts 8 bytes into a 16-byte line of the row arrays) a lane reads and writes its two rows
// as two 8-byte accesses instead of joining half lines with its neighbour.  The provider columns of iteration i+1
// are on their way while the two AES blocks of iteration i are computed.
template <int HW>
__device__ __forceinline__ void fu_fast_unit(const FusedParams &P, const AesCtrKey &K, uint32_t lb, FuPair<uint64_t> &S, uint64_t k,
	uint32_t lox, uint32_t loy, AesCtrGroup &gx, AesCtrGroup &gy) {
	const uint64_t *pa = S.a + k, *pb = S.b + k;
	uint64_t *pmb = S.mb + k, *pma = S.ma + k;
	uint64_t an[2], bn[2];
	if(HW == 0) {
		FuLine<uint64_t>::load(pa, an);
		FuLine<uint64_t>::load(pb, bn);
	} else {
		an[0] = __ldg(pa); an[1] = __ldg(pa + 1);
		bn[0] = __ldg(pb); bn[1] = __ldg(pb + 1);
	}
	aes_ctr_group(K, lb, lox, gx);          // the unit is one group of the x stream
#pragma unroll 1
	for(int i = 0; i < FU_U; i++) {
		const uint64_t a0 = an[0], a1 = an[1], b0 = bn[0], b1 = bn[1];
		if(i + 1 < FU_U) {
			pa += 64; pb += 64;                  // 32 blocks = 512 bytes on
			if(HW == 0) {
				FuLine<uint64_t>::load(pa, an);
				FuLine<uint64_t>::load(pb, bn);
			} else {
				an[0] = __ldg(pa); an[1] = __ldg(pa + 1);
				bn[0] = __ldg(pb); bn[1] = __ldg(pb + 1);
			}
		}
		uint32_t xs[4], ys[4];
		aes_ctr_block_in_group(K, lb, lox, gx, xs);
		if((loy >> 8) != gy.id) aes_ctr_group(K, lb, loy, gy);
		aes_ctr_block_in_group(K, lb, loy, gy, ys);
		lox += 32; loy += 32;
		const uint64_t x0 = ((uint64_t) xs[1] << 32) | xs[0], x1 = ((uint64_t) xs[3] << 32) | xs[2];
		const uint64_t y0 = ((uint64_t) ys[1] << 32) | ys[0], y1 = ((uint64_t) ys[3] << 32) | ys[2];
		const uint64_t mb0 = b0 + x0, mb1 = b1 + x1, ma0 = a0 - y0, ma1 = a1 - y1;
		S.dxy += x0 * y0 + x1 * y1;
		S.da += mb0 * y0 + mb1 * y1;
		S.db += ma0 * b0 + ma1 * b1;
		if(HW == 0) {
			__stcs(reinterpret_cast<ulonglong2 *>(pmb), make_ulonglong2(mb0, mb1));
			__stcs(reinterpret_cast<ulonglong2 *>(pma), make_ulonglong2(ma0, ma1));
		} else {
			__stcs(pmb, mb0); __stcs(pmb + 1, mb1);
			__stcs(pma, ma0); __stcs(pma + 1, ma1);
		}
		pmb += 64; pma += 64;
	}
}

template <typename W, bool ALIGNED, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_masking_fused(const FusedParams P) {
	extern __shared__ __align__(256) uint8_t smem[];
	constexpr int w = sizeof(W);
	uint8_t *sT = smem + aes_table_offset(smem, 0, 0);
	aes_load_tables(sT, P.T0);
	__syncthreads();
	const uint32_t lane = threadIdx.x & 31, lb = aes_lane_base(sT, lane);
	const uint32_t warps_total = gridDim.x * (blockDim.x >> 5);
	const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint32_t units = P.count * P.units_per_pair;
	const uint64_t nbytes = P.n_rows * w;
	const uint64_t ydelta = P.n_total * w;
	const uint32_t ysh = (uint32_t) (ydelta & 15);
	AesCtrGroup gx, gy, gy1;
	gx.id = gy.id = gy1.id = 0xffffffffu;
	uint32_t last_hi = P.key.hi;      // high counter word the group caches were computed for

	// A warp takes `chunk` consecutive units at a time: they mostly belong to one pair, so the three running sums are
	// reduced over the warp and added to memory once per chunk instead of once per unit, and the y-stream group of a
	// unit's last blocks is still cached when the next unit starts.
	// Whole rounds of chunks (every warp gets the same number) first, the remaining units one at a time, so that the
	// last round of a launch is no longer than one unit.
	const uint32_t chunk = P.chunk;
	const uint32_t full_chunks = units / (chunk * warps_total) * warps_total;     // chunks in whole rounds
	const uint32_t U1 = full_chunks * chunk, items = full_chunks + (units - U1);
	uint32_t cur_p = 0xffffffffu;
	uint64_t sxy = 0, sa = 0, sb = 0;
	auto flush = [&]() {
		if(cur_p == 0xffffffffu) return;
		sxy = p1_warp_sum(sxy);
		sa = p1_warp_sum(sa);
		sb = p1_warp_sum(sb);
		if(lane == 0) {
			if(sxy) p1_red_add(P.xy + cur_p, sxy);
			if(sa) p1_red_add(P.acc + P.slots_a[cur_p], sa);
			if(sb) p1_red_add(P.acc + P.slots_b[cur_p], sb);
		}
		sxy = sa = sb = 0;
	};
	for(uint32_t it = gw; it < items; it += warps_total)
	for(uint32_t u = it < full_chunks ? it * chunk : U1 + (it - full_chunks), u_end = it < full_chunks ? u + chunk : u + 1; u < u_end; u++) {
		const uint32_t p = u / P.units_per_pair, uu = u - p * P.units_per_pair;
		if(p != cur_p) {
			flush();
			cur_p = p;
		}
		const uint64_t W0 = (P.q0 + p) * (2 * P.n_total + 1);
		const uint64_t X0 = (W0 + P.row0) * w;
		const uint64_t b_first = X0 >> 4;
		FuPair<W> S;
		S.a = (const W *) P.xt + (uint64_t) P.cols_a[p] * P.ld_x;
		S.b = (const W *) P.xt + (uint64_t) P.cols_b[p] * P.ld_x;
		S.mb = (W *) P.msg_b + (uint64_t) p * P.ld;
		S.ma = (W *) P.msg_a + (uint64_t) p * P.ld;
		S.xo = P.x ? (W *) P.x + (uint64_t) p * P.ld : nullptr;
		S.yo = P.y ? (W *) P.y + (uint64_t) p * P.ld : nullptr;
		S.dxy = sxy; S.da = sa; S.db = sb;
		uint32_t cx[4] = {0, 0, 0, 0}, cy[4] = {0, 0, 0, 0};
		bool carry = false;
		const uint64_t B0 = (b_first & ~(uint64_t) 255) + (uint64_t) uu * (32 * FU_U);
		// offset of the block group in the shard's mask; the phase (o & 15) is the same for every block of the pair
		const uint32_t hw = ((uint32_t) X0 & 15u) >> 2;   // words of a block that complete the line begun by the block before
		if constexpr(w == 8 && ALIGNED) {
			// warp-uniform: the whole unit inside the mask, one high counter word for both streams - the one the launch
			// started in (key) or the next (keyb: a launch of many pairs may cross a multiple of 2^32 blocks) -, masks not kept
			const int64_t o0 = (int64_t) (B0 << 4) - (int64_t) X0;
			const uint64_t By0 = B0 + (ydelta >> 4);
			const uint32_t hiw = (uint32_t) (B0 >> 32);
			if(!P.x && !P.y && P.opt_fast && o0 >= 0 && o0 + 16 * 32 * FU_U <= (int64_t) nbytes
				&& (hiw == P.key.hi || hiw == P.keyb.hi) && (uint32_t) ((B0 + 32 * FU_U - 1) >> 32) == hiw
				&& (uint32_t) (By0 >> 32) == hiw && (uint32_t) ((By0 + 32 * FU_U - 1) >> 32) == hiw) {
				const uint64_t k = ((uint64_t) o0 >> 3) + 2 * lane;
				// the group caches belong to one (key, hi): drop them when the unit uses the other
				if(hiw != last_hi) { gx.id = gy.id = gy1.id = 0xffffffffu; last_hi = hiw; }
				if(hiw == P.key.hi) {
					if(hw == 0) fu_fast_unit<0>(P, P.key, lb, S, k, (uint32_t) B0 + lane, (uint32_t) By0 + lane, gx, gy);
					else fu_fast_unit<2>(P, P.key, lb, S, k, (uint32_t) B0 + lane, (uint32_t) By0 + lane, gx, gy);
				} else {
					if(hw == 0) fu_fast_unit<0>(P, P.keyb, lb, S, k, (uint32_t) B0 + lane, (uint32_t) By0 + lane, gx, gy);
					else fu_fast_unit<2>(P, P.keyb, lb, S, k, (uint32_t) B0 + lane, (uint32_t) By0 + lane, gx, gy);
				}
				goto unit_done;
			}
			// the general loop caches groups of key.hi only
			if(last_hi != P.key.hi) { gx.id = gy.id = gy1.id = 0xffffffffu; last_hi = P.key.hi; }
		}
#pragma unroll 1
		for(int i = 0; i < FU_U; i++) {
			const uint64_t B = B0 + i * 32 + lane;
			const int64_t o = (int64_t) (B << 4) - (int64_t) X0;
			const int64_t o0 = (int64_t) ((B0 + i * 32) << 4) - (int64_t) X0;            // lane 0's offset
			// warp-uniform: all 32 blocks of this iteration / of the next one lie inside the shard's mask
			const bool full = o0 >= 0 && o0 + 512 <= (int64_t) nbytes;
			const bool next_full = i + 1 < FU_U && o0 + 512 >= 0 && o0 + 1024 <= (int64_t) nbytes;
			if(o <= -16 || o >= (int64_t) nbytes) { carry = false; continue; }
			uint32_t xs[4], ys[4];
			aes_ctr_block_g(P.key, lb, B, gx, xs);
			const uint64_t By = ((B << 4) + ydelta) >> 4;
			if(ALIGNED) {
				aes_ctr_block_g(P.key, lb, By, gy, ys);
			} else {
				uint32_t ya[4], yb[4];
				aes_ctr_block_g(P.key, lb, By, gy, ya);
				aes_ctr_block_g(P.key, lb, By + 1, gy1, yb);
				const uint32_t ws = ysh >> 2;
#pragma unroll
				for(int c = 0; c < 4; c++) {
					const uint32_t idx = c + ws;
					uint32_t v = 0;
#pragma unroll
					for(int z = 0; z < 8; z++) if(idx == (uint32_t) z) v = z < 4 ? ya[z & 3] : yb[z & 3];
					ys[c] = v;
				}
			}
			if(full) {
				if(hw == 0) fu_full<W, 0>(S, o, xs, ys, cx, cy, carry, !next_full, lane);
				else if(hw == 2) fu_full<W, 2>(S, o, xs, ys, cx, cy, carry, !next_full, lane);
				else if(w == 4 && hw == 1) fu_full<W, 1>(S, o, xs, ys, cx, cy, carry, !next_full, lane);
				else if(w == 4) fu_full<W, 3>(S, o, xs, ys, cx, cy, carry, !next_full, lane);
				continue;
			}
			carry = false;
			// shard edge: row by row, with bounds
#pragma unroll
			for(int e = 0; e < 16 / w; e++) {
				const int64_t ob = o + e * w;
				if(ob < 0 || ob + w > (int64_t) nbytes) continue;
				W xv, yv;
				if(w == 8) {
					xv = (W) (((uint64_t) xs[2 * e + 1] << 32) | xs[2 * e]);
					yv = (W) (((uint64_t) ys[2 * e + 1] << 32) | ys[2 * e]);
				} else {
					xv = (W) xs[e];
					yv = (W) ys[e];
				}
				fu_elem<W>(S, (uint64_t) ob / w, xv, yv);
			}
		}
	unit_done:
		sxy = S.dxy; sa = S.da; sb = S.db;
		if(P.r && uu == 0) {
			// r = S[W0 + 2 n_total] (src/phase1.c:273)
			const uint64_t Rb = (W0 + 2 * P.n_total) * w;
			uint32_t rs[4];
			aes_ctr_block(P.key, lb, Rb >> 4, rs);
			const uint32_t wi = (uint32_t) (Rb & 15) >> 2;
			uint32_t lo = 0, hi = 0;
#pragma unroll
			for(int z = 0; z < 4; z++) {
				if(wi == (uint32_t) z) lo = rs[z];
				if(w == 8 && wi + 1 == (uint32_t) z) hi = rs[z];
			}
			if(lane == 0) P.r[p] = ((uint64_t) hi << 32) | lo;
		}
	}
	flush();
}

template <typename W, bool ALIGNED, int THREADS>
static int launch_fused_t(p1_ctx *ctx, const FusedParams &P) {
	const size_t smem = AES_SMEM_BYTES;
	P1_CUDA(cudaFuncSetAttribute(k_masking_fused<W, ALIGNED, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	const uint64_t units = (uint64_t) P.count * P.units_per_pair;
	P1_CHECK(units < 0xffffffffull, "p1_masking_fused: batch too large (%llu work units)", (unsigned long long) units);
	uint64_t grid = (uint64_t) ctx->sm_count;
	const uint64_t need = (units + THREADS / 32 - 1) / (THREADS / 32);   // small launches: one unit per warp
	if(grid > need) grid = need;
	p1_prof_begin(ctx, P1_K_MASKING_FUSED);
	k_masking_fused<W, ALIGNED, THREADS><<<(unsigned) grid, THREADS, smem, ctx->stream>>>(P);
	P1_LAUNCHED(ctx);
	return 0;
}

// CTA size: 512 threads by default; option "fused_threads" picks one of the other compiled sizes (each with the
// register budget of its own launch bound)
template <typename W>
static int launch_fused(p1_ctx *ctx, const FusedParams &P, bool aligned) {
	const long t = ctx->opt_fused_threads ? ctx->opt_fused_threads : FU_THREADS;
#define FU_CASE(T) if(t == T) return aligned ? launch_fused_t<W, true, T>(ctx, P) : launch_fused_t<W, false, T>(ctx, P);
	FU_CASE(384) FU_CASE(512) FU_CASE(640) FU_CASE(768) FU_CASE(1024)
#undef FU_CASE
	P1_CHECK(0, "p1_masking_fused: no kernel for %ld threads per CTA (384, 512, 640, 768, 1024)", t);
}

extern "C" int p1_masking_fused(p1_ctx *ctx, const uint8_t key[16], uint64_t q0, size_t count, size_t n_total, size_t row0,
	const p1_data *data, const int32_t *cols_a, const int32_t *cols_b, const int64_t *slots_a, const int64_t *slots_b,
	void *msg_b_dev, void *msg_a_dev, size_t ld_words, void *x_dev, void *y_dev,
	uint64_t *xy_dev, uint64_t *r_dev, uint64_t *acc_dev) {
	P1_ENTER(ctx);
	if(!count) return 0;
	P1_CHECK(key && data && cols_a && cols_b && slots_a && slots_b && msg_b_dev && msg_a_dev && xy_dev && acc_dev,
		"p1_masking_fused: arguments may not be null");
	P1_CHECK(data->bits == ctx->bits_p1, "p1_masking_fused: data was encoded for a %d-bit ring", data->bits);
	const size_t n_rows = data->n;
	P1_CHECK(n_rows > 0 && row0 + n_rows <= n_total, "p1_masking_fused: rows [%zu,%zu) outside [0,%zu)", row0, row0 + n_rows, n_total);
	const int w = ctx->bits_p1 / 8;
	P1_CHECK(ld_words >= n_rows && (ld_words * w) % 16 == 0, "p1_masking_fused: ld_words %zu must be >= n_rows and a multiple of 16 bytes", ld_words);
	P1_CHECK(count <= 0xffffffffu, "p1_masking_fused: batch too large");
	FusedParams P;
	if(aes_device_T0(ctx, &P.T0)) return 1;
	const uint32_t hi0 = (uint32_t) ((((uint64_t) q0 * (2 * (uint64_t) n_total + 1) + row0) * w) >> 36);
	aes_host_ctr_key(key, hi0, &P.key);
	aes_host_ctr_key(key, hi0 + 1, &P.keyb);
	const void *ca = NULL, *cb = NULL, *sa = NULL, *sb = NULL;
	if(p1_small_to_dev(ctx, cols_a, count * sizeof(int32_t), 0, &ca)) return 1;
	if(p1_small_to_dev(ctx, cols_b, count * sizeof(int32_t), 1, &cb)) return 1;
	if(p1_small_to_dev(ctx, slots_a, count * sizeof(int64_t), 2, &sa)) return 1;
	if(p1_small_to_dev(ctx, slots_b, count * sizeof(int64_t), 3, &sb)) return 1;
	P.q0 = q0;
	P.count = (uint32_t) count;
	P.n_total = n_total;
	P.row0 = row0;
	P.n_rows = n_rows;
	P.xt = data->xt;
	P.ld_x = data->ld;
	P.cols_a = (const int32_t *) ca;
	P.cols_b = (const int32_t *) cb;
	P.slots_a = (const int64_t *) sa;
	P.slots_b = (const int64_t *) sb;
	P.msg_b = msg_b_dev;
	P.msg_a = msg_a_dev;
	P.ld = ld_words;
	P.x = x_dev;
	P.y = y_dev;
	P.xy = xy_dev;
	P.r = r_dev;
	P.acc = acc_dev;
	P.opt_fast = ctx->opt_fused_fast ? 1u : 0u;
	P.chunk = ctx->opt_fused_chunk > 0 ? (uint32_t) ctx->opt_fused_chunk : 8u;
	const uint64_t nblk = ((uint64_t) n_rows * w + 15 + 15) / 16;
	P.units_per_pair = (uint32_t) ((nblk + 255 + 32 * FU_U - 1) / (32 * FU_U));
	P1_CUDA(cudaMemsetAsync(xy_dev, 0, count * sizeof(uint64_t), ctx->stream));
	const bool aligned = ((uint64_t) n_total * w) % 16 == 0;
	if(w == 8) return launch_fused<uint64_t>(ctx, P, aligned);
	return launch_fused<uint32_t>(ctx, P, aligned);
}
