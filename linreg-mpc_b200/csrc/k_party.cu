// Party side of the TI masking protocol (src/phase1.c:173-224), batched over cross
// pairs.  Pure HBM streaming: 128-bit loads/stores, warp-shuffle + one atomic per CTA.
//   k_party_b_mask : out = col + T                        (:203-205)
//   k_party_a_step : out = col - T ; acc += <in, T>       (:188-190,195)
//   k_party_b_share: acc += <in, col>                     (:221)
#include "p1_common.cuh"

#define PK_THREADS 256
#define PK_CHUNK 8192 // words per CTA

template <typename W> struct Vec;  // 16-byte vector of ring words
template <> struct Vec<uint64_t> { typedef ulonglong2 t; static constexpr int N = 2; };
template <> struct Vec<uint32_t> { typedef uint4 t; static constexpr int N = 4; };

template <typename W> __device__ __forceinline__ void unpack(const typename Vec<W>::t &v, W *o);
template <> __device__ __forceinline__ void unpack<uint64_t>(const ulonglong2 &v, uint64_t *o) { o[0] = v.x; o[1] = v.y; }
template <> __device__ __forceinline__ void unpack<uint32_t>(const uint4 &v, uint32_t *o) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
template <typename W> __device__ __forceinline__ typename Vec<W>::t pack(const W *o);
template <> __device__ __forceinline__ ulonglong2 pack<uint64_t>(const uint64_t *o) { return make_ulonglong2(o[0], o[1]); }
template <> __device__ __forceinline__ uint4 pack<uint32_t>(const uint32_t *o) { return make_uint4(o[0], o[1], o[2], o[3]); }

template <typename T> __device__ __forceinline__ T ld_stream(const T *p) { return __ldcs(p); }
template <typename T> __device__ __forceinline__ void st_stream(T *p, const T &v) { __stcs(p, v); }

__device__ __forceinline__ void block_add(uint64_t v, uint64_t *dst) {
	__shared__ uint64_t part[PK_THREADS / 32];
	v = p1_warp_sum(v);
	if((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = v;
	__syncthreads();
	if(threadIdx.x < 32) {
		uint64_t s = threadIdx.x < PK_THREADS / 32 ? part[threadIdx.x] : 0;
		s = p1_warp_sum(s);
		if(threadIdx.x == 0 && s) p1_red_add(dst, s);
	}
}

// MODE 0: b mask, 1: a step, 2: b share.   grid (chunks, pairs)
// All vectors are 16-byte aligned (ld multiples of 2/4 words, column starts at ld*c).
template <typename W, int MODE>
__global__ void __launch_bounds__(PK_THREADS) k_party(const W *__restrict__ xt, size_t ld, size_t n,
	const int32_t *__restrict__ cols, const int64_t *__restrict__ slots,
	const W *__restrict__ T, size_t ld_T, const W *__restrict__ in, size_t ld_in,
	W *__restrict__ out, size_t ld_out, uint64_t *__restrict__ acc) {
	typedef typename Vec<W>::t V;
	constexpr int N = Vec<W>::N;
	const size_t p = blockIdx.y;
	const W *col = xt + (size_t) cols[p] * ld;
	const W *Tp = MODE != 2 ? T + p * ld_T : nullptr;
	const W *inp = MODE != 0 ? in + p * ld_in : nullptr;
	W *outp = MODE != 2 ? out + p * ld_out : nullptr;
	size_t k_lo = (size_t) blockIdx.x * PK_CHUNK, k_hi = k_lo + PK_CHUNK;
	if(k_hi > n) k_hi = n;
	uint64_t dot = 0;
	size_t nv_hi = k_hi / N; // whole vectors
	for(size_t v = k_lo / N + threadIdx.x; v < nv_hi; v += PK_THREADS) {
		W c[N], t[N], i[N], o[N];
		unpack<W>(__ldg(reinterpret_cast<const V *>(col) + v), c); // columns are re-read across pairs: keep cached
		if(MODE != 2) unpack<W>(ld_stream(reinterpret_cast<const V *>(Tp) + v), t);
		if(MODE != 0) unpack<W>(ld_stream(reinterpret_cast<const V *>(inp) + v), i);
#pragma unroll
		for(int e = 0; e < N; e++) {
			if(MODE == 0) o[e] = c[e] + t[e];
			if(MODE == 1) { o[e] = c[e] - t[e]; dot += (uint64_t) i[e] * (uint64_t) t[e]; }
			if(MODE == 2) dot += (uint64_t) i[e] * (uint64_t) c[e];
		}
		if(MODE != 2) st_stream(reinterpret_cast<V *>(outp) + v, pack<W>(o));
	}
	// ragged tail (n not a multiple of the vector width): last CTA, scalar
	if(k_hi == n && (n % N) && threadIdx.x == 0) {
		for(size_t k = n / N * N; k < n; k++) {
			W c = col[k];
			if(MODE == 0) outp[k] = c + Tp[k];
			if(MODE == 1) { outp[k] = c - Tp[k]; dot += (uint64_t) inp[k] * (uint64_t) Tp[k]; }
			if(MODE == 2) dot += (uint64_t) inp[k] * (uint64_t) c;
		}
	}
	if(MODE != 0) block_add(dot, acc + slots[p]);
}

__global__ void k_acc_sub(uint64_t *__restrict__ acc, const int64_t *__restrict__ slots, const uint64_t *__restrict__ value, size_t count) {
	size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t < count) p1_red_add(acc + slots[t], (uint64_t) 0 - value[t]);
}

__global__ void k_ti_values(const uint64_t *__restrict__ xy, const uint64_t *__restrict__ r, size_t count, uint64_t mask,
	uint64_t *__restrict__ va, uint64_t *__restrict__ vb) {
	size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t < count) {
		va[t] = (xy[t] - r[t]) & mask;
		vb[t] = r[t] & mask;
	}
}

template <int MODE>
static int launch_party(p1_ctx *ctx, const p1_data *data, const int32_t *cols, const int64_t *slots, size_t count,
	const void *T, size_t ld_T, const void *in, size_t ld_in, void *out, size_t ld_out, uint64_t *acc) {
	if(!count) return 0;
	P1_CHECK(data, "null data");
	P1_CHECK(data->bits == ctx->bits_p1, "data was encoded for a %d-bit ring", data->bits);
	size_t align = 16 / (data->bits / 8);
	P1_CHECK((MODE == 2 || (ld_T % align == 0 && ld_out % align == 0)) && (MODE == 0 || ld_in % align == 0),
		"vector strides must be multiples of %zu words (16 bytes)", align);
	P1_CHECK(count <= 65535, "at most 65535 pairs per batch");
	const void *cols_d = NULL, *slots_d = NULL;
	if(p1_small_to_dev(ctx, cols, count * sizeof(int32_t), 0, &cols_d)) return 1;
	if(MODE != 0 && p1_small_to_dev(ctx, slots, count * sizeof(int64_t), 1, &slots_d)) return 1;
	dim3 grid((unsigned) ((data->n + PK_CHUNK - 1) / PK_CHUNK), (unsigned) count);
	p1_prof_begin(ctx, MODE == 0 ? P1_K_PARTY_B_MASK : MODE == 1 ? P1_K_PARTY_A_STEP : P1_K_PARTY_B_SHARE);
	if(data->bits == 64) {
		k_party<uint64_t, MODE><<<grid, PK_THREADS, 0, ctx->stream>>>((const uint64_t *) data->xt, data->ld, data->n,
			(const int32_t *) cols_d, (const int64_t *) slots_d, (const uint64_t *) T, ld_T, (const uint64_t *) in, ld_in,
			(uint64_t *) out, ld_out, acc);
	} else {
		k_party<uint32_t, MODE><<<grid, PK_THREADS, 0, ctx->stream>>>((const uint32_t *) data->xt, data->ld, data->n,
			(const int32_t *) cols_d, (const int64_t *) slots_d, (const uint32_t *) T, ld_T, (const uint32_t *) in, ld_in,
			(uint32_t *) out, ld_out, acc);
	}
	P1_LAUNCHED(ctx);
	return 0;
}


extern "C" int p1_party_b_mask(p1_ctx *ctx, const p1_data *data, const int32_t *cols, size_t count,
	const void *T_dev, size_t ld_T, void *out_dev, size_t ld_out) {
	P1_ENTER(ctx);
	P1_CHECK(cols && T_dev && out_dev, "p1_party_b_mask: arguments may not be null");
	return launch_party<0>(ctx, data, cols, NULL, count, T_dev, ld_T, NULL, 0, out_dev, ld_out, NULL);
}

extern "C" int p1_party_a_step(p1_ctx *ctx, const p1_data *data, const int32_t *cols, const int64_t *slots, size_t count,
	const void *T_dev, size_t ld_T, const void *in_dev, size_t ld_in, void *out_dev, size_t ld_out, uint64_t *acc_dev) {
	P1_ENTER(ctx);
	P1_CHECK(cols && slots && T_dev && in_dev && out_dev && acc_dev, "p1_party_a_step: arguments may not be null");
	return launch_party<1>(ctx, data, cols, slots, count, T_dev, ld_T, in_dev, ld_in, out_dev, ld_out, acc_dev);
}

extern "C" int p1_party_b_share(p1_ctx *ctx, const p1_data *data, const int32_t *cols, const int64_t *slots, size_t count,
	const void *in_dev, size_t ld_in, uint64_t *acc_dev) {
	P1_ENTER(ctx);
	P1_CHECK(cols && slots && in_dev && acc_dev, "p1_party_b_share: arguments may not be null");
	return launch_party<2>(ctx, data, cols, slots, count, NULL, 0, in_dev, ld_in, NULL, 0, acc_dev);
}

extern "C" int p1_acc_sub_values(p1_ctx *ctx, uint64_t *acc_dev, const int64_t *slots, const uint64_t *value_dev, size_t count) {
	P1_ENTER(ctx);
	if(!count) return 0;
	P1_CHECK(acc_dev && slots && value_dev, "p1_acc_sub_values: arguments may not be null");
	const void *slots_d = NULL;
	if(p1_small_to_dev(ctx, slots, count * sizeof(int64_t), 1, &slots_d)) return 1;
	p1_prof_begin(ctx, P1_K_MISC);
	k_acc_sub<<<(unsigned) ((count + 255) / 256), 256, 0, ctx->stream>>>(acc_dev, (const int64_t *) slots_d, value_dev, count);
	P1_LAUNCHED(ctx);
	return 0;
}

extern "C" int p1_ti_values(p1_ctx *ctx, const uint64_t *xy_dev, const uint64_t *r_dev, size_t count,
	uint64_t *value_a_dev, uint64_t *value_b_dev) {
	P1_ENTER(ctx);
	if(!count) return 0;
	P1_CHECK(xy_dev && r_dev && value_a_dev && value_b_dev, "p1_ti_values: arguments may not be null");
	uint64_t mask = ctx->bits_p1 == 64 ? ~0ull : 0xffffffffull;
	p1_prof_begin(ctx, P1_K_MISC);
	k_ti_values<<<(unsigned) ((count + 255) / 256), 256, 0, ctx->stream>>>(xy_dev, r_dev, count, mask, value_a_dev, value_b_dev);
	P1_LAUNCHED(ctx);
	return 0;
}
