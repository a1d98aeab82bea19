// Text ingestion (src/linear.c:27-102): the reference reads X and y with one fscanf("%lf")
// per element.  Here the whitespace-separated tokens of the whole buffer are found with a
// prefix count (a token starts at a non-blank byte that follows a blank one, so token k is
// the start with k starts before it) and every token is converted by one thread:
//   k_tok_count  : token starts per 8 KB tile
//   k_tok_scan1/2/3 : exclusive scan of the tile counts (any number of tiles)
//   k_tok_parse  : decimal -> binary64 with the Eisel-Lemire algorithm (128-bit product with a
//                  truncated power of five, pow5_table.inc; exact for mantissas of up to 19
//                  digits, N. Mushtak, D. Lemire, "Fast number parsing without fallback"), i.e.
//                  the correctly rounded double that glibc's strtod returns.
// Tokens the grammar does not cover are stored as NaN and reported for the caller's strtod.
#include "p1_common.cuh"
#include <algorithm>

#define TX_THREADS 256
#define TX_BPT 32
#define TX_TILE (TX_THREADS * TX_BPT)
#define TX_MAXTOK 256                    // longest token converted on the device
#define TX_SCAN_PER 4096                 // tile counts per CTA of the scan kernels (256 threads x 16)


__device__ static const ulonglong2 pow5_tab[651] = {
#include "pow5_table.inc"
};

// binary64 bits (sign excluded) of w * 10^q, 0 < w < 2^64
__device__ __forceinline__ uint64_t el_bits(uint64_t w, int q) {
	if(q < -342) return 0;
	if(q > 308) return 0x7ffull << 52;
	const int lz = __clzll((long long) w);
	w <<= lz;
	const ulonglong2 t = pow5_tab[q + 342];
	uint64_t hi = __umul64hi(w, t.x), lo = w * t.x;
	if((hi & 0x1ffu) == 0x1ffu) {
		const uint64_t second_hi = __umul64hi(w, t.y);
		lo += second_hi;
		if(second_hi > lo) hi++;
	}
	const int upper = (int) (hi >> 63);
	const int shift = upper + 64 - 52 - 3;
	uint64_t m = hi >> shift;
	int power2 = (((152170 + 65536) * q) >> 16) + 63 + upper - lz + 1023;
	if(power2 <= 0) {                       // subnormal
		if(-power2 + 1 >= 64) return 0;
		m >>= -power2 + 1;
		m += m & 1;
		m >>= 1;
		return m;                           // bit 52 set = the smallest normal
	}
	if(lo <= 1 && q >= -4 && q <= 23 && (m & 3) == 1 && (m << shift) == hi) m &= ~1ull;   // exactly halfway: to even
	m += m & 1;
	m >>= 1;
	if(m >= (2ull << 52)) { m = 1ull << 52; power2++; }
	m &= ~(1ull << 52);
	if(power2 >= 0x7ff) return 0x7ffull << 52;
	return m | ((uint64_t) power2 << 52);
}

__device__ __forceinline__ bool tx_is_space(uint32_t c) { return c == 0x20u || (c - 9u) <= 4u; }   // isspace() of the C locale

// bit i: byte i of the thread's 32 bytes is white space
__device__ __forceinline__ uint32_t tx_space_mask(const uint4 a, const uint4 b) {
	const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
	uint32_t m = 0;
#pragma unroll
	for(int i = 0; i < 8; i++) {
#pragma unroll
		for(int k = 0; k < 4; k++) m |= (tx_is_space((w[i] >> (8 * k)) & 0xffu) ? 1u : 0u) << (4 * i + k);
	}
	return m;
}

// 16 bytes at offset `off` of the text, blanks outside [0, bytes)
__device__ __forceinline__ uint4 tx_load16(const uint8_t *__restrict__ text, int64_t off, uint64_t bytes) {
	if(off >= 0 && (uint64_t) off + 16 <= bytes) return __ldcs(reinterpret_cast<const uint4 *>(text + off));
	uint32_t w[4] = {0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u};
	for(int i = 0; i < 16; i++) {
		const int64_t o = off + i;
		if(o >= 0 && (uint64_t) o < bytes) w[i >> 2] = (w[i >> 2] & ~(0xffu << (8 * (i & 3)))) | ((uint32_t) text[o] << (8 * (i & 3)));
	}
	return make_uint4(w[0], w[1], w[2], w[3]);
}

struct TextParams {
	const uint8_t *text;
	uint64_t bytes;
	uint64_t tiles;
	double *out;
	uint64_t max_tokens;
	uint32_t *tile_cnt;     // [tiles]
	uint64_t *tile_off;     // [tiles]
	uint64_t *part;         // [ceil(tiles / TX_SCAN_PER)] sums, then exclusive offsets
	uint64_t *counters;     // [0] = tokens, [1] = deferred
	uint64_t *deferred;     // [cap]
	uint64_t cap;
};

__global__ void __launch_bounds__(TX_THREADS) k_tok_count(const TextParams P) {
	__shared__ uint32_t s_warp[TX_THREADS / 32];
	const uint64_t base = (uint64_t) blockIdx.x * TX_TILE + threadIdx.x * TX_BPT;
	uint32_t cnt = 0;
	if(base < P.bytes) {
		const uint4 a = tx_load16(P.text, (int64_t) base, P.bytes), b = tx_load16(P.text, (int64_t) base + 16, P.bytes);
		const uint32_t sp = tx_space_mask(a, b);
		const uint32_t prev = base == 0 ? 1u : (tx_is_space(P.text[base - 1]) ? 1u : 0u);
		cnt = __popc(~sp & ((sp << 1) | prev));
	}
#pragma unroll
	for(int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
	if((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
	__syncthreads();
	if(threadIdx.x == 0) {
		uint32_t t = 0;
		for(int i = 0; i < TX_THREADS / 32; i++) t += s_warp[i];
		P.tile_cnt[blockIdx.x] = t;
	}
}

__global__ void __launch_bounds__(TX_THREADS) k_tok_scan1(const TextParams P) {
	__shared__ uint32_t s_warp[TX_THREADS / 32 + 1];
	const uint64_t first = (uint64_t) blockIdx.x * TX_SCAN_PER + threadIdx.x * 16;
	uint32_t sum = 0;
	for(int e = 0; e < 16; e++) if(first + e < P.tiles) sum += P.tile_cnt[first + e];
	uint32_t total;
	p1_block_excl_scan256(sum, s_warp, &total);
	if(threadIdx.x == 0) P.part[blockIdx.x] = total;
}

__global__ void k_tok_scan2(const TextParams P, uint64_t nparts) {
	if(threadIdx.x || blockIdx.x) return;
	uint64_t run = 0;
	for(uint64_t i = 0; i < nparts; i++) { const uint64_t c = P.part[i]; P.part[i] = run; run += c; }
	P.counters[0] = run;
}

__global__ void __launch_bounds__(TX_THREADS) k_tok_scan3(const TextParams P) {
	__shared__ uint32_t s_warp[TX_THREADS / 32 + 1];
	const uint64_t first = (uint64_t) blockIdx.x * TX_SCAN_PER + threadIdx.x * 16;
	uint32_t c[16], sum = 0;
	for(int e = 0; e < 16; e++) { c[e] = first + e < P.tiles ? P.tile_cnt[first + e] : 0; sum += c[e]; }
	uint32_t total;
	uint64_t run = P.part[blockIdx.x] + p1_block_excl_scan256(sum, s_warp, &total);
	for(int e = 0; e < 16; e++) if(first + e < P.tiles) { P.tile_off[first + e] = run; run += c[e]; }
}

// one token: [+-]digits[.digits][(e|E)[+-]digits] up to the next blank.  Returns false if the
// token is not of that form (or too long, or its >19-digit mantissa is ambiguous).
__device__ __forceinline__ bool tx_convert(const uint8_t *s, uint32_t avail, double *out) {
	uint32_t i = 0;
	uint32_t c = s[0];
	const bool neg = c == '-';
	if(c == '-' || c == '+') c = s[++i];
	uint64_t w = 0;
	int nd = 0, exp10 = 0;
	bool any = false, truncated = false;
	for(; c - '0' <= 9u && i < avail; c = s[++i]) {
		any = true;
		if(w == 0 && c == '0') continue;
		if(nd < 19) { w = w * 10 + (c - '0'); nd++; }
		else { exp10++; truncated |= c != '0'; }
	}
	if(c == '.') {
		c = s[++i];
		for(; c - '0' <= 9u && i < avail; c = s[++i]) {
			any = true;
			if(w == 0 && c == '0') { exp10--; continue; }
			if(nd < 19) { w = w * 10 + (c - '0'); nd++; exp10--; }
			else truncated |= c != '0';
		}
	}
	if(!any) return false;
	if((c | 0x20u) == 'e') {
		c = s[++i];
		const bool eneg = c == '-';
		if(c == '-' || c == '+') c = s[++i];
		if(c - '0' > 9u) return false;
		int e = 0;
		for(; c - '0' <= 9u && i < avail; c = s[++i]) if(e < 100000) e = e * 10 + (int) (c - '0');
		exp10 += eneg ? -e : e;
	}
	if(i >= avail || !tx_is_space(c)) return false;      // trailing characters, or longer than the window
	uint64_t bits = 0;
	if(w) {
		bits = el_bits(w, exp10);
		if(truncated && el_bits(w + 1, exp10) != bits) return false;
	}
	*out = __longlong_as_double((long long) (bits | ((uint64_t) neg << 63)));
	return true;
}

__global__ void __launch_bounds__(TX_THREADS) k_tok_parse(const TextParams P) {
	__shared__ __align__(16) uint8_t sb[16 + TX_TILE + TX_MAXTOK + 16];
	__shared__ uint32_t s_warp[TX_THREADS / 32 + 1];
	const uint64_t t0 = (uint64_t) blockIdx.x * TX_TILE;
	for(uint32_t c = threadIdx.x; c < sizeof(sb) / 16; c += TX_THREADS)
		*reinterpret_cast<uint4 *>(sb + c * 16) = tx_load16(P.text, (int64_t) t0 - 16 + (int64_t) c * 16, P.bytes);
	__syncthreads();
	const uint32_t lb = 16 + threadIdx.x * TX_BPT;
	const uint4 a = *reinterpret_cast<const uint4 *>(sb + lb), b = *reinterpret_cast<const uint4 *>(sb + lb + 16);
	const uint32_t sp = tx_space_mask(a, b);
	const uint32_t prev = (t0 == 0 && threadIdx.x == 0) ? 1u : (tx_is_space(sb[lb - 1]) ? 1u : 0u);
	const uint32_t starts = ~sp & ((sp << 1) | prev);
	uint32_t total;
	uint64_t k = P.tile_off[blockIdx.x] + p1_block_excl_scan256(__popc(starts), s_warp, &total);
	for(uint32_t s = starts; s; s &= s - 1, k++) {
		const uint32_t i = __ffs(s) - 1;
		double v;
		const bool ok = tx_convert(sb + lb + i, sizeof(sb) - (lb + i) - 4, &v);   // 3 bytes of slack for the unchecked steps over '.', 'e' and a sign
		if(!ok) {
			v = __longlong_as_double(0x7ff8000000000000ll);
			const uint64_t slot = atomicAdd(reinterpret_cast<unsigned long long *>(P.counters + 1), 1ull);
			if(slot < P.cap) P.deferred[slot] = t0 + threadIdx.x * TX_BPT + i;
		}
		if(k < P.max_tokens) P.out[k] = v;
	}
}

extern "C" int p1_parse_doubles(p1_ctx *ctx, const char *text, size_t bytes, double *out_dev, size_t max_tokens,
	uint64_t *n_tokens, uint64_t *n_deferred, uint64_t *deferred, size_t cap) {
	P1_ENTER(ctx);
	P1_CHECK(n_tokens && n_deferred && (deferred || !cap) && (out_dev || !max_tokens), "p1_parse_doubles: arguments may not be null");
	*n_tokens = *n_deferred = 0;
	if(!bytes) return 0;
	P1_CHECK(text, "p1_parse_doubles: null text");
	const bool on_dev = p1_is_device_ptr(text);
	P1_CHECK(!on_dev || ((uintptr_t) text & 15) == 0, "p1_parse_doubles: a device text buffer must be 16-byte aligned");
	const uint64_t tiles = (bytes + TX_TILE - 1) / TX_TILE;
	const uint64_t nparts = (tiles + TX_SCAN_PER - 1) / TX_SCAN_PER;
	P1_CHECK(tiles < (1ull << 31), "p1_parse_doubles: buffer too large for one call (%zu bytes)", bytes);
	// one allocation: [text copy][tile_off][part][counters][deferred][tile_cnt]
	size_t o_text = 0, o_off = o_text + (on_dev ? 0 : p1_round_up(bytes, 256));
	size_t o_part = o_off + tiles * 8, o_cnt = o_part + nparts * 8, o_def = o_cnt + 16, o_tc = o_def + cap * 8;
	const size_t total = o_tc + tiles * 4;
	p1_tmp tmp;
	P1_CUDA(tmp.alloc(total, ctx->stream));
	char *ws = (char *) tmp.p;
	TextParams P;
	P.text = (const uint8_t *) text;
	if(!on_dev) {
		P1_CUDA(cudaMemcpyAsync(ws + o_text, text, bytes, cudaMemcpyHostToDevice, ctx->stream));
		P.text = (const uint8_t *) ws + o_text;
	}
	P.bytes = bytes; P.tiles = tiles; P.out = out_dev; P.max_tokens = max_tokens;
	P.tile_off = (uint64_t *) (ws + o_off); P.part = (uint64_t *) (ws + o_part);
	P.counters = (uint64_t *) (ws + o_cnt); P.deferred = (uint64_t *) (ws + o_def); P.cap = cap;
	P.tile_cnt = (uint32_t *) (ws + o_tc);
	P1_CUDA(cudaMemsetAsync(P.counters, 0, 16, ctx->stream));
	p1_prof_begin(ctx, P1_K_TEXT);
	k_tok_count<<<(unsigned) tiles, TX_THREADS, 0, ctx->stream>>>(P);
	k_tok_scan1<<<(unsigned) nparts, TX_THREADS, 0, ctx->stream>>>(P);
	k_tok_scan2<<<1, 32, 0, ctx->stream>>>(P, nparts);
	k_tok_scan3<<<(unsigned) nparts, TX_THREADS, 0, ctx->stream>>>(P);
	k_tok_parse<<<(unsigned) tiles, TX_THREADS, 0, ctx->stream>>>(P);
	ctx->launches += 4;
	P1_LAUNCHED(ctx);
	uint64_t counters[2] = {0, 0};
	P1_CUDA(cudaMemcpyAsync(counters, P.counters, 16, cudaMemcpyDeviceToHost, ctx->stream));
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	*n_tokens = counters[0];
	*n_deferred = counters[1];
	const size_t got = counters[1] < cap ? (size_t) counters[1] : cap;
	if(got) {
		P1_CUDA(cudaMemcpy(deferred, P.deferred, got * 8, cudaMemcpyDeviceToHost));
		std::sort(deferred, deferred + got);
	}
	return 0;
}
