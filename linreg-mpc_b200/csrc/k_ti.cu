// Trusted-initializer mask generation (src/phase1.c:270-278): AES-128-CTR expansion of
// the masks x, y of a batch of cross pairs, fused with <x,y> and the draw of r.
// Persistent kernel (one CTA set sized to the SM count; the replicated T-table is built
// once per CTA), work unit = 32 lanes x TI_U consecutive stream blocks of one pair.
#include "p1_common.cuh"
#include "aes_dev.cuh"
#include <string.h>

// ------------------------------------------------------------------ host AES helpers

static uint8_t h_sbox[256];
static uint32_t h_T0[256];
static bool h_ready = false;
static std::mutex h_mu;

static uint8_t gmul(uint8_t a, uint8_t b) {
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

static void host_init(void) {
	std::lock_guard<std::mutex> l(h_mu);
	if(h_ready) return;
	for(int x = 0; x < 256; x++) {
		uint8_t inv = 0;
		if(x) {
			uint8_t r = 1, b = (uint8_t) x;
			for(int e = 254; e; e >>= 1) {
				if(e & 1) r = gmul(r, b);
				b = gmul(b, b);
			}
			inv = r;
		}
		uint8_t s = inv;
		for(int i = 1; i <= 4; i++) s ^= (uint8_t) ((inv << i) | (inv >> (8 - i)));
		h_sbox[x] = s ^ 0x63;
	}
	for(int x = 0; x < 256; x++) {
		uint32_t s = h_sbox[x];
		h_T0[x] = (uint32_t) gmul((uint8_t) s, 2) | (s << 8) | (s << 16) | ((uint32_t) gmul((uint8_t) s, 3) << 24);
	}
	h_ready = true;
}

void aes_host_expand(const uint8_t key[16], AesKey *out) {
	host_init();
	uint8_t rk[176];
	memcpy(rk, key, 16);
	uint8_t rcon = 1;
	for(int r = 1; r <= 10; r++) {
		const uint8_t *p = rk + 16 * (r - 1);
		uint8_t *q = rk + 16 * r;
		uint8_t t[4] = {(uint8_t) (h_sbox[p[13]] ^ rcon), h_sbox[p[14]], h_sbox[p[15]], h_sbox[p[12]]};
		rcon = gmul(rcon, 2);
		for(int i = 0; i < 4; i++) q[i] = p[i] ^ t[i];
		for(int i = 4; i < 16; i++) q[i] = p[i] ^ q[i - 4];
	}
	for(int i = 0; i < 44; i++) {
		out->rk[i] = (uint32_t) rk[4 * i] | ((uint32_t) rk[4 * i + 1] << 8) | ((uint32_t) rk[4 * i + 2] << 16) | ((uint32_t) rk[4 * i + 3] << 24);
	}
}

const uint32_t *aes_host_T0(void) {
	host_init();
	return h_T0;
}

static inline uint32_t rotl8(uint32_t v, int k) { return k ? (v << (8 * k)) | (v >> (32 - 8 * k)) : v; }
static inline uint32_t bswap32(uint32_t v) { return (v >> 24) | ((v >> 8) & 0xff00u) | ((v << 8) & 0xff0000u) | (v << 24); }

// round 1 of the counter block 0^64 || BE32(hi) || BE32(lo) without the lookups of the
// low counter word: T_k[x] = rotl(T0[x], 8k)
void aes_host_ctr_key(const uint8_t key[16], uint32_t hi, AesCtrKey *out) {
	aes_host_expand(key, &out->k);
	const uint32_t *rk = out->k.rk;
	const uint32_t s0 = rk[0], s1 = rk[1], s2 = bswap32(hi) ^ rk[2];
	out->hi = hi;
	out->c1[0] = h_T0[s0 & 0xff] ^ rotl8(h_T0[(s1 >> 8) & 0xff], 1) ^ rotl8(h_T0[(s2 >> 16) & 0xff], 2) ^ rk[4];
	out->c1[1] = h_T0[s1 & 0xff] ^ rotl8(h_T0[(s2 >> 8) & 0xff], 1) ^ rotl8(h_T0[s0 >> 24], 3) ^ rk[5];
	out->c1[2] = h_T0[s2 & 0xff] ^ rotl8(h_T0[(s0 >> 16) & 0xff], 2) ^ rotl8(h_T0[s1 >> 24], 3) ^ rk[6];
	out->c1[3] = rotl8(h_T0[(s0 >> 8) & 0xff], 1) ^ rotl8(h_T0[(s1 >> 16) & 0xff], 2) ^ rotl8(h_T0[s2 >> 24], 3) ^ rk[7];
}

static uint32_t *d_T0[64] = {NULL};

int aes_device_T0(p1_ctx *ctx, const uint32_t **dT0) {
	const uint32_t *t = aes_host_T0();
	std::lock_guard<std::mutex> l(h_mu);
	P1_CHECK(ctx->device < 64, "device index too large");
	if(!d_T0[ctx->device]) {
		uint32_t *p = NULL;
		P1_CUDA(cudaMalloc(&p, sizeof(h_T0)));
		P1_CUDA(cudaMemcpy(p, t, sizeof(h_T0), cudaMemcpyHostToDevice));
		d_T0[ctx->device] = p;
	}
	*dT0 = d_T0[ctx->device];
	return 0;
}

// ------------------------------------------------------------------ kernel

#define TI_U 8 // stream blocks per lane per work unit: a unit is one 256-block group of the x stream (aes_dev.cuh, AesCtrGroup)

#define TI_THREADS 512          // launch bound (128 registers); the launch uses TI_THREADS_DEFAULT or the ti_threads knob
#define TI_THREADS_DEFAULT 512  // leaves registers and thread slots on every SM for the HBM-bound party kernels of the other pipe

struct TiParams {
	AesCtrKey kx, ky, kr;   // keys of the x, y and r words (one key three times in the reference's stream, p1_ti_generate)
	uint64_t q0;
	uint32_t count;
	uint64_t n_total, row0, n_rows;
	void *x, *y;       // may be NULL
	uint64_t ld;       // words between pairs
	uint64_t *xy, *r;  // may be NULL
	uint32_t units_per_pair;
	const uint32_t *T0;
};

// 16 keystream bytes per lane, 512 consecutive bytes per warp, to p (4-byte aligned; p + 4*WS is
// 16-byte aligned): one aligned 16-byte store per lane assembled with the low WS words of the
// next lane's block; the first lane's head and the last lane's tail go out as words.  WS is a
// template constant so that the word selection costs no instructions.
template <int WS>
__device__ __forceinline__ void ti_store_row_ws(char *p, const uint32_t s[4], uint32_t lane) {
	if(WS == 0) {
		__stcs(reinterpret_cast<uint4 *>(p), make_uint4(s[0], s[1], s[2], s[3]));
		return;
	}
	uint32_t c[4];
#pragma unroll
	for(int i = 0; i < 4; i++) c[i] = i + WS < 4 ? s[(i + WS) & 3] : __shfl_down_sync(0xffffffffu, s[(i + WS) & 3], 1);
	uint32_t *q = reinterpret_cast<uint32_t *>(p);
	if(lane < 31) __stcs(reinterpret_cast<uint4 *>(p + 4 * WS), make_uint4(c[0], c[1], c[2], c[3]));
	else {
#pragma unroll
		for(int i = WS; i < 4; i++) q[i] = s[i];
	}
	if(lane == 0) {
#pragma unroll
		for(int i = 0; i < WS; i++) q[i] = s[i];
	}
}
// sh = bytes from p to the next 16-byte line, uniform over the warp
__device__ __forceinline__ void ti_store_row(char *p, const uint32_t s[4], uint32_t sh, uint32_t lane) {
	switch(sh >> 2) {
	case 0: ti_store_row_ws<0>(p, s, lane); break;
	case 1: ti_store_row_ws<1>(p, s, lane); break;
	case 2: ti_store_row_ws<2>(p, s, lane); break;
	default: ti_store_row_ws<3>(p, s, lane); break;
	}
}

template <typename W, bool ALIGNED>
__global__ void __launch_bounds__(TI_THREADS, 1) k_ti_generate(const TiParams P) {
	extern __shared__ __align__(256) uint8_t smem[];
	constexpr int w = sizeof(W);
	constexpr int EPB = 16 / w; // elements per AES block
	uint8_t *sT = smem + aes_table_offset(smem, 0, 0);
	aes_load_tables(sT, P.T0);
	__syncthreads();
	const uint32_t lane = threadIdx.x & 31, lb = aes_lane_base(sT, lane);
	const uint32_t warps_total = gridDim.x * (blockDim.x >> 5);
	const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint32_t units = P.count * P.units_per_pair;   // < 2^32, checked by the launcher
	const bool need_x = P.x || P.xy, need_y = P.y || P.xy;
	const uint64_t nbytes = P.n_rows * w;        // bytes of one mask on this shard
	const uint64_t ydelta = P.n_total * w;       // stream distance x -> y
	const uint32_t ysh = (uint32_t) (ydelta & 15);
	AesCtrGroup gx, gy, gy1;                     // round-1/2 caches of the streams this lane walks
	gx.id = gy.id = gy1.id = 0xffffffffu;

	for(uint32_t u = gw; u < units; u += warps_total) {
		const uint32_t p = u / P.units_per_pair, uu = u - p * P.units_per_pair;
		const uint64_t W0 = (P.q0 + p) * (2 * P.n_total + 1);
		const uint64_t X0 = (W0 + P.row0) * w;       // stream byte of this shard's x[0]
		const uint64_t b_first = X0 >> 4;
		W *xo = P.x ? (W *) P.x + (uint64_t) p * P.ld : nullptr;
		W *yo = P.y ? (W *) P.y + (uint64_t) p * P.ld : nullptr;
		uint64_t dot = 0;
#pragma unroll 1
		for(int i = 0; i < TI_U; i++) {
			// units are the 256-aligned block groups of the x stream that overlap the shard's mask
			const uint64_t B = (b_first & ~(uint64_t) 255) + (uint64_t) uu * (32 * TI_U) + i * 32 + lane;
			const int64_t o = (int64_t) (B << 4) - (int64_t) X0; // byte offset of this block in the shard's mask
			// uniform over the warp: all 32 blocks lie inside the shard's mask (voted before any lane drops out)
			const bool full = __all_sync(0xffffffffu, o >= 0 && o + 16 <= (int64_t) nbytes);
			if(o <= -16 || o >= (int64_t) nbytes) continue;      // before the start / past the end
			uint32_t xs[4] = {0, 0, 0, 0}, ys[4] = {0, 0, 0, 0};
			if(need_x) aes_ctr_block_g(P.kx, lb, B, gx, xs);
			if(need_y) {
				const uint64_t By = ((B << 4) + ydelta) >> 4;
				if(ALIGNED) {
					aes_ctr_block_g(P.ky, lb, By, gy, ys);
				} else {
					// y bytes for these positions straddle two stream blocks
					uint32_t ya[4], yb[4];
					aes_ctr_block_g(P.ky, lb, By, gy, ya);
					aes_ctr_block_g(P.ky, lb, By + 1, gy1, yb);
					const uint32_t ws = ysh >> 2; // whole 32-bit words (ysh is a multiple of w >= 4)
#pragma unroll
					for(int c = 0; c < 4; c++) {
						uint32_t idx = c + ws;
						uint32_t v = 0;
#pragma unroll
						for(int z = 0; z < 8; z++) if(idx == (uint32_t) z) v = z < 4 ? ya[z & 3] : yb[z & 3];
						ys[c] = v;
					}
				}
			}
			// the warp's 32 blocks are 512 consecutive mask bytes: when all of them lie inside the shard
			// they leave as aligned 16-byte stores (full sectors), whatever the phase of the mask in
			// the stream (blocks of an odd pair start 8 bytes into a 16-byte line)
			if(full) {
#pragma unroll
				for(int e = 0; e < EPB; e++) {
					if(w == 8) dot += (((uint64_t) xs[2 * e + 1] << 32) | xs[2 * e]) * (((uint64_t) ys[2 * e + 1] << 32) | ys[2 * e]);
					else dot += (uint64_t) xs[e] * (uint64_t) ys[e];
				}
				const uint32_t sh = (16u - ((uint32_t) o & 15u)) & 15u;   // uniform: bytes to the next 16-byte line
				if(xo) ti_store_row(reinterpret_cast<char *>(xo) + o, xs, sh, lane);
				if(yo) ti_store_row(reinterpret_cast<char *>(yo) + o, ys, sh, lane);
				continue;
			}
#pragma unroll
			for(int e = 0; e < EPB; e++) {
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
				const uint64_t k = (uint64_t) ob / w;
				if(xo) xo[k] = xv;
				if(yo) yo[k] = yv;
				dot += (uint64_t) xv * (uint64_t) yv;
			}
		}
		if(P.xy) {
			dot = p1_warp_sum(dot);
			if(lane == 0 && dot) p1_red_add(P.xy + p, dot);
		}
		if(P.r && uu == 0) {
			// r = S[W0 + 2 n_total] (src/phase1.c:273)
			const uint64_t Rb = (W0 + 2 * P.n_total) * w;
			uint32_t rs[4];
			aes_ctr_block(P.kr, lb, Rb >> 4, rs);
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
}

template <typename W>
static int launch_ti(p1_ctx *ctx, const TiParams &P, bool aligned) {
	const size_t smem = AES_SMEM_BYTES;
	if(aligned) P1_CUDA(cudaFuncSetAttribute(k_ti_generate<W, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	else P1_CUDA(cudaFuncSetAttribute(k_ti_generate<W, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	// persistent: one 1024-thread CTA per SM (the 128 KB table allows no more)
	uint64_t units = (uint64_t) P.count * P.units_per_pair;
	P1_CHECK(units < 0xffffffffull, "p1_ti_generate: batch too large (%llu work units)", (unsigned long long) units);
	const unsigned threads = ctx->opt_ti_threads ? (unsigned) ctx->opt_ti_threads : TI_THREADS_DEFAULT;
	uint64_t grid = (uint64_t) ctx->sm_count;
	uint64_t need = (units + threads / 32 - 1) / (threads / 32);
	if(grid > need) grid = need;
	p1_prof_begin(ctx, P1_K_TI);
	if(aligned) k_ti_generate<W, true><<<(unsigned) grid, threads, smem, ctx->stream>>>(P);
	else k_ti_generate<W, false><<<(unsigned) grid, threads, smem, ctx->stream>>>(P);
	P1_LAUNCHED(ctx);
	return 0;
}

extern "C" int p1_ti_generate_keys(p1_ctx *ctx, const uint8_t key_x[16], const uint8_t key_y[16], const uint8_t key_r[16],
	uint64_t q0, size_t count, size_t n_total, size_t row0, size_t n_rows, void *x_dev, void *y_dev, size_t ld_words,
	uint64_t *xy_dev, uint64_t *r_dev) {
	P1_CHECK(ctx, "null context");
	std::lock_guard<std::mutex> lock_(ctx->mu);
	P1_CUDA(cudaSetDevice(ctx->device));
	if(!count) return 0;
	const bool need_x = x_dev || xy_dev, need_y = y_dev || xy_dev;
	P1_CHECK((!need_x || key_x) && (!need_y || key_y) && (!r_dev || key_r), "p1_ti_generate: null key for a requested output");
	P1_CHECK(n_rows > 0 && row0 + n_rows <= n_total, "p1_ti_generate: rows [%zu,%zu) outside [0,%zu)", row0, row0 + n_rows, n_total);
	P1_CHECK((!x_dev && !y_dev) || ld_words >= n_rows, "p1_ti_generate: ld_words %zu < n_rows %zu", ld_words, n_rows);
	P1_CHECK(count <= 0xffffffffu, "p1_ti_generate: batch too large");
	const int w = ctx->bits_p1 / 8;
	TiParams P;
	if(aes_device_T0(ctx, &P.T0)) return 1;
	// most of a launch shares the high counter word of its first stream block
	const uint32_t hi = (uint32_t) ((((uint64_t) q0 * (2 * (uint64_t) n_total + 1) + row0) * w) >> 36);
	static const uint8_t zero_key[16] = {0};
	aes_host_ctr_key(key_x ? key_x : zero_key, hi, &P.kx);
	if(key_y == key_x || (key_x && key_y && !memcmp(key_x, key_y, 16))) P.ky = P.kx;
	else aes_host_ctr_key(key_y ? key_y : zero_key, hi, &P.ky);
	if(key_r == key_x || (key_x && key_r && !memcmp(key_x, key_r, 16))) P.kr = P.kx;
	else aes_host_ctr_key(key_r ? key_r : zero_key, hi, &P.kr);
	P.q0 = q0;
	P.count = (uint32_t) count;
	P.n_total = n_total;
	P.row0 = row0;
	P.n_rows = n_rows;
	P.x = x_dev;
	P.y = y_dev;
	P.ld = ld_words;
	P.xy = xy_dev;
	P.r = r_dev;
	// stream blocks that can overlap one shard mask: bytes + up to 15 of misalignment
	uint64_t nblk = ((uint64_t) n_rows * w + 15 + 15) / 16;
	// 256-aligned groups of stream blocks that can overlap the mask (its first block sits anywhere in a group)
	P.units_per_pair = (uint32_t) ((nblk + 255 + 32 * TI_U - 1) / (32 * TI_U));
	if(xy_dev) P1_CUDA(cudaMemsetAsync(xy_dev, 0, count * sizeof(uint64_t), ctx->stream));
	bool aligned = ((uint64_t) n_total * w) % 16 == 0;
	if(w == 8) return launch_ti<uint64_t>(ctx, P, aligned);
	return launch_ti<uint32_t>(ctx, P, aligned);
}

extern "C" int p1_ti_generate(p1_ctx *ctx, const uint8_t key[16], uint64_t q0, size_t count,
	size_t n_total, size_t row0, size_t n_rows, void *x_dev, void *y_dev, size_t ld_words,
	uint64_t *xy_dev, uint64_t *r_dev) {
	P1_CHECK(ctx, "null context");
	P1_CHECK(key || !count, "p1_ti_generate: null key");
	return p1_ti_generate_keys(ctx, key, key, key, q0, count, n_total, row0, n_rows, x_dev, y_dev, ld_words, xy_dev, r_dev);
}
