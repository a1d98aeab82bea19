// Internal declarations shared by the translation units of libp1gpu.so.
// Public contract: include/p1gpu.h.  Target: sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <mutex>
#include <vector>
#include "../../include/p1gpu.h"

#define P1_SQ_LIMIT_BITS 26   // |word| < 2^26  =>  v*v exact in binary64 (p1gpu.h, p1_finalize)

struct p1_ctx {
	int device;
	int bits_p1, bits_p2;
	int sm_count;
	cudaStream_t stream;
	cudaEvent_t ev;
	cudaEvent_t marks[8];               // p1_ctx_mark / p1_ctx_mark_sync
	std::mutex mu;
	uint64_t launches;
	// small pinned + device scratch for index arrays passed from the host
	void *pin;      size_t pin_bytes;
	void *scratch;  size_t scratch_bytes;
	// staging for p1_encode from host memory (two buffers, events)
	void *stage[2]; size_t stage_bytes; cudaEvent_t stage_ev[2];
	// tcgen05 Gram workspace (limb planes), grown on demand
	void *limb_ws;  size_t limb_ws_bytes;
	// tile sums / offsets of the wire-format kernels (k_wire.cu), grown on demand
	void *wire_ws;  size_t wire_ws_bytes;
	long opt_tc_max_slices, opt_tc_chunk_blocks, opt_tc_min_cols, opt_tc_whole, opt_fused_fast, opt_fused_chunk;
	long opt_ti_threads;              // threads per CTA of k_ti_generate (0 = default)
	long opt_fused_threads;           // threads per CTA of k_masking_fused (0 = default)
	long opt_encode_async;            // encode through the cp.async pipeline (k_encode_async) when the layout allows
	long opt_encode_planes;           // p1_encode also writes the digit planes of the tcgen05 Gram (0 = off)
	// per-kernel-class event timing (p1_ctx_profile)
	bool prof;
	int prof_open;                       // class of the launch being bracketed, -1 if none
	cudaEvent_t prof_a;
	struct ProfRec { int cls; cudaEvent_t a, b; };
	std::vector<ProfRec> prof_recs;
	std::vector<cudaEvent_t> prof_pool;
	double prof_ms[P1_K_COUNT];
	uint64_t prof_n[P1_K_COUNT];
};

// retired blocks kept per device / for page-locked host memory (p1_api.cu).  take: NULL if the driver has no
// memory left even after the kept blocks were released; give: the caller guarantees nothing on the device still
// uses the block (it synchronised, as a cudaFree would have)
void *p1_pool_dev_take(int device, size_t bytes);
void p1_pool_dev_give(int device, void *p, size_t bytes);
void p1_pool_dev_give_secret(int device, void *p, size_t bytes);   // zeroed first (P1GPU_POOL_SCRUB, p1_api.cu)
void *p1_pool_host_take(size_t bytes);
void p1_pool_host_give(void *p, size_t bytes);

// call right before a kernel launch; P1_LAUNCHED closes the bracket
void p1_prof_begin(p1_ctx *ctx, int cls);
void p1_prof_end(p1_ctx *ctx);

struct p1_data {
	int device, bits;
	size_t n, d;        // rows, feature columns
	bool has_y;         // column d holds the target
	size_t ld;          // words between consecutive columns (n rounded up to 256)
	void *xt;           // (d+1) x ld ring words, column-major; rows >= n are zero
	uint64_t *stats;    // (d+1) x 4 : sq_lo, sq_hi, viol, maxabs
	size_t bytes, stat_bytes;   // allocation sizes of xt and stats
	// digit planes written by the encode pass for the tcgen05 Gram (k_gram_tc.cu): P[row/128][l][column 0..d][128],
	// Lg balanced base-256 digits per word; NULL when the encode pass did not produce them
	uint8_t *planes; int Lg; size_t plane_bytes;
};

// error plumbing (check() of the reference: message to stderr, return 1)
int p1_fail(const char *fmt, ...);
#define P1_CHECK(cond, ...) do { if(!(cond)) return p1_fail(__VA_ARGS__); } while(0)
#define P1_CUDA(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) \
	return p1_fail("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while(0)
#define P1_LAUNCHED(ctx) do { (ctx)->launches++; p1_prof_end(ctx); cudaError_t e_ = cudaGetLastError(); if(e_ != cudaSuccess) \
	return p1_fail("%s:%d: kernel launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); } while(0)

// every entry point: context checked, calls on one context serialised, device selected
#define P1_ENTER(ctx) P1_CHECK(ctx, "null context"); std::lock_guard<std::mutex> lock_((ctx)->mu); P1_CUDA(cudaSetDevice((ctx)->device))

// stream-ordered temporary (cudaMallocAsync) released on every return path of an entry point
struct p1_tmp {
	void *p = nullptr;
	cudaStream_t st = nullptr;
	cudaError_t alloc(size_t bytes, cudaStream_t stream) { st = stream; return cudaMallocAsync(&p, bytes, stream); }
	~p1_tmp() { if(p) cudaFreeAsync(p, st); }
};

static inline size_t p1_round_up(size_t v, size_t m) { return (v + m - 1) / m * m; }
static inline size_t p1_nA(size_t d) { return d * (d + 1) / 2; }

// host or device pointer?
bool p1_is_device_ptr(const void *p);
// returns a device copy of a small host array (or the pointer itself if already on the device)
int p1_small_to_dev(p1_ctx *ctx, const void *src, size_t bytes, size_t slot, const void **dev);

// ---- launchers implemented in the kernel translation units ----
int p1_tc_digits(uint64_t maxabs, int ring_bytes);
int p1_launch_encode(p1_ctx *ctx, const double *X_dev, size_t rows, size_t d, size_t row_off,
	const double *y_dev, double normalizer, int precision, p1_data *out);
int p1_launch_words_in(p1_ctx *ctx, const void *Xw_dev, const void *yw_dev, p1_data *out);
int p1_launch_words_out(p1_ctx *ctx, const p1_data *data, void *Xw_dev, void *yw_dev);
int p1_launch_gram_imad(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, uint64_t *acc_dev);
int p1_launch_gram_tc(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, uint64_t *acc_dev);
int p1_launch_diag_stats(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, uint64_t *acc_dev);
int p1_launch_diag_seq(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int precision, double *xy_dev);
int p1_launch_finalize(p1_ctx *ctx, const uint64_t *acc_dev, size_t num, size_t d, const int32_t *col_start_dev, int precision,
	int precision_p2, double normalizer2, const double *diag_xy_dev, int32_t *need_dev, uint64_t *out_dev);

// device helpers
#ifdef __CUDACC__
template <typename W> struct p1_word;
template <> struct p1_word<uint64_t> { typedef int64_t s; static constexpr int bits = 64; };
template <> struct p1_word<uint32_t> { typedef int32_t s; static constexpr int bits = 32; };

__device__ __forceinline__ uint64_t p1_warp_sum(uint64_t v) {
#pragma unroll
	for(int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}
__device__ __forceinline__ void p1_red_add(uint64_t *p, uint64_t v) {
	atomicAdd(reinterpret_cast<unsigned long long *>(p), static_cast<unsigned long long>(v));
}
// exclusive scan of one 32-bit value per thread over a 256-thread CTA (s_warp: 9 words of shared
// memory); *total = sum over the CTA.  Ends with a barrier, so s_warp may be reused at once.
__device__ __forceinline__ uint32_t p1_block_excl_scan256(uint32_t v, uint32_t *s_warp, uint32_t *total) {
	const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for(int o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
		if(lane >= (uint32_t) o) inc += t;
	}
	if(lane == 31) s_warp[wid] = inc;
	__syncthreads();
	if(threadIdx.x == 0) {
		uint32_t run = 0;
		for(int i = 0; i < 8; i++) { const uint32_t c = s_warp[i]; s_warp[i] = run; run += c; }
		s_warp[8] = run;
	}
	__syncthreads();
	const uint32_t r = s_warp[wid] + inc - v;
	*total = s_warp[8];
	__syncthreads();
	return r;
}
#endif
