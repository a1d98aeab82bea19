// Dense int8 tensor-core peak probe: MEASURED_PEAKS.json has no int8 entry, so the
// roofline of the tcgen05 Gram (k_gram_tc.cu) is taken against this kernel's rate, measured on the same
// box in the same run.  Same instruction and operand path as the Gram kernel - tcgen05.mma
// cta_group::1 kind::i8, M = 128, K = 32 per instruction, both operands K-major SWIZZLE_128B tiles in
// shared memory, s32 accumulators in TMEM - but nothing else: one CTA per SM, operands written once
// (no TMA, no HBM traffic, no epilogue), one thread issuing back-to-back MMAs into two alternating
// TMEM regions.  What it reports is therefore the ceiling a kernel of this shape can reach at the
// clocks the chip holds under this load.
#include "p1_common.cuh"
#define TC_BM 128
#include "tc_ptx.cuh"

#define PROBE_THREADS 128
#define PROBE_KBLK 128            // bytes of K per operand row in smem (one swizzle row) = 4 MMAs of K = 32

// N = accumulator columns per MMA (64 .. 256)
__global__ void __launch_bounds__(PROBE_THREADS, 1) k_probe_i8(int N, int iters, uint32_t seed, uint32_t *sink) {
	extern __shared__ __align__(1024) uint8_t smem_raw[];
	__shared__ __align__(8) uint64_t bar_done;
	__shared__ uint32_t tmem_base_s;
	uint8_t *smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	// operands: A 128 x 128 bytes, B 256 x 128 bytes of pseudo-random s8 (zeros would flatter the clocks)
	uint32_t *w32 = reinterpret_cast<uint32_t *>(smem);
	uint32_t st = seed ^ (blockIdx.x * 2654435761u) ^ (threadIdx.x * 40503u);
	for(int i = threadIdx.x; i < (128 + 256) * PROBE_KBLK / 4; i += blockDim.x) {
		st = st * 1664525u + 1013904223u;
		w32[i] = st ^ (st >> 13);
	}
	if(threadIdx.x == 0) {
		mbar_init(&bar_done, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if(warp == 0) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the tensor core
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = tmem_base_s;
	if(warp == 0 && lane == 0) {
		const uint32_t a_lo = ((smem_u32(smem) >> 4) & 0x3fffu) | (1u << 16);
		const uint32_t b_lo = a_lo + ((128 * PROBE_KBLK) >> 4);
		const uint32_t idesc = tc_idesc(N);
		const uint32_t span = (uint32_t) N <= 256u ? 256u : 512u;
		for(int it = 0; it < iters; it++) {
			const uint32_t d = tmem + ((it & 1) ? span : 0u);       // two accumulator regions, so consecutive MMAs do not chain on one
#pragma unroll
			for(int k4 = 0; k4 < PROBE_KBLK / 32; k4++)
				tc_mma_i8(d, tc_desc(a_lo + (uint32_t) (k4 * 2)), tc_desc(b_lo + (uint32_t) (k4 * 2)), idesc, it > 1 || k4 > 0 ? 1u : 0u);
		}
		tc_commit(&bar_done);
		mbar_wait(&bar_done, 0);
		tc_fence_after();
	}
	__syncthreads();
	if(warp == 0) {
		// keep the accumulators observable
		uint32_t v;
		asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(tmem));
		asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
		if(lane == 0 && sink) atomicAdd(sink, v);
		tc_fence_before();
		__syncwarp();
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
	}
}

extern "C" int p1_probe_tensor_i8(p1_ctx *ctx, int n_cols, double min_ms, double *tops, double *ms_out, uint64_t *ops_out) {
	P1_ENTER(ctx);
	P1_CHECK(n_cols >= 16 && n_cols <= 256 && n_cols % 16 == 0, "p1_probe_tensor_i8: N must be a multiple of 16 in [16,256]");
	const size_t smem = (128 + 256) * PROBE_KBLK + 1024;
	P1_CUDA(cudaFuncSetAttribute(k_probe_i8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
	uint32_t *sink = (uint32_t *) ctx->scratch;
	cudaEvent_t a, b;
	P1_CUDA(cudaEventCreate(&a));
	P1_CUDA(cudaEventCreate(&b));
	int iters = 2000;
	float ms = 0;
	for(int attempt = 0; attempt < 8; attempt++) {
		k_probe_i8<<<ctx->sm_count, PROBE_THREADS, smem, ctx->stream>>>(n_cols, 200, 1u, sink);   // warm-up
		P1_CUDA(cudaEventRecord(a, ctx->stream));
		k_probe_i8<<<ctx->sm_count, PROBE_THREADS, smem, ctx->stream>>>(n_cols, iters, 7u + attempt, sink);
		ctx->launches += 2;
		P1_CUDA(cudaEventRecord(b, ctx->stream));
		P1_CUDA(cudaEventSynchronize(b));
		P1_CUDA(cudaGetLastError());
		P1_CUDA(cudaEventElapsedTime(&ms, a, b));
		if(ms >= min_ms || iters >= (1 << 26)) break;
		double scale = ms > 0.01 ? min_ms / ms * 1.2 : 16.0;
		if(scale > 16.0) scale = 16.0;
		iters = (int) (iters * scale) + 1;
	}
	cudaEventDestroy(a);
	cudaEventDestroy(b);
	const double ops = 2.0 * 128.0 * n_cols * 32.0 * (PROBE_KBLK / 32) * (double) iters * ctx->sm_count;
	if(tops) *tops = ops / (ms * 1e-3) / 1e12;
	if(ms_out) *ms_out = ms;
	if(ops_out) *ops_out = (uint64_t) ops;
	return 0;
}
