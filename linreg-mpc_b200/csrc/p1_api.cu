// C-ABI entry points of libp1gpu.so (include/p1gpu.h): context, memory, encode
// orchestration, local-block dispatch, finalize.  Kernels live in the k_*.cu files.
#include "p1_common.cuh"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <unordered_map>

static thread_local char g_err[512] = "";

int p1_fail(const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
	fprintf(stderr, "p1gpu: %s\n", g_err);
	return 1;
}

extern "C" const char *p1_last_error(void) { return g_err; }

bool p1_is_device_ptr(const void *p) {
	cudaPointerAttributes a;
	if(cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static const size_t SMALL_SLOTS = 4;

int p1_small_to_dev(p1_ctx *ctx, const void *src, size_t bytes, size_t slot, const void **dev) {
	if(p1_is_device_ptr(src)) {
		*dev = src;
		return 0;
	}
	size_t per = ctx->pin_bytes / SMALL_SLOTS;
	P1_CHECK(bytes <= per && slot < SMALL_SLOTS, "index array of %zu bytes exceeds the %zu-byte staging slot; upload it once and pass the device pointer", bytes, per);
	// the pinned slot may still be in flight from an earlier call: wait for the stream
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	char *h = (char *) ctx->pin + slot * per, *d = (char *) ctx->scratch + slot * per;
	memcpy(h, src, bytes);
	P1_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
	*dev = d;
	return 0;
}

// Freed blocks are kept for the next request of the same size (device: per device; host: page-locked), up to
// P1_POOL_MAX bytes each: a party that serves one run after another - and every test - asks for the same
// buffers again, and cudaHostAlloc / cudaMalloc / cudaFree cost milliseconds per 100 MB and synchronise the device.
// Caps and scrubbing are process-wide settings read once from the environment:
//   P1GPU_POOL_MAX_MB  bytes kept per device and of page-locked memory, in MB (default 8192; 0 = keep nothing, every
//                      free goes back to the driver)
//   P1GPU_POOL_SCRUB   1 (default): device blocks are zeroed before they are kept or released - OT round keys, pads,
//                      TI masks and whatever a caller's p1_dev_alloc block held do not outlive their owner - and so are
//                      page-locked blocks of up to 1 MB (keys and values of seeded-mask messages); 2: page-locked blocks
//                      of any size as well (a host memset of the whole wire buffer per free); 0: nothing is zeroed
static size_t pool_env(const char *name, size_t dflt) {
	const char *v = getenv(name);
	if(!v || !*v) return dflt;
	char *end = NULL;
	const unsigned long long x = strtoull(v, &end, 10);
	return end && *end == 0 ? (size_t) x : dflt;
}
static size_t pool_max_bytes() {
	static const size_t v = pool_env("P1GPU_POOL_MAX_MB", 8192) << 20;
	return v;
}
static int pool_scrub_level() {
	static const int v = (int) pool_env("P1GPU_POOL_SCRUB", 1);
	return v;
}
#define P1_POOL_MAX pool_max_bytes()
struct BlockPool {
	std::mutex mu;
	std::multimap<size_t, void *> blocks;
	size_t held = 0;
	void *take(size_t bytes) {
		std::lock_guard<std::mutex> l(mu);
		auto it = blocks.find(bytes);
		if(it == blocks.end()) return NULL;
		void *p = it->second;
		blocks.erase(it);
		held -= bytes;
		return p;
	}
	bool give(size_t bytes, void *p) {
		std::lock_guard<std::mutex> l(mu);
		if(held + bytes > P1_POOL_MAX) return false;
		blocks.emplace(bytes, p);
		held += bytes;
		return true;
	}
	template <typename F> void drain(F release) {
		std::lock_guard<std::mutex> l(mu);
		for(auto &b : blocks) release(b.second);
		blocks.clear();
		held = 0;
	}
};
static BlockPool g_dev_pool[64], g_host_pool;
static std::mutex g_size_mu;
static std::unordered_map<const void *, size_t> g_block_size;   // bytes of every live p1_dev_alloc / p1_host_alloc block

static void block_size_set(const void *p, size_t bytes) {
	std::lock_guard<std::mutex> l(g_size_mu);
	g_block_size[p] = bytes;
}
static size_t block_size_take(const void *p) {
	std::lock_guard<std::mutex> l(g_size_mu);
	auto it = g_block_size.find(p);
	if(it == g_block_size.end()) return 0;
	size_t b = it->second;
	g_block_size.erase(it);
	return b;
}

extern "C" void p1_pool_trim(void) {
	int cur = 0;
	cudaGetDevice(&cur);
	for(int d = 0; d < 64; d++) {
		bool any;
		{
			std::lock_guard<std::mutex> l(g_dev_pool[d].mu);
			any = !g_dev_pool[d].blocks.empty();
		}
		if(!any) continue;
		cudaSetDevice(d);
		g_dev_pool[d].drain([](void *p) { cudaFree(p); });
	}
	cudaSetDevice(cur);
	g_host_pool.drain([](void *p) { cudaFreeHost(p); });
}

void *p1_pool_dev_take(int device, size_t bytes) {
	if(!bytes) bytes = 1;
	void *p = g_dev_pool[device & 63].take(bytes);
	if(p) return p;
	if(cudaMalloc(&p, bytes) == cudaSuccess) return p;
	cudaGetLastError();
	p1_pool_trim();                     // the retired blocks may be what is in the way
	if(cudaMalloc(&p, bytes) == cudaSuccess) return p;
	cudaGetLastError();
	return NULL;
}
void p1_pool_dev_give(int device, void *p, size_t bytes) {
	if(!p) return;
	if(!bytes) bytes = 1;
	if(!g_dev_pool[device & 63].give(bytes, p)) cudaFree(p);
}
// the same for a block that held key, pad or mask material: zeroed first (device `device` is current and idle, as for
// p1_pool_dev_give; the memset runs on the legacy stream and is waited for, so the next owner on any stream sees zeros)
void p1_pool_dev_give_secret(int device, void *p, size_t bytes) {
	if(!p) return;
	if(pool_scrub_level() >= 1 && bytes) {
		if(cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) cudaGetLastError();
	}
	p1_pool_dev_give(device, p, bytes);
}
void *p1_pool_host_take(size_t bytes) {
	if(!bytes) bytes = 1;
	void *p = g_host_pool.take(bytes);
	if(p) return p;
	if(cudaHostAlloc(&p, bytes, cudaHostAllocPortable) == cudaSuccess) return p;
	cudaGetLastError();
	return NULL;
}
void p1_pool_host_give(void *p, size_t bytes) {
	if(!p) return;
	if(!bytes) bytes = 1;
	if(!g_host_pool.give(bytes, p)) cudaFreeHost(p);
}

extern "C" int p1_ctx_create(int device, int ring_bits_p1, int ring_bits_p2, p1_ctx **out) {
	P1_CHECK(out, "p1_ctx_create: out may not be null");
	*out = NULL;
	P1_CHECK((ring_bits_p1 == 64 && (ring_bits_p2 == 64 || ring_bits_p2 == 32)) || (ring_bits_p1 == 32 && ring_bits_p2 == 32),
		"unsupported ring widths (%d,%d): supported are (64,64), (32,32), (64,32)", ring_bits_p1, ring_bits_p2);
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	P1_CHECK(e == cudaSuccess && count > 0, "no CUDA device available (%s); libp1gpu has no CPU fallback", cudaGetErrorString(e));
	P1_CHECK(device >= 0 && device < count, "device %d out of range (%d devices)", device, count);
	P1_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	P1_CUDA(cudaGetDeviceProperties(&prop, device));
	P1_CHECK(prop.major == 10, "device %d is sm_%d%d; libp1gpu is built for sm_100a only", device, prop.major, prop.minor);
	{
		// temporaries come from the device's stream-ordered pool (p1_tmp): let it keep what it has handed out once
		// instead of returning the memory to the driver at every synchronisation
		cudaMemPool_t mp;
		uint64_t keep = UINT64_MAX;
		if(cudaDeviceGetDefaultMemPool(&mp, device) == cudaSuccess) cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &keep);
		cudaGetLastError();
	}
	p1_ctx *ctx = new p1_ctx();
	ctx->device = device;
	ctx->bits_p1 = ring_bits_p1;
	ctx->bits_p2 = ring_bits_p2;
	ctx->sm_count = prop.multiProcessorCount;
	ctx->launches = 0;
	for(int i = 0; i < 8; i++) ctx->marks[i] = NULL;
	ctx->pin_bytes = ctx->scratch_bytes = 4u << 20;
	ctx->stage[0] = ctx->stage[1] = NULL;
	ctx->stage_bytes = 0;
	ctx->limb_ws = NULL;
	ctx->limb_ws_bytes = 0;
	ctx->wire_ws = NULL;
	ctx->wire_ws_bytes = 0;
	ctx->opt_tc_max_slices = ctx->opt_tc_chunk_blocks = ctx->opt_tc_min_cols = 0;
	ctx->opt_tc_whole = 1;
	ctx->opt_ti_threads = 0;
	ctx->opt_fused_threads = 0;
	ctx->opt_fused_fast = 1;
	ctx->opt_fused_chunk = 0;
	ctx->opt_encode_planes = 0;
	ctx->opt_encode_async = 1;
	ctx->prof = false;
	ctx->prof_open = -1;
	for(int i = 0; i < P1_K_COUNT; i++) { ctx->prof_ms[i] = 0; ctx->prof_n[i] = 0; }
	if(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
	   cudaEventCreateWithFlags(&ctx->ev, cudaEventDisableTiming) != cudaSuccess ||
	   !(ctx->pin = p1_pool_host_take(ctx->pin_bytes)) ||
	   !(ctx->scratch = p1_pool_dev_take(device, ctx->scratch_bytes))) {
		delete ctx;
		return p1_fail("p1_ctx_create: %s", cudaGetErrorString(cudaGetLastError()));
	}
	*out = ctx;
	return 0;
}

extern "C" void p1_ctx_destroy(p1_ctx *ctx) {
	if(!ctx) return;
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	for(int i = 0; i < 2; i++) {
		if(ctx->stage[i]) {
			p1_pool_dev_give(ctx->device, ctx->stage[i], ctx->stage_bytes);
			cudaEventDestroy(ctx->stage_ev[i]);
		}
	}
	p1_pool_dev_give(ctx->device, ctx->limb_ws, ctx->limb_ws_bytes);
	p1_pool_dev_give(ctx->device, ctx->wire_ws, ctx->wire_ws_bytes);
	for(auto &r : ctx->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
	for(auto e : ctx->prof_pool) cudaEventDestroy(e);
	p1_pool_dev_give(ctx->device, ctx->scratch, ctx->scratch_bytes);
	p1_pool_host_give(ctx->pin, ctx->pin_bytes);
	for(int i = 0; i < 8; i++) if(ctx->marks[i]) cudaEventDestroy(ctx->marks[i]);
	cudaEventDestroy(ctx->ev);
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}


extern "C" int p1_ctx_sync(p1_ctx *ctx) {
	P1_ENTER(ctx);
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

extern "C" int p1_ctx_wait(p1_ctx *waiter, p1_ctx *signaler) {
	P1_CHECK(waiter && signaler, "null context");
	if(waiter == signaler) return 0;
	{
		std::lock_guard<std::mutex> l(signaler->mu);
		P1_CUDA(cudaSetDevice(signaler->device));
		P1_CUDA(cudaEventRecord(signaler->ev, signaler->stream));
	}
	std::lock_guard<std::mutex> l(waiter->mu);
	P1_CUDA(cudaSetDevice(waiter->device));
	P1_CUDA(cudaStreamWaitEvent(waiter->stream, signaler->ev, 0));
	return 0;
}

extern "C" int p1_ctx_mark(p1_ctx *ctx, int slot) {
	P1_ENTER(ctx);
	P1_CHECK(slot >= 0 && slot < 8, "p1_ctx_mark: slot %d out of range [0,8)", slot);
	if(!ctx->marks[slot]) P1_CUDA(cudaEventCreateWithFlags(&ctx->marks[slot], cudaEventDisableTiming));
	P1_CUDA(cudaEventRecord(ctx->marks[slot], ctx->stream));
	return 0;
}

extern "C" int p1_ctx_mark_sync(p1_ctx *ctx, int slot) {
	P1_CHECK(ctx && slot >= 0 && slot < 8 && ctx->marks[slot], "p1_ctx_mark_sync: no mark %d on this context", slot);
	// no context lock: another thread may keep enqueueing work while this one waits
	P1_CUDA(cudaEventSynchronize(ctx->marks[slot]));
	return 0;
}

extern "C" int p1_ctx_set_option(p1_ctx *ctx, const char *name, long value) {
	P1_ENTER(ctx);
	P1_CHECK(name && value >= 0, "p1_ctx_set_option: invalid arguments");
	if(!strcmp(name, "gram_tc_max_slices")) ctx->opt_tc_max_slices = value;
	else if(!strcmp(name, "gram_tc_chunk_blocks")) ctx->opt_tc_chunk_blocks = value;
	else if(!strcmp(name, "gram_tc_min_cols")) ctx->opt_tc_min_cols = value;
	else if(!strcmp(name, "gram_tc_whole")) ctx->opt_tc_whole = value;
	else if(!strcmp(name, "fused_fast")) ctx->opt_fused_fast = value;
	else if(!strcmp(name, "fused_chunk")) {
		P1_CHECK(value >= 0 && value <= 4096, "fused_chunk must be in [0, 4096] (0 = default)");
		ctx->opt_fused_chunk = value;
	}
	else if(!strcmp(name, "ti_threads")) {
		P1_CHECK(value == 0 || (value % 32 == 0 && value >= 32 && value <= 512), "ti_threads must be a multiple of 32 in [32,512]");
		ctx->opt_ti_threads = value;
	}
	else if(!strcmp(name, "fused_threads")) {
		P1_CHECK(value == 0 || value == 384 || value == 512 || value == 640 || value == 768 || value == 1024, "fused_threads must be 0 (default), 384, 512, 640, 768 or 1024");
		ctx->opt_fused_threads = value;
	}
	else if(!strcmp(name, "encode_planes")) ctx->opt_encode_planes = value;
	else if(!strcmp(name, "encode_async")) ctx->opt_encode_async = value;
	else return p1_fail("p1_ctx_set_option: unknown option '%s'", name);
	return 0;
}

extern "C" void *p1_ctx_stream(p1_ctx *ctx) { return ctx ? (void *) ctx->stream : NULL; }
extern "C" uint64_t p1_ctx_launches(p1_ctx *ctx) { return ctx ? ctx->launches : 0; }

// ------------------------------------------------------------------ profiling

static cudaEvent_t prof_event(p1_ctx *ctx) {
	cudaEvent_t e = NULL;
	if(!ctx->prof_pool.empty()) {
		e = ctx->prof_pool.back();
		ctx->prof_pool.pop_back();
	} else {
		cudaEventCreate(&e);
	}
	return e;
}

void p1_prof_begin(p1_ctx *ctx, int cls) {
	if(!ctx->prof) return;
	ctx->prof_open = cls;
	ctx->prof_a = prof_event(ctx);
	cudaEventRecord(ctx->prof_a, ctx->stream);
}

void p1_prof_end(p1_ctx *ctx) {
	if(!ctx->prof || ctx->prof_open < 0) return;
	cudaEvent_t b = prof_event(ctx);
	cudaEventRecord(b, ctx->stream);
	ctx->prof_recs.push_back({ctx->prof_open, ctx->prof_a, b});
	ctx->prof_open = -1;
}

static const char *const K_NAMES[P1_K_COUNT] = {"encode", "layout", "diag", "gram_imad", "gram_limbs", "gram_tc", "ti_generate",
	"party_b_mask", "party_a_step", "party_b_share", "cot_recv", "cot_send", "cot_finish", "misc", "msg_pack", "msg_unpack", "share_bits", "text", "masking_fused"};

extern "C" const char *p1_kernel_class_name(int c) { return c >= 0 && c < P1_K_COUNT ? K_NAMES[c] : NULL; }

extern "C" int p1_ctx_profile(p1_ctx *ctx, int enable) {
	P1_ENTER(ctx);
	ctx->prof = enable != 0;
	ctx->prof_open = -1;
	return 0;
}

extern "C" int p1_ctx_profile_read(p1_ctx *ctx, int cls, double *ms_total, uint64_t *launches) {
	P1_ENTER(ctx);
	P1_CHECK(cls >= 0 && cls < P1_K_COUNT, "p1_ctx_profile_read: unknown kernel class %d", cls);
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	for(auto &r : ctx->prof_recs) {
		float ms = 0;
		if(cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
			ctx->prof_ms[r.cls] += ms;
			ctx->prof_n[r.cls]++;
		}
		ctx->prof_pool.push_back(r.a);
		ctx->prof_pool.push_back(r.b);
	}
	ctx->prof_recs.clear();
	if(ms_total) *ms_total = ctx->prof_ms[cls];
	if(launches) *launches = ctx->prof_n[cls];
	return 0;
}

extern "C" int p1_dev_alloc(p1_ctx *ctx, size_t bytes, void **dev) {
	P1_ENTER(ctx);
	P1_CHECK(dev, "null output");
	*dev = NULL;
	if(!bytes) bytes = 1;
	void *p = p1_pool_dev_take(ctx->device, bytes);
	P1_CHECK(p, "p1_dev_alloc: cannot allocate %zu bytes on device %d", bytes, ctx->device);
	block_size_set(p, bytes);
	*dev = p;
	return 0;
}
extern "C" int p1_dev_free(p1_ctx *ctx, void *dev) {
	P1_ENTER(ctx);
	if(!dev) return 0;
	P1_CUDA(cudaDeviceSynchronize());           // as cudaFree would: no stream of the device still uses the block
	const size_t bytes = block_size_take(dev);
	if(bytes) p1_pool_dev_give_secret(ctx->device, dev, bytes);    // callers keep masks and messages in these blocks
	else P1_CUDA(cudaFree(dev));
	return 0;
}
extern "C" int p1_dev_zero(p1_ctx *ctx, void *dev, size_t bytes) {
	P1_ENTER(ctx);
	P1_CUDA(cudaMemsetAsync(dev, 0, bytes, ctx->stream));
	return 0;
}
extern "C" int p1_host_alloc(p1_ctx *ctx, size_t bytes, void **host) {
	P1_ENTER(ctx);
	P1_CHECK(host, "p1_host_alloc: null output");
	*host = NULL;
	if(!bytes) bytes = 1;
	void *p = p1_pool_host_take(bytes);
	P1_CHECK(p, "p1_host_alloc: cannot page-lock %zu bytes", bytes);
	block_size_set(p, bytes);
	*host = p;
	return 0;
}
extern "C" int p1_host_free(p1_ctx *ctx, void *host) {
	P1_ENTER(ctx);
	if(!host) return 0;
	const size_t bytes = block_size_take(host);
	if(bytes && (pool_scrub_level() >= 2 || (pool_scrub_level() == 1 && bytes <= ((size_t) 1 << 20)))) {
		P1_CUDA(cudaStreamSynchronize(ctx->stream));              // no copy of this context still targets the block
		memset(host, 0, bytes);
	}
	if(bytes) p1_pool_host_give(host, bytes);
	else P1_CUDA(cudaFreeHost(host));
	return 0;
}
extern "C" int p1_copy_h2d(p1_ctx *ctx, void *dst, const void *src, size_t bytes) {
	P1_ENTER(ctx);
	P1_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
	return 0;
}
extern "C" int p1_copy_d2h_async(p1_ctx *ctx, void *dst, const void *src, size_t bytes) {
	P1_ENTER(ctx);
	P1_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	return 0;
}
extern "C" int p1_copy_d2h(p1_ctx *ctx, void *dst, const void *src, size_t bytes) {
	P1_ENTER(ctx);
	P1_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	P1_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

// ------------------------------------------------------------------ data

// The column store comes from, and goes back to, the per-device block pool: a provider re-encoding per run, every
// test and bench.py's first steps ask for the same multi-GB buffers again (cudaMalloc/cudaFree of those cost
// milliseconds and synchronise the device).
__global__ void k_zero_pads(unsigned char *__restrict__ xt, size_t col_bytes, size_t live_bytes, size_t cols) {
	// zero bytes [live_bytes, col_bytes) of every column (rows >= n), 16 bytes per thread
	const size_t pad = col_bytes - live_bytes, per = (pad + 15) / 16;
	size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if(t >= per * cols) return;
	size_t c = t / per, o = (t % per) * 16;
	unsigned char *p = xt + c * col_bytes + live_bytes + o;
	size_t k = pad - o < 16 ? pad - o : 16;
	for(size_t i = 0; i < k; i++) p[i] = 0;
}

static int data_alloc(p1_ctx *ctx, size_t n, size_t d, bool has_y, p1_data **out, int Lg = 0) {
	p1_data *dt = new p1_data();
	dt->device = ctx->device;
	dt->bits = ctx->bits_p1;
	dt->n = n;
	dt->d = d;
	dt->has_y = has_y;
	dt->ld = p1_round_up(n ? n : 1, 256);
	dt->xt = NULL;
	dt->stats = NULL;
	dt->planes = NULL;
	dt->Lg = 0;
	dt->plane_bytes = 0;
	const size_t w = (size_t) (dt->bits / 8);
	const size_t plane_bytes = (size_t) Lg * (d + 1) * dt->ld;
	size_t bytes = (d + 1) * dt->ld * w, stat_bytes = (d + 1) * 4 * sizeof(uint64_t);
	if(plane_bytes) dt->planes = (uint8_t *) p1_pool_dev_take(ctx->device, plane_bytes);   // no room for the planes: the Gram falls back to k_limb_split
	if(dt->planes) {
		dt->Lg = Lg;
		dt->plane_bytes = plane_bytes;
		// rows >= n of the last K-blocks must hold zero digits (an absent target column is never read:
		// p1_local_block refuses owns_target without one)
		const size_t kb0 = n / 128, per_kb = (size_t) Lg * (d + 1) * 128;
		if(kb0 * 128 < dt->ld) cudaMemsetAsync(dt->planes + kb0 * per_kb, 0, (dt->ld / 128 - kb0) * per_kb, ctx->stream);
	}
	dt->xt = p1_pool_dev_take(ctx->device, bytes);
	dt->stats = dt->xt ? (uint64_t *) p1_pool_dev_take(ctx->device, stat_bytes) : NULL;
	if(!dt->stats) {
		p1_pool_dev_give(ctx->device, dt->xt, bytes);
		p1_pool_dev_give(ctx->device, dt->planes, plane_bytes);
		delete dt;
		return p1_fail("p1_encode: cannot allocate %zu bytes on device %d", bytes, ctx->device);
	}
	dt->bytes = bytes;
	dt->stat_bytes = stat_bytes;
	// zero the padding rows of every column, the absent target column and the statistics
	if(dt->ld > n) {
		size_t per = ((dt->ld - n) * w + 15) / 16, threads = per * (d + 1);
		k_zero_pads<<<(unsigned) ((threads + 255) / 256), 256, 0, ctx->stream>>>((unsigned char *) dt->xt, dt->ld * w, n * w, d + 1);
	}
	if(!has_y) cudaMemsetAsync((char *) dt->xt + d * dt->ld * w, 0, dt->ld * w, ctx->stream);
	cudaMemsetAsync(dt->stats, 0, stat_bytes, ctx->stream);
	*out = dt;
	return 0;
}

extern "C" void p1_data_free(p1_data *dt) {
	if(!dt) return;
	cudaSetDevice(dt->device);
	cudaDeviceSynchronize();
	p1_pool_dev_give(dt->device, dt->xt, dt->bytes);
	p1_pool_dev_give(dt->device, dt->stats, dt->stat_bytes);
	p1_pool_dev_give(dt->device, dt->planes, dt->plane_bytes);
	delete dt;
}

extern "C" size_t p1_data_rows(const p1_data *dt) { return dt ? dt->n : 0; }
extern "C" size_t p1_data_cols(const p1_data *dt) { return dt ? dt->d : 0; }
extern "C" const void *p1_data_colmajor_dev(const p1_data *dt, size_t *ld) {
	if(!dt) return NULL;
	if(ld) *ld = dt->ld;
	return dt->xt;
}

static int ensure_stage(p1_ctx *ctx, size_t bytes) {
	if(ctx->stage_bytes >= bytes) return 0;
	for(int i = 0; i < 2; i++) {
		if(ctx->stage[i]) {
			P1_CUDA(cudaDeviceSynchronize());
			p1_pool_dev_give(ctx->device, ctx->stage[i], ctx->stage_bytes);
		} else {
			P1_CUDA(cudaEventCreateWithFlags(&ctx->stage_ev[i], cudaEventDisableTiming));
		}
		ctx->stage[i] = p1_pool_dev_take(ctx->device, bytes);
		if(!ctx->stage[i]) {
			ctx->stage_bytes = 0;
			return p1_fail("p1_encode: cannot allocate a %zu-byte staging buffer", bytes);
		}
	}
	ctx->stage_bytes = bytes;
	return 0;
}

// X / y (host or device) -> the column store of dt, on the context's stream
static int encode_fill(p1_ctx *ctx, p1_data *dt, const double *X, const double *y, size_t n, size_t d, double normalizer, int precision) {
	int st = 0;
	if(p1_is_device_ptr(X)) {
		st = p1_launch_encode(ctx, X, n, d, 0, NULL, normalizer, precision, dt);
	} else {
		// host rows -> device staging (double-buffered) -> encode+transpose kernel
		size_t row_bytes = d * sizeof(double);
		size_t chunk_rows = (256u << 20) / row_bytes;
		chunk_rows = chunk_rows / 256 * 256;
		if(chunk_rows < 256) chunk_rows = 256;
		if(chunk_rows > n) chunk_rows = n;
		st = ensure_stage(ctx, chunk_rows * row_bytes);
		int buf = 0;
		for(size_t r0 = 0; r0 < n && !st; r0 += chunk_rows, buf ^= 1) {
			size_t rows = n - r0 < chunk_rows ? n - r0 : chunk_rows;
			// stage[buf] was last read by the encode launched two chunks ago on the same stream: ordered
			if(cudaMemcpyAsync(ctx->stage[buf], X + r0 * d, rows * row_bytes, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) {
				st = p1_fail("p1_encode: H2D copy failed: %s", cudaGetErrorString(cudaGetLastError()));
				break;
			}
			st = p1_launch_encode(ctx, (const double *) ctx->stage[buf], rows, d, r0, NULL, normalizer, precision, dt);
		}
	}
	if(!st && y) {
		const double *y_dev = y;
		if(!p1_is_device_ptr(y)) {
			st = ensure_stage(ctx, n * sizeof(double));
			if(!st && cudaMemcpyAsync(ctx->stage[0], y, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) {
				st = p1_fail("p1_encode: H2D copy of y failed: %s", cudaGetErrorString(cudaGetLastError()));
			}
			y_dev = (const double *) ctx->stage[0];
		}
		if(!st) st = p1_launch_encode(ctx, NULL, n, d, 0, y_dev, normalizer, precision, dt);
	}
	return st;
}

// digit planes for the tensor-core Gram, written in the encode pass (saves k_limb_split's re-read of every
// word).  The digit count must be fixed before the data has been seen: enough for |value| <= 16 at this
// scale; if a block turns out to need more (statistics of this very pass), its Gram splits again itself.
static int encode_plane_digits(p1_ctx *ctx, size_t n, double normalizer, int precision) {
	int Lg = 0;
	if(ctx->opt_encode_planes && n >= 4096) {
		const double bound = 16.0 * (double) (1ll << precision) / normalizer;
		if(bound < 9.0e15) Lg = p1_tc_digits((uint64_t) bound + 1, ctx->bits_p1 / 8);
		if(Lg > 4) Lg = 0;
	}
	return Lg;
}

extern "C" int p1_encode(p1_ctx *ctx, const double *X, const double *y, size_t n, size_t d,
	double normalizer, int precision, p1_data **out) {
	P1_ENTER(ctx);
	P1_CHECK(out && X, "p1_encode: arguments may not be null");
	*out = NULL;
	P1_CHECK(n > 0 && d > 0, "p1_encode: empty matrix (%zu x %zu)", n, d);
	P1_CHECK(precision >= 0 && precision < 63, "p1_encode: precision %d out of range", precision);
	p1_data *dt = NULL;
	if(data_alloc(ctx, n, d, y != NULL, &dt, encode_plane_digits(ctx, n, normalizer, precision))) return 1;
	if(encode_fill(ctx, dt, X, y, n, d, normalizer, precision)) {
		p1_data_free(dt);
		return 1;
	}
	*out = dt;
	return 0;
}

extern "C" int p1_encode_again(p1_ctx *ctx, p1_data *dt, const double *X, const double *y, double normalizer, int precision) {
	P1_ENTER(ctx);
	P1_CHECK(dt && X, "p1_encode_again: arguments may not be null");
	P1_CHECK(dt->device == ctx->device && dt->bits == ctx->bits_p1, "p1_encode_again: the store belongs to another device or ring");
	P1_CHECK((y != NULL) == dt->has_y, "p1_encode_again: the store was allocated %s a target column", dt->has_y ? "with" : "without");
	P1_CHECK(precision >= 0 && precision < 63, "p1_encode_again: precision %d out of range", precision);
	P1_CHECK(!dt->planes || dt->Lg == encode_plane_digits(ctx, dt->n, normalizer, precision), "p1_encode_again: the scale needs another digit count than the store holds");
	P1_CUDA(cudaMemsetAsync(dt->stats, 0, dt->stat_bytes, ctx->stream));
	return encode_fill(ctx, dt, X, y, dt->n, dt->d, normalizer, precision);
}

extern "C" int p1_data_from_words(p1_ctx *ctx, const void *Xw, const void *yw, size_t n, size_t d, p1_data **out) {
	P1_ENTER(ctx);
	P1_CHECK(out && Xw, "p1_data_from_words: arguments may not be null");
	*out = NULL;
	P1_CHECK(n > 0 && d > 0, "p1_data_from_words: empty matrix");
	p1_data *dt = NULL;
	if(data_alloc(ctx, n, d, yw != NULL, &dt)) return 1;
	size_t w = (size_t) ctx->bits_p1 / 8;
	void *xd = NULL, *yd = NULL;
	bool xfree = false, yfree = false;
	int st = 0;
	if(p1_is_device_ptr(Xw)) xd = (void *) Xw;
	else {
		if(cudaMalloc(&xd, n * d * w) != cudaSuccess) st = p1_fail("p1_data_from_words: out of device memory");
		else {
			xfree = true;
			cudaMemcpyAsync(xd, Xw, n * d * w, cudaMemcpyHostToDevice, ctx->stream);
		}
	}
	if(!st && yw) {
		if(p1_is_device_ptr(yw)) yd = (void *) yw;
		else {
			if(cudaMalloc(&yd, n * w) != cudaSuccess) st = p1_fail("p1_data_from_words: out of device memory");
			else {
				yfree = true;
				cudaMemcpyAsync(yd, yw, n * w, cudaMemcpyHostToDevice, ctx->stream);
			}
		}
	}
	if(!st) st = p1_launch_words_in(ctx, xd, yd, dt);
	cudaStreamSynchronize(ctx->stream);
	if(xfree) cudaFree(xd);
	if(yfree) cudaFree(yd);
	if(st) {
		p1_data_free(dt);
		return 1;
	}
	*out = dt;
	return 0;
}

extern "C" int p1_data_download(p1_ctx *ctx, const p1_data *dt, void *Xw, void *yw) {
	P1_ENTER(ctx);
	P1_CHECK(dt, "null data");
	size_t w = (size_t) dt->bits / 8;
	void *xd = NULL, *yd = NULL;
	if(Xw) P1_CUDA(cudaMalloc(&xd, dt->n * dt->d * w));
	if(yw) P1_CUDA(cudaMalloc(&yd, dt->n * w));
	int st = p1_launch_words_out(ctx, dt, xd, yd);
	if(!st && Xw && cudaMemcpyAsync(Xw, xd, dt->n * dt->d * w, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) st = p1_fail("D2H failed");
	if(!st && yw && cudaMemcpyAsync(yw, yd, dt->n * w, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) st = p1_fail("D2H failed");
	cudaStreamSynchronize(ctx->stream);
	cudaFree(xd);
	cudaFree(yd);
	return st;
}

// ------------------------------------------------------------------ local block

extern "C" size_t p1_acc_len(size_t d) { return p1_nA(d) + 4 * d; }

extern "C" int p1_local_block(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target,
	uint64_t *acc_dev, int algo) {
	P1_ENTER(ctx);
	P1_CHECK(data && acc_dev, "p1_local_block: arguments may not be null");
	P1_CHECK(col_lo <= col_hi && col_hi <= data->d, "p1_local_block: invalid column range [%zu,%zu) for d=%zu", col_lo, col_hi, data->d);
	P1_CHECK(!owns_target || data->has_y, "p1_local_block: owns_target set but the data holds no target vector");
	P1_CHECK(data->bits == ctx->bits_p1, "p1_local_block: data was encoded for a %d-bit ring", data->bits);
	if(col_lo == col_hi) return 0;
	if(p1_launch_diag_stats(ctx, data, col_lo, col_hi, acc_dev)) return 1;
	size_t m = col_hi - col_lo + (owns_target ? 1 : 0);
	if(m < 2) return 0;
	if(algo == P1_GRAM_AUTO) {
		// crossover measured on B200 (profiles/r1_kbench.txt): at n = 1M the limb GEMM (digit split included) already
		// wins 3x at 32 columns; tiny blocks and tiny row counts stay on the CUDA-core kernel
		size_t min_cols = ctx->opt_tc_min_cols > 0 ? (size_t) ctx->opt_tc_min_cols : 16;
		algo = (m >= min_cols && data->n >= 4096) ? P1_GRAM_TCGEN05 : P1_GRAM_IMAD;
	}
	if(algo == P1_GRAM_TCGEN05) return p1_launch_gram_tc(ctx, data, col_lo, col_hi, owns_target, acc_dev);
	P1_CHECK(algo == P1_GRAM_IMAD, "p1_local_block: unknown algo %d", algo);
	return p1_launch_gram_imad(ctx, data, col_lo, col_hi, owns_target, acc_dev);
}

extern "C" int p1_diag_sequential(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int precision, double *xy_inout) {
	P1_ENTER(ctx);
	P1_CHECK(data && xy_inout, "p1_diag_sequential: arguments may not be null");
	P1_CHECK(col_lo <= col_hi && col_hi <= data->d, "p1_diag_sequential: invalid column range");
	size_t m = col_hi - col_lo;
	if(!m) return 0;
	double *xy_dev = NULL;
	P1_CUDA(cudaMalloc(&xy_dev, m * sizeof(double)));
	int st = 0;
	if(cudaMemcpyAsync(xy_dev, xy_inout, m * sizeof(double), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) st = p1_fail("H2D failed");
	if(!st) st = p1_launch_diag_seq(ctx, data, col_lo, col_hi, precision, xy_dev);
	if(!st && cudaMemcpyAsync(xy_inout, xy_dev, m * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) st = p1_fail("D2H failed");
	cudaStreamSynchronize(ctx->stream);
	cudaFree(xy_dev);
	return st;
}

// ------------------------------------------------------------------ finalize

// (fixed_t)/(fixed_p1_t) cast of a double as x86-64 cvttsd2si does it (src/fixed.c:3-5,11-13)
static int64_t cast_x86(double v, int bits) {
	if(bits == 64) {
		if(!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) return INT64_MIN;
		return (int64_t) v;
	}
	if(!(v > -2147483649.0 && v < 2147483648.0)) return (int64_t) INT32_MIN;
	return (int64_t) (int32_t) v;
}

extern "C" int p1_finalize(p1_ctx *ctx, const uint64_t *acc, size_t d, size_t col_lo, size_t col_hi,
	int precision, int precision_p2, double normalizer2, const double *diag_xy, int *need_sequential,
	void **res_A, void **res_b) {
	P1_ENTER(ctx);
	P1_CHECK(acc, "p1_finalize: null accumulator");
	P1_CHECK(col_lo <= col_hi && col_hi <= d, "p1_finalize: invalid column range");
	if(res_A) *res_A = NULL;
	if(res_b) *res_b = NULL;
	size_t nA = p1_nA(d), len = nA + 4 * d;
	uint64_t *h = (uint64_t *) malloc(len * sizeof(uint64_t));
	P1_CHECK(h, "out of memory");
	if(p1_is_device_ptr(acc)) {
		cudaError_t e = cudaMemcpyAsync(h, acc, len * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream);
		if(e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
		if(e != cudaSuccess) {
			free(h);
			return p1_fail("p1_finalize: D2H failed: %s", cudaGetErrorString(e));
		}
	} else {
		memcpy(h, acc, len * sizeof(uint64_t));
	}
	const int b1 = ctx->bits_p1, b2 = ctx->bits_p2;
	// diagonal words (src/phase1.c:562-567)
	int missing = 0;
	for(size_t c = col_lo; c < col_hi; c++) {
		uint64_t lo = h[nA + d + c], hi = h[nA + 2 * d + c], viol = h[nA + 3 * d + c];
		unsigned __int128 S = ((unsigned __int128) hi << 32) + lo;
		double xy;
		bool exact = viol == 0 && S <= ((unsigned __int128) 1 << 53);
		if(need_sequential) need_sequential[c - col_lo] = exact ? 0 : 1;
		if(exact) {
			xy = ldexp((double) (uint64_t) S, -precision);
		} else if(diag_xy && !isnan(diag_xy[c - col_lo])) {
			xy = diag_xy[c - col_lo];
		} else {
			missing++;
			continue;
		}
		double scaled = (xy / normalizer2) * (double) (1ll << precision);
		h[c * (c + 1) / 2 + c] = (uint64_t) cast_x86(scaled, b1);
	}
	if(missing) {
		free(h);
		return p1_fail("p1_finalize: %d diagonal column(s) need the sequential double sum (p1_diag_sequential)", missing);
	}
	size_t wout = (size_t) b2 / 8;
	void *A = res_A ? calloc(nA ? nA : 1, b1 == 64 && b2 == 32 ? 8 : wout) : NULL; // reference over-allocates in the 64->32 case (src/phase1.c:611)
	void *B = res_b ? calloc(d, b1 == 64 && b2 == 32 ? 8 : wout) : NULL;
	if((res_A && !A) || (res_b && !B)) {
		free(A);
		free(B);
		free(h);
		return p1_fail("p1_finalize: out of memory");
	}
	for(size_t k = 0; k < nA + d; k++) {
		void *dst = k < nA ? A : B;
		size_t i = k < nA ? k : k - nA;
		if(!dst) continue;
		uint64_t v = h[k];
		if(b1 == 64 && b2 == 64) ((uint64_t *) dst)[i] = v;
		else if(b1 == 32) ((uint32_t *) dst)[i] = (uint32_t) v;
		else ((uint32_t *) dst)[i] = (uint32_t) (((int64_t) v) >> (precision - precision_p2)); // src/phase1.c:617,633
	}
	free(h);
	if(res_A) *res_A = A;
	if(res_b) *res_b = B;
	return 0;
}

extern "C" int p1_finalize_async(p1_ctx *ctx, const uint64_t *acc_dev, size_t num, size_t d, const int32_t *col_start_dev,
	int precision, int precision_p2, double normalizer2, const double *diag_xy_dev, int32_t *need_dev, uint64_t *out_dev) {
	P1_ENTER(ctx);
	P1_CHECK(acc_dev && col_start_dev && need_dev && out_dev, "p1_finalize_async: arguments may not be null");
	P1_CHECK(p1_is_device_ptr(acc_dev) && p1_is_device_ptr(col_start_dev) && p1_is_device_ptr(out_dev), "p1_finalize_async: device pointers expected");
	P1_CHECK(precision >= 0 && precision < 63 && precision_p2 <= precision, "p1_finalize_async: invalid precision");
	return p1_launch_finalize(ctx, acc_dev, num, d, col_start_dev, precision, precision_p2, normalizer2, diag_xy_dev, need_dev, out_dev);
}
