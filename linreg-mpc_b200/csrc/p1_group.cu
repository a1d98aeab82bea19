// Row sharding behind the C-ABI: a p1_group owns one context per GPU of the box
// inside ONE process, so that the reference's single-process callers (run_party /
// run_trusted_initializer, src/phase1.h:14-29) use every GPU without knowing about ranks.
//   - p1_group_encode scatters row slabs [row0_g, row0_g + rows_g) of X, y over the devices
//     (256-row aligned equal slices, as phase1.shard_rows);
//   - every batched call fans out per device on that device's rows: AES-CTR and the OT extension are
//     random access, so each device expands only its slice of every mask / PRG stream
//     (row0 / n_rows of p1_ti_generate, p1_cot_*);
//   - every share word is a sum over rows in Z_2^64, so the partial accumulators combine exactly with a
//     wrapping u64 sum: k_group_reduce on device 0 reads the peers' accumulators over NVLink (P2P
//     loads, one launch), or ncclAllReduce(ncclUint64, ncclSum) from the system libnccl when asked
//     for (option "reduce" = 1; SURVEY.md D.4(v): both measured, tools/kbench.py group);
//   - the one order-dependent quantity, the sequential double diagonal (src/phase1.c:563-566), is
//     chained over the shards in device order;
//   - vectors of the masking protocol (TI masks, party messages) live as row slices on the devices
//     (p1_gvec); whole vectors for the wire format are gathered to / scattered from device 0 with
//     peer copies.
// Ordering: device 0's stream is the group's main stream.  Every group call forks (each device's
// stream waits for an event on the main stream) and joins (the main stream waits for every device),
// so group calls and plain calls on p1_group_ctx(grp, 0) execute in program order.
#include "p1_common.cuh"
#include <dlfcn.h>
#include <math.h>
#include <string.h>
#include <string>
#include <thread>

struct p1_group {
	int ndev;
	int bits_p1, bits_p2;
	std::vector<p1_ctx *> ctx;
	std::vector<cudaEvent_t> ev;        // per device: end of its part of the current call
	cudaEvent_t ev_main;
	size_t acc_len;
	std::vector<uint64_t *> acc;        // per device partial accumulators (acc_len words)
	std::vector<uint64_t *> part;       // per device small partial sums (<x,y> of a batch), part_len words
	size_t part_len;
	uint64_t **peer_tab_dev;            // on device 0: table of per-device pointers for k_group_reduce
	int reduce_mode;                    // 0 = P2P reduce kernel, 1 = NCCL
	void *nccl_lib;
	std::vector<void *> nccl_comm;
	bool reduced;                       // acc[0] already holds the sum over devices
};

struct p1_gdata {
	p1_group *grp;
	size_t n, d;
	bool has_y;
	std::vector<p1_data *> shard;
	std::vector<size_t> row0, rows;
};

struct p1_gvec {
	p1_group *grp;
	size_t n_total, batch;
	std::vector<void *> buf;
	std::vector<size_t> ld, row0, rows;
};

struct p1_gcot {
	p1_group *grp;
	int role;
	std::vector<p1_cot *> s;
	std::vector<void *> u, c;           // per device wire slices
	size_t n_total;
};

// ------------------------------------------------------------------ helpers

static void group_rows(int ndev, size_t n_total, int g, size_t *row0, size_t *rows) {
	// whole 256-row blocks, spread evenly (the first blocks % ndev devices hold one more): nobody is empty once
	// there are at least ndev blocks
	const size_t blocks = (n_total + 255) / 256, q = blocks / ndev, r = blocks % ndev;
	size_t r0 = ((size_t) g * q + ((size_t) g < r ? (size_t) g : r)) * 256;
	size_t r1 = r0 + (q + ((size_t) g < r ? 1 : 0)) * 256;
	if(r0 > n_total) r0 = n_total;
	if(r1 > n_total) r1 = n_total;
	*row0 = r0;
	*rows = r1 - r0;
}

// runs fn(g) for every device, on one host thread per device when there are several (host-side
// synchronisations inside a call - staging copies, statistics reads - then overlap across devices)
template <typename F> static int fan_out(p1_group *grp, F fn) {
	if(grp->ndev == 1) return fn(0);
	std::vector<int> st(grp->ndev, 0);
	std::vector<std::string> msg(grp->ndev);
	std::vector<std::thread> th;
	for(int g = 0; g < grp->ndev; g++) {
		th.emplace_back([&, g]() {
			st[g] = fn(g);
			if(st[g]) msg[g] = p1_last_error();
		});
	}
	for(auto &t : th) t.join();
	for(int g = 0; g < grp->ndev; g++) if(st[g]) return p1_fail("device %d: %s", grp->ctx[g]->device, msg[g].c_str());
	return 0;
}

// every device's stream waits for what the main stream (device 0) has enqueued so far
static int group_fork(p1_group *grp) {
	if(grp->ndev == 1) return 0;
	P1_CUDA(cudaSetDevice(grp->ctx[0]->device));
	P1_CUDA(cudaEventRecord(grp->ev_main, grp->ctx[0]->stream));
	for(int g = 1; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaStreamWaitEvent(grp->ctx[g]->stream, grp->ev_main, 0));
	}
	return 0;
}
// the main stream waits for every device's part
static int group_join(p1_group *grp) {
	if(grp->ndev == 1) return 0;
	for(int g = 1; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaEventRecord(grp->ev[g], grp->ctx[g]->stream));
	}
	P1_CUDA(cudaSetDevice(grp->ctx[0]->device));
	for(int g = 1; g < grp->ndev; g++) P1_CUDA(cudaStreamWaitEvent(grp->ctx[0]->stream, grp->ev[g], 0));
	return 0;
}

// dst[i] += sum over the other devices' src_g[i] (wrapping u64): P2P loads over NVLink.  clear: the peers' words
// are zeroed once they have been added (P2P stores), so that a partial sum is never counted twice
__global__ void k_group_reduce(uint64_t *__restrict__ dst, uint64_t *const *__restrict__ src, int nsrc, size_t len, int clear) {
	for(size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (size_t) gridDim.x * blockDim.x) {
		uint64_t s = dst[i];
		for(int g = 1; g < nsrc; g++) {
			s += src[g][i];
			if(clear) src[g][i] = 0;
		}
		dst[i] = s;
	}
}

// The kernel reads the peers' addresses (entries 1..ndev-1) from a table on device 0.  The two tables - accumulators
// in slot 0, <x,y> partials in slot 1 - are written when those buffers are allocated (stage_peer_table, a blocking
// copy right after the cudaMalloc), so a reduce itself moves nothing from the host and needs no host round trip.
#define PEER_TAB_STRIDE 64
static int stage_peer_table(p1_group *grp, int slot, const std::vector<uint64_t *> &bufs) {
	P1_CUDA(cudaSetDevice(grp->ctx[0]->device));
	P1_CUDA(cudaStreamSynchronize(grp->ctx[0]->stream));      // no reduce that reads the old table is still running
	P1_CUDA(cudaMemcpy(grp->peer_tab_dev + slot * PEER_TAB_STRIDE, bufs.data(), grp->ndev * sizeof(uint64_t *), cudaMemcpyHostToDevice));
	return 0;
}

static int reduce_p2p(p1_group *grp, int slot, uint64_t *dst, size_t len, bool clear = false) {
	if(grp->ndev == 1 || !len) return 0;
	p1_ctx *c0 = grp->ctx[0];
	P1_CUDA(cudaSetDevice(c0->device));
	unsigned grid = (unsigned) ((len + 255) / 256);
	if(grid > (unsigned) c0->sm_count * 4) grid = c0->sm_count * 4;
	p1_prof_begin(c0, P1_K_MISC);
	k_group_reduce<<<grid, 256, 0, c0->stream>>>(dst, grp->peer_tab_dev + slot * PEER_TAB_STRIDE, grp->ndev, len, clear ? 1 : 0);
	P1_LAUNCHED(c0);
	return 0;
}

// ---- NCCL (system libnccl.so.2, loaded on demand; prototypes of nccl.h 2.x)
typedef struct { char internal[128]; } p1_ncclUniqueId;
typedef int (*nccl_comm_init_all_fn)(void **comms, int ndev, const int *devlist);
typedef int (*nccl_all_reduce_fn)(const void *send, void *recv, size_t count, int dtype, int op, void *comm, cudaStream_t stream);
typedef int (*nccl_group_fn)(void);
typedef int (*nccl_comm_destroy_fn)(void *comm);
static nccl_comm_init_all_fn f_init_all;
static nccl_all_reduce_fn f_all_reduce;
static nccl_group_fn f_group_start, f_group_end;
static nccl_comm_destroy_fn f_comm_destroy;

static int nccl_load(p1_group *grp) {
	if(grp->nccl_lib) return 0;
	void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
	P1_CHECK(h, "p1_group: option reduce=1 needs libnccl.so.2 (%s)", dlerror());
	f_init_all = (nccl_comm_init_all_fn) dlsym(h, "ncclCommInitAll");
	f_all_reduce = (nccl_all_reduce_fn) dlsym(h, "ncclAllReduce");
	f_group_start = (nccl_group_fn) dlsym(h, "ncclGroupStart");
	f_group_end = (nccl_group_fn) dlsym(h, "ncclGroupEnd");
	f_comm_destroy = (nccl_comm_destroy_fn) dlsym(h, "ncclCommDestroy");
	P1_CHECK(f_init_all && f_all_reduce && f_group_start && f_group_end && f_comm_destroy, "p1_group: libnccl.so.2 lacks the expected symbols");
	std::vector<int> devs;
	for(auto c : grp->ctx) devs.push_back(c->device);
	grp->nccl_comm.assign(grp->ndev, NULL);
	int rc = f_init_all(grp->nccl_comm.data(), grp->ndev, devs.data());
	P1_CHECK(rc == 0, "ncclCommInitAll failed (%d)", rc);
	grp->nccl_lib = h;
	return 0;
}

static int reduce_nccl(p1_group *grp, std::vector<uint64_t *> &bufs, size_t len, bool clear = false) {
	if(grp->ndev == 1 || !len) return 0;
	if(nccl_load(grp)) return 1;
	const int ncclUint64 = 5, ncclSum = 0;      // nccl.h: ncclDataType_t / ncclRedOp_t
	if(group_fork(grp)) return 1;
	P1_CHECK(f_group_start() == 0, "ncclGroupStart failed");
	for(int g = 0; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		int rc = f_all_reduce(bufs[g], bufs[g], len, ncclUint64, ncclSum, grp->nccl_comm[g], grp->ctx[g]->stream);
		P1_CHECK(rc == 0, "ncclAllReduce failed (%d)", rc);
	}
	P1_CHECK(f_group_end() == 0, "ncclGroupEnd failed");
	// the all-reduce leaves the sum everywhere; only device 0 keeps it, or the next reduce would add it ndev times
	if(clear) for(int g = 1; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaMemsetAsync(bufs[g], 0, len * sizeof(uint64_t), grp->ctx[g]->stream));
	}
	return group_join(grp);
}

static int group_reduce(p1_group *grp, std::vector<uint64_t *> &bufs, size_t len, bool clear) {
	return grp->reduce_mode == 1 ? reduce_nccl(grp, bufs, len, clear) : reduce_p2p(grp, 0, bufs[0], len, clear);   // only called on grp->acc
}

// ------------------------------------------------------------------ group

extern "C" void p1_group_destroy(p1_group *grp) {
	if(!grp) return;
	for(int g = 0; g < (int) grp->ctx.size(); g++) {
		if(!grp->ctx[g]) continue;
		cudaSetDevice(grp->ctx[g]->device);
		cudaStreamSynchronize(grp->ctx[g]->stream);
		if(g < (int) grp->acc.size() && grp->acc[g]) cudaFree(grp->acc[g]);
		if(g < (int) grp->part.size() && grp->part[g]) cudaFree(grp->part[g]);
		if(g < (int) grp->ev.size() && grp->ev[g]) cudaEventDestroy(grp->ev[g]);
	}
	if(grp->nccl_lib) for(auto c : grp->nccl_comm) if(c) f_comm_destroy(c);
	if(!grp->ctx.empty() && grp->ctx[0]) {
		cudaSetDevice(grp->ctx[0]->device);
		if(grp->peer_tab_dev) cudaFree(grp->peer_tab_dev);
		if(grp->ev_main) cudaEventDestroy(grp->ev_main);
	}
	for(auto c : grp->ctx) p1_ctx_destroy(c);
	delete grp;
}

extern "C" int p1_group_create(const int *devices, int ndev, int ring_bits_p1, int ring_bits_p2, p1_group **out) {
	P1_CHECK(out && devices && ndev >= 1 && ndev <= 64, "p1_group_create: invalid arguments");
	*out = NULL;
	p1_group *grp = new p1_group();
	grp->ndev = ndev;
	grp->bits_p1 = ring_bits_p1;
	grp->bits_p2 = ring_bits_p2;
	grp->acc_len = grp->part_len = 0;
	grp->peer_tab_dev = NULL;
	grp->reduce_mode = 0;
	grp->nccl_lib = NULL;
	grp->ev_main = NULL;
	grp->reduced = false;
	grp->ctx.assign(ndev, NULL);
	grp->ev.assign(ndev, NULL);
	grp->acc.assign(ndev, NULL);
	grp->part.assign(ndev, NULL);
	// The first touch of a GPU in a process (primary context, loading this library's kernels onto it) takes the better
	// part of a second, and enabling peer access costs time per ordered pair: one host thread per device does both,
	// in two rounds (every context exists before any peer mapping is asked for), instead of one device after another.
	{
		std::vector<int> st(ndev, 0);
		std::vector<std::thread> th;
		for(int g = 0; g < ndev; g++) th.emplace_back([&, g] {
			if(p1_ctx_create(devices[g], ring_bits_p1, ring_bits_p2, &grp->ctx[g])) { st[g] = 1; return; }
			if(cudaEventCreateWithFlags(&grp->ev[g], cudaEventDisableTiming) != cudaSuccess) st[g] = 2;
		});
		for(auto &t : th) t.join();
		for(int g = 0; g < ndev; g++) if(st[g]) {
			p1_group_destroy(grp);
			return st[g] == 1 ? 1 : p1_fail("p1_group_create: cudaEventCreate failed");
		}
	}
	// peer access between every pair of distinct devices (NVLink / NVSwitch on the box)
	{
		std::vector<int> bad(ndev, -1);
		std::vector<cudaError_t> err(ndev, cudaSuccess);
		std::vector<std::thread> th;
		for(int a = 0; a < ndev; a++) {
			bool first = true;                     // a device listed twice (tests) is handled once
			for(int k = 0; k < a; k++) if(devices[k] == devices[a]) first = false;
			if(!first) continue;
			th.emplace_back([&, a] {
				cudaSetDevice(devices[a]);
				for(int b = 0; b < ndev; b++) {
					if(devices[a] == devices[b]) continue;
					int can = 0;
					cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
					if(!can) { bad[a] = b; err[a] = cudaSuccess; return; }
					cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
					cudaGetLastError();
					if(e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { bad[a] = b; err[a] = e; return; }
				}
			});
		}
		for(auto &t : th) t.join();
		for(int a = 0; a < ndev; a++) if(bad[a] >= 0) {
			p1_group_destroy(grp);
			if(err[a] == cudaSuccess) return p1_fail("p1_group_create: device %d cannot access device %d (no P2P)", devices[a], devices[bad[a]]);
			return p1_fail("cudaDeviceEnablePeerAccess(%d -> %d): %s", devices[a], devices[bad[a]], cudaGetErrorString(err[a]));
		}
	}
	cudaSetDevice(devices[0]);
	if(cudaEventCreateWithFlags(&grp->ev_main, cudaEventDisableTiming) != cudaSuccess ||
	   cudaMalloc(&grp->peer_tab_dev, 2 * PEER_TAB_STRIDE * sizeof(uint64_t *)) != cudaSuccess) { p1_group_destroy(grp); return p1_fail("p1_group_create: out of memory"); }
	*out = grp;
	return 0;
}

extern "C" int p1_group_size(const p1_group *grp) { return grp ? grp->ndev : 0; }
extern "C" p1_ctx *p1_group_ctx(p1_group *grp, int g) { return grp && g >= 0 && g < grp->ndev ? grp->ctx[g] : NULL; }
extern "C" void p1_group_rows(const p1_group *grp, size_t n_total, int g, size_t *row0, size_t *n_rows) {
	size_t r0 = 0, r = 0;
	if(grp && g >= 0 && g < grp->ndev) group_rows(grp->ndev, n_total, g, &r0, &r);
	if(row0) *row0 = r0;
	if(n_rows) *n_rows = r;
}

extern "C" int p1_group_set_option(p1_group *grp, const char *name, long value) {
	P1_CHECK(grp && name, "p1_group_set_option: invalid arguments");
	if(!strcmp(name, "reduce")) {
		P1_CHECK(value == 0 || value == 1, "reduce: 0 = P2P reduce kernel, 1 = NCCL all-reduce");
		grp->reduce_mode = (int) value;
		return 0;
	}
	for(auto c : grp->ctx) if(p1_ctx_set_option(c, name, value)) return 1;
	return 0;
}

extern "C" int p1_group_sync(p1_group *grp) {
	P1_CHECK(grp, "null group");
	for(auto c : grp->ctx) if(p1_ctx_sync(c)) return 1;
	return 0;
}

extern "C" uint64_t p1_group_launches(p1_group *grp) {
	uint64_t s = 0;
	if(grp) for(auto c : grp->ctx) s += p1_ctx_launches(c);
	return s;
}

// ------------------------------------------------------------------ accumulators

static int ensure_acc(p1_group *grp, size_t d) {
	const size_t len = p1_acc_len(d);
	if(grp->acc_len == len) return 0;
	for(int g = 0; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		if(grp->acc[g]) {
			P1_CUDA(cudaStreamSynchronize(grp->ctx[g]->stream));
			P1_CUDA(cudaFree(grp->acc[g]));
			grp->acc[g] = NULL;
		}
		P1_CUDA(cudaMalloc(&grp->acc[g], len * sizeof(uint64_t)));
		P1_CUDA(cudaMemsetAsync(grp->acc[g], 0, len * sizeof(uint64_t), grp->ctx[g]->stream));
	}
	grp->acc_len = len;
	grp->reduced = false;
	return stage_peer_table(grp, 0, grp->acc);
}

static int ensure_part(p1_group *grp, size_t len) {
	if(grp->part_len >= len) return 0;
	for(int g = 0; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		if(grp->part[g]) {
			P1_CUDA(cudaStreamSynchronize(grp->ctx[g]->stream));
			P1_CUDA(cudaFree(grp->part[g]));
			grp->part[g] = NULL;
		}
		P1_CUDA(cudaMalloc(&grp->part[g], len * sizeof(uint64_t)));
	}
	grp->part_len = len;
	return stage_peer_table(grp, 1, grp->part);
}

extern "C" int p1_group_acc_reset(p1_group *grp, size_t d) {
	P1_CHECK(grp, "null group");
	if(ensure_acc(grp, d)) return 1;
	if(group_fork(grp)) return 1;
	for(int g = 0; g < grp->ndev; g++) {
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaMemsetAsync(grp->acc[g], 0, grp->acc_len * sizeof(uint64_t), grp->ctx[g]->stream));
	}
	grp->reduced = false;
	return group_join(grp);
}

extern "C" uint64_t *p1_group_acc_dev(p1_group *grp, int g) { return grp && g >= 0 && g < grp->ndev ? grp->acc[g] : NULL; }

// sum of the devices' partial accumulators, left in device 0's accumulator; the other devices' accumulators are
// zero afterwards, so accumulating further and reducing again adds only what is new (device 0 + new partials)
extern "C" int p1_group_reduce(p1_group *grp) {
	P1_CHECK(grp && grp->acc_len, "p1_group_reduce: no accumulator");
	if(grp->reduced) return 0;
	if(group_reduce(grp, grp->acc, grp->acc_len, true)) return 1;
	grp->reduced = true;
	return 0;
}

// ------------------------------------------------------------------ data

extern "C" void p1_gdata_free(p1_gdata *gd) {
	if(!gd) return;
	for(auto s : gd->shard) if(s) p1_data_free(s);
	delete gd;
}

extern "C" int p1_group_encode(p1_group *grp, const double *X, const double *y, size_t n, size_t d,
	double normalizer, int precision, p1_gdata **out) {
	P1_CHECK(grp && X && out, "p1_group_encode: arguments may not be null");
	*out = NULL;
	P1_CHECK(n >= (size_t) grp->ndev * 256 || grp->ndev == 1, "p1_group_encode: %zu rows are too few for %d devices (256-row slices)", n, grp->ndev);
	p1_gdata *gd = new p1_gdata();
	gd->grp = grp;
	gd->n = n;
	gd->d = d;
	gd->has_y = y != NULL;
	gd->shard.assign(grp->ndev, NULL);
	gd->row0.assign(grp->ndev, 0);
	gd->rows.assign(grp->ndev, 0);
	for(int g = 0; g < grp->ndev; g++) group_rows(grp->ndev, n, g, &gd->row0[g], &gd->rows[g]);
	for(int g = 0; g < grp->ndev; g++) if(!gd->rows[g]) { delete gd; return p1_fail("p1_group_encode: device %d would hold no rows", g); }
	if(group_fork(grp)) { delete gd; return 1; }
	// X, y on the host: every device pulls its slab over its own PCIe link; on a device (the text parser's
	// output on device 0): the encode kernels of the other devices read the peer's memory over NVLink
	int st = fan_out(grp, [&](int g) {
		return p1_encode(grp->ctx[g], X + gd->row0[g] * d, y ? y + gd->row0[g] : NULL, gd->rows[g], d, normalizer, precision, &gd->shard[g]);
	});
	if(!st) st = group_join(grp);
	if(st) { p1_gdata_free(gd); return 1; }
	*out = gd;
	return 0;
}

extern "C" const p1_data *p1_gdata_shard(const p1_gdata *gd, int g) { return gd && g >= 0 && g < (int) gd->shard.size() ? gd->shard[g] : NULL; }

extern "C" int p1_group_local_block(p1_group *grp, const p1_gdata *gd, size_t col_lo, size_t col_hi, int owns_target, int algo) {
	P1_CHECK(grp && gd && gd->grp->ndev == grp->ndev, "p1_group_local_block: invalid arguments");
	if(ensure_acc(grp, gd->d)) return 1;
	if(group_fork(grp)) return 1;
	grp->reduced = false;
	if(fan_out(grp, [&](int g) { return p1_local_block(grp->ctx[g], gd->shard[g], col_lo, col_hi, owns_target, grp->acc[g], algo); })) return 1;
	return group_join(grp);
}

// ------------------------------------------------------------------ row-sliced vectors

extern "C" void p1_gvec_free(p1_gvec *v) {
	if(!v) return;
	for(int g = 0; g < (int) v->buf.size(); g++) if(v->buf[g]) p1_dev_free(v->grp->ctx[g], v->buf[g]);
	delete v;
}

extern "C" int p1_gvec_alloc(p1_group *grp, size_t n_total, size_t batch, p1_gvec **out) {
	P1_CHECK(grp && out && batch, "p1_gvec_alloc: invalid arguments");
	*out = NULL;
	p1_gvec *v = new p1_gvec();
	v->grp = grp;
	v->n_total = n_total;
	v->batch = batch;
	v->buf.assign(grp->ndev, NULL);
	v->ld.assign(grp->ndev, 0);
	v->row0.assign(grp->ndev, 0);
	v->rows.assign(grp->ndev, 0);
	const size_t w = (size_t) grp->bits_p1 / 8;
	for(int g = 0; g < grp->ndev; g++) {
		group_rows(grp->ndev, n_total, g, &v->row0[g], &v->rows[g]);
		v->ld[g] = p1_round_up(v->rows[g] ? v->rows[g] : 1, 256);
		if(p1_dev_alloc(grp->ctx[g], batch * v->ld[g] * w, &v->buf[g])) { p1_gvec_free(v); return 1; }
	}
	*out = v;
	return 0;
}

extern "C" void *p1_gvec_dev(const p1_gvec *v, int g, size_t *ld_words) {
	if(!v || g < 0 || g >= (int) v->buf.size()) return NULL;
	if(ld_words) *ld_words = v->ld[g];
	return v->buf[g];
}

// whole vectors (count x n_total words, ld0 words apart) on device 0  <->  row slices on the devices
static int gvec_move(p1_group *grp, p1_gvec *v, size_t count, void *full_dev0, size_t ld0, bool gather) {
	P1_CHECK(grp && v && v->grp == grp && full_dev0 && count <= v->batch && ld0 >= v->n_total, "p1_gvec_gather/scatter: invalid arguments");
	const size_t w = (size_t) grp->bits_p1 / 8;
	if(group_fork(grp)) return 1;
	for(int g = 0; g < grp->ndev; g++) {
		if(!v->rows[g]) continue;
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		char *full = (char *) full_dev0 + v->row0[g] * w;
		if(gather) P1_CUDA(cudaMemcpy2DAsync(full, ld0 * w, v->buf[g], v->ld[g] * w, v->rows[g] * w, count, cudaMemcpyDefault, grp->ctx[g]->stream));
		else P1_CUDA(cudaMemcpy2DAsync(v->buf[g], v->ld[g] * w, full, ld0 * w, v->rows[g] * w, count, cudaMemcpyDefault, grp->ctx[g]->stream));
	}
	return group_join(grp);
}
extern "C" int p1_gvec_gather(p1_group *grp, const p1_gvec *v, size_t count, void *full_dev0, size_t ld0) {
	return gvec_move(grp, (p1_gvec *) v, count, full_dev0, ld0, true);
}
extern "C" int p1_gvec_scatter(p1_group *grp, const void *full_dev0, size_t ld0, size_t count, p1_gvec *v) {
	return gvec_move(grp, v, count, (void *) full_dev0, ld0, false);
}

// ------------------------------------------------------------------ masking protocol, fanned out

extern "C" int p1_group_ti_generate_keys(p1_group *grp, const uint8_t key_x[16], const uint8_t key_y[16], const uint8_t key_r[16],
	uint64_t q0, size_t count, size_t n_total, p1_gvec *x, p1_gvec *y, uint64_t *xy_dev0, uint64_t *r_dev0) {
	P1_CHECK(grp, "p1_group_ti_generate: invalid arguments");
	P1_CHECK((!x || (x->grp == grp && x->n_total == n_total && count <= x->batch)) && (!y || (y->grp == grp && y->n_total == n_total && count <= y->batch)),
		"p1_group_ti_generate: vectors do not match the call");
	if(!count) return 0;
	if(xy_dev0 && ensure_part(grp, count)) return 1;
	if(group_fork(grp)) return 1;
	int st = fan_out(grp, [&](int g) {
		size_t row0, rows;
		group_rows(grp->ndev, n_total, g, &row0, &rows);
		if(!rows) {
			if(xy_dev0 && g) { cudaSetDevice(grp->ctx[g]->device); cudaMemsetAsync(grp->part[g], 0, count * 8, grp->ctx[g]->stream); }
			return 0;
		}
		uint64_t *xy = !xy_dev0 ? NULL : (g == 0 ? xy_dev0 : grp->part[g]);
		return p1_ti_generate_keys(grp->ctx[g], key_x, key_y, key_r, q0, count, n_total, row0, rows, x ? x->buf[g] : NULL, y ? y->buf[g] : NULL,
			x ? x->ld[g] : (y ? y->ld[g] : rows), xy, g == 0 ? r_dev0 : NULL);
	});
	if(st || group_join(grp)) return 1;
	if(xy_dev0 && grp->ndev > 1) {
		return reduce_p2p(grp, 1, xy_dev0, count);
	}
	return 0;
}

extern "C" int p1_group_ti_generate(p1_group *grp, const uint8_t key[16], uint64_t q0, size_t count, size_t n_total,
	p1_gvec *x, p1_gvec *y, uint64_t *xy_dev0, uint64_t *r_dev0) {
	P1_CHECK(grp && key, "p1_group_ti_generate: invalid arguments");
	return p1_group_ti_generate_keys(grp, key, key, key, q0, count, n_total, x, y, xy_dev0, r_dev0);
}

static int check_party(p1_group *grp, const p1_gdata *gd, const p1_gvec *a, const p1_gvec *b, const p1_gvec *c, size_t count) {
	P1_CHECK(grp && gd && gd->grp->ndev == grp->ndev, "invalid group data");
	for(const p1_gvec *v : {a, b, c}) P1_CHECK(!v || (v->grp == grp && v->n_total == gd->n && count <= v->batch), "vector set does not match the data");
	return ensure_acc(grp, gd->d);
}

extern "C" int p1_group_party_b_mask(p1_group *grp, const p1_gdata *gd, const int32_t *cols, size_t count, const p1_gvec *T, p1_gvec *out) {
	P1_CHECK(cols && T && out, "p1_group_party_b_mask: arguments may not be null");
	if(check_party(grp, gd, T, out, NULL, count) || group_fork(grp)) return 1;
	if(fan_out(grp, [&](int g) { return p1_party_b_mask(grp->ctx[g], gd->shard[g], cols, count, T->buf[g], T->ld[g], out->buf[g], out->ld[g]); })) return 1;
	return group_join(grp);
}

extern "C" int p1_group_party_a_step(p1_group *grp, const p1_gdata *gd, const int32_t *cols, const int64_t *slots, size_t count,
	const p1_gvec *T, const p1_gvec *in, p1_gvec *out) {
	P1_CHECK(cols && slots && T && in && out, "p1_group_party_a_step: arguments may not be null");
	if(check_party(grp, gd, T, in, out, count) || group_fork(grp)) return 1;
	grp->reduced = false;
	if(fan_out(grp, [&](int g) {
		return p1_party_a_step(grp->ctx[g], gd->shard[g], cols, slots, count, T->buf[g], T->ld[g], in->buf[g], in->ld[g], out->buf[g], out->ld[g], grp->acc[g]);
	})) return 1;
	return group_join(grp);
}

extern "C" int p1_group_party_b_share(p1_group *grp, const p1_gdata *gd, const int32_t *cols, const int64_t *slots, size_t count, const p1_gvec *in) {
	P1_CHECK(cols && slots && in, "p1_group_party_b_share: arguments may not be null");
	if(check_party(grp, gd, in, NULL, NULL, count) || group_fork(grp)) return 1;
	grp->reduced = false;
	if(fan_out(grp, [&](int g) { return p1_party_b_share(grp->ctx[g], gd->shard[g], cols, slots, count, in->buf[g], in->ld[g], grp->acc[g]); })) return 1;
	return group_join(grp);
}

// share -= value: once, on device 0's accumulator (src/phase1.c:196,222)
extern "C" int p1_group_acc_sub_values(p1_group *grp, size_t d, const int64_t *slots, const uint64_t *value_dev0, size_t count) {
	P1_CHECK(grp, "null group");
	if(ensure_acc(grp, d)) return 1;
	return p1_acc_sub_values(grp->ctx[0], grp->acc[0], slots, value_dev0, count);
}

// ------------------------------------------------------------------ --use_ot, fanned out

extern "C" void p1_group_cot_free(p1_gcot *s) {
	if(!s) return;
	for(int g = 0; g < (int) s->s.size(); g++) {
		if(s->s[g]) p1_cot_free(s->s[g]);
		if(s->u[g]) p1_dev_free(s->grp->ctx[g], s->u[g]);
		if(s->c[g]) p1_dev_free(s->grp->ctx[g], s->c[g]);
	}
	delete s;
}

static int gcot_new(p1_group *grp, int role, const uint8_t *delta, const uint8_t (*sa)[16], const uint8_t (*sb)[16], p1_gcot **out) {
	P1_CHECK(grp && out, "p1_group_cot_*_new: invalid arguments");
	*out = NULL;
	p1_gcot *s = new p1_gcot();
	s->grp = grp;
	s->role = role;
	s->n_total = 0;
	s->s.assign(grp->ndev, NULL);
	s->u.assign(grp->ndev, NULL);
	s->c.assign(grp->ndev, NULL);
	for(int g = 0; g < grp->ndev; g++) {
		int st = role ? p1_cot_sender_new(grp->ctx[g], delta, sa, &s->s[g]) : p1_cot_receiver_new(grp->ctx[g], sa, sb, &s->s[g]);
		if(st) { p1_group_cot_free(s); return 1; }
	}
	*out = s;
	return 0;
}
extern "C" int p1_group_cot_sender_new(p1_group *grp, const uint8_t delta[16], const uint8_t seeds[128][16], p1_gcot **s) {
	P1_CHECK(delta && seeds, "p1_group_cot_sender_new: arguments may not be null");
	return gcot_new(grp, 1, delta, seeds, NULL, s);
}
extern "C" int p1_group_cot_receiver_new(p1_group *grp, const uint8_t seeds0[128][16], const uint8_t seeds1[128][16], p1_gcot **r) {
	P1_CHECK(seeds0 && seeds1, "p1_group_cot_receiver_new: arguments may not be null");
	return gcot_new(grp, 0, NULL, seeds0, seeds1, r);
}

// wire slices of device g for an inner product over n_total rows: U is 2048 B per 128 OTs, C one ring word per OT
static void cot_slice(const p1_group *grp, size_t n_total, int g, size_t *rows, size_t *u_off, size_t *u_len, size_t *c_off, size_t *c_len, size_t *row0_out) {
	size_t row0, r;
	group_rows(grp->ndev, n_total, g, &row0, &r);
	const size_t bits = (size_t) grp->bits_p1, w = bits / 8;
	const size_t tiles = (r * bits + 127) / 128;
	*rows = r;
	*row0_out = row0;
	*u_off = row0 * bits / 128 * 2048;
	*u_len = tiles * 2048;
	*c_off = row0 * bits * w;
	*c_len = tiles * 128 * w;
}

static int gcot_bufs(p1_gcot *s, size_t n_total) {
	if(s->n_total == n_total) return 0;
	p1_group *grp = s->grp;
	for(int g = 0; g < grp->ndev; g++) {
		size_t rows, uo, ul, co, cl, r0;
		cot_slice(grp, n_total, g, &rows, &uo, &ul, &co, &cl, &r0);
		if(s->u[g]) p1_dev_free(grp->ctx[g], s->u[g]);
		if(s->c[g]) p1_dev_free(grp->ctx[g], s->c[g]);
		s->u[g] = s->c[g] = NULL;
		if(p1_dev_alloc(grp->ctx[g], ul ? ul : 16, &s->u[g]) || p1_dev_alloc(grp->ctx[g], cl ? cl : 16, &s->c[g])) return 1;
	}
	s->n_total = n_total;
	return 0;
}

// receiver, first half: every device extends its rows and sends its slice of U to u_host over its own PCIe link
extern "C" int p1_group_cot_receiver_extend(p1_group *grp, p1_gcot *r, const p1_gdata *gd, int32_t col, void *u_host) {
	P1_CHECK(grp && r && gd && u_host && r->grp == grp && gd->grp->ndev == grp->ndev && r->role == 0, "p1_group_cot_receiver_extend: invalid arguments");
	if(gcot_bufs(r, gd->n) || group_fork(grp)) return 1;
	int st = fan_out(grp, [&](int g) {
		size_t rows, uo, ul, co, cl, r0;
		cot_slice(grp, gd->n, g, &rows, &uo, &ul, &co, &cl, &r0);
		if(p1_cot_receiver_extend(grp->ctx[g], r->s[g], gd->shard[g], col, gd->n, r0, r->u[g])) return 1;
		P1_CUDA(cudaMemcpyAsync((char *) u_host + uo, r->u[g], ul, cudaMemcpyDeviceToHost, grp->ctx[g]->stream));
		return 0;
	});
	if(st || group_join(grp)) return 1;
	return p1_group_sync(grp);          // u_host is about to go on the wire
}

extern "C" int p1_group_cot_sender_step(p1_group *grp, p1_gcot *s, const p1_gdata *gd, int32_t col, const void *u_host, void *c_host, int64_t slot) {
	P1_CHECK(grp && s && gd && u_host && c_host && s->grp == grp && gd->grp->ndev == grp->ndev && s->role == 1, "p1_group_cot_sender_step: invalid arguments");
	if(ensure_acc(grp, gd->d) || gcot_bufs(s, gd->n) || group_fork(grp)) return 1;
	grp->reduced = false;
	int st = fan_out(grp, [&](int g) {
		size_t rows, uo, ul, co, cl, r0;
		cot_slice(grp, gd->n, g, &rows, &uo, &ul, &co, &cl, &r0);
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaMemcpyAsync(s->u[g], (const char *) u_host + uo, ul, cudaMemcpyHostToDevice, grp->ctx[g]->stream));
		if(p1_cot_sender_step(grp->ctx[g], s->s[g], gd->shard[g], col, gd->n, r0, s->u[g], s->c[g], grp->acc[g], slot)) return 1;
		size_t live = rows * (size_t) grp->bits_p1 * ((size_t) grp->bits_p1 / 8);      // C bytes of real OTs (the last tile may be partial)
		P1_CUDA(cudaMemcpyAsync((char *) c_host + co, s->c[g], live < cl ? live : cl, cudaMemcpyDeviceToHost, grp->ctx[g]->stream));
		return 0;
	});
	if(st || group_join(grp)) return 1;
	return p1_group_sync(grp);
}

extern "C" int p1_group_cot_receiver_finish(p1_group *grp, p1_gcot *r, const p1_gdata *gd, const void *c_host, int64_t slot) {
	P1_CHECK(grp && r && gd && c_host && r->grp == grp && r->role == 0, "p1_group_cot_receiver_finish: invalid arguments");
	if(ensure_acc(grp, gd->d) || group_fork(grp)) return 1;
	grp->reduced = false;
	int st = fan_out(grp, [&](int g) {
		size_t rows, uo, ul, co, cl, r0;
		cot_slice(grp, gd->n, g, &rows, &uo, &ul, &co, &cl, &r0);
		size_t live = rows * (size_t) grp->bits_p1 * ((size_t) grp->bits_p1 / 8);
		P1_CUDA(cudaSetDevice(grp->ctx[g]->device));
		P1_CUDA(cudaMemcpyAsync(r->c[g], (const char *) c_host + co, live < cl ? live : cl, cudaMemcpyHostToDevice, grp->ctx[g]->stream));
		return p1_cot_receiver_finish(grp->ctx[g], r->s[g], r->c[g], grp->acc[g], slot);
	});
	if(st) return 1;
	return group_join(grp);
}

// ------------------------------------------------------------------ finalize

// Tail of run_party (src/phase1.c:605-641) for the provider owning [col_lo, col_hi): ring reduction of
// the devices' accumulators, the diagonal (exact shortcut, or the literal sequential loop chained over the
// row slabs in device order), 64 -> 32 shift, calloc'ed outputs.
extern "C" int p1_group_finalize(p1_group *grp, const p1_gdata *gd, const uint64_t *acc_host, size_t col_lo, size_t col_hi, int precision, int precision_p2,
	double normalizer2, void **res_A, void **res_b) {
	P1_CHECK(grp && gd && gd->grp->ndev == grp->ndev, "p1_group_finalize: invalid arguments");
	if(res_A) *res_A = NULL;
	if(res_b) *res_b = NULL;
	const size_t d = gd->d, nA = p1_nA(d), len = p1_acc_len(d);
	std::vector<uint64_t> h(len);
	if(acc_host) {
		memcpy(h.data(), acc_host, len * sizeof(uint64_t));      // already combined by the caller (per-peer OT threads)
	} else {
		if(ensure_acc(grp, d)) return 1;
		if(p1_group_reduce(grp)) return 1;
		if(p1_copy_d2h(grp->ctx[0], h.data(), grp->acc[0], len * sizeof(uint64_t))) return 1;
	}
	bool need = false;
	for(size_t c = col_lo; c < col_hi; c++) {
		const uint64_t lo = h[nA + d + c], hi = h[nA + 2 * d + c], viol = h[nA + 3 * d + c];
		if(viol || hi >= (1ull << 22) || ((hi << 32) + lo) > (1ull << 53)) need = true;
	}
	std::vector<double> xy;
	if(need) {
		// rank-ordered chain: device g continues the running sums of device g-1 over its own rows
		xy.assign(col_hi - col_lo, 0.0);
		for(int g = 0; g < grp->ndev; g++)
			if(p1_diag_sequential(grp->ctx[g], gd->shard[g], col_lo, col_hi, precision, xy.data())) return 1;
	}
	return p1_finalize(grp->ctx[0], h.data(), d, col_lo, col_hi, precision, precision_p2, normalizer2, need ? xy.data() : NULL, NULL, res_A, res_b);
}
