/* phase1_gpu.c — replaces src/phase1.c of schoppmp/linreg-mpc in the reference's own build:
 * the same two entry points (src/phase1.h:14-29), the same callers (src/cmd/linreg.c:116-123,
 * src/cmd/secure_multiplication.c:86-89), the same messages on the same sockets, with every
 * arithmetic block of phase 1 executed by libp1gpu.so (include/p1gpu.h) on the B200s of the box.
 *
 *   reference block (src/phase1.c)                 here
 *   read_matrix / read_vector (:473-477)           fread + p1_parse_doubles + p1_group_encode
 *   TI loop body (:270-284)                        p1_group_ti_generate, p1_ti_values, p1_msg_pack
 *   inner_product_ti (:148-236)                    p1_msg_unpack, p1_group_party_b_mask / _a_step / _b_share,
 *                                                  p1_msg_pack, p1_group_acc_sub_values
 *   local pairs and diagonal (:562-571, :359-383)  p1_group_local_block (+ the sequential diagonal chain)
 *   run_party_ot_thread (:342-450)                 p1_group_cot_* with U / C exchanged over the peer socket
 *   tail (:605-641)                                p1_group_finalize
 *
 * It compiles against the reference's headers (phase1.h, node.h, config.h, fixed.h: the ring
 * widths are the reference's compile-time BIT_WIDTH_32_P1 / _P2) and needs nothing of Obliv-C
 * beyond ProtocolDesc / osend / orecv.  Messages keep the framing of send_pmsg / recv_pmsg
 * (:100-145): raw size_t length, proto2 bytes, zero-length orecv flush.
 *
 * GPUs.  One process per party as in the reference; inside it a p1_group (include/p1gpu.h) spreads the
 * rows of the party's columns over the devices listed in P1GPU_DEVICES ("0,1,2,3"; default: device
 * P1GPU_DEVICE or 0): encode, the local Gram, mask expansion, the party kernels and the OT extension run
 * on every device's row slab; the partial shares are summed over NVLink inside p1_group_finalize.
 * Device 0 of the list is the I/O device: whole message vectors are (de)serialised there and moved to /
 * from the slabs with peer copies.
 *
 * Batching.  The reference walks the pairs one by one.  Here consecutive cross pairs between the
 * SAME two providers (a run: j sweeps one provider's columns for a fixed row i) are processed
 * together, up to `batch` pairs: the TI sends the run's 2B messages in the reference's order
 * (a, b, a, b, ...), party b sends its B masked columns, party a answers with B, so every socket
 * carries exactly the reference's byte stream.  All parties visit the runs in the global pair
 * order, hence the earliest unfinished run can always progress (no new deadlocks) PROVIDED every
 * party uses the same batch size: it is derived from n alone (pick_batch); the P1GPU_BATCH override
 * (tests) must be given to all parties alike.
 *
 * Pipelining.  Every host buffer that crosses PCIe is page-locked.  The TI prepares run k+1 (mask
 * expansion, values, serialisation, device->host copy - all asynchronous on the stream) while it writes
 * run k to the sockets from the other half of a double buffer (p1_ctx_mark / p1_ctx_mark_sync).  A party
 * reads message p+1 from its socket while message p travels to the device, and brings a batch of
 * serialised messages back with one asynchronous copy.
 *
 * Seeded masks (P1GPU_SEEDED_MASKS=1 in the environment of the TI and of every provider; default off =
 * the reference's byte stream).  The TI derives one key per provider, K_p = SHA-256(label || K || p)[0..16), and sends
 * it to provider p once.  For cross pair q the masks keep their positions in the stream of p1gpu.h (x = S[W0 + k],
 * y = S[W0 + n + k], r = S[W0 + 2n]) but x is read from the keystream of K_b (b = owner of row j), y from that of K_a
 * and r from the TI's own K (p1_ti_generate_keys): party a expands y, party b expands x on its own GPUs, neither can
 * expand the other's mask, and the TI's message of :277-284 shrinks to {vector = (empty), value}: the same proto2
 * message with no vector field, 2-11 bytes instead of ~9.5 n.  Message count and order on every socket are the
 * reference's; the party <-> party messages are unchanged.
 *
 * Two things the reference obtains inside Obliv-C are hooks here (weak defaults below):
 *   p1host_ti_key   : the TI's PRG key (newBCipherRandomGen draws a random one, :243)
 *   p1host_base_ot  : the kappa = 128 base OTs of honestOTExt{Sender,Recver}New (:394,397)
 * There is no CPU fallback: without a GPU p1_group_create fails and so do both entry points.
 */
#define _POSIX_C_SOURCE 201605L
#define _DEFAULT_SOURCE
#include <ctype.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <sys/random.h>

#include "phase1.h"
#include "p1gpu.h"

#define GCHECK(call) check((call) == 0, "%s:%d: %s", __FILE__, __LINE__, p1_last_error())
#define MAX_DEVICES 16

/* ------------------------------------------------------------------ hooks */

__attribute__((weak)) int p1host_ti_key(uint8_t key[16]) {
	return getrandom(key, 16, 0) == 16 ? 0 : 1;
}

/* role: 1 = this party is the OT-extension sender (gets delta and seeds0 = k_i^{delta_i}),
 * 0 = receiver (gets seeds0 = k_i^0 and seeds1 = k_i^1).  Parties are 0-based node indices.
 * The default is in host/base_ot.c (Chou-Orlandi base OTs over the peer socket); a build against
 * Obliv-C may override it with the npot glue of INTEGRATION.md. */
int p1host_base_ot(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]);

/* devices of this party's group: P1GPU_DEVICES = "0,1,.." or the single device P1GPU_DEVICE (default 0) */
__attribute__((weak)) int p1host_devices(int *devs, int max) {
	const char *e = getenv("P1GPU_DEVICES");
	int n = 0;
	if(e && *e) {
		while(*e && n < max) {
			char *end = NULL;
			long v = strtol(e, &end, 10);
			if(end == e) break;
			devs[n++] = (int) v;
			e = *end == ',' ? end + 1 : end;
		}
	}
	if(!n) {
		const char *one = getenv("P1GPU_DEVICE");
		devs[n++] = one ? atoi(one) : 0;
	}
	return n;
}

/* ------------------------------------------------------------------ small helpers */

/* owner of a row as a 0-based node index; the target (row d) belongs to the last party (:25-33) */
/* P1GPU_TRACE=1: one line per party on stderr with the seconds spent per stage (where does a run's time go) */
static double now_s(void) {
	struct timespec t;
	clock_gettime(CLOCK_MONOTONIC, &t);
	return (double) t.tv_sec + 1e-9 * (double) t.tv_nsec;
}
static int trace_on(void) {
	const char *e = getenv("P1GPU_TRACE");
	return e && atoi(e) > 0;
}

/* ------------------------------------------------------------------ seeded masks (see the header) */

void p1host_sha256(const uint8_t *msg, size_t len, uint8_t out[32]);     /* host/base_ot.c */

static int seeded_masks(void) {
	const char *e = getenv("P1GPU_SEEDED_MASKS");
	return e && atoi(e) > 0;
}

/* K_p = SHA-256("linreg-mpc p1gpu seeded mask key" || K || LE32(p))[0..16), p = 0-based node index */
static void seeded_party_key(const uint8_t ti_key[16], int party, uint8_t out[16]) {
	static const char label[32] = "linreg-mpc p1gpu seeded mask key";
	uint8_t msg[32 + 16 + 4], dig[32];
	memcpy(msg, label, 32);
	memcpy(msg + 32, ti_key, 16);
	for(int k = 0; k < 4; k++) msg[48 + k] = (uint8_t) ((uint32_t) party >> (8 * k));
	p1host_sha256(msg, sizeof(msg), dig);
	memcpy(out, dig, 16);
}

/* SecureMultiplication.Msg with an empty vector: field 2 (value), wire type 0 */
static size_t value_msg_encode(uint64_t value, uint8_t out[16]) {
	size_t k = 0;
	out[k++] = 0x10;
	do {
		uint8_t b = (uint8_t) (value & 0x7f);
		value >>= 7;
		out[k++] = (uint8_t) (b | (value ? 0x80 : 0));
	} while(value);
	return k;
}
static int value_msg_decode(const uint8_t *in, size_t size, uint64_t *value) {
	if(size < 2 || in[0] != 0x10) return 1;
	uint64_t v = 0;
	size_t k = 1;
	for(int shift = 0; k < size && shift < 64; shift += 7) {
		const uint8_t b = in[k++];
		v |= (uint64_t) (b & 0x7f) << shift;
		if(!(b & 0x80)) {
			*value = v;
			return k == size ? 0 : 1;
		}
	}
	return 1;
}

#ifdef P1_WITH_BASE_OT_SELFTEST
/* test hooks (the loop-back build of oracle/Makefile only; tests/test_host.py) */
size_t p1host_selftest_value_msg(uint64_t value, uint8_t out[16], uint64_t *back, int *decode_status) {
	const size_t len = value_msg_encode(value, out);
	*decode_status = value_msg_decode(out, len, back);
	return len;
}
int p1host_selftest_value_msg_decode(const uint8_t *in, size_t size, uint64_t *value) { return value_msg_decode(in, size, value); }
void p1host_selftest_party_key(const uint8_t ti_key[16], int party, uint8_t out[16]) { seeded_party_key(ti_key, party, out); }
#endif

static int owner_of(size_t row, const config *c) {
	int party = 0;
	for(; party + 1 < c->num_parties && c->index_owned[party + 1] <= (ssize_t) row; party++);
	return party;
}

/* framing of send_pmsg (:131-141) for bytes that were serialised on the GPU */
static int send_wire(ProtocolDesc *pd, const uint8_t *buf, size_t size) {
	check(osend(pd, 0, &size, sizeof(size)) >= 0, "osend: %s", strerror(errno));
	check(osend(pd, 0, buf, size) >= 0, "osend: %s", strerror(errno));
	orecv(pd, 0, NULL, 0);
	return 0;
error:
	return 1;
}

/* framing of recv_pmsg (:115-128) */
static int recv_wire(ProtocolDesc *pd, uint8_t *buf, size_t cap, size_t *size) {
	size_t msg_size = 0;
	check(orecv(pd, 0, &msg_size, sizeof(msg_size)) == sizeof(msg_size), "orecv: %s", strerror(errno));
	check(msg_size <= cap, "message of %zu bytes exceeds the %zu-byte receive buffer", msg_size, cap);
	check((size_t) orecv(pd, 0, buf, msg_size) == msg_size, "orecv: %s", strerror(errno));
	*size = msg_size;
	return 0;
error:
	return 1;
}

/* Pairs per batch: ~64 MB of ring words per vector buffer, at most 32 (a run is d/p pairs long).  A function
 * of n ALONE, so that every party of an instance batches alike (see the header); P1GPU_BATCH overrides it
 * for all parties of a test. */
static size_t pick_batch(size_t n) {
	size_t b = ((size_t) 64 << 20) / (n * sizeof(ufixed_p1_t) + 1);
	const char *e = getenv("P1GPU_BATCH");
	if(e && atoi(e) > 0) b = (size_t) atoi(e);
	return b < 1 ? 1 : (b > 32 ? 32 : b);
}

/* the party's GPUs and the buffers one side of the masking protocol needs for a batch */
typedef struct {
	p1_group *grp;
	p1_ctx *ctx;                   /* context of the I/O device = group member 0 */
	int ndev;
	size_t n, ld, stride, batch;
	void *vec[3];                  /* batch x ld ring words on the I/O device: TI masks / received / produced */
	p1_gvec *gv[3];                /* the same as row slices (ndev > 1) */
	int sets;                      /* wire buffers: sets x 2 classes x batch messages */
	uint8_t *wire_dev, *wire_host;
	uint64_t *sizes_dev, *sizes_host;
	uint64_t *vals_dev, *vals2_dev, *zero_dev;
	int32_t *status_dev, *status_host;
} gpu_state;

static uint8_t *wire_dev_at(gpu_state *g, int set, int cls) { return g->wire_dev + ((size_t) (set * 2 + cls)) * g->batch * g->stride; }
static uint8_t *wire_host_at(gpu_state *g, int set, int cls) { return g->wire_host + ((size_t) (set * 2 + cls)) * g->batch * g->stride; }
static uint64_t *sizes_dev_at(gpu_state *g, int set, int cls) { return g->sizes_dev + ((size_t) (set * 2 + cls)) * g->batch; }
static uint64_t *sizes_host_at(gpu_state *g, int set, int cls) { return g->sizes_host + ((size_t) (set * 2 + cls)) * g->batch; }

static void gpu_free(gpu_state *g) {
	if(!g->grp) return;
	for(int k = 0; k < 3; k++) {
		if(g->gv[k]) p1_gvec_free(g->gv[k]);
		if(g->vec[k]) p1_dev_free(g->ctx, g->vec[k]);
	}
	void *dv[] = {g->wire_dev, g->sizes_dev, g->vals_dev, g->vals2_dev, g->zero_dev, g->status_dev};
	for(size_t k = 0; k < sizeof(dv) / sizeof(dv[0]); k++) if(dv[k]) p1_dev_free(g->ctx, dv[k]);
	if(g->wire_host) p1_host_free(g->ctx, g->wire_host);
	if(g->sizes_host) p1_host_free(g->ctx, g->sizes_host);
	if(g->status_host) p1_host_free(g->ctx, g->status_host);
	p1_group_destroy(g->grp);
	memset(g, 0, sizeof(*g));
}

/* nvec vector buffers, `sets` wire double-buffer halves (0: no masking-protocol buffers at all) */
static int gpu_init(gpu_state *g, size_t n, int nvec, int sets) {
	memset(g, 0, sizeof(*g));
	int devs[MAX_DEVICES];
	int ndev = p1host_devices(devs, MAX_DEVICES);
	if(ndev > 1 && n < (size_t) ndev * 256) ndev = 1;          /* slabs are 256-row aligned: tiny instances stay on one GPU */
	GCHECK(p1_group_create(devs, ndev, FIXED_BIT_SIZE, FIXED_BIT_SIZE_P2, &g->grp));
	g->ndev = ndev;
	g->ctx = p1_group_ctx(g->grp, 0);
	g->n = n;
	g->ld = (n + 255) / 256 * 256;
	g->batch = pick_batch(n);
	g->stride = p1_msg_max_bytes(g->ctx, n);
	g->sets = sets;
	for(int k = 0; k < nvec; k++) {
		GCHECK(p1_dev_alloc(g->ctx, g->batch * g->ld * sizeof(ufixed_p1_t), &g->vec[k]));
		if(ndev > 1) GCHECK(p1_gvec_alloc(g->grp, n, g->batch, &g->gv[k]));
	}
	if(sets < 0) {            /* the TI with seeded masks: values only */
		g->sets = 0;
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals_dev));
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals2_dev));
	} else if(sets) {
		const size_t msgs = (size_t) sets * 2 * g->batch;
		GCHECK(p1_dev_alloc(g->ctx, msgs * g->stride, (void **) &g->wire_dev));
		GCHECK(p1_dev_alloc(g->ctx, msgs * 8, (void **) &g->sizes_dev));
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals_dev));
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals2_dev));
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->zero_dev));
		GCHECK(p1_dev_zero(g->ctx, g->zero_dev, g->batch * 8));
		GCHECK(p1_dev_alloc(g->ctx, g->batch * 4, (void **) &g->status_dev));
		GCHECK(p1_host_alloc(g->ctx, msgs * g->stride, (void **) &g->wire_host));
		GCHECK(p1_host_alloc(g->ctx, msgs * 8, (void **) &g->sizes_host));
		GCHECK(p1_host_alloc(g->ctx, g->batch * 4, (void **) &g->status_host));
	}
	return 0;
error:
	gpu_free(g);
	return 1;
}

/* serialise `count` messages {vector = vec[p], value = values_dev[p]} into half (set, cls) of the wire
 * buffer and start their copy to the host; nothing waits (wire_ready does) */
static int pack_start(gpu_state *g, int set, int cls, const void *vec, const uint64_t *values_dev, size_t count) {
	GCHECK(p1_msg_pack(g->ctx, vec, g->ld, g->n, count, values_dev, wire_dev_at(g, set, cls), g->stride, sizes_dev_at(g, set, cls)));
	GCHECK(p1_copy_d2h_async(g->ctx, sizes_host_at(g, set, cls), sizes_dev_at(g, set, cls), count * 8));
	GCHECK(p1_copy_d2h_async(g->ctx, wire_host_at(g, set, cls), wire_dev_at(g, set, cls), count * g->stride));
	return 0;
error:
	return 1;
}
static int wire_mark(gpu_state *g, int set) { return p1_ctx_mark(g->ctx, set); }
static int wire_ready(gpu_state *g, int set) { return p1_ctx_mark_sync(g->ctx, set); }

/* receive `count` messages from pd and parse them on the GPU into vec / values_dev; message p+1 is read from
 * the socket while message p is on its way to the device */
static int recv_to_device(gpu_state *g, ProtocolDesc *pd, size_t count, void *vec, uint64_t *values_dev) {
	uint8_t *wh = wire_host_at(g, 0, 0), *wd = wire_dev_at(g, 0, 0);
	uint64_t *sizes = sizes_host_at(g, 0, 0);
	for(size_t p = 0; p < count; p++) {
		size_t size = 0;
		check(!recv_wire(pd, wh + p * g->stride, g->stride, &size), "could not receive message %zu of the batch", p);
		sizes[p] = size;
		GCHECK(p1_copy_h2d(g->ctx, wd + p * g->stride, wh + p * g->stride, (size + 15) / 16 * 16));
	}
	GCHECK(p1_msg_unpack(g->ctx, wd, g->stride, sizes, count, g->n, vec, g->ld, values_dev, g->status_dev));
	GCHECK(p1_copy_d2h(g->ctx, g->status_host, g->status_dev, count * 4));          /* also: the buffers above are free again */
	for(size_t p = 0; p < count; p++) check(g->status_host[p] == 0, "msg__unpack: malformed message (flags %d)", g->status_host[p]);
	return 0;
error:
	return 1;
}

/* party kernels on the party's data: one device directly, several through the row slices */
/* t_sliced: the TI vector is already in the row slices gv[0] (seeded masks are expanded there) */
static int do_b_mask(gpu_state *g, const p1_gdata *gd, const int32_t *cols, size_t count, int t_sliced) {      /* vec[2] = col + vec[0] */
	if(g->ndev == 1) return p1_party_b_mask(g->ctx, p1_gdata_shard(gd, 0), cols, count, g->vec[0], g->ld, g->vec[2], g->ld);
	if(!t_sliced && p1_gvec_scatter(g->grp, g->vec[0], g->ld, count, g->gv[0])) return 1;
	if(p1_group_party_b_mask(g->grp, gd, cols, count, g->gv[0], g->gv[2])) return 1;
	return p1_gvec_gather(g->grp, g->gv[2], count, g->vec[2], g->ld);
}
static int do_a_step(gpu_state *g, const p1_gdata *gd, const int32_t *cols, const int64_t *slots, size_t count, int t_sliced) {
	/* vec[2] = col - vec[0]; share += <vec[1], vec[0]> */
	if(g->ndev == 1) return p1_party_a_step(g->ctx, p1_gdata_shard(gd, 0), cols, slots, count, g->vec[0], g->ld, g->vec[1], g->ld,
		g->vec[2], g->ld, p1_group_acc_dev(g->grp, 0));
	if(!t_sliced && p1_gvec_scatter(g->grp, g->vec[0], g->ld, count, g->gv[0])) return 1;
	if(p1_gvec_scatter(g->grp, g->vec[1], g->ld, count, g->gv[1])) return 1;
	if(p1_group_party_a_step(g->grp, gd, cols, slots, count, g->gv[0], g->gv[1], g->gv[2])) return 1;
	return p1_gvec_gather(g->grp, g->gv[2], count, g->vec[2], g->ld);
}
static int do_b_share(gpu_state *g, const p1_gdata *gd, const int32_t *cols, const int64_t *slots, size_t count) {
	/* share += <vec[1], col> */
	if(g->ndev == 1) return p1_party_b_share(g->ctx, p1_gdata_shard(gd, 0), cols, slots, count, g->vec[1], g->ld, p1_group_acc_dev(g->grp, 0));
	if(p1_gvec_scatter(g->grp, g->vec[1], g->ld, count, g->gv[1])) return 1;
	return p1_group_party_b_share(g->grp, gd, cols, slots, count, g->gv[1]);
}

/* ------------------------------------------------------------------ pair schedule
 * Enumeration of src/phase1.c:257-258 / :534,544: i = 0..d, j = 0..min(i, d-1).  next_run()
 * returns the next maximal run of consecutive CROSS pairs with the same owners (same i,
 * j = j0 .. j0+count-1), at most `batch` long; q0 = index of its first pair among the cross
 * pairs (the TI's stream position). */
typedef struct {
	size_t i, j;      /* cursor */
	uint64_t q;       /* cross pairs before the cursor */
} pair_cursor;

typedef struct {
	size_t i, j0, count;
	uint64_t q0;
	int owner_i, owner_j;
} pair_run;

static int next_run(pair_cursor *pc, const config *c, size_t batch, pair_run *r) {
	const size_t d = c->d;
	for(; pc->i <= d; pc->i++, pc->j = 0) {
		const int oi = owner_of(pc->i, c);
		while(pc->j <= pc->i && pc->j < d) {
			const int oj = owner_of(pc->j, c);
			if(oi == oj) { pc->j++; continue; }
			size_t k = 0;
			while(k < batch && pc->j + k <= pc->i && pc->j + k < d && owner_of(pc->j + k, c) == oj) k++;
			r->i = pc->i; r->j0 = pc->j; r->count = k; r->q0 = pc->q; r->owner_i = oi; r->owner_j = oj;
			pc->j += k;
			pc->q += k;
			return 1;
		}
	}
	return 0;
}

/* ------------------------------------------------------------------ trusted initializer */

/* A second sending thread: the TI's messages to party A and to party B of a run travel on different connections,
 * so the two streams go out side by side (each connection still carries its messages in the reference's order;
 * only the interleaving ACROSS connections, which no peer can observe, differs from :281-284). */
typedef struct {
	pthread_t th;
	pthread_mutex_t mu;
	pthread_cond_t cv;
	int started, state;            /* 0 idle, 1 job posted, 2 quit */
	ProtocolDesc *pd;
	const uint8_t *base;
	const uint64_t *sizes;
	size_t count, stride;
	int status;
} wire_sender;

static void *wire_sender_main(void *arg) {
	wire_sender *w = arg;
	pthread_mutex_lock(&w->mu);
	for(;;) {
		while(w->state == 0) pthread_cond_wait(&w->cv, &w->mu);
		if(w->state == 2) break;
		pthread_mutex_unlock(&w->mu);
		int st = 0;
		for(size_t p = 0; p < w->count && !st; p++) st = send_wire(w->pd, w->base + p * w->stride, w->sizes[p]);
		pthread_mutex_lock(&w->mu);
		w->status = st;
		w->state = 0;
		pthread_cond_broadcast(&w->cv);
	}
	pthread_mutex_unlock(&w->mu);
	return NULL;
}
static int wire_sender_start(wire_sender *w) {
	memset(w, 0, sizeof(*w));
	pthread_mutex_init(&w->mu, NULL);
	pthread_cond_init(&w->cv, NULL);
	if(pthread_create(&w->th, NULL, wire_sender_main, w)) return 1;
	w->started = 1;
	return 0;
}
static void wire_sender_post(wire_sender *w, ProtocolDesc *pd, const uint8_t *base, const uint64_t *sizes, size_t count, size_t stride) {
	pthread_mutex_lock(&w->mu);
	w->pd = pd; w->base = base; w->sizes = sizes; w->count = count; w->stride = stride;
	w->state = 1;
	pthread_cond_broadcast(&w->cv);
	pthread_mutex_unlock(&w->mu);
}
static int wire_sender_wait(wire_sender *w) {
	pthread_mutex_lock(&w->mu);
	while(w->state == 1) pthread_cond_wait(&w->cv, &w->mu);
	const int st = w->status;
	pthread_mutex_unlock(&w->mu);
	return st;
}
static void wire_sender_stop(wire_sender *w) {
	if(!w->started) return;
	pthread_mutex_lock(&w->mu);
	while(w->state == 1) pthread_cond_wait(&w->cv, &w->mu);
	w->state = 2;
	pthread_cond_broadcast(&w->cv);
	pthread_mutex_unlock(&w->mu);
	pthread_join(w->th, NULL);
	pthread_mutex_destroy(&w->mu);
	pthread_cond_destroy(&w->cv);
	w->started = 0;
}

/* everything of one run up to the host copy of its 2 x count messages, asynchronously, into wire half `set`:
 * x, y, r, <x,y> (:270-274), xy - r and r (:277-278), msg_a = {y, xy - r} (class 0), msg_b = {x, r} (class 1) */
static int ti_prepare(gpu_state *g, const uint8_t key[16], const config *c, const pair_run *r, int set, uint64_t *xy_dev, uint64_t *r_dev) {
	if(g->ndev == 1) {
		GCHECK(p1_ti_generate(g->ctx, key, r->q0, r->count, c->n, 0, c->n, g->vec[0], g->vec[1], g->ld, xy_dev, r_dev));
	} else {
		GCHECK(p1_group_ti_generate(g->grp, key, r->q0, r->count, c->n, g->gv[0], g->gv[1], xy_dev, r_dev));
		GCHECK(p1_gvec_gather(g->grp, g->gv[0], r->count, g->vec[0], g->ld));
		GCHECK(p1_gvec_gather(g->grp, g->gv[1], r->count, g->vec[1], g->ld));
	}
	GCHECK(p1_ti_values(g->ctx, xy_dev, r_dev, r->count, g->vals_dev, g->vals2_dev));
	check(!pack_start(g, set, 0, g->vec[1], g->vals_dev, r->count), "could not serialise the messages for party A");
	check(!pack_start(g, set, 1, g->vec[0], g->vals2_dev, r->count), "could not serialise the messages for party B");
	GCHECK(wire_mark(g, set));
	return 0;
error:
	return 1;
}

/* seeded masks: only <x,y> and r are computed (x under the key of b = owner_j, y under that of a = owner_i, r under
 * the TI's own key); the two values of every pair go to the host half `set` of vals_host */
static int ti_prepare_seeded(gpu_state *g, const uint8_t key[16], const config *c, const pair_run *r, int set,
	uint64_t *xy_dev, uint64_t *r_dev, uint64_t *vals_host) {
	uint8_t ka[16], kb[16];
	seeded_party_key(key, r->owner_i, ka);
	seeded_party_key(key, r->owner_j, kb);
	if(g->ndev == 1) {
		GCHECK(p1_ti_generate_keys(g->ctx, kb, ka, key, r->q0, r->count, c->n, 0, c->n, NULL, NULL, g->ld, xy_dev, r_dev));
	} else {
		GCHECK(p1_group_ti_generate_keys(g->grp, kb, ka, key, r->q0, r->count, c->n, NULL, NULL, xy_dev, r_dev));
	}
	GCHECK(p1_ti_values(g->ctx, xy_dev, r_dev, r->count, g->vals_dev, g->vals2_dev));
	uint64_t *vh = vals_host + (size_t) set * 2 * g->batch;
	GCHECK(p1_copy_d2h_async(g->ctx, vh, g->vals_dev, r->count * 8));
	GCHECK(p1_copy_d2h_async(g->ctx, vh + g->batch, g->vals2_dev, r->count * 8));
	GCHECK(wire_mark(g, set));
	return 0;
error:
	return 1;
}

int run_trusted_initializer(node *self, config *c, int precision, bool use_ot) {
	(void) precision;
	gpu_state g;
	memset(&g, 0, sizeof(g));
	uint64_t *xy_dev = NULL, *r_dev = NULL;
	int status = 1;
	if(use_ot) return 0;                 /* the TI takes no part in the OT variant (:256) */
	double t_wait = 0, t_send = 0;
	const double t_begin = now_s();
	wire_sender to_b;
	memset(&to_b, 0, sizeof(to_b));
	const int seeded = seeded_masks();
	uint64_t *vals_host = NULL;
	check(!gpu_init(&g, c->n, seeded ? 0 : 2, seeded ? -1 : 2), "could not set up the GPU");
	check(!wire_sender_start(&to_b), "could not start the second sending thread");
	uint8_t key[16];
	check(!p1host_ti_key(key), "could not draw the PRG key");
	GCHECK(p1_dev_alloc(g.ctx, g.batch * 8, (void **) &xy_dev));
	GCHECK(p1_dev_alloc(g.ctx, g.batch * 8, (void **) &r_dev));
	if(seeded) {
		GCHECK(p1_host_alloc(g.ctx, (size_t) 2 * 2 * g.batch * 8, (void **) &vals_host));
		for(int p = 2; p < self->num_parties; p++) {          /* every provider's mask key, once */
			uint8_t kp[16];
			seeded_party_key(key, p, kp);
			check(!send_wire(self->peer[p], kp, sizeof(kp)), "Could not send the mask key to party %d", p + 1);
		}
	}

	pair_cursor pc = {0, 0, 0};
	pair_run cur, nxt;
	int have = next_run(&pc, c, g.batch, &cur), set = 0;
	if(have) check(!(seeded ? ti_prepare_seeded(&g, key, c, &cur, set, xy_dev, r_dev, vals_host) : ti_prepare(&g, key, c, &cur, set, xy_dev, r_dev)),
		"could not prepare the first run");
	while(have) {
		check(cur.owner_i >= 2 && cur.owner_j >= 2, "Invalid owners %d, %d for rows %zd, %zd", cur.owner_i, cur.owner_j, cur.i, cur.j0);
		/* the GPU works on the next run while this one goes out on the sockets */
		const int have_next = next_run(&pc, c, g.batch, &nxt);
		if(have_next) check(!(seeded ? ti_prepare_seeded(&g, key, c, &nxt, set ^ 1, xy_dev, r_dev, vals_host) : ti_prepare(&g, key, c, &nxt, set ^ 1, xy_dev, r_dev)),
			"could not prepare a run");
		double t0 = now_s();
		GCHECK(wire_ready(&g, set));
		t_wait += now_s() - t0;
		t0 = now_s();
		if(seeded) {
			/* {vector = (empty), value}: a first, then b, per pair (:281-284) */
			const uint64_t *va = vals_host + (size_t) set * 2 * g.batch, *vb = va + g.batch;
			for(size_t p = 0; p < cur.count; p++) {
				uint8_t m[16];
				size_t len = value_msg_encode((ufixed_p1_t) va[p], m);        /* the field has the ring's width */
				check(!send_wire(self->peer[cur.owner_i], m, len), "Could not send message to party A (%d)", cur.owner_i);
				len = value_msg_encode((ufixed_p1_t) vb[p], m);
				check(!send_wire(self->peer[cur.owner_j], m, len), "Could not send message to party B (%d)", cur.owner_j);
			}
			t_send += now_s() - t0;
			cur = nxt;
			have = have_next;
			set ^= 1;
			continue;
		}
		const uint64_t *sa = sizes_host_at(&g, set, 0), *sb = sizes_host_at(&g, set, 1);
		const uint8_t *wa = wire_host_at(&g, set, 0), *wb = wire_host_at(&g, set, 1);
		wire_sender_post(&to_b, self->peer[cur.owner_j], wb, sb, cur.count, g.stride);
		for(size_t p = 0; p < cur.count; p++) {
			status = send_wire(self->peer[cur.owner_i], wa + p * g.stride, sa[p]);
			if(status) break;
		}
		const int status_b = wire_sender_wait(&to_b);     /* also: the buffers of this half are free again */
		check(!status, "Could not send message to party A (%d)", cur.owner_i);
		check(!status_b, "Could not send message to party B (%d)", cur.owner_j);
		status = 1;
		t_send += now_s() - t0;
		cur = nxt;
		have = have_next;
		set ^= 1;
	}
	if(trace_on()) fprintf(stderr, "p1gpu trace TI: waiting for the GPU %.3f  sending %.3f  total %.3f s, batch %zu\n", t_wait, t_send, now_s() - t_begin, g.batch);
	status = 0;
error:
	wire_sender_stop(&to_b);
	if(g.ctx) {
		p1_group_sync(g.grp);
		if(xy_dev) p1_dev_free(g.ctx, xy_dev);
		if(r_dev) p1_dev_free(g.ctx, r_dev);
		if(vals_host) p1_host_free(g.ctx, vals_host);
	}
	gpu_free(&g);
	return status;
}

/* ------------------------------------------------------------------ data provider: inputs */

/* read_matrix + read_vector (src/linear.c:27-102) for the rest of c->input: the text goes to the
 * GPU as it is, is tokenised and converted there (p1_parse_doubles) and encoded from the device
 * buffer (p1_group_encode; the other devices of the group read it over NVLink).  Tokens: "n m",
 * n*m values, "l", l values.  Tokens the GPU parser defers (inf, nan, hex floats, more than 19 digits with
 * an ambiguous rounding, > 256 characters) are converted with strtod, as fscanf("%lf") would, and patched in. */
static int load_inputs(gpu_state *g, config *c, int precision, p1_gdata **data) {
	p1_ctx *ctx = g->ctx;
	char *text = NULL;
	double *tok_dev = NULL;
	uint64_t *deferred = NULL;
	int status = 1;
	*data = NULL;
	const long pos = ftell(c->input);
	check(pos >= 0 && fseek(c->input, 0, SEEK_END) == 0, "fseek: %s", strerror(errno));
	const long end = ftell(c->input);
	check(end >= pos && fseek(c->input, pos, SEEK_SET) == 0, "fseek: %s", strerror(errno));
	const size_t bytes = (size_t) (end - pos);
	GCHECK(p1_host_alloc(ctx, bytes + 1, (void **) &text));      /* page-locked: the copy to the GPU runs at link rate */
	check(fread(text, 1, bytes, c->input) == bytes, "fread: %s", strerror(errno));
	text[bytes] = 0;
	const size_t n = c->n, d = c->d, need = 2 + n * d + 1 + n;
	GCHECK(p1_dev_alloc(ctx, need * sizeof(double), (void **) &tok_dev));
	uint64_t n_tokens = 0, n_deferred = 0;
	size_t cap = 64;
	deferred = malloc(cap * sizeof(*deferred));
	check(deferred, "Out of memory");
	GCHECK(p1_parse_doubles(ctx, text, bytes, tok_dev, need, &n_tokens, &n_deferred, deferred, cap));
	check(n_tokens >= need, "fscanf: input ends after %llu of %zu values", (unsigned long long) n_tokens, need);
	if(n_deferred > cap) {
		cap = (size_t) n_deferred;
		free(deferred);
		deferred = malloc(cap * sizeof(*deferred));
		check(deferred, "Out of memory");
		GCHECK(p1_parse_doubles(ctx, text, bytes, tok_dev, need, &n_tokens, &n_deferred, deferred, cap));
	}
	if(n_deferred) {
		/* token index of every deferred offset: one pass over the text (offsets ascend) */
		size_t tok = 0, k = 0, o = 0;
		while(o < bytes && k < n_deferred) {
			while(o < bytes && isspace((unsigned char) text[o])) o++;
			if(o >= bytes) break;
			if(o == deferred[k]) {
				char *endp = NULL;
				errno = 0;
				const double v = strtod(text + o, &endp);
				check(endp != text + o, "fscanf: the token at byte %zu of the data section is not a number", o);
				if(tok < need) GCHECK(p1_copy_h2d(ctx, tok_dev + tok, &v, sizeof(v)));
				GCHECK(p1_ctx_sync(ctx));       /* v lives on this stack frame */
				k++;
			}
			while(o < bytes && !isspace((unsigned char) text[o])) o++;
			tok++;
		}
		check(k == n_deferred, "internal: deferred token offsets do not line up with the text");
	}
	double hdr[2], len;
	GCHECK(p1_copy_d2h(ctx, hdr, tok_dev, sizeof(hdr)));
	GCHECK(p1_copy_d2h(ctx, &len, tok_dev + 2 + n * d, sizeof(len)));
	check((size_t) len == n && (size_t) hdr[1] == d && (size_t) hdr[0] == n,
		"Input dimensions invalid: (%zd, %zd), %zd", (size_t) hdr[0], (size_t) hdr[1], (size_t) len);
	const double normalizer = sqrt(pow(2, precision) * c->normalizer1);   /* :473 */
	GCHECK(p1_group_encode(g->grp, tok_dev + 2, tok_dev + 2 + n * d + 1, n, d, normalizer, precision, data));
	GCHECK(p1_group_sync(g->grp));
	status = 0;
error:
	if(tok_dev) p1_dev_free(ctx, tok_dev);
	free(deferred);
	if(text) p1_host_free(ctx, text);
	return status;
}

/* ------------------------------------------------------------------ data provider: --use_ot threads */

typedef struct {
	node *self;
	config *c;
	const p1_gdata *data;
	int peer;                 /* 0-based node index this thread talks to (own index: local block) */
	uint64_t *acc_host;       /* p1_acc_len(d) words, this thread's private result (:498-503) */
	struct timespec wait_total;
	int status;
} ot_thread_args;

static void *run_party_ot_thread_gpu(void *vargs) {
	ot_thread_args *args = vargs;
	node *self = args->self;
	config *c = args->c;
	const int me = self->party - 1, P = self->num_parties;
	const size_t d = c->d, n = c->n, nA = d * (d + 1) / 2;
	gpu_state g;                                   /* a group of its own: group calls belong to one thread at a time */
	p1_gcot *sess = NULL;
	uint8_t *u_host = NULL, *c_host = NULL;
	args->status = 1;
	memset(&g, 0, sizeof(g));
	check(!gpu_init(&g, n, 0, 0), "could not set up the GPU");
	const size_t acc_len = p1_acc_len(d);
	GCHECK(p1_group_acc_reset(g.grp, d));

	if(me == args->peer) {
		/* local pairs, <col, y> for the last provider, diagonal statistics (:359-383) */
		const size_t lo = (size_t) c->index_owned[me], hi = me < P - 1 ? (size_t) c->index_owned[me + 1] : d;
		GCHECK(p1_group_local_block(g.grp, args->data, lo, hi, me == P - 1, P1_GRAM_AUTO));
	} else {
		ProtocolDesc *pd = self->peer[args->peer];
		int party_i, party_j;
		orecv(pd, 0, 0, 0);
		/* sender iff the indices are congruent mod 2 and ours is smaller, or neither (:392) */
		const int sender = ((me % 2 == args->peer % 2) == (me < args->peer));
		uint8_t seeds0[128][16], seeds1[128][16], delta[16];
		if(sender) { party_i = me; party_j = args->peer; } else { party_j = me; party_i = args->peer; }
		check(!p1host_base_ot(pd, party_i, party_j, sender, seeds0, seeds1, delta), "base OTs failed");
		if(sender) {
			GCHECK(p1_group_cot_sender_new(g.grp, delta, seeds0, &sess));
		} else {
			GCHECK(p1_group_cot_receiver_new(g.grp, seeds0, seeds1, &sess));
		}
		/* the wire messages of one inner product, page-locked: every device moves its slice over its own PCIe link */
		const size_t u_bytes = p1_cot_u_bytes(g.ctx, n), c_bytes = p1_cot_c_bytes(g.ctx, n);
		GCHECK(p1_host_alloc(g.ctx, u_bytes, (void **) &u_host));
		GCHECK(p1_host_alloc(g.ctx, c_bytes, (void **) &c_host));
		const size_t i_lo = (size_t) c->index_owned[party_i], i_max = party_i < P - 1 ? (size_t) c->index_owned[party_i + 1] : d;
		const size_t j_lo = (size_t) c->index_owned[party_j], j_max = party_j < P - 1 ? (size_t) c->index_owned[party_j + 1] : d;
		struct timespec wait_start, wait_end;
		clock_gettime(CLOCK_MONOTONIC, &wait_start);
		/* inner products in the order of :404-438; column d = the target vector */
		for(size_t pass = 0; pass < 2; pass++) {
			const size_t a_lo = pass == 0 ? i_lo : d, a_hi = pass == 0 ? i_max : (party_i == P - 1 ? d + 1 : d);
			for(size_t i = a_lo; i < a_hi; i++) {
				const size_t b_hi = (pass == 0 && party_j == P - 1) ? j_max + 1 : j_max;
				for(size_t jj = j_lo; jj < b_hi; jj++) {
					const size_t j = jj < j_max ? jj : d;        /* the extra step of pass 0: (i, target) */
					const int64_t slot = i == d ? (int64_t) (nA + j) : (j == d ? (int64_t) (nA + i) : (int64_t) idx(i, j));
					orecv(pd, 0, 0, 0);
					if(sender) {
						check((size_t) orecv(pd, 0, u_host, u_bytes) == u_bytes, "orecv: %s", strerror(errno));
						GCHECK(p1_group_cot_sender_step(g.grp, sess, args->data, (int32_t) i, u_host, c_host, slot));
						check(osend(pd, 0, c_host, c_bytes) >= 0, "osend: %s", strerror(errno));
					} else {
						GCHECK(p1_group_cot_receiver_extend(g.grp, sess, args->data, (int32_t) j, u_host));
						check(osend(pd, 0, u_host, u_bytes) >= 0, "osend: %s", strerror(errno));
						check((size_t) orecv(pd, 0, c_host, c_bytes) == c_bytes, "orecv: %s", strerror(errno));
						GCHECK(p1_group_cot_receiver_finish(g.grp, sess, args->data, c_host, slot));
					}
				}
			}
		}
		orecv(pd, 0, 0, 0);
		clock_gettime(CLOCK_MONOTONIC, &wait_end);
		args->wait_total.tv_sec += (wait_end.tv_sec - wait_start.tv_sec);
		args->wait_total.tv_nsec += (wait_end.tv_nsec - wait_start.tv_nsec);
	}
	GCHECK(p1_group_reduce(g.grp));
	GCHECK(p1_copy_d2h(g.ctx, args->acc_host, p1_group_acc_dev(g.grp, 0), acc_len * 8));
	args->status = 0;
error:
	if(g.grp) p1_group_sync(g.grp);
	if(sess) p1_group_cot_free(sess);
	if(g.ctx) {
		if(u_host) p1_host_free(g.ctx, u_host);
		if(c_host) p1_host_free(g.ctx, c_host);
	}
	gpu_free(&g);
	return NULL;
}

/* ------------------------------------------------------------------ data provider */

int run_party(node *self, config *c, int precision, int precision_p2, struct timespec *wait_total,
	ufixed_t **res_A, ufixed_t **res_b, bool use_ot) {
	gpu_state g;
	memset(&g, 0, sizeof(g));
	p1_gdata *data = NULL;
	uint64_t *acc_host = NULL;
	int32_t *cols = NULL;
	int64_t *slots = NULL;
	uint64_t *vals_host = NULL;
	uint8_t my_key[16];
	const int seeded = !use_ot && seeded_masks();
	int status = 1;
	if(res_A) *res_A = NULL;
	if(res_b) *res_b = NULL;
	if(wait_total) wait_total->tv_sec = wait_total->tv_nsec = 0;
	const int me = c->party - 1, P = self->num_parties;
	const size_t d = c->d, nA = d * (d + 1) / 2;
	check(me >= 2 && me < P, "run_party: party %d is not a data provider", c->party);
	const size_t lo = (size_t) c->index_owned[me], hi = me < P - 1 ? (size_t) c->index_owned[me + 1] : d;

	double t_ti = 0, t_peer = 0, t_gpu = 0, t_send = 0;
	const double t_begin = now_s();
	check(!gpu_init(&g, c->n, use_ot ? 0 : 3, use_ot ? 0 : 1), "could not set up the GPU");
	const double t_init = now_s();
	status = load_inputs(&g, c, precision, &data);
	check(!status, "Could not read data");
	status = 1;
	const double t_loaded = now_s();
	const size_t acc_len = p1_acc_len(d);

	if(use_ot) {
		/* one thread per provider, ourselves included, each with a private result (:494-529) */
		const int T = P - 2;
		acc_host = calloc(acc_len, 8);
		pthread_t *th = malloc((size_t) T * sizeof(*th));
		ot_thread_args *ta = calloc((size_t) T, sizeof(*ta));
		int bad = !th || !ta || !acc_host, started = 0;
		for(int peer = 2; peer < P && !bad; peer++) {
			ta[peer - 2] = (ot_thread_args) {.self = self, .c = c, .data = data, .peer = peer,
				.acc_host = calloc(acc_len, 8), .wait_total = {0, 0}, .status = 1};
			if(!ta[peer - 2].acc_host || pthread_create(&th[peer - 2], NULL, run_party_ot_thread_gpu, &ta[peer - 2])) {
				fprintf(stderr, "run_party: could not start the thread for party %d\n", peer + 1);
				free(ta[peer - 2].acc_host);
				bad = 1;
				break;
			}
			started++;
		}
		for(int k = 0; k < started; k++) {
			pthread_join(th[k], NULL);
			for(size_t w = 0; w < acc_len; w++) acc_host[w] += ta[k].acc_host[w];   /* ring sum (:515-521) */
			free(ta[k].acc_host);
			bad |= ta[k].status;
			if(wait_total) {
				wait_total->tv_sec += ta[k].wait_total.tv_sec;
				wait_total->tv_nsec += ta[k].wait_total.tv_nsec;
			}
		}
		free(th);
		free(ta);
		check(!bad, "an OT thread failed");
	} else {
		GCHECK(p1_group_acc_reset(g.grp, d));
		if(seeded) {
			size_t size = 0;
			check(!recv_wire(self->peer[0], my_key, sizeof(my_key), &size) && size == sizeof(my_key), "Could not receive the mask key from the TI");
			GCHECK(p1_host_alloc(g.ctx, g.batch * 8, (void **) &vals_host));
		}
		/* pairs we own both rows of, and the diagonal statistics (:562-571) */
		GCHECK(p1_group_local_block(g.grp, data, lo, hi, me == P - 1, P1_GRAM_AUTO));
		cols = malloc(g.batch * sizeof(*cols));
		slots = malloc(g.batch * sizeof(*slots));
		check(cols && slots, "malloc: %s", strerror(errno));
		pair_cursor pc = {0, 0, 0};
		pair_run r;
		struct timespec wait_start, wait_end;
		while(next_run(&pc, c, g.batch, &r)) {
			check(r.owner_i >= 2 && r.owner_j >= 2, "Invalid owners %d, %d for rows %zd, %zd", r.owner_i, r.owner_j, r.i, r.j0);
			if(r.owner_i != me && r.owner_j != me) continue;        /* we own neither row: skip (:556) */
			const size_t count = r.count;
			for(size_t p = 0; p < count; p++) {
				const size_t j = r.j0 + p;
				slots[p] = r.i < d ? (int64_t) idx(r.i, j) : (int64_t) (nA + j);   /* :581-585 */
				cols[p] = r.owner_i == me ? (int32_t) r.i : (int32_t) j;           /* column d = the target vector */
			}
			/* random vectors and values from the TI (:170) */
			double t0 = now_s(), t1;
			if(seeded) {
				/* the values arrive, the vector is expanded here: y (party a) or x (party b) of the run's pairs under our key */
				for(size_t p = 0; p < count; p++) {
					uint8_t m[16];
					size_t size = 0;
					check(!recv_wire(self->peer[0], m, sizeof(m), &size), "Could not receive message from TI");
					check(!value_msg_decode(m, size, &vals_host[p]), "msg__unpack: malformed value message from the TI");
				}
				GCHECK(p1_copy_h2d(g.ctx, g.vals_dev, vals_host, count * 8));
				const uint8_t *kx = r.owner_i == me ? NULL : my_key, *ky = r.owner_i == me ? my_key : NULL;
				if(g.ndev == 1) {
					GCHECK(p1_ti_generate_keys(g.ctx, kx, ky, NULL, r.q0, count, c->n, 0, c->n, kx ? g.vec[0] : NULL, ky ? g.vec[0] : NULL, g.ld, NULL, NULL));
				} else {
					GCHECK(p1_group_ti_generate_keys(g.grp, kx, ky, NULL, r.q0, count, c->n, kx ? g.gv[0] : NULL, ky ? g.gv[0] : NULL, NULL, NULL));
				}
			} else {
				check(!recv_to_device(&g, self->peer[0], count, g.vec[0], g.vals_dev), "Could not receive message from TI");
			}
			t_ti += (t1 = now_s()) - t0;
			if(r.owner_i == me) {                                  /* party a (:173-196) */
				ProtocolDesc *pb = self->peer[r.owner_j];
				clock_gettime(CLOCK_MONOTONIC, &wait_start);
				status = recv_to_device(&g, pb, count, g.vec[1], g.vals2_dev);
				clock_gettime(CLOCK_MONOTONIC, &wait_end);
				check(!status, "Could not receive message from party B (%d)", r.owner_j);
				status = 1;
				t_peer += (t0 = now_s()) - t1;
				GCHECK(do_a_step(&g, data, cols, slots, count, seeded));           /* out = a - y, share += <b + x, y> */
				check(!pack_start(&g, 0, 1, g.vec[2], g.zero_dev, count), "could not serialise the messages for party B");
				GCHECK(wire_mark(&g, 0));
				GCHECK(wire_ready(&g, 0));
				t_gpu += (t1 = now_s()) - t0;
				for(size_t p = 0; p < count; p++)
					check(!send_wire(pb, wire_host_at(&g, 0, 1) + p * g.stride, sizes_host_at(&g, 0, 1)[p]), "Could not send message to party B (%d)", r.owner_j);
				t_send += now_s() - t1;
			} else {                                             /* party b (:198-224) */
				ProtocolDesc *pa = self->peer[r.owner_i];
				GCHECK(do_b_mask(&g, data, cols, count, seeded));                  /* out = b + x */
				check(!pack_start(&g, 0, 1, g.vec[2], g.zero_dev, count), "could not serialise the messages for party A");
				GCHECK(wire_mark(&g, 0));
				GCHECK(wire_ready(&g, 0));
				t_gpu += (t0 = now_s()) - t1;
				for(size_t p = 0; p < count; p++)
					check(!send_wire(pa, wire_host_at(&g, 0, 1) + p * g.stride, sizes_host_at(&g, 0, 1)[p]), "Could not send message to party A (%d)", r.owner_i);
				t_send += (t1 = now_s()) - t0;
				clock_gettime(CLOCK_MONOTONIC, &wait_start);
				status = recv_to_device(&g, pa, count, g.vec[1], g.vals2_dev);
				clock_gettime(CLOCK_MONOTONIC, &wait_end);
				check(!status, "Could not receive message from party A (%d)", r.owner_i);
				status = 1;
				t_peer += now_s() - t1;
				GCHECK(do_b_share(&g, data, cols, slots, count));          /* share += <a - y, b> */
			}
			GCHECK(p1_group_acc_sub_values(g.grp, d, slots, g.vals_dev, count));      /* share -= value (:196,222) */
			if(wait_total) {
				wait_total->tv_sec += (wait_end.tv_sec - wait_start.tv_sec);
				wait_total->tv_nsec += (wait_end.tv_nsec - wait_start.tv_nsec);
			}
		}
	}

	/* tail (:605-641): ring sum over the devices, diagonal words, 64 -> 32 shift, calloc'ed outputs the caller frees */
	{
		void *A = NULL, *B = NULL;
		GCHECK(p1_group_finalize(g.grp, data, acc_host, lo, hi, precision, precision_p2, (double) c->normalizer2, &A, &B));
		if(res_A) *res_A = A; else free(A);
		if(res_b) *res_b = B; else free(B);
	}
	if(trace_on()) fprintf(stderr, "p1gpu trace party %d: setup %.3f  inputs %.3f  protocol %.3f (from the TI %.3f, from the peer %.3f, "
		"GPU + D2H %.3f, send %.3f)  total %.3f s, %d device(s), batch %zu\n", c->party, t_init - t_begin, t_loaded - t_init,
		now_s() - t_loaded, t_ti, t_peer, t_gpu, t_send, now_s() - t_begin, g.ndev, g.batch);
	status = 0;
error:
	if(g.grp) p1_group_sync(g.grp);
	if(data) p1_gdata_free(data);
	if(vals_host) p1_host_free(g.ctx, vals_host);
	free(acc_host);
	free(cols);
	free(slots);
	gpu_free(&g);
	return status;
}
