/* phase1_gpu.c — replaces src/phase1.c of schoppmp/linreg-mpc in the reference's own build:
 * the same two entry points (src/phase1.h:14-29), the same callers (src/cmd/linreg.c:116-123,
 * src/cmd/secure_multiplication.c:86-89), the same messages on the same sockets, with every
 * arithmetic block of phase 1 executed by libp1gpu.so (include/p1gpu.h) on a B200.
 *
 *   reference block (src/phase1.c)                 here
 *   read_matrix / read_vector (:473-477)           fread + p1_parse_doubles + p1_encode
 *   TI loop body (:270-284)                        p1_ti_generate, p1_ti_values, p1_msg_pack
 *   inner_product_ti (:148-236)                    p1_msg_unpack, p1_party_b_mask / _a_step / _b_share,
 *                                                  p1_msg_pack, p1_acc_sub_values
 *   local pairs and diagonal (:562-571, :359-383)  p1_local_block (+ p1_diag_sequential)
 *   run_party_ot_thread (:342-450)                 p1_cot_* with U / C exchanged over the peer socket
 *   tail (:605-641)                                p1_finalize
 *
 * It compiles against the reference's headers (phase1.h, node.h, config.h, fixed.h: the ring
 * widths are the reference's compile-time BIT_WIDTH_32_P1 / _P2) and needs nothing of Obliv-C
 * beyond ProtocolDesc / osend / orecv.  Messages keep the framing of send_pmsg / recv_pmsg
 * (:100-145): raw size_t length, proto2 bytes, zero-length orecv flush.
 *
 * Batching.  The reference walks the pairs one by one.  Here consecutive cross pairs between the
 * SAME two providers (a run: j sweeps one provider's columns for a fixed row i) are processed
 * together, up to `batch` pairs: the TI sends the run's 2B messages in the reference's order
 * (a, b, a, b, ...), party b sends its B masked columns, party a answers with B, so every socket
 * carries exactly the reference's byte stream.  All parties visit the runs in the global pair
 * order, hence the earliest unfinished run can always progress (no new deadlocks).
 *
 * Two things the reference obtains inside Obliv-C are hooks here (weak defaults below):
 *   p1host_ti_key   : the TI's PRG key (newBCipherRandomGen draws a random one, :243)
 *   p1host_base_ot  : the kappa = 128 base OTs of honestOTExt{Sender,Recver}New (:394,397)
 * There is no CPU fallback: without a GPU p1_ctx_create fails and so do both entry points.
 */
#define _POSIX_C_SOURCE 201605L
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <sys/random.h>

#include "phase1.h"
#include "p1gpu.h"

#define GCHECK(call) check((call) == 0, "%s:%d: %s", __FILE__, __LINE__, p1_last_error())

/* ------------------------------------------------------------------ hooks */

__attribute__((weak)) int p1host_ti_key(uint8_t key[16]) {
	return getrandom(key, 16, 0) == 16 ? 0 : 1;
}

/* role: 1 = this party is the OT-extension sender (gets delta and seeds0 = k_i^{delta_i}),
 * 0 = receiver (gets seeds0 = k_i^0 and seeds1 = k_i^1).  Parties are 0-based node indices. */
__attribute__((weak)) int p1host_base_ot(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]) {
	(void) pd; (void) sender_party; (void) receiver_party; (void) role; (void) seeds0; (void) seeds1; (void) delta;
	fprintf(stderr, "p1host_base_ot: link the base-OT provider (Obliv-C's npot sender/receiver) to use --use_ot\n");
	return 1;
}

__attribute__((weak)) int p1host_device(void) {
	const char *e = getenv("P1GPU_DEVICE");
	return e ? atoi(e) : 0;
}

/* ------------------------------------------------------------------ small helpers */

/* owner of a row as a 0-based node index; the target (row d) belongs to the last party (:25-33) */
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

static size_t pick_batch(size_t n) {
	/* ~64 MB of ring words per vector buffer, at most 32 pairs (a run is d/p pairs long) */
	size_t b = ((size_t) 64 << 20) / (n * sizeof(ufixed_p1_t) + 1);
	const char *e = getenv("P1GPU_BATCH");
	if(e && atoi(e) > 0) b = (size_t) atoi(e);
	return b < 1 ? 1 : (b > 32 ? 32 : b);
}

/* device + host buffers one side of the masking protocol needs for a batch */
typedef struct {
	p1_ctx *ctx;
	size_t n, ld, stride, batch;
	void *vec[3];                 /* batch x ld ring words: TI masks / received / produced */
	uint8_t *wire_dev, *wire_host; /* batch x stride bytes */
	uint64_t *sizes_dev, *vals_dev, *vals2_dev, *zero_dev;
	int32_t *status_dev;
	uint64_t *sizes_host;
	int32_t *status_host;
} gpu_bufs;

static void bufs_free(gpu_bufs *g) {
	if(!g->ctx) return;
	for(int k = 0; k < 3; k++) if(g->vec[k]) p1_dev_free(g->ctx, g->vec[k]);
	void *dv[] = {g->wire_dev, g->sizes_dev, g->vals_dev, g->vals2_dev, g->zero_dev, g->status_dev};
	for(size_t k = 0; k < sizeof(dv) / sizeof(dv[0]); k++) if(dv[k]) p1_dev_free(g->ctx, dv[k]);
	if(g->wire_host) p1_host_free(g->ctx, g->wire_host);
	free(g->sizes_host);
	free(g->status_host);
	p1_ctx_destroy(g->ctx);
	memset(g, 0, sizeof(*g));
}

static int bufs_init(gpu_bufs *g, size_t n, int nvec) {
	memset(g, 0, sizeof(*g));
	GCHECK(p1_ctx_create(p1host_device(), FIXED_BIT_SIZE, FIXED_BIT_SIZE_P2, &g->ctx));
	g->n = n;
	g->ld = (n + 255) / 256 * 256;
	g->batch = pick_batch(n);
	g->stride = p1_msg_max_bytes(g->ctx, n);
	for(int k = 0; k < nvec; k++) GCHECK(p1_dev_alloc(g->ctx, g->batch * g->ld * sizeof(ufixed_p1_t), &g->vec[k]));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * g->stride, (void **) &g->wire_dev));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->sizes_dev));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals_dev));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->vals2_dev));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * 8, (void **) &g->zero_dev));
	GCHECK(p1_dev_zero(g->ctx, g->zero_dev, g->batch * 8));
	GCHECK(p1_dev_alloc(g->ctx, g->batch * 4, (void **) &g->status_dev));
	GCHECK(p1_host_alloc(g->ctx, g->batch * g->stride, (void **) &g->wire_host));
	g->sizes_host = calloc(g->batch, 8);
	g->status_host = calloc(g->batch, 4);
	check(g->sizes_host && g->status_host, "malloc: %s", strerror(errno));
	return 0;
error:
	bufs_free(g);
	return 1;
}

/* serialise `count` messages {vector = vec[p], value = values_dev[p]} and bring them to the host */
static int pack_to_host(gpu_bufs *g, const void *vec, const uint64_t *values_dev, size_t count) {
	GCHECK(p1_msg_pack(g->ctx, vec, g->ld, g->n, count, values_dev, g->wire_dev, g->stride, g->sizes_dev));
	GCHECK(p1_copy_d2h(g->ctx, g->sizes_host, g->sizes_dev, count * 8));
	for(size_t p = 0; p < count; p++)
		GCHECK(p1_copy_d2h(g->ctx, g->wire_host + p * g->stride, g->wire_dev + p * g->stride, g->sizes_host[p]));
	return 0;
error:
	return 1;
}

/* receive `count` messages from pd and parse them on the GPU into vec / values_dev */
static int recv_to_device(gpu_bufs *g, ProtocolDesc *pd, size_t count, void *vec, uint64_t *values_dev) {
	for(size_t p = 0; p < count; p++) {
		size_t size = 0;
		check(!recv_wire(pd, g->wire_host + p * g->stride, g->stride, &size), "could not receive message %zu of the batch", p);
		g->sizes_host[p] = size;
		GCHECK(p1_copy_h2d(g->ctx, g->wire_dev + p * g->stride, g->wire_host + p * g->stride, (size + 15) / 16 * 16));
	}
	GCHECK(p1_msg_unpack(g->ctx, g->wire_dev, g->stride, g->sizes_host, count, g->n, vec, g->ld, values_dev, g->status_dev));
	GCHECK(p1_copy_d2h(g->ctx, g->status_host, g->status_dev, count * 4));
	for(size_t p = 0; p < count; p++) check(g->status_host[p] == 0, "msg__unpack: malformed message (flags %d)", g->status_host[p]);
	return 0;
error:
	return 1;
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

static int next_run(pair_cursor *pc, const config *c, size_t batch,
	size_t *ri, size_t *rj0, size_t *count, uint64_t *q0, int *owner_i, int *owner_j) {
	const size_t d = c->d;
	for(; pc->i <= d; pc->i++, pc->j = 0) {
		const int oi = owner_of(pc->i, c);
		while(pc->j <= pc->i && pc->j < d) {
			const int oj = owner_of(pc->j, c);
			if(oi == oj) { pc->j++; continue; }
			size_t k = 0;
			while(k < batch && pc->j + k <= pc->i && pc->j + k < d && owner_of(pc->j + k, c) == oj) k++;
			*ri = pc->i; *rj0 = pc->j; *count = k; *q0 = pc->q; *owner_i = oi; *owner_j = oj;
			pc->j += k;
			pc->q += k;
			return 1;
		}
	}
	return 0;
}

/* ------------------------------------------------------------------ trusted initializer */

int run_trusted_initializer(node *self, config *c, int precision, bool use_ot) {
	(void) precision;
	gpu_bufs g;
	memset(&g, 0, sizeof(g));
	uint8_t *wire_b = NULL;
	uint64_t *sizes_b = NULL, *xy_dev = NULL, *r_dev = NULL;
	int status = 1;
	if(use_ot) return 0;                 /* the TI takes no part in the OT variant (:256) */
	check(!bufs_init(&g, c->n, 2), "could not set up the GPU");
	uint8_t key[16];
	check(!p1host_ti_key(key), "could not draw the PRG key");
	wire_b = malloc(g.batch * g.stride);
	sizes_b = calloc(g.batch, 8);
	check(wire_b && sizes_b, "malloc: %s", strerror(errno));
	GCHECK(p1_dev_alloc(g.ctx, g.batch * 8, (void **) &xy_dev));
	GCHECK(p1_dev_alloc(g.ctx, g.batch * 8, (void **) &r_dev));

	pair_cursor pc = {0, 0, 0};
	size_t i, j0, count;
	uint64_t q0;
	int party_a, party_b;
	while(next_run(&pc, c, g.batch, &i, &j0, &count, &q0, &party_a, &party_b)) {
		check(party_a >= 2 && party_b >= 2, "Invalid owners %d, %d for rows %zd, %zd", party_a, party_b, i, j0);
		/* x, y, r of the run's pairs, <x,y> (:270-274); the shard is the whole vector here */
		GCHECK(p1_ti_generate(g.ctx, key, q0, count, c->n, 0, c->n, g.vec[0], g.vec[1], g.ld, xy_dev, r_dev));
		GCHECK(p1_ti_values(g.ctx, xy_dev, r_dev, count, g.vals_dev, g.vals2_dev));      /* xy - r, r (:277-278) */
		/* msg_b = {x, r} first into the spare host buffer, then msg_a = {y, xy - r} */
		check(!pack_to_host(&g, g.vec[0], g.vals2_dev, count), "could not serialise the messages for party B");
		memcpy(sizes_b, g.sizes_host, count * 8);
		for(size_t p = 0; p < count; p++) memcpy(wire_b + p * g.stride, g.wire_host + p * g.stride, sizes_b[p]);
		check(!pack_to_host(&g, g.vec[1], g.vals_dev, count), "could not serialise the messages for party A");
		for(size_t p = 0; p < count; p++) {          /* the reference's order on the wire: a, b per pair (:281-284) */
			status = send_wire(self->peer[party_a], g.wire_host + p * g.stride, g.sizes_host[p]);
			check(!status, "Could not send message to party A (%d)", party_a);
			status = send_wire(self->peer[party_b], wire_b + p * g.stride, sizes_b[p]);
			check(!status, "Could not send message to party B (%d)", party_b);
		}
	}
	status = 0;
error:
	if(g.ctx) {
		if(xy_dev) p1_dev_free(g.ctx, xy_dev);
		if(r_dev) p1_dev_free(g.ctx, r_dev);
	}
	free(wire_b);
	free(sizes_b);
	bufs_free(&g);
	return status;
}

/* ------------------------------------------------------------------ data provider: inputs */

/* read_matrix + read_vector (src/linear.c:27-102) for the rest of c->input: the text goes to the
 * GPU as it is, is tokenised and converted there (p1_parse_doubles) and encoded from the device
 * buffer (p1_encode).  Tokens: "n m", n*m values, "l", l values. */
static int load_inputs(p1_ctx *ctx, config *c, int precision, p1_data **data) {
	char *text = NULL;
	double *tok_dev = NULL;
	int status = 1;
	*data = NULL;
	const long pos = ftell(c->input);
	check(pos >= 0 && fseek(c->input, 0, SEEK_END) == 0, "fseek: %s", strerror(errno));
	const long end = ftell(c->input);
	check(end >= pos && fseek(c->input, pos, SEEK_SET) == 0, "fseek: %s", strerror(errno));
	const size_t bytes = (size_t) (end - pos);
	text = malloc(bytes + 1);
	check(text, "Out of memory");
	check(fread(text, 1, bytes, c->input) == bytes, "fread: %s", strerror(errno));
	const size_t n = c->n, d = c->d, need = 2 + n * d + 1 + n;
	GCHECK(p1_dev_alloc(ctx, need * sizeof(double), (void **) &tok_dev));
	uint64_t n_tokens = 0, n_deferred = 0, deferred[4] = {0, 0, 0, 0};
	GCHECK(p1_parse_doubles(ctx, text, bytes, tok_dev, need, &n_tokens, &n_deferred, deferred, 4));
	check(n_tokens >= need, "fscanf: input ends after %llu of %zu values", (unsigned long long) n_tokens, need);
	check(n_deferred == 0, "fscanf: the token at byte %llu of the data section is not a plain decimal number",
		(unsigned long long) deferred[0]);
	double hdr[2], len;
	GCHECK(p1_copy_d2h(ctx, hdr, tok_dev, sizeof(hdr)));
	GCHECK(p1_copy_d2h(ctx, &len, tok_dev + 2 + n * d, sizeof(len)));
	check((size_t) len == n && (size_t) hdr[1] == d && (size_t) hdr[0] == n,
		"Input dimensions invalid: (%zd, %zd), %zd", (size_t) hdr[0], (size_t) hdr[1], (size_t) len);
	const double normalizer = sqrt(pow(2, precision) * c->normalizer1);   /* :473 */
	GCHECK(p1_encode(ctx, tok_dev + 2, tok_dev + 2 + n * d + 1, n, d, normalizer, precision, data));
	GCHECK(p1_ctx_sync(ctx));
	status = 0;
error:
	if(tok_dev) p1_dev_free(ctx, tok_dev);
	free(text);
	return status;
}

/* ------------------------------------------------------------------ data provider: --use_ot threads */

typedef struct {
	node *self;
	config *c;
	const p1_data *data;
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
	p1_ctx *ctx = NULL;
	p1_cot *sess = NULL;
	uint64_t *acc = NULL;
	void *u_dev = NULL, *c_dev = NULL;
	uint8_t *u_host = NULL, *c_host = NULL;
	args->status = 1;
	GCHECK(p1_ctx_create(p1host_device(), FIXED_BIT_SIZE, FIXED_BIT_SIZE_P2, &ctx));
	const size_t acc_len = p1_acc_len(d);
	GCHECK(p1_dev_alloc(ctx, acc_len * 8, (void **) &acc));
	GCHECK(p1_dev_zero(ctx, acc, acc_len * 8));

	if(me == args->peer) {
		/* local pairs, <col, y> for the last provider, diagonal statistics (:359-383) */
		const size_t lo = (size_t) c->index_owned[me], hi = me < P - 1 ? (size_t) c->index_owned[me + 1] : d;
		GCHECK(p1_local_block(ctx, args->data, lo, hi, me == P - 1, acc, P1_GRAM_AUTO));
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
			GCHECK(p1_cot_sender_new(ctx, delta, seeds0, &sess));
		} else {
			GCHECK(p1_cot_receiver_new(ctx, seeds0, seeds1, &sess));
		}
		const size_t u_bytes = p1_cot_u_bytes(ctx, n), c_bytes = p1_cot_c_bytes(ctx, n);
		GCHECK(p1_dev_alloc(ctx, u_bytes, &u_dev));
		GCHECK(p1_dev_alloc(ctx, c_bytes, &c_dev));
		u_host = malloc(u_bytes);
		c_host = malloc(c_bytes);
		check(u_host && c_host, "malloc: %s", strerror(errno));
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
						GCHECK(p1_copy_h2d(ctx, u_dev, u_host, u_bytes));
						GCHECK(p1_cot_sender_step(ctx, sess, args->data, (int32_t) i, n, 0, u_dev, c_dev, acc, slot));
						GCHECK(p1_copy_d2h(ctx, c_host, c_dev, c_bytes));
						check(osend(pd, 0, c_host, c_bytes) >= 0, "osend: %s", strerror(errno));
					} else {
						GCHECK(p1_cot_receiver_extend(ctx, sess, args->data, (int32_t) j, n, 0, u_dev));
						GCHECK(p1_copy_d2h(ctx, u_host, u_dev, u_bytes));
						check(osend(pd, 0, u_host, u_bytes) >= 0, "osend: %s", strerror(errno));
						check((size_t) orecv(pd, 0, c_host, c_bytes) == c_bytes, "orecv: %s", strerror(errno));
						GCHECK(p1_copy_h2d(ctx, c_dev, c_host, c_bytes));
						GCHECK(p1_cot_receiver_finish(ctx, sess, c_dev, acc, slot));
					}
				}
			}
		}
		orecv(pd, 0, 0, 0);
		clock_gettime(CLOCK_MONOTONIC, &wait_end);
		args->wait_total.tv_sec += (wait_end.tv_sec - wait_start.tv_sec);
		args->wait_total.tv_nsec += (wait_end.tv_nsec - wait_start.tv_nsec);
	}
	GCHECK(p1_copy_d2h(ctx, args->acc_host, acc, acc_len * 8));
	args->status = 0;
error:
	if(sess) p1_cot_free(sess);
	if(ctx) {
		if(u_dev) p1_dev_free(ctx, u_dev);
		if(c_dev) p1_dev_free(ctx, c_dev);
		if(acc) p1_dev_free(ctx, acc);
		p1_ctx_destroy(ctx);
	}
	free(u_host);
	free(c_host);
	return NULL;
}

/* ------------------------------------------------------------------ data provider */

int run_party(node *self, config *c, int precision, int precision_p2, struct timespec *wait_total,
	ufixed_t **res_A, ufixed_t **res_b, bool use_ot) {
	gpu_bufs g;
	memset(&g, 0, sizeof(g));
	p1_data *data = NULL;
	uint64_t *acc = NULL, *acc_host = NULL;
	int32_t *cols = NULL;
	int64_t *slots = NULL;
	double *diag_xy = NULL;
	int status = 1;
	if(res_A) *res_A = NULL;
	if(res_b) *res_b = NULL;
	if(wait_total) wait_total->tv_sec = wait_total->tv_nsec = 0;
	const int me = c->party - 1, P = self->num_parties;
	const size_t d = c->d, nA = d * (d + 1) / 2;
	check(me >= 2 && me < P, "run_party: party %d is not a data provider", c->party);
	const size_t lo = (size_t) c->index_owned[me], hi = me < P - 1 ? (size_t) c->index_owned[me + 1] : d;

	check(!bufs_init(&g, c->n, use_ot ? 0 : 3), "could not set up the GPU");
	status = load_inputs(g.ctx, c, precision, &data);
	check(!status, "Could not read data");
	status = 1;
	const size_t acc_len = p1_acc_len(d);
	acc_host = calloc(acc_len, 8);
	check(acc_host, "malloc: %s", strerror(errno));

	if(use_ot) {
		/* one thread per provider, ourselves included, each with a private result (:494-529) */
		const int T = P - 2;
		pthread_t *th = malloc((size_t) T * sizeof(*th));
		ot_thread_args *ta = calloc((size_t) T, sizeof(*ta));
		int bad = !th || !ta;
		for(int peer = 2; peer < P && !bad; peer++) {
			ta[peer - 2] = (ot_thread_args) {.self = self, .c = c, .data = data, .peer = peer,
				.acc_host = calloc(acc_len, 8), .wait_total = {0, 0}, .status = 1};
			pthread_create(&th[peer - 2], NULL, run_party_ot_thread_gpu, &ta[peer - 2]);
		}
		for(int peer = 2; peer < P && !bad; peer++) {
			pthread_join(th[peer - 2], NULL);
			for(size_t k = 0; k < acc_len; k++) acc_host[k] += ta[peer - 2].acc_host[k];   /* ring sum (:515-521) */
			free(ta[peer - 2].acc_host);
			if(wait_total) {
				wait_total->tv_sec += ta[peer - 2].wait_total.tv_sec;
				wait_total->tv_nsec += ta[peer - 2].wait_total.tv_nsec;
			}
		}
		for(int peer = 2; peer < P && !bad; peer++) bad |= ta[peer - 2].status;
		free(th);
		free(ta);
		check(!bad, "an OT thread failed");
	} else {
		GCHECK(p1_dev_alloc(g.ctx, acc_len * 8, (void **) &acc));
		GCHECK(p1_dev_zero(g.ctx, acc, acc_len * 8));
		/* pairs we own both rows of, and the diagonal statistics (:562-571) */
		GCHECK(p1_local_block(g.ctx, data, lo, hi, me == P - 1, acc, P1_GRAM_AUTO));
		cols = malloc(g.batch * sizeof(*cols));
		slots = malloc(g.batch * sizeof(*slots));
		check(cols && slots, "malloc: %s", strerror(errno));
		pair_cursor pc = {0, 0, 0};
		size_t i, j0, count;
		uint64_t q0;
		int owner_i, owner_j;
		struct timespec wait_start, wait_end;
		while(next_run(&pc, c, g.batch, &i, &j0, &count, &q0, &owner_i, &owner_j)) {
			check(owner_i >= 2 && owner_j >= 2, "Invalid owners %d, %d for rows %zd, %zd", owner_i, owner_j, i, j0);
			if(owner_i != me && owner_j != me) continue;        /* we own neither row: skip (:556) */
			for(size_t p = 0; p < count; p++) {
				const size_t j = j0 + p;
				slots[p] = i < d ? (int64_t) idx(i, j) : (int64_t) (nA + j);   /* :581-585 */
				cols[p] = owner_i == me ? (int32_t) i : (int32_t) j;           /* column d = the target vector */
			}
			/* random vectors and values from the TI (:170) */
			check(!recv_to_device(&g, self->peer[0], count, g.vec[0], g.vals_dev), "Could not receive message from TI");
			if(owner_i == me) {                                  /* party a (:173-196) */
				ProtocolDesc *pb = self->peer[owner_j];
				clock_gettime(CLOCK_MONOTONIC, &wait_start);
				status = recv_to_device(&g, pb, count, g.vec[1], g.vals2_dev);
				clock_gettime(CLOCK_MONOTONIC, &wait_end);
				check(!status, "Could not receive message from party B (%d)", owner_j);
				status = 1;
				/* out = a - y, share += <b + x, y> */
				GCHECK(p1_party_a_step(g.ctx, data, cols, slots, count, g.vec[0], g.ld, g.vec[1], g.ld, g.vec[2], g.ld, acc));
				check(!pack_to_host(&g, g.vec[2], g.zero_dev, count), "could not serialise the messages for party B");
				for(size_t p = 0; p < count; p++)
					check(!send_wire(pb, g.wire_host + p * g.stride, g.sizes_host[p]), "Could not send message to party B (%d)", owner_j);
			} else {                                             /* party b (:198-224) */
				ProtocolDesc *pa = self->peer[owner_i];
				GCHECK(p1_party_b_mask(g.ctx, data, cols, count, g.vec[0], g.ld, g.vec[2], g.ld));   /* out = b + x */
				check(!pack_to_host(&g, g.vec[2], g.zero_dev, count), "could not serialise the messages for party A");
				for(size_t p = 0; p < count; p++)
					check(!send_wire(pa, g.wire_host + p * g.stride, g.sizes_host[p]), "Could not send message to party A (%d)", owner_i);
				clock_gettime(CLOCK_MONOTONIC, &wait_start);
				status = recv_to_device(&g, pa, count, g.vec[1], g.vals2_dev);
				clock_gettime(CLOCK_MONOTONIC, &wait_end);
				check(!status, "Could not receive message from party A (%d)", owner_i);
				status = 1;
				GCHECK(p1_party_b_share(g.ctx, data, cols, slots, count, g.vec[1], g.ld, acc));      /* share += <a - y, b> */
			}
			GCHECK(p1_acc_sub_values(g.ctx, acc, slots, g.vals_dev, count));                         /* share -= value (:196,222) */
			if(wait_total) {
				wait_total->tv_sec += (wait_end.tv_sec - wait_start.tv_sec);
				wait_total->tv_nsec += (wait_end.tv_nsec - wait_start.tv_nsec);
			}
		}
		GCHECK(p1_copy_d2h(g.ctx, acc_host, acc, acc_len * 8));
	}

	/* tail (:605-641): diagonal words, 64 -> 32 shift, calloc'ed outputs the caller frees */
	{
		int need = 0;
		for(size_t col = lo; col < hi; col++) {
			const uint64_t s_lo = acc_host[nA + d + col], s_hi = acc_host[nA + 2 * d + col], viol = acc_host[nA + 3 * d + col];
			if(viol || s_hi >= (1ull << 22) || ((s_hi << 32) + s_lo) > (1ull << 53)) need = 1;
		}
		if(need) {      /* the exact shortcut does not apply: the literal sequential loop (:563-566) */
			diag_xy = calloc(hi - lo, sizeof(double));
			check(diag_xy, "malloc: %s", strerror(errno));
			GCHECK(p1_diag_sequential(g.ctx, data, lo, hi, precision, diag_xy));
		}
		void *A = NULL, *B = NULL;
		GCHECK(p1_finalize(g.ctx, acc_host, d, lo, hi, precision, precision_p2, (double) c->normalizer2, diag_xy, NULL, &A, &B));
		if(res_A) *res_A = A; else free(A);
		if(res_b) *res_b = B; else free(B);
	}
	status = 0;
error:
	if(g.ctx && acc) p1_dev_free(g.ctx, acc);
	if(data) p1_data_free(data);
	free(acc_host);
	free(cols);
	free(slots);
	free(diag_xy);
	bufs_free(&g);
	return status;
}
