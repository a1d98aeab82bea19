/* TEST INFRASTRUCTURE (oracle/_ref build only). Not part of the product path.
 *
 * Runs the reference's own phase 1 — run_trusted_initializer / run_party from
 * /root/reference/src/phase1.c, compiled unmodified — with every party as a thread
 * of this process over the loopback transport of p1shim.c.  It plays the role of
 * /root/reference/src/cmd/secure_multiplication.c:68-104 (config_new, node_new,
 * barrier, timed run_*), minus TCP.  Outputs every data provider's share words.
 */
#define _POSIX_C_SOURCE 201605L
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <math.h>

#include "phase1.h"
#include "p1shim_internal.h"

typedef struct {
	const char *path;
	int party; /* 1-based */
	int precision, precision_p2, use_ot;
	node *self;
	pthread_barrier_t *bar;
	ufixed_t *res_A, *res_b;
	double wall, cpu, wait;
	int status;
} party_args;

static double ts_diff(struct timespec a, struct timespec b) {
	return (double) (b.tv_sec - a.tv_sec) + 1e-9 * (double) (b.tv_nsec - a.tv_nsec);
}

static void *party_main(void *v) {
	party_args *pa = v;
	config *c = NULL;
	struct timespec w0, w1, c0, c1, wait_total = {0, 0};
	pa->status = config_new(&c, pa->path);
	if(!pa->status) c->party = pa->party;
	pthread_barrier_wait(pa->bar);
	clock_gettime(CLOCK_MONOTONIC, &w0);
	clock_gettime(CLOCK_THREAD_CPUTIME_ID, &c0);
	if(!pa->status) {
		if(pa->party == 1) {
			pa->status = run_trusted_initializer(pa->self, c, pa->precision, pa->use_ot);
		} else if(pa->party > 2) {
			pa->status = run_party(pa->self, c, pa->precision, pa->precision_p2, &wait_total,
				&pa->res_A, &pa->res_b, pa->use_ot);
		}
	}
	clock_gettime(CLOCK_MONOTONIC, &w1);
	clock_gettime(CLOCK_THREAD_CPUTIME_ID, &c1);
	pa->wall = ts_diff(w0, w1);
	pa->cpu = ts_diff(c0, c1);
	pa->wait = (double) wait_total.tv_sec + 1e-9 * (double) wait_total.tv_nsec;
	config_destroy(&c);
	return NULL;
}

int p1ref_bits_p1(void) { return FIXED_BIT_SIZE; }
int p1ref_bits_p2(void) { return FIXED_BIT_SIZE_P2; }

/* the reference's scalar codec and index map, exported for unit-level pinning */
int64_t p1ref_double_to_fixed(double d, int p) { return (int64_t) double_to_fixed(d, p); }
int64_t p1ref_double_to_fixed_p1(double d, int p) { return (int64_t) double_to_fixed_p1(d, p); }
double p1ref_fixed_to_double_p1(int64_t f, int p) { return fixed_to_double_p1((fixed_p1_t) f, p); }
size_t p1ref_idx(size_t i, size_t j) { return idx(i, j); }

/* reference text parse + encode (linear.c:27-102) of the X and y blocks of an instance file */
int p1ref_read_inputs(const char *path, int precision, int64_t *X_out, int64_t *y_out, double *seconds) {
	config *c = NULL;
	matrix_t data = {.value = NULL};
	vector_t target = {.value = NULL};
	struct timespec t0, t1;
	if(config_new(&c, path)) return 1;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	double normalizer = sqrt(pow(2, precision) * c->normalizer1); /* phase1.c:473 */
	int st = read_matrix(c->input, &data, precision, true, normalizer);
	if(!st) st = read_vector(c->input, &target, precision, true, normalizer);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	if(seconds) *seconds = ts_diff(t0, t1);
	if(!st) {
		if(X_out) for(size_t i = 0; i < data.d[0] * data.d[1]; i++) X_out[i] = data.value[i];
		if(y_out) for(size_t i = 0; i < target.len; i++) y_out[i] = target.value[i];
	}
	free(data.value);
	free(target.value);
	config_destroy(&c);
	return st;
}

/* shares_out: num_dp x (d(d+1)/2 + d) entries, widened to uint64_t.
 * stats_out : num_parties x 3 doubles (wall, cpu, wait seconds), may be NULL.
 * bytes_out : num_parties x num_parties bytes sent [from][to], may be NULL. */
int p1ref_run(const char *path, int precision, int precision_p2, int use_ot,
	const uint8_t ti_key[16], const uint8_t ot_key[16],
	uint64_t *shares_out, double *stats_out, uint64_t *bytes_out) {
	config *c0 = NULL;
	if(config_new(&c0, path)) return 1;
	int P = c0->num_parties;
	size_t d = c0->d;
	config_destroy(&c0);
	size_t nA = d * (d + 1) / 2;

	p1shim_set_keys(ti_key, ot_key);

	/* full mesh of directed pipes (node.c:26-62 builds the same mesh over TCP) */
	p1shim_pipe **pipes = calloc((size_t) P * P, sizeof(*pipes));
	for(int a = 0; a < P; a++) for(int b = 0; b < P; b++) if(a != b) pipes[a * P + b] = p1shim_pipe_new(1 << 22);
	node *nodes = calloc(P, sizeof(node));
	for(int a = 0; a < P; a++) {
		nodes[a].party = a + 1;
		nodes[a].num_parties = P;
		nodes[a].peer = calloc(P, sizeof(ProtocolDesc *));
		for(int b = 0; b < P; b++) if(a != b) nodes[a].peer[b] = p1shim_pd_new(pipes[a * P + b], pipes[b * P + a], a, b);
	}

	pthread_barrier_t bar;
	pthread_barrier_init(&bar, NULL, P);
	party_args *pa = calloc(P, sizeof(*pa));
	pthread_t *th = calloc(P, sizeof(*th));
	for(int a = 0; a < P; a++) {
		pa[a] = (party_args) {.path = path, .party = a + 1, .precision = precision, .precision_p2 = precision_p2,
			.use_ot = use_ot, .self = &nodes[a], .bar = &bar};
		pthread_create(&th[a], NULL, party_main, &pa[a]);
	}
	int status = 0;
	for(int a = 0; a < P; a++) {
		pthread_join(th[a], NULL);
		status |= pa[a].status;
	}
	for(int a = 2; a < P && !status; a++) {
		uint64_t *row = shares_out + (size_t) (a - 2) * (nA + d);
		for(size_t i = 0; i < nA; i++) row[i] = pa[a].res_A[i];
		for(size_t i = 0; i < d; i++) row[nA + i] = pa[a].res_b[i];
	}
	for(int a = 0; a < P; a++) {
		if(stats_out) {
			stats_out[3 * a] = pa[a].wall;
			stats_out[3 * a + 1] = pa[a].cpu;
			stats_out[3 * a + 2] = pa[a].wait;
		}
		free(pa[a].res_A);
		free(pa[a].res_b);
		for(int b = 0; b < P; b++) {
			if(a == b) continue;
			if(bytes_out) bytes_out[a * P + b] = p1shim_pipe_total(pipes[a * P + b]);
			free(nodes[a].peer[b]);
		}
		free(nodes[a].peer);
	}
	for(int i = 0; i < P * P; i++) p1shim_pipe_free(pipes[i]);
	pthread_barrier_destroy(&bar);
	free(pipes);
	free(nodes);
	free(pa);
	free(th);
	return status;
}

/* ---- self-test of the drop-in's default base-OT provider (linreg-mpc_b200/host/base_ot.c), present only in
 * the gpuhost build: both roles as two threads over one loop-back connection */
#ifdef P1_WITH_BASE_OT_SELFTEST
int p1host_base_ot_co15(ProtocolDesc *pd, int sender_party, int receiver_party, int role,
	uint8_t seeds0[128][16], uint8_t seeds1[128][16], uint8_t delta[16]);

typedef struct {
	ProtocolDesc *pd;
	int role, status;
	uint8_t (*s0)[16], (*s1)[16], *delta;
} bot_args;

static void *bot_main(void *v) {
	bot_args *a = v;
	a->status = p1host_base_ot_co15(a->pd, 3, 2, a->role, a->s0, a->s1, a->delta);
	return NULL;
}

/* recv_s0 / recv_s1: the extension receiver's k_i^0, k_i^1; send_s: the extension sender's k_i^{Delta_i}; delta: its choice string */
int p1ref_base_ot_selftest(uint8_t recv_s0[128][16], uint8_t recv_s1[128][16], uint8_t send_s[128][16], uint8_t delta[16]) {
	p1shim_pipe *ab = p1shim_pipe_new(1 << 16), *ba = p1shim_pipe_new(1 << 16);
	ProtocolDesc *pa = p1shim_pd_new(ab, ba, 0, 1), *pb = p1shim_pd_new(ba, ab, 1, 0);
	uint8_t dummy_delta[16], dummy_s1[128][16];
	bot_args ra = {.pd = pa, .role = 0, .s0 = recv_s0, .s1 = recv_s1, .delta = dummy_delta};
	bot_args sa = {.pd = pb, .role = 1, .s0 = send_s, .s1 = dummy_s1, .delta = delta};
	pthread_t t0, t1;
	pthread_create(&t0, NULL, bot_main, &ra);
	pthread_create(&t1, NULL, bot_main, &sa);
	pthread_join(t0, NULL);
	pthread_join(t1, NULL);
	free(pa);
	free(pb);
	p1shim_pipe_free(ab);
	p1shim_pipe_free(ba);
	return ra.status | sa.status;
}
#endif
