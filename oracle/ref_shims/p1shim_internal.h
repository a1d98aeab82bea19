/* TEST INFRASTRUCTURE (oracle/_ref build only): internals shared by p1shim.c and ref_driver.c */
#pragma once
#include "p1shim.h"

typedef struct p1shim_pipe p1shim_pipe;
p1shim_pipe *p1shim_pipe_new(size_t cap);
void p1shim_pipe_free(p1shim_pipe *p);
uint64_t p1shim_pipe_total(p1shim_pipe *p);

struct ProtocolDesc {
	p1shim_pipe *tx, *rx;
	int self, peer; /* 0-based party indices (party-1) */
};
ProtocolDesc *p1shim_pd_new(p1shim_pipe *tx, p1shim_pipe *rx, int self, int peer);

void p1shim_set_keys(const uint8_t ti_key[16], const uint8_t ot_key[16]);
