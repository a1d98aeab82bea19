/* TEST INFRASTRUCTURE (oracle/_ref build only). Not part of the product path.
 *
 * Stand-ins for the external symbols that /root/reference/src/phase1.c pulls from
 * Obliv-C (obliv.h, obliv_common.h, obliv_types.h, obliv_bits.h, bcrandom.h) and
 * protobuf-c — neither is vendored under /root/reference.
 * They let the reference's phase1.c / linear.c / fixed.c / config.c compile
 * UNMODIFIED, from where they lie, into oracle/_ref/ with all parties running as
 * threads of one process:
 *   - ProtocolDesc/osend/orecv : in-process byte-stream loopback (blocking, bounded)
 *   - BCipherRandomGen         : AES-128-CTR byte stream (oracle/aes128.h), key injectable
 *   - honest*OTExt*            : correlated-OT extension defined in oracle/cot_ext.h
 *   - SecureMultiplication__Msg: proto2 wire format (packed varints), hand-written
 */
#pragma once
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <sys/types.h>

/* ---- transport ---- */
typedef struct ProtocolDesc ProtocolDesc;
int osend(ProtocolDesc *pd, int dest, const void *data, size_t len);
int orecv(ProtocolDesc *pd, int src, void *data, size_t len);

/* ---- PRG (bcrandom.h) ---- */
typedef struct BCipherRandomGen BCipherRandomGen;
BCipherRandomGen *newBCipherRandomGen(void);
void randomizeBuffer(BCipherRandomGen *gen, char *dest, size_t len);
void releaseBCipherRandomGen(BCipherRandomGen *gen);

/* ---- OT extension ---- */
struct HonestOTExtSender;
struct HonestOTExtRecver;
struct HonestOTExtSender *honestOTExtSenderNew(ProtocolDesc *pd, int destparty);
struct HonestOTExtRecver *honestOTExtRecverNew(ProtocolDesc *pd, int srcparty);
void honestOTExtSenderRelease(struct HonestOTExtSender *s);
void honestOTExtRecverRelease(struct HonestOTExtRecver *r);
typedef void (*OcOtCorrelator)(char *, const char *, int, void *);
void honestCorrelatedOTExtSend1Of2(struct HonestOTExtSender *s, char *opt0, char *opt1,
	int n, int len, OcOtCorrelator corr, void *corrArg);
void honestCorrelatedOTExtRecv1Of2(struct HonestOTExtRecver *r, char *dest, const bool *sel,
	int n, int len);
void dhRandomInit(void);
