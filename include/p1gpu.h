/* p1gpu.h — C-ABI of libp1gpu.so: B200 (sm_100a) kernels for linreg-mpc's phase-1
 * secure aggregation.  This is the drop-in boundary: the reference's host code keeps
 * its own entry points (run_trusted_initializer / run_party, src/phase1.h:14-29) and
 * calls the functions below where src/phase1.c, src/linear.c and src/fixed.c do the
 * arithmetic today.  Each entry cites the reference block it replaces (paths are
 * under the reference tree).  No torch types, plain pointers and sizes only.
 *
 * Conventions (mirroring src/check_error.h:5-9 and src/phase1.c:644-655):
 *   - every call returns 0 on success, 1 on failure; on failure a message is printed
 *     to stderr and kept for p1_last_error(); outputs are left NULL/untouched.
 *   - there is NO CPU fallback: without a CUDA device every call fails.
 *   - ring width is chosen at context creation (the reference's compile-time
 *     BIT_WIDTH_32_P1 / BIT_WIDTH_32_P2, src/fixed.h:4-30): (p1,p2) in
 *     {(64,64),(32,32),(64,32)}.
 *   - "ring word" = ufixed_p1_t: uint64_t when ring_bits_p1 == 64, else uint32_t.
 *   - pointers named *_dev must be device pointers of the context's GPU; pointers
 *     documented "host or device" are classified with cudaPointerGetAttributes.
 *   - share accumulators are always uint64_t (a Z_2^32 share is the low half), so a
 *     wrapping 64-bit sum all-reduce (NCCL ncclUint64/ncclSum) combines row shards
 *     exactly for both rings.
 *   - a context owns one CUDA stream; calls are asynchronous on it unless they return
 *     host data.  Calls on one context are serialised by an internal mutex, so the
 *     per-peer pthreads of OT mode (src/phase1.c:503-513) may share a context.
 */
#ifndef P1GPU_H
#define P1GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct p1_ctx p1_ctx;
typedef struct p1_data p1_data;       /* encoded X (and y), resident in HBM, column-major */
typedef struct p1_cot p1_cot;         /* one correlated-OT session with one peer */

/* ------------------------------------------------------------------ context */

int p1_ctx_create(int device, int ring_bits_p1, int ring_bits_p2, p1_ctx **ctx);
void p1_ctx_destroy(p1_ctx *ctx);
const char *p1_last_error(void);
int p1_ctx_sync(p1_ctx *ctx);
/* make `waiter`'s stream wait for everything enqueued so far on `signaler`'s stream
 * (a message hand-over between two parties that live on the same GPU) */
int p1_ctx_wait(p1_ctx *waiter, p1_ctx *signaler);
/* the context's cudaStream_t, for callers that time with CUDA events */
void *p1_ctx_stream(p1_ctx *ctx);
/* number of kernels this context has launched so far */
uint64_t p1_ctx_launches(p1_ctx *ctx);

/* Per-kernel-class device timing for bench.py's roofline line: while enabled, every
 * kernel launch of this context is bracketed by CUDA events on the context's stream.
 * p1_ctx_profile_read synchronises the stream, adds the elapsed times recorded since
 * the last read to the per-class totals and returns the totals of `kernel_class`. */
enum {
	P1_K_ENCODE = 0,      /* k_rows_to_cols / k_vec_encode */
	P1_K_LAYOUT,          /* k_cols_to_rows, word import/export */
	P1_K_DIAG,            /* k_diag_stats / k_diag_seq */
	P1_K_GRAM_IMAD,       /* k_gram_imad */
	P1_K_GRAM_LIMBS,      /* k_limb_split (operand planes of the tensor-core Gram) */
	P1_K_GRAM_TC,         /* k_gram_tc (tcgen05 kind::i8) */
	P1_K_TI,              /* k_ti_generate */
	P1_K_PARTY_B_MASK,    /* k_party<0> */
	P1_K_PARTY_A_STEP,    /* k_party<1> */
	P1_K_PARTY_B_SHARE,   /* k_party<2> */
	P1_K_COT_RECV,        /* k_cot_tiles<receiver> */
	P1_K_COT_SEND,        /* k_cot_tiles<sender> */
	P1_K_COT_FINISH,      /* k_cot_finish */
	P1_K_MISC,            /* k_acc_sub, k_ti_values */
	P1_K_MSG_PACK,        /* k_pack_len / k_pack_scan / k_pack_write */
	P1_K_MSG_UNPACK,      /* k_unpack_count / k_unpack_scan / k_unpack_write */
	P1_K_SHARE_BITS,      /* k_share_bits */
	P1_K_TEXT,            /* k_tok_count / k_tok_scan* / k_tok_parse */
	P1_K_MASKING_FUSED,   /* k_masking_fused */
	P1_K_COUNT
};
int p1_ctx_profile(p1_ctx *ctx, int enable);
int p1_ctx_profile_read(p1_ctx *ctx, int kernel_class, double *ms_total, uint64_t *launches);
const char *p1_kernel_class_name(int kernel_class);

/* Tuning knobs (0 = built-in choice).  Used by the parity tests to force the multi-chunk
 * and multi-slice paths at small sizes, and by bench.py sweeps.
 *   "gram_tc_max_slices"   : upper bound on the row slices (grid.y) of k_gram_tc
 *   "gram_tc_chunk_blocks" : 128-row blocks accumulated in s32 before a drain (only lowered)
 *   "gram_tc_min_cols"     : P1_GRAM_AUTO picks the tcgen05 kernel from this block width on
 *   "ti_threads"           : threads per CTA of the persistent mask-expansion kernel (one CTA per SM)
 *   "fused_threads"        : CTA size of the fused masking-loop kernel (p1_masking_fused): 0 = default (512), or one
 *                            of the compiled sizes 384, 512, 640, 768, 1024 (each with its own register budget)
 *   "fused_fast"           : 0 = every work unit of p1_masking_fused through the general loop (default 1: interior
 *                            units of a 64-bit ring with n_total even take the straight loop; identical results)
 *   "fused_chunk"          : consecutive work units a warp of p1_masking_fused takes at a time (0 = default 8)
 *   "gram_tc_whole"        : 0 = one-digit blocks of <= 256 columns also use the tiled tcgen05 kernel (default 1:
 *                            k_gram_whole, every CTA multiplies the whole block over its rows; identical results)
 *   "encode_async"         : 0 = never use the cp.async pipeline kernel of p1_encode (default 1: used whenever d
 *                            is even and the source is 16-byte aligned; identical results)
 *   "encode_planes"        : 1 = p1_encode also writes, in the same pass, the base-256 digit planes of ALL
 *                            columns that the tcgen05 Gram reads (no digit split before p1_local_block; +3..4
 *                            bytes of HBM per word).  Worth it when one p1_data serves every provider's block
 *                            (a box that holds all columns); a single party multiplies only its own columns
 *                            and leaves it off (default 0).  Results are identical either way. */
int p1_ctx_set_option(p1_ctx *ctx, const char *name, long value);

/* p1_dev_free / p1_host_free keep the block for the next request of the same size (up to P1GPU_POOL_MAX_MB, default
 * 8192, MB per device and as much page-locked memory per process; environment, read once; 0 = keep nothing): a party
 * that serves run after run asks for the same buffers again.  Blocks are zeroed before they are kept or released
 * (P1GPU_POOL_SCRUB: 1 = device blocks and page-locked blocks up to 1 MB, the default; 2 = page-locked blocks of any
 * size too; 0 = off), and so are the round keys and pads of a p1_cot session when it is freed.  p1_dev_free waits for
 * the whole device, as cudaFree does.  p1_pool_trim returns everything that is kept to the driver. */
int p1_dev_alloc(p1_ctx *ctx, size_t bytes, void **dev);
int p1_dev_free(p1_ctx *ctx, void *dev);
void p1_pool_trim(void);
int p1_dev_zero(p1_ctx *ctx, void *dev, size_t bytes);
/* page-locked host memory for buffers that cross PCIe every message (full-rate, truly asynchronous copies) */
int p1_host_alloc(p1_ctx *ctx, size_t bytes, void **host);
int p1_host_free(p1_ctx *ctx, void *host);
int p1_copy_h2d(p1_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes);
int p1_copy_d2h(p1_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes); /* synchronises */
/* device -> host without waiting (dst_host page-locked for a truly asynchronous copy); the data is there after
 * p1_ctx_sync, or after p1_ctx_mark_sync on a mark recorded behind the copy */
int p1_copy_d2h_async(p1_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes);
/* markers on the context's stream (slot 0..7): p1_ctx_mark records one behind everything enqueued so far,
 * p1_ctx_mark_sync blocks the calling host thread until the work before the mark is done - while later work
 * (the next batch) keeps running.  This is what lets the host file send batch k while the GPU prepares k+1. */
int p1_ctx_mark(p1_ctx *ctx, int slot);
int p1_ctx_mark_sync(p1_ctx *ctx, int slot);

/* ------------------------------------------------------------------ a2-a4: encode
 * Replaces the arithmetic of read_matrix / read_vector (src/linear.c:46-51,83-87):
 *     val /= normalizer;  word = (fixed_p1_t) double_to_fixed(val, precision)
 * with double_to_fixed = (fixed_t)(val * (1ll << precision)) (src/fixed.c:3-5), C
 * truncation with x86 cvttsd2si semantics (out of range / NaN -> INT_MIN of the
 * PHASE-2 width), then sign extension into the phase-1 ring.  Text parsing stays on
 * the host.  normalizer = sqrt(pow(2,precision) * n_total) (src/phase1.c:473).
 *   X : n x d row-major doubles (value[i*d+j], src/linear.c:51), host or device.
 *   y : n doubles or NULL (only the last provider needs the target).
 * The result lives on the GPU column-major (one contiguous n-vector per feature; the
 * target is column d), which is what every consumer below streams.  Per-column
 * statistics (max |word|, exact sum of squares) are gathered in the same pass. */
int p1_encode(p1_ctx *ctx, const double *X, const double *y, size_t n, size_t d,
              double normalizer, int precision, p1_data **out);
/* a provider that runs the protocol again on new values of the same shape (bench.py's step; a service that
 * re-fits as rows are corrected): the same pass into the store that an earlier call made - same n, d and
 * presence of y - with no allocation and no host synchronisation, ordered on the context's stream */
int p1_encode_again(p1_ctx *ctx, p1_data *data, const double *X, const double *y, double normalizer, int precision);
/* same, from ring words that are already encoded (row-major n x d, and n) */
int p1_data_from_words(p1_ctx *ctx, const void *X_words, const void *y_words, size_t n, size_t d,
                       p1_data **out);
void p1_data_free(p1_data *data);
/* encoded words back in the reference's layout (matrix_t.value row-major, vector_t.value) */
int p1_data_download(p1_ctx *ctx, const p1_data *data, void *X_words, void *y_words);
size_t p1_data_rows(const p1_data *data);
size_t p1_data_cols(const p1_data *data);
/* device pointer and leading dimension (in words) of the column-major store */
const void *p1_data_colmajor_dev(const p1_data *data, size_t *ld_words);

/* ------------------------------------------------------------------ share accumulator
 * layout of the uint64_t accumulator every call below adds into (length p1_acc_len(d)):
 *     [0, nA)            share_A, packed lower triangle, idx(i,j) = i(i+1)/2 + j (src/linear.c:11-16)
 *     [nA, nA+d)         share_b
 *     [nA+d, nA+2d)      diagonal: sum of squares, low part  } two sums of 32-bit halves of partial sums;
 *     [nA+2d, nA+3d)     diagonal: sum of squares, high part } (high << 32) + low = sum_k word_k^2 exactly
 *                                                              whenever the next entry is 0
 *     [nA+3d, nA+4d)     diagonal: number of rows with |word| >= 2^26
 * All entries are sums over rows, so row shards combine by a wrapping u64 all-reduce. */
size_t p1_acc_len(size_t d);

/* ------------------------------------------------------------------ a7/a8: local block
 * Replaces the "own both" branch and the diagonal branch of the pair loop
 * (src/phase1.c:562-571; OT-mode copy :359-383) for the provider owning columns
 * [col_lo, col_hi): for every pair col_lo <= j < i < col_hi
 *     acc[idx(i,j)] += sum_k X[k][i] * X[k][j]            (inner_product_local, :14-20)
 * and, if owns_target, acc[nA + j] += sum_k y[k] * X[k][j] for every owned j (:540-543),
 * plus the diagonal statistics of the owned columns (see p1_finalize).
 * algo: 0 = choose by size, 1 = 64-bit IMAD CUDA-core kernel, 2 = tcgen05 kind::i8
 * signed-limb GEMM (TMA -> smem -> tensor core -> TMEM). */
#define P1_GRAM_AUTO 0
#define P1_GRAM_IMAD 1
#define P1_GRAM_TCGEN05 2
int p1_local_block(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target,
                   uint64_t *acc_dev, int algo);

/* Number of balanced base-256 digits per word the tcgen05 path uses for this block (from
 * the max |word| statistics of the encode pass): the kernel executes 2*L*L int8 ops per
 * ring MAC.  Reporting aid for rooflines; synchronises the stream. */
int p1_gram_digits(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int owns_target, int *digits);

/* Sequential double-precision diagonal, the literal loop of src/phase1.c:563-566:
 *     for k: xy += pow(fixed_to_double_p1(word_k, precision), 2) * pow(2, precision)
 * continued from xy_inout[c] over this shard's rows for columns col_lo..col_hi-1
 * (host array of col_hi-col_lo doubles).  Only needed when p1_finalize reports that
 * the exact integer shortcut does not apply; row shards chain it in rank order. */
int p1_diag_sequential(p1_ctx *ctx, const p1_data *data, size_t col_lo, size_t col_hi, int precision,
                       double *xy_inout);

/* ------------------------------------------------------------------ a9: trusted initializer
 * Replaces the body of the TI loop (src/phase1.c:270-278) for cross pairs
 * q0 .. q0+count-1 (q = running index of the pair in the loop order of :257-267).
 * Mask stream: AES-128-CTR keystream S under `key`, zero IV, 128-bit big-endian block
 * counter; the TI's randomizeBuffer calls read S consecutively, so pair q owns the words
 *     x = S[W0 .. W0+n), y = S[W0+n .. W0+2n), r = S[W0+2n],  W0 = q*(2*n_total+1)
 * (words = keystream bytes in host little-endian order).  Only rows
 * [row0, row0+n_rows) are produced (row sharding; CTR is random access).
 *   x_dev, y_dev : count vectors of n_rows words, ld_words apart; either may be NULL
 *                  (a provider expanding only its own mask from a seed, or the TI
 *                  computing only the values).
 *   xy_dev       : count u64, WRITTEN with sum over the shard's rows of x*y; NULL to skip
 *                  (needs both masks generated, stored or not).
 *   r_dev        : count u64, written with r (identical on every shard); NULL to skip.
 * The messages of :277-278 are then {vector = y, value = xy - r} and {vector = x, value = r}. */
int p1_ti_generate(p1_ctx *ctx, const uint8_t key[16], uint64_t q0, size_t count,
                   size_t n_total, size_t row0, size_t n_rows,
                   void *x_dev, void *y_dev, size_t ld_words,
                   uint64_t *xy_dev, uint64_t *r_dev);
/* The same with one key per word class: x = S_kx[W0 + k], y = S_ky[W0 + n + k], r = S_kr[W0 + 2n] (same positions,
 * three keystreams).  This is what seed-instead-of-vector TI messages need (host/phase1_gpu.c with
 * P1GPU_SEEDED_MASKS=1): party b shares kx with the TI and expands x itself, party a shares ky and expands y, so
 * neither learns the other's mask; a key may be NULL when its words are not requested.  With kx = ky = kr this is
 * p1_ti_generate. */
int p1_ti_generate_keys(p1_ctx *ctx, const uint8_t key_x[16], const uint8_t key_y[16], const uint8_t key_r[16],
                        uint64_t q0, size_t count, size_t n_total, size_t row0, size_t n_rows,
                        void *x_dev, void *y_dev, size_t ld_words, uint64_t *xy_dev, uint64_t *r_dev);
/* value_a[p] = xy[p] - r[p], value_b[p] = r[p] (src/phase1.c:277-278), after xy was all-reduced */
int p1_ti_values(p1_ctx *ctx, const uint64_t *xy_dev, const uint64_t *r_dev, size_t count,
                 uint64_t *value_a_dev, uint64_t *value_b_dev);

/* ------------------------------------------------------------------ a10: inner_product_ti
 * Batched party side of the masking protocol (src/phase1.c:173-224).  For pair p of
 * the batch: cols[p] = this party's row of the pair (column index into `data`; d means
 * the target vector), slots[p] = accumulator index (idx(i,j), or nA + j when i == d,
 * src/phase1.c:581-585).  cols/slots: host or device arrays.
 *   party b, step 1 (:202-205)  out[p][k]  = col[k] + T[p][k]                 (T = x)
 *   party a         (:187-195)  out[p][k]  = col[k] - T[p][k]                 (T = y)
 *                               acc[slot] += sum_k in[p][k] * T[p][k]
 *   party b, step 2 (:221)      acc[slot] += sum_k in[p][k] * col[k]
 *   both            (:196,222)  acc[slot] -= value[p]      (once, not per row shard)
 * Vectors are n_rows(data) words long, ld_* words apart. */
int p1_party_b_mask(p1_ctx *ctx, const p1_data *data, const int32_t *cols, size_t count,
                    const void *T_dev, size_t ld_T, void *out_dev, size_t ld_out);
int p1_party_a_step(p1_ctx *ctx, const p1_data *data, const int32_t *cols, const int64_t *slots, size_t count,
                    const void *T_dev, size_t ld_T, const void *in_dev, size_t ld_in,
                    void *out_dev, size_t ld_out, uint64_t *acc_dev);
int p1_party_b_share(p1_ctx *ctx, const p1_data *data, const int32_t *cols, const int64_t *slots, size_t count,
                     const void *in_dev, size_t ld_in, uint64_t *acc_dev);
int p1_acc_sub_values(p1_ctx *ctx, uint64_t *acc_dev, const int64_t *slots, const uint64_t *value_dev, size_t count);

/* ------------------------------------------------------------------ a9 + a10 fused: one box, all parties
 * The body of the TI loop (src/phase1.c:270-278) and both parties' inner_product_ti
 * (:187-195 and :202-205,221) for cross pairs q0 .. q0+count-1 whose two owners both live in
 * `data` (all columns of the instance on this GPU: a single-box deployment, or one row shard
 * of it).  One persistent kernel: the masks x, y of a pair are expanded (same stream as
 * p1_ti_generate) and consumed in registers, so they never touch HBM; the two party messages
 * are the protocol's data plane and ARE written:
 *     msg_b[p][k] = col_b[k] + x[k]        (b -> a, :203-205)
 *     msg_a[p][k] = col_a[k] - y[k]        (a -> b, :188-190)
 *     xy[p]             = sum_k x[k] * y[k]               (WRITTEN; the TI's value_a = xy - r)
 *     acc[slots_a[p]]  += sum_k msg_b[p][k] * y[k]        (:195)
 *     acc[slots_b[p]]  += sum_k msg_a[p][k] * col_b[k]    (:221)
 *     r[p]              = S[W0 + 2 n_total]               (:273; NULL to skip)
 * cols_a[p] / cols_b[p] = row i / row j of the pair (column index into `data`, d = target);
 * slots_*: absolute accumulator indices (the providers' accumulators may be slices of one
 * buffer).  The four index arrays: host or device.  x_dev / y_dev (count vectors, ld_words
 * apart) may be given to ALSO materialise the masks (tests: identical words to the split
 * kernels), normally NULL.  `- value` is applied afterwards as for the split path
 * (p1_ti_values, p1_acc_sub_values).  Rows: data's shard [row0, row0 + rows(data)) of n_total. */
int p1_masking_fused(p1_ctx *ctx, const uint8_t key[16], uint64_t q0, size_t count, size_t n_total, size_t row0,
                     const p1_data *data, const int32_t *cols_a, const int32_t *cols_b,
                     const int64_t *slots_a, const int64_t *slots_b,
                     void *msg_b_dev, void *msg_a_dev, size_t ld_words, void *x_dev, void *y_dev,
                     uint64_t *xy_dev, uint64_t *r_dev, uint64_t *acc_dev);

/* ------------------------------------------------------------------ a12/a13: --use_ot
 * Correlated-OT extension behind inner_product_ot_sender / inner_product_ot_recver
 * (src/phase1.c:53-96).  The external Obliv-C primitive is replaced by the semi-honest
 * IKNP-style construction defined in oracle/cot_ext.h (kappa = 128, AES-128-CTR column
 * PRG, 128x128 bit transposes, fixed-key AES row hash); the base-OT outputs enter as
 * seeds.  One inner product of n_rows words = n_rows * W OTs, OT index ni = k*W + bit
 * (:45-46).  Wire messages stay on the device:
 *   receiver -> sender : U, 2048 bytes per 128 OTs
 *   sender -> receiver : C, one ring word per OT
 * sender   (:43-51,66-68): opt0 = H(q), opt1 = (1<<bit)*x[k] + opt0, C = opt1 ^ H(q^Delta),
 *                          acc[slot] += -sum opt0
 * receiver (:78-83,90-92): choice = bit `bit` of a[k], dest = H(t) ^ (choice ? C : 0),
 *                          acc[slot] += +sum dest */
int p1_cot_sender_new(p1_ctx *ctx, const uint8_t delta[16], const uint8_t seeds[128][16], p1_cot **s);
int p1_cot_receiver_new(p1_ctx *ctx, const uint8_t seeds0[128][16], const uint8_t seeds1[128][16], p1_cot **r);
void p1_cot_free(p1_cot *session);
/* wire sizes for an inner product over n_rows rows of this ring (row shard: row0 and
 * n_rows must be multiples of 128/W words unless the shard is the last one) */
size_t p1_cot_u_bytes(const p1_ctx *ctx, size_t n_rows);
size_t p1_cot_c_bytes(const p1_ctx *ctx, size_t n_rows);
/* receiver, first half: expands both column PRGs for the OTs of rows [row0, row0+n_rows)
 * of an n_total-row inner product, writes U, keeps the hashed pads for _finish, and
 * advances the session cursor by the whole inner product (n_total rows). */
int p1_cot_receiver_extend(p1_ctx *ctx, p1_cot *r, const p1_data *data, int32_t col,
                           size_t n_total, size_t row0, void *u_dev);
/* sender: consumes U, writes C, adds -sum(opt0) into acc_dev[slot], advances the cursor */
int p1_cot_sender_step(p1_ctx *ctx, p1_cot *s, const p1_data *data, int32_t col,
                       size_t n_total, size_t row0, const void *u_dev,
                       void *c_dev, uint64_t *acc_dev, int64_t slot);
/* Moves the session cursor past `count` inner products of n_total rows without computing them:
 * two sessions built from the same base-OT outputs can then take alternate inner products of
 * one peer session on two streams (each receiver keeps its own pads), every one at the cursor
 * the single-session order gives it. */
int p1_cot_advance(p1_cot *session, size_t n_total, size_t count);
/* Sets the cursor outright (PRG block, OT counter): resuming a session, and the tests that start
 * beyond 2^32 blocks. */
int p1_cot_set_cursor(p1_cot *session, uint64_t prg_block, uint64_t ot_counter);
/* receiver, second half: adds +sum(dest) into acc_dev[slot] */
int p1_cot_receiver_finish(p1_ctx *ctx, p1_cot *r, const void *c_dev, uint64_t *acc_dev, int64_t slot);

/* Dense int8 tensor-core rate of this GPU, measured (the roofline denominator of the tcgen05 Gram;
 * MEASURED_PEAKS.json has no int8 entry).  One CTA per SM issues back-to-back
 * tcgen05.mma cta_group::1 kind::i8 instructions of shape M = 128, N = n_cols, K = 32 on operands that
 * stay in shared memory (no TMA, no epilogue) for at least min_ms milliseconds; *tops = executed int8
 * ops (2 per MAC) / time / 1e12.  n_cols: multiple of 16 in [16, 256] (192 = the Gram kernel's
 * instruction for 3-digit words). */
int p1_probe_tensor_i8(p1_ctx *ctx, int n_cols, double min_ms, double *tops, double *ms, uint64_t *ops);

/* ------------------------------------------------------------------ a14: finalize
 * Tail of run_party (src/phase1.c:605-641).  `acc` (host or device, p1_acc_len(d)
 * words, already summed over row shards) becomes the caller-owned calloc'ed arrays
 * *res_A (d(d+1)/2 ufixed_t) and *res_b (d ufixed_t), which the caller free()s
 * (src/cmd/linreg.c:203-204).  ufixed_t is uint64_t or uint32_t per ring_bits_p2.
 *   - off-diagonal and share_b words: the ring word, and for (p1,p2) = (64,32) the
 *     arithmetic shift (ufixed_t)(((fixed_p1_t) share) >> (precision - precision_p2)).
 *   - diagonal words of owned columns [col_lo, col_hi) (src/phase1.c:562-567):
 *     share = double_to_fixed_p1(xy / normalizer2, precision), where xy is the
 *     sequential double sum.  When no row of the column has |word| >= 2^26 and the
 *     exact integer S = sum word^2 (from the accumulator) is <= 2^53, every partial
 *     sum of the reference's loop is exactly representable, so xy = S * 2^-precision
 *     exactly, in any summation order.  Otherwise the column's xy must be supplied in
 *     diag_xy[c - col_lo] (from p1_diag_sequential); if diag_xy is NULL or the entry is
 *     NaN the call fails and lists the column in need_sequential (if not NULL, an
 *     array of col_hi - col_lo ints set to 0/1). */
int p1_finalize(p1_ctx *ctx, const uint64_t *acc, size_t d, size_t col_lo, size_t col_hi,
                int precision, int precision_p2, double normalizer2,
                const double *diag_xy, int *need_sequential,
                void **res_A, void **res_b);

/* The same tail for `num` providers at once, stream-ordered and without a host round trip (the short steps of
 * cfg4 / cfg5, where a D2H plus a host loop per provider costs more than the kernels).  acc_dev: num accumulators
 * of p1_acc_len(d) words, already combined over row shards; provider k owns the columns
 * [col_start_dev[k], col_start_dev[k + 1]) (device array of num + 1 int32).  out_dev: num x (d(d+1)/2 + d) uint64
 * words, res_A | res_b of provider k as ufixed_t widened to 64 bits (zero extended).  diag_xy_dev: d doubles
 * indexed by column, or NULL; *need_dev counts the diagonal words whose exact shortcut does not apply and for
 * which no xy was supplied (they are left 0 - the caller then falls back to p1_diag_sequential + the host call). */
int p1_finalize_async(p1_ctx *ctx, const uint64_t *acc_dev, size_t num, size_t d, const int32_t *col_start_dev,
                      int precision, int precision_p2, double normalizer2, const double *diag_xy_dev,
                      int32_t *need_dev, uint64_t *out_dev);

/* ------------------------------------------------------------------ f2: messages on the wire
 * SecureMultiplication__Msg (src/protobuf/secure_multiplication{,_32}.proto):
 *     message msg { repeated uintW vector = 1 [packed=true]; required uintW value = 2; }
 * as protobuf-c's pack_to_buffer writes it in send_pmsg (src/phase1.c:131-139) and
 * secure_multiplication__msg__unpack reads it in recv_pmsg (:111), for a batch of `count`
 * messages whose vectors live on the GPU (the x / y masks of the TI, the party messages).
 * proto2 wire bytes of one message:
 *     [0x0A][varint payload_bytes][varint(vector[0]) .. varint(vector[n-1])]   (absent when n == 0)
 *     [0x10][varint(value)]
 * Message p occupies out_dev + p*stride_bytes; its length (the size_t that send_pmsg ships
 * first, :133) is written to sizes_dev[p].  stride_bytes must be a multiple of 16 and at
 * least p1_msg_max_bytes(ctx, n_words); buffers must be 16-byte aligned; vectors are ld_words
 * apart (ld_words * word size a multiple of 16).  values: count uint64_t, host or device
 * (the low W bits are used). */
size_t p1_msg_max_bytes(const p1_ctx *ctx, size_t n_words);
int p1_msg_pack(p1_ctx *ctx, const void *vec_dev, size_t ld_words, size_t n_words, size_t count,
                const uint64_t *values, uint8_t *out_dev, size_t stride_bytes, uint64_t *sizes_dev);
/* p1_msg_pack for callers that must not allocate.  (No counterpart in the reference: it is the same
 * pack_to_buffer replacement as above; only where the kernels keep their working tables differs.)
 * p1_msg_pack takes 12 bytes of device scratch per 2048-word tile from the library's block pool, which
 * falls through to cudaMalloc the first time a size is asked for.  Hosts that cannot have that happen when
 * a message is due - a provider whose own arena already holds all of the GPU's memory, the per-peer threads
 * of --use_ot mode (an allocation waits for the whole device and so stalls every other peer's session on
 * that GPU), a party under an operator's memory cap that might refuse the request in mid-protocol - call
 * this instead.  It makes no cudaMalloc / cudaFree / pool request and stages nothing: the scratch is a fixed
 * array inside libp1gpu.so's device image (P1_MSG_PACK_STATIC_MAX_TILES * 12 bytes = 768 KB, resident on
 * every GPU that loads the library until the process exits).
 *   - arguments, output bytes and sizes exactly as p1_msg_pack, except that `values_dev` must be a DEVICE
 *     array (a host array would have to be staged);
 *   - limit: count * max(1, ceil(n_words / 2048)) <= P1_MSG_PACK_STATIC_MAX_TILES, i.e. up to 134 M ring
 *     words per call; a larger call fails with a message that says so - split the batch, or use p1_msg_pack;
 *   - the first call on each GPU looks up the array's address on that device (cudaGetSymbolAddress, with the
 *     context's device current) and keeps it; make that first call before starting any stream capture;
 *   - every call on a GPU, from any context, works in the same array.  Nothing in it needs clearing between
 *     calls (each entry is written before it is read within a call), but two calls whose kernels may run at the
 *     same time on one GPU - different contexts, hence different streams - overwrite each other's tables and
 *     produce garbage: order them (p1_ctx_wait) or keep to one context per GPU for this entry point;
 *   - not covered by p1_ctx_profile (the event brackets allocate). */
#define P1_MSG_PACK_STATIC_MAX_TILES 65536
int p1_msg_pack_static(p1_ctx *ctx, const void *vec_dev, size_t ld_words, size_t n_words, size_t count,
                       const uint64_t *values_dev, uint8_t *out_dev, size_t stride_bytes, uint64_t *sizes_dev);
/* Inverse.  sizes: count uint64_t message lengths, host or device (each <= stride_bytes; the
 * whole stride must be readable).  Every message must carry exactly n_words vector elements
 * in the canonical form above (what protobuf-c emits); status_dev[p] = 0 on success, else a
 * bit set of P1_MSG_* flags, and the caller falls back to its own parser for that message.
 * Varints longer than W bits are truncated as protobuf parsers do. */
#define P1_MSG_BAD_HEADER 1   /* first tag / payload length / trailer malformed or past the end */
#define P1_MSG_BAD_COUNT 2    /* number of vector elements != n_words */
#define P1_MSG_BAD_VARINT 4   /* an element longer than 10 bytes or cut off by the payload end */
int p1_msg_unpack(p1_ctx *ctx, const uint8_t *in_dev, size_t stride_bytes, const uint64_t *sizes, size_t count,
                  size_t n_words, void *vec_dev, size_t ld_words, uint64_t *values_dev, int32_t *status_dev);

/* ------------------------------------------------------------------ f3: hand-off to phase 2
 * The choice bits dcsSendIntArray builds from the shares (src/input.c:37-41):
 *     sel[i*intsize + j] = (input[i] >> j) & 1,   intsize = FIXED_BIT_SIZE_P2
 * one byte (C bool) per bit.  shares: n ufixed_t words (uint64_t or uint32_t per
 * ring_bits_p2), host or device; sel: n * ring_bits_p2 bytes, host or device. */
int p1_share_bits(p1_ctx *ctx, const void *shares, size_t n, uint8_t *sel);

/* ------------------------------------------------------------------ f4: text ingestion
 * read_matrix / read_vector (src/linear.c:27-102) fscanf one "%lf" token at a time.  Here the
 * whitespace-separated tokens of a text buffer are located with a prefix count and converted on
 * the GPU (Eisel-Lemire decimal -> binary64: correctly rounded, i.e. the double glibc's strtod
 * returns in the default rounding mode).
 *   text       : `bytes` bytes, host or device (a device buffer must be 16-byte aligned)
 *   out_dev    : max_tokens doubles on the device; token k -> out_dev[k]; tokens beyond
 *                max_tokens are counted but not stored
 *   n_tokens   : (host, out) number of tokens in the buffer
 *   n_deferred : (host, out) number of tokens the kernel did not convert; they are stored as NaN
 *                and the byte offsets of up to `cap` of them are written to `deferred`
 *                (host, ascending; call again with a larger list if n_deferred > cap).  Deferred: anything that is not
 *                [+-]digits[.digits][(e|E)[+-]digits] over the whole token (inf, nan, hex floats,
 *                "localhost:1234"), tokens longer than 256 bytes, and mantissas of more than 19
 *                digits whose truncation is ambiguous.  The caller converts those with its own
 *                strtod and patches out_dev.
 * The result feeds p1_encode directly.  For the reference's instance files, positioned at the X
 * header (src/config.c:20): token 0, 1 = "n m", then n*m values of X (row-major), then "l", then
 * l values of y.  Synchronises the stream. */
int p1_parse_doubles(p1_ctx *ctx, const char *text, size_t bytes, double *out_dev, size_t max_tokens,
                     uint64_t *n_tokens, uint64_t *n_deferred, uint64_t *deferred, size_t cap);

/* ------------------------------------------------------------------ 8b/8e: row sharding behind the boundary
 * A p1_group owns one context per GPU of the box inside ONE process, so that the reference's callers
 * (run_party / run_trusted_initializer, src/phase1.h:14-29, one process per party) use every GPU.
 * Rows of the provider's X, y are split into equal 256-row aligned slabs (device g holds rows
 * [row0_g, row0_g + rows_g), p1_group_rows); every call below fans out over the devices on their rows
 * and is asynchronous like its single-device counterpart, ordered with plain calls on
 * p1_group_ctx(grp, 0) (device 0's stream is the group's main stream).  Partial shares combine exactly
 * by the wrapping u64 sum p1_group_reduce performs (peers' accumulators read over NVLink by one kernel
 * on device 0; option "reduce" = 1: ncclAllReduce(ncclUint64, ncclSum) from the system libnccl.so.2).
 * The devices must have peer access to each other.  Listing one device twice is allowed (tests).
 * A group and everything made from it belong to one host thread at a time; the per-peer threads of OT
 * mode (src/phase1.c:503-513) each create their own group and may share one p1_gdata, which they only read. */
typedef struct p1_group p1_group;
typedef struct p1_gdata p1_gdata;     /* encoded X, y as one p1_data per device */
typedef struct p1_gvec p1_gvec;       /* a batch of n-row vectors as row slices on the devices */
typedef struct p1_gcot p1_gcot;       /* one correlated-OT session, one p1_cot per device */

int p1_group_create(const int *devices, int ndev, int ring_bits_p1, int ring_bits_p2, p1_group **grp);
void p1_group_destroy(p1_group *grp);
int p1_group_size(const p1_group *grp);
p1_ctx *p1_group_ctx(p1_group *grp, int g);
void p1_group_rows(const p1_group *grp, size_t n_total, int g, size_t *row0, size_t *n_rows);
/* "reduce" (see above); every other name is forwarded to p1_ctx_set_option of each context */
int p1_group_set_option(p1_group *grp, const char *name, long value);
int p1_group_sync(p1_group *grp);
uint64_t p1_group_launches(p1_group *grp);

/* p1_encode over the group: X (n x d row-major doubles) and y on the host - every device pulls its slab
 * over its own PCIe link - or on a device of the group (the other devices read it over NVLink). */
int p1_group_encode(p1_group *grp, const double *X, const double *y, size_t n, size_t d,
                    double normalizer, int precision, p1_gdata **out);
void p1_gdata_free(p1_gdata *data);
const p1_data *p1_gdata_shard(const p1_gdata *data, int g);

/* the group's accumulators: p1_acc_len(d) words per device, zeroed by p1_group_acc_reset; every call
 * below adds into them; p1_group_reduce leaves their sum in device 0's (p1_group_acc_dev(grp, 0)) and ZEROES
 * the other devices' (in both reduce modes), so the sequence reduce -> more pairs -> reduce is allowed and
 * equals one reduce at the end: no partial sum is ever counted twice.  Between a reduce and the next
 * accumulating call, p1_group_acc_dev(grp, g >= 1) therefore reads as zeros. */
int p1_group_acc_reset(p1_group *grp, size_t d);
uint64_t *p1_group_acc_dev(p1_group *grp, int g);
int p1_group_reduce(p1_group *grp);

/* p1_local_block on every slab (src/phase1.c:562-571) */
int p1_group_local_block(p1_group *grp, const p1_gdata *data, size_t col_lo, size_t col_hi, int owns_target, int algo);

/* batch x n_total ring words as row slices; gather / scatter move whole vectors (ld0 words apart) on
 * device 0 - where p1_msg_pack / p1_msg_unpack work - from / to the slices with peer copies */
int p1_gvec_alloc(p1_group *grp, size_t n_total, size_t batch, p1_gvec **out);
void p1_gvec_free(p1_gvec *v);
void *p1_gvec_dev(const p1_gvec *v, int g, size_t *ld_words);
int p1_gvec_gather(p1_group *grp, const p1_gvec *v, size_t count, void *full_dev0, size_t ld0);
int p1_gvec_scatter(p1_group *grp, const void *full_dev0, size_t ld0, size_t count, p1_gvec *v);

/* p1_ti_generate / p1_party_* / p1_acc_sub_values over the group; xy (sum over ALL rows), r and the
 * values live on device 0; cols / slots: host or device-0... host arrays (each context stages its copy) */
int p1_group_ti_generate(p1_group *grp, const uint8_t key[16], uint64_t q0, size_t count, size_t n_total,
                         p1_gvec *x, p1_gvec *y, uint64_t *xy_dev0, uint64_t *r_dev0);
int p1_group_ti_generate_keys(p1_group *grp, const uint8_t key_x[16], const uint8_t key_y[16], const uint8_t key_r[16],
                              uint64_t q0, size_t count, size_t n_total, p1_gvec *x, p1_gvec *y, uint64_t *xy_dev0, uint64_t *r_dev0);
int p1_group_party_b_mask(p1_group *grp, const p1_gdata *data, const int32_t *cols, size_t count, const p1_gvec *T, p1_gvec *out);
int p1_group_party_a_step(p1_group *grp, const p1_gdata *data, const int32_t *cols, const int64_t *slots, size_t count,
                          const p1_gvec *T, const p1_gvec *in, p1_gvec *out);
int p1_group_party_b_share(p1_group *grp, const p1_gdata *data, const int32_t *cols, const int64_t *slots, size_t count, const p1_gvec *in);
int p1_group_acc_sub_values(p1_group *grp, size_t d, const int64_t *slots, const uint64_t *value_dev0, size_t count);

/* --use_ot over the group: every device extends / answers the OTs of its rows and moves its slice of
 * the wire messages U (p1_cot_u_bytes(n) bytes) and C (p1_cot_c_bytes(n)) from / to the HOST buffers over
 * its own PCIe link (page-locked buffers from p1_host_alloc make the copies concurrent).  The calls
 * return when the host buffer they produce is complete. */
int p1_group_cot_sender_new(p1_group *grp, const uint8_t delta[16], const uint8_t seeds[128][16], p1_gcot **s);
int p1_group_cot_receiver_new(p1_group *grp, const uint8_t seeds0[128][16], const uint8_t seeds1[128][16], p1_gcot **r);
void p1_group_cot_free(p1_gcot *session);
int p1_group_cot_receiver_extend(p1_group *grp, p1_gcot *r, const p1_gdata *data, int32_t col, void *u_host);
int p1_group_cot_sender_step(p1_group *grp, p1_gcot *s, const p1_gdata *data, int32_t col, const void *u_host, void *c_host, int64_t slot);
int p1_group_cot_receiver_finish(p1_group *grp, p1_gcot *r, const p1_gdata *data, const void *c_host, int64_t slot);

/* p1_finalize for the group (src/phase1.c:605-641): p1_group_reduce, the diagonal of the owned columns
 * [col_lo, col_hi) (exact shortcut, or the sequential double loop chained over the slabs in device
 * order), the 64 -> 32 shift, calloc'ed outputs the caller free()s.  acc_host: NULL = the group's own
 * accumulators; else p1_acc_len(d) words on the host that are already the sum over devices and threads (the
 * per-peer threads of OT mode keep private results that the caller adds up, src/phase1.c:514-523). */
int p1_group_finalize(p1_group *grp, const p1_gdata *data, const uint64_t *acc_host, size_t col_lo, size_t col_hi,
                      int precision, int precision_p2, double normalizer2, void **res_A, void **res_b);

#ifdef __cplusplus
}
#endif
#endif /* P1GPU_H */
