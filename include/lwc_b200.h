/*
 * lwc_b200.h -- C ABI of the B200-native scoring hot path for llm-weighted-consensus.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  A Rust caller
 * binds these symbols through `extern "C"` (rust/lwc-b200-sys) exactly where the reference
 * crate computes the same quantities on the CPU; INTEGRATION.md shows the binding.
 *
 * What each entry point replaces in the reference (paths relative to /root/reference):
 *
 *   lwc_score_batch ........ the inline tally / normalise / voter-agreement block of
 *                            score::completions::Client::create_streaming,
 *                            src/score/completions/client.rs:384-456, batched over R
 *                            requests, plus the embeddings-similarity stages that the
 *                            training-table weight slot carries the data for
 *                            (src/score/completions/weight.rs:15-18,66-74;
 *                            src/embeddings/response.rs:4-24).  With
 *                            similarity = LWC_SIM_IDENTITY it is the reference tally and
 *                            nothing else.
 *   lwc_merge_error_codes .. the all-votes-failed status merge, client.rs:386-409,459-463
 *                            and src/score/completions/error.rs:41.
 *   lwc_stage_f64,
 *   lwc_stage_f64_bf16,
 *   lwc_score_batch_typed .. the Vec<f64> -> device fp32 hop for
 *                            embeddings::response::Embedding.embedding
 *                            (src/embeddings/response.rs:19-24): on the host (fp32 or bf16
 *                            transport) or on the device (f64 transport).
 *   lwc_pool_consensus ..... client.rs:384-456 for ONE request whose pool is too large for a
 *                            batch slot (BASELINE.json config 5: 262,144 candidates), rows
 *                            sharded over the GPUs of a node; no other reference counterpart.
 *   lwc_vote_from_logprobs . the numeric tail of get_vote, client.rs:1765-1791: exp(logprob) of
 *                            the sibling-leaf alternatives, accumulated per choice, divided by
 *                            the probability sum.
 *   lwc_table_load /
 *   lwc_table_weights ...... what weight::Fetcher<CTX, TrainingTableData> would compute -- the slot
 *                            the reference ships as unimplemented!(), src/score/completions/weight.rs:99-117
 *                            (config src/score/model/mod.rs:278-306, limits src/score/llm/mod.rs:663-688);
 *                            build-defined policy, no reference arithmetic to match.
 *   lwc_score_batches, lwc_pool_consensus_static, lwc_comm_*, lwc_pool_p2p_*, lwc_host_* / lwc_device_* /
 *   lwc_memcpy, lwc_ctx_* .. no reference counterpart (batch lists, allocation-free variant, multi-GPU
 *                            plumbing, lifecycle).
 *
 * Error convention mirrors error::ResponseError { code: u16, .. } (src/error.rs:8-13):
 * every function returns 0 on success or an HTTP-style status -- 400 for a malformed
 * argument (shape, null pointer, misalignment), 500 for a CUDA/NCCL failure, 503 when no
 * usable device exists.  lwc_last_error_message() gives the text for the `message` field.
 *
 * Threading: a context may be used from many host threads; calls on one context are
 * serialised internally.  Create one context per device.
 */
#ifndef LWC_B200_H
#define LWC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LWC_OK 0
#define LWC_ERR_INVALID 400
#define LWC_ERR_INTERNAL 500
#define LWC_ERR_UNAVAILABLE 503

#define LWC_MAX_CHOICES 128 /* per-request candidates handled by lwc_score_batch */
#define LWC_MAX_VOTERS 128  /* score::model: 1 <= LLMs <= 128, src/score/model/mod.rs:23-35 */

typedef struct lwc_ctx lwc_ctx;

/* lwc_config.flags */
#define LWC_CTX_INPUTS_RESIDENT 1u /* LWC_MEM_DEVICE calls only: the caller guarantees that E, V and w were
                                      completely written before the call is issued and are not being written by
                                      earlier work of the same stream (a batch kept resident in HBM, or uploaded
                                      and synchronised beforehand).  Consecutive lwc_score_batch launches on one
                                      stream may then overlap (programmatic dependent launch): the next launch
                                      streams its inputs while the previous one drains; its result stores still
                                      wait for the previous launch to finish. */

typedef struct {
    int32_t device;          /* CUDA ordinal this context drives */
    int32_t chunk_requests;  /* host-memory calls are pipelined in chunks of this many requests; 0 = auto */
    uint64_t flags;          /* LWC_CTX_* */
    uint64_t reserved[3];
} lwc_config;

enum { LWC_SIM_IDENTITY = 0, /* S = I: the reference tally, client.rs:410-415 */
       LWC_SIM_COSINE = 1 /* S = cosine similarity of the candidates' embeddings */ };

enum { LWC_PREC_FP32 = 0,   /* fp32-grade result (confidences within 1e-5 relative); the Gram runs
                               as error-compensated 3xTF32 on tensor cores or as fp32 FMA, by shape */
       LWC_PREC_TF32X3 = 1, /* force the compensated tensor-core route */
       LWC_PREC_TF32 = 2,   /* single-pass TF32 inputs, fp32 accumulate (confidences ~1e-3 relative) */
       LWC_PREC_BF16 = 3 /* bf16 inputs, fp32 accumulate (pool path; confidences ~1e-2 relative) */ };

enum { LWC_MEM_HOST = 0,  /* every pointer is host memory; the call returns with outputs written */
       LWC_MEM_DEVICE = 1 /* every pointer is device memory on ctx's device; the call is
                             asynchronous on `stream` */ };

typedef struct {
    int32_t R;             /* requests in the batch */
    int32_t N;             /* candidate responses (choices) per request, 1..LWC_MAX_CHOICES (the
                              reference itself requires >= 2, client.rs:115-120) */
    int32_t D;             /* embedding dimension (ignored for LWC_SIM_IDENTITY) */
    int32_t M;             /* voters (LLMs) per request, 0..LWC_MAX_VOTERS */
    int32_t similarity;    /* LWC_SIM_* */
    int32_t precision;     /* LWC_PREC_* */
    int32_t memory;        /* LWC_MEM_* */
    int32_t w_per_request; /* 1: w is [R][M]; 0: one [M] vector shared by all requests */
} lwc_batch_desc;

/* replaces: nothing (lifecycle).  Binds the context to cfg->device, creates its streams. */
int32_t lwc_ctx_create(const lwc_config *cfg, lwc_ctx **out);
void lwc_ctx_destroy(lwc_ctx *ctx);
/* Text of the last failure on this context (or of the last failed lwc_ctx_create when ctx is
 * NULL); maps to ResponseError.message, src/error.rs:11. Never NULL. */
const char *lwc_last_error_message(const lwc_ctx *ctx);
/* Library / device identification: "lwc_b200 <version> sm_100a ..." */
const char *lwc_version(void);

/* Pinned host memory for the batch packer (the Rust host lays request embeddings out into
 * one of these, contiguous [R][N][D]); freed with lwc_host_free. */
int32_t lwc_host_alloc(lwc_ctx *ctx, size_t bytes, void **out);
int32_t lwc_host_free(lwc_ctx *ctx, void *p);
/* Plain device memory on the context's device, for callers that keep batches resident. */
int32_t lwc_device_alloc(lwc_ctx *ctx, size_t bytes, void **out);
int32_t lwc_device_free(lwc_ctx *ctx, void *p);
/* Blocking copies for such callers (kind: 0 = host->device, 1 = device->host). */
int32_t lwc_memcpy(lwc_ctx *ctx, void *dst, const void *src, size_t bytes, int32_t kind);
/* The context's compute stream as a cudaStream_t (so a caller can record its own events on it). */
void *lwc_ctx_stream(lwc_ctx *ctx);
/* Waits for everything queued on `stream` (NULL = the context's own stream). */
int32_t lwc_stream_synchronize(lwc_ctx *ctx, void *stream);

/*
 * Score a batch.  replaces client.rs:384-456 (see the table above).
 *
 *   E      [R][N][D] fp32, row-major, 16-byte aligned        (unused for LWC_SIM_IDENTITY)
 *   V      [R][M][N] fp32 vote rows (Delta.vote, src/score/completions/response.rs:189);
 *          a voter without a vote (vote == None, client.rs:410) is a row whose FIRST element
 *          is NaN
 *   w      [R][M] or [M] fp32 voter weights (choice.weight); NaN = None -> 0
 *          (unwrap_or(ZERO), client.rs:412-413)
 *   conf   [R][N] confidence  = cw / sum(cw) if sum > 0 else 0      (client.rs:428-437)
 *   cw     [R][N] choice weight                                     (client.rs:410-415)
 *   agree  [R][M] voter confidence sum_i conf_i * V_mi; NaN for absent voters
 *                                                                  (client.rs:438-449)
 *   winner [R]    argmax_j conf_j, ties -> lowest j, -1 when sum(cw) is not > 0
 *   rank   [R][N] choice indices by confidence, descending, stable
 * Any output pointer may be NULL.  `stream` is a cudaStream_t (NULL = the context's stream)
 * and is honoured for LWC_MEM_DEVICE; LWC_MEM_HOST calls run on the context's own streams,
 * chunked so that host->device copies overlap the kernels, and return when outputs are home.
 */
int32_t lwc_score_batch(lwc_ctx *ctx, const lwc_batch_desc *desc, const float *E, const float *V,
                        const float *w, float *conf, float *cw, float *agree, int32_t *winner,
                        int32_t *rank, void *stream);

/* replaces: the element-type hop of Embedding.embedding (Vec<f64>, src/embeddings/response.rs:21; JSON floats on
 * the wire) on the way to the kernels.  lwc_score_batch for embeddings that are not fp32 when they reach the
 * library: E is [R][N][D] of `e_type`, 16-byte aligned, host or device memory per desc->memory; everything else as
 * in lwc_score_batch.
 *   LWC_ELEM_F64   the reference's own element type.  8 bytes per element cross PCIe and a streaming kernel rounds
 *                  them to fp32 (nearest even) on the device: the results are the bits of converting on the host
 *                  (lwc_stage_f64) and calling lwc_score_batch.  For hosts whose f64 rows already sit in pinned
 *                  memory; a host that has to touch every row anyway should convert while it copies (fp32 halves
 *                  the PCIe time, which is what bounds a host-memory call: DESIGN.md 3.9).
 *   LWC_ELEM_BF16  OPT-IN reduced transport: 2 bytes per element cross PCIe (lwc_stage_f64_bf16 converts on the
 *                  host), expanded exactly on the device and scored with fp32 accumulate -- the tensor core
 *                  reads a bf16 value exactly, so one pass suffices whatever desc->precision says.  The inputs
 *                  carry 8 mantissa bits: stated tolerance 1e-2 of a request's top confidence, winner exact
 *                  where the oracle's margin exceeds that.
 * Device-memory calls convert into one buffer owned by the context: issue them on one stream. */
enum { LWC_ELEM_F32 = 0, LWC_ELEM_F64 = 1, LWC_ELEM_BF16 = 2 };
int32_t lwc_score_batch_typed(lwc_ctx *ctx, const lwc_batch_desc *desc, int32_t e_type, const void *E, const float *V,
                              const float *w, float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank,
                              void *stream);
/* Host helper for LWC_ELEM_BF16: n doubles -> n bf16 (round to nearest even).  Like lwc_stage_f64 it is a plain
 * loop meant to be called by many threads on their own rows. */
int32_t lwc_stage_f64_bf16(lwc_ctx *ctx, const double *src, size_t n, uint16_t *dst);

/* No counterpart in the reference (which scores one request per task): lwc_score_batch for a LIST of
 * device-resident batches, enqueued back to back on `stream` inside one call.  descs [n_batches]; every
 * other argument is an array of n_batches pointers with lwc_score_batch's meaning (an output array, or an
 * entry of one, may be NULL).  Every batch must be LWC_MEM_DEVICE.  The list is validated before the first
 * launch: a 400 means nothing was enqueued.  Results are bit-identical to n_batches single calls.  What
 * changes: (1) one lock, one argument check and one library call per list instead of per batch; (2) a RUN
 * of consecutive entries with the same descriptor that take the small-pool ring kernel (N <= 32, cosine,
 * more than 56 KB of embeddings per request -- smaller requests run one CTA each and launch batch by batch)
 * is scored by ONE launch per 32 batches -- its persistent CTAs walk the whole run, so launch latency, the
 * first bytes' latency and the last requests' tail are paid once per run instead of once per batch, and
 * nothing depends on the host staying a launch ahead of a ~60 us kernel (cfg2: 63.1 us per batch = 0.99 of
 * the measured HBM peak at 20 batches per call, against 66.5 us = 0.94 launched one by one).  Entries of
 * one list must not share output buffers (inside a fused launch their order is undefined); the
 * environment variable LWC_LIST_FUSION=0 restores one launch per batch.  Asynchronous on `stream` like a
 * device-memory lwc_score_batch. */
int32_t lwc_score_batches(lwc_ctx *ctx, int32_t n_batches, const lwc_batch_desc *descs, const float *const *E,
                          const float *const *V, const float *const *w, float *const *conf, float *const *cw,
                          float *const *agree, int32_t *const *winner, int32_t *const *rank, void *stream);

/* Name of the kernel lwc_score_batch would launch for this descriptor ("small16_tf32x3", ...);
 * empty string for an invalid descriptor.  For logs, tests and the bench's roofline record. */
const char *lwc_kernel_for(const lwc_ctx *ctx, const lwc_batch_desc *desc);

/* Number of kernels this library has launched on behalf of ctx since creation. */
uint64_t lwc_launch_count(const lwc_ctx *ctx);

/* replaces: client.rs:386-409,459-463.  codes[m] = 0 for a voter without an error, else its
 * status code.  Returns 0 if any voter succeeded, else the status AllVotesFailed maps to. */
int32_t lwc_merge_error_codes(const uint16_t *codes, int32_t n_voters);

/* replaces: the f64 -> fp32 hop for Embedding.embedding (embeddings/response.rs:21).
 * Converts n doubles at src (host) to floats at dst (host, typically from lwc_host_alloc). */
int32_t lwc_stage_f64(lwc_ctx *ctx, const double *src, size_t n, float *dst);

/* replaces: client.rs:1765-1791 (the producer of Delta.vote, response.rs:189), batched.
 *   logprob [n_rows][n_top] fp32  TopLogprob.logprob of each alternative (NaN = logprob None)
 *   leaf    [n_rows][n_top] i32   choice index the alternative's character selects in the lowest
 *                                 prefix-tree branch, or -1 when it is not a sibling leaf
 *   V       [n_rows][n_choices]   vote rows out; a row whose probability sum is 0 (unreachable!() in
 *                                 the reference) is written as NaN = "voter without a vote"
 * `memory` as in lwc_batch_desc; device calls are asynchronous on `stream`. */
int32_t lwc_vote_from_logprobs(lwc_ctx *ctx, int32_t n_rows, int32_t n_top, int32_t n_choices, const float *logprob,
                               const int32_t *leaf, float *V, int32_t memory, void *stream);

/* ---- one giant pool (all-pairs consensus), optionally row-sharded over the ranks of a node ---- */

typedef struct {
    int64_t N;         /* candidate rows THIS rank holds (every rank must hold the same number) */
    int32_t D;         /* embedding dimension, D % 8 == 0 */
    int32_t M;         /* voters, 0..LWC_MAX_VOTERS */
    int32_t precision; /* LWC_PREC_BF16: normalised rows are rounded to bf16 for the tcgen05 pass
                          (confidences within 1e-2 of the top confidence); other values -> 400 */
    int32_t memory;    /* LWC_MEM_* for every pointer below */
} lwc_pool_desc;

/* replaces: client.rs:384-456 applied to one request of N * nranks candidates.
 *   E [N][D] own rows; V [M][N] the vote entries of own rows (a voter without a vote: NaN row on every
 *   rank); w [M].  Outputs: conf, cw [N] for own rows; agree [M] and *winner (global row index, -1 if
 *   the weight sum is not > 0) identical on every rank.  rank order is not produced for a pool.
 * With a communicator attached (lwc_comm_init) the normalised rows and the tally vector are
 * all-gathered over NVLink (NCCL) -- the path's only exchange step -- and three scalars per rank are
 * reduced; without one the call is single-GPU.  Blocking: returns with outputs written. */
int32_t lwc_pool_consensus(lwc_ctx *ctx, const lwc_pool_desc *desc, const float *E, const float *V, const float *w,
                           float *conf, float *cw, float *agree, int64_t *winner, void *stream);

/* No counterpart in the reference: this entry point exists for hosts that must not allocate.
 * A serving process that has given all of a GPU's memory to its own pool, or whose scoring thread may
 * not take the driver's allocation lock (cudaMalloc/cudaFree synchronise the whole device and would
 * stall every other stream, e.g. the batcher threads of the other requests), can still score one pool
 * with this call.  It is lwc_pool_consensus for a single GPU with every pointer in DEVICE memory
 * (desc->memory must be LWC_MEM_DEVICE, no communicator attached, `winner` is a device int64), and it
 * takes its working memory from a fixed array that lives inside the library's device image instead of
 * from cudaMalloc: LWC_STATIC_SCRATCH_BYTES per GPU, held from the moment the library's pool module is
 * loaded on that GPU until the process exits.  CUDA loads modules lazily by default, so that moment is
 * the first pool call on the GPU -- at the latest the first call of this function.  A host that is about
 * to hand ALL of a GPU's memory to its own pool must therefore make one warm-up call of this function
 * on that GPU first (or run with CUDA_MODULE_LOADING=EAGER); afterwards no call needs memory.
 *   Limits: N <= LWC_STATIC_POOL_MAX_ROWS and N * D <= LWC_STATIC_POOL_MAX_ELEMS; anything larger is
 *   refused with 400 and a message naming the limit (use lwc_pool_consensus then).
 *   The call makes no cudaMalloc/cudaFree or other allocating call.  It is asynchronous on `stream`
 *   (NULL = the context's stream): outputs are valid once the stream has reached this point.
 *   The first call on each device looks up the array's device address (once, with that device
 *   current); make that first call before starting any stream capture.
 *   All calls on a device share the one array, which still holds the previous call's intermediate
 *   data: the few words that must start at zero are cleared by each call, everything else is
 *   overwritten before it is read.  Two calls whose GPU work overlaps (different streams, or
 *   different contexts on the same device) race on the array and give wrong results: run them one
 *   after the other. */
#define LWC_STATIC_SCRATCH_BYTES (13u * 1024u * 1024u)
#define LWC_STATIC_POOL_MAX_ROWS 16384
#define LWC_STATIC_POOL_MAX_ELEMS 4194304
int32_t lwc_pool_consensus_static(lwc_ctx *ctx, const lwc_pool_desc *desc, const float *E, const float *V,
                                  const float *w, float *conf, float *cw, float *agree, int64_t *winner,
                                  void *stream);

/* NCCL plumbing for the sharded pool: rank 0 creates a 128-byte id, every rank calls lwc_comm_init
 * with it (collective).  NCCL is loaded with dlopen("libnccl.so.2") on first use; 503 if absent. */
int32_t lwc_comm_unique_id(void *id128);
int32_t lwc_comm_init(lwc_ctx *ctx, int32_t nranks, int32_t rank, const void *id128);
int32_t lwc_comm_destroy(lwc_ctx *ctx);

/* Optional: fuse the pool path's all-gather INTO the tensor pass.  After lwc_comm_init, every rank
 * calls lwc_pool_p2p_export for the shard shape it will score (it sizes the context's buffers and
 * returns two 64-byte CUDA IPC handles), the caller all-gathers the 128-byte records in rank order
 * (any transport), and every rank calls lwc_pool_p2p_attach with the nranks * 128 bytes.  From then on
 * lwc_pool_consensus calls with that shape (N % 256 == 0) skip the NCCL all-gather: a gather warp in
 * every CTA of the tensor kernel pulls the peers' normalised rows over NVLink (peer loads) while the
 * tensor cores already work on the own shard, and each peer shard is consumed as soon as it has landed.
 * Only a one-word NCCL all-reduce remains as the cross-rank barrier.  One process per GPU (CUDA IPC). */
int32_t lwc_pool_p2p_export(lwc_ctx *ctx, int64_t n_own, int32_t D, void *handles_out /* 128 bytes */);
int32_t lwc_pool_p2p_attach(lwc_ctx *ctx, const void *all_handles /* nranks * 128 bytes */);

/* Training-table weights.  Fills the slot the reference leaves `unimplemented!()`:
 * weight::Fetcher for WeightType::TrainingTable (src/score/completions/weight.rs:99-117), configured by
 * `WeightTrainingTable { embeddings, top }` (src/score/model/mod.rs:278-306) and the per-LLM
 * `base_weight / min_weight / max_weight` (src/score/llm/mod.rs:663-688).  The reference holds no
 * arithmetic for it, so the policy is BUILD-DEFINED and its parity is unpinned.
 *
 * lwc_table_load: copies a table of `rows` past prompts to the GPU and keeps it resident in the
 *   context: T [rows][D] prompt embeddings, Q [rows][M] each LLM's recorded quality on that prompt
 *   (>= 0; e.g. the confidence it gave the winning choice).  Replaces a previously loaded table.
 * lwc_table_weights: for each of n_prompts new prompt embeddings q [n_prompts][D], finds
 *   the `top` most cosine-similar table rows (ties -> lowest row) and writes
 *     weights[p][m] = clamp(base[m] * rho_m / mean_m'(rho_m'), minw[m], maxw[m]),
 *     rho_m = sum_k max(sim_k, 0) Q[idx_k][m] / sum_k max(sim_k, 0);   base[m] when no neighbour has sim > 0.
 *   400 unless 0 < minw <= base <= maxw for every LLM (the reference's validate()) and 1 <= top <= 64.
 *   top_idx / top_sim ([n_prompts][top], -1 / 0 past the table's size) may be NULL.
 *   memory = LWC_MEM_HOST: q, weights, top_idx, top_sim are host pointers and the call blocks.
 *   memory = LWC_MEM_DEVICE: they are device pointers and the call is asynchronous on the context's stream
 *   (lwc_ctx_stream): `weights` can be passed as `w` (w_per_request = 1) to an lwc_score_batch call on
 *   that stream without touching the host -- also under LWC_CTX_INPUTS_RESIDENT: the context remembers that a
 *   producer of weights is queued and makes the next scoring launch wait for it before reading its inputs.
 *   base / minw / maxw are host arrays either way. */
int32_t lwc_table_load(lwc_ctx *ctx, const float *T, const float *Q, int64_t rows, int32_t D, int32_t M);
int32_t lwc_table_weights(lwc_ctx *ctx, const float *q, int32_t n_prompts, int32_t top, const float *base,
                          const float *minw, const float *maxw, float *weights, int32_t *top_idx, float *top_sim,
                          int32_t memory);
/* A table too large for one GPU is split by rows over n_parts contexts (ranks): each part runs lwc_table_weights on
 * its shard and reports top_idx / top_sim; the host looks up the quality rows Q[top_idx] of its shard, all-gathers the
 * n_parts * n_prompts * top records (any transport: they are tiny) and calls this HOST-ONLY function, which keeps the
 * `top` best of the n_parts * top candidates per prompt (ties -> lowest global row = row_offset[part] + local row) and
 * applies the same policy with the same arithmetic in the same order as the device path: a split table gives the
 * bits of the unsplit one.  idx / sim: [n_parts][n_prompts][top]; q_rows: [n_parts][n_prompts][top][M] (rows of
 * entries with idx < 0 are ignored); top_row ([n_prompts][top], global rows, -1 past the end) and top_sim may be NULL. */
int32_t lwc_table_merge(int32_t n_parts, int32_t n_prompts, int32_t top, int32_t M, const int64_t *row_offset,
                        const int32_t *idx, const float *sim, const float *q_rows, const float *base, const float *minw,
                        const float *maxw, float *weights, int64_t *top_row, float *top_sim);

#ifdef __cplusplus
}
#endif
#endif /* LWC_B200_H */
