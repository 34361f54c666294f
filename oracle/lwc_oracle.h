/*
 * lwc_oracle -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the scoring hot path of
 * ObjectiveAI/llm-weighted-consensus, used as the parity checker (tests/, smoke()) and as
 * the timed CPU baseline (bench.py cpu_baseline / --impl reference).  The product library
 * (liblwc_b200.so) never links, loads or calls anything in this directory.
 *
 * PARITY UNPINNED: the reference is Rust-only (no cargo/rustc here), ships no tests, golden
 * vectors or fixtures for this path, and its arithmetic lives in rust_decimal 1.39.0, whose
 * source is absent from /root/reference.  The Decimal tally below restates
 * src/score/completions/client.rs:384-463 line by line; the normalise / cosine /
 * similarity-weighted-vote / argmax stages do not exist in the reference at all
 *  and follow the build-defined spec of SURVEY.md 8.0.
 */
#ifndef LWC_ORACLE_H
#define LWC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { LWC_ORACLE_SIM_IDENTITY = 0, LWC_ORACLE_SIM_COSINE = 1 };

/* One decimal operator on strings: op 0 '+', 1 '-', 2 '*', 3 '/', 4 compare (out = "-1|0|1").
 * out receives the value-canonical string (trailing zeros stripped). Returns a DEC96_* code. */
int lwc_oracle_dec_op(const char *a, const char *b, int op, char *out, size_t outlen);

/* Decimal string -> nearest double; float -> shortest-round-trip decimal string. */
double lwc_oracle_dec_to_double(const char *a);
int lwc_oracle_dec_from_f32(float x, char *out, size_t outlen);

/*
 * The reference tally, in Decimal (client.rs:384-449).
 *   votes_csv  : voters separated by ';', each voter "v0,v1,...,v{n-1}" or "-" when that
 *                voter has no vote (choice.delta.vote == None, client.rs:410).
 *   weights_csv: one entry per voter, "-" when choice.weight == None (unwrap_or(ZERO),
 *                client.rs:412-413).
 * Outputs (comma separated, value-canonical decimals): choice weight, choice confidence,
 * per-voter confidence ("-" for a voter without a vote: its confidence stays None,
 * client.rs:438-449).
 */
int lwc_oracle_tally_decimal(int n_choices, int n_voters, const char *votes_csv, const char *weights_csv,
                             char *weight_out, char *confidence_out, char *voter_conf_out, size_t outlen);

/* All-votes-failed status merge (client.rs:386-409, 459-463; error.rs:41).
 * codes[m] = 0 when voter m carries no error, else its HTTP-style code.
 * Returns 0 when at least one voter succeeded, else the merged status the reference
 * would answer with (AllVotesFailed(None) -> 500). */
int lwc_oracle_merge_error_codes(const uint16_t *codes, int n_voters);

/*
 * Full path in f64 on the fp32 inputs the device sees.
 *   E [R][N][D] fp32, V [R][M][N] fp32 (voter absent <=> V[r][m][0] is NaN),
 *   w [R][M] or [M] fp32 (NaN <=> weight None -> 0).
 * Outputs: conf,cw [R][N]; agree [R][M] (NaN for absent voters); winner [R] (-1 when the
 * weight sum is not > 0); rank [R][N] (confidence descending, stable). Any output may be NULL.
 * nthreads <= 0 -> all cores (OpenMP), requests are the parallel unit.
 */
int lwc_oracle_score_f64(int R, int N, int D, int M, int similarity, const float *E, const float *V,
                         const float *w, int w_per_request, double *conf, double *cw, double *agree,
                         int32_t *winner, int32_t *rank, int nthreads);

/*
 * The CPU baseline "port": what the reference does on its own types, per request --
 * votes and weights as Decimal (tally client.rs:385-416 in dec96), embeddings as f64
 * (embeddings/response.rs:21), then the build-defined stages in f64 and the Decimal-style
 * normalise of client.rs:422-449 carried out in f64.  Same I/O contract as score_f64.
 */
int lwc_oracle_score_port(int R, int N, int D, int M, int similarity, const float *E, const float *V,
                          const float *w, int w_per_request, double *conf, double *cw, double *agree,
                          int32_t *winner, int32_t *rank, int nthreads);

/*
 * One big pool (cfg5 shape): votes are dense N-vectors.  method 0 = explicit pairwise
 * matrix O(N^2 D); method 1 = centroid identity cw_j = <E^_j, sum_i t_i E^_i>, O(N D)
 * (SURVEY.md 8.0 note) -- the only feasible oracle at N = 262,144.
 */
int lwc_oracle_pool_f64(int64_t N, int D, int M, int method, const float *E, const float *V, const float *w,
                        double *conf, double *cw, double *agree, int64_t *winner, int nthreads);

/* Numeric tail of get_vote (client.rs:1765-1791): p_k = exp(logprob_k) accumulated on
 * leaf[k], divided by the sum.  leaf[k] < 0 means "alternative is not a sibling leaf". */
int lwc_oracle_vote_from_logprobs(int n_choices, int n_top, const double *logprob, const int32_t *leaf,
                                  double *vote);

/* f64 restatement of the BUILD-DEFINED training-table weight policy (include/lwc_b200.h,
 * lwc_table_weights; the reference slot weight.rs:99-117 is unimplemented, so this checks the CUDA
 * path against its own specification, not against the reference).  Returns 1 on invalid limits. */
int lwc_oracle_table_weights(int64_t rows, int D, int M, const float *T, const float *Q, int n_prompts,
                             const float *q, int top, const float *base, const float *minw, const float *maxw,
                             double *weights, int32_t *top_idx, double *top_sim);

/* Seconds the last lwc_oracle_score_port call spent in its compute phase (the fp32 -> Decimal
 * conversion of the synthetic inputs excluded: in the reference those arrive as Decimals). */
double lwc_oracle_port_last_seconds(void);

int lwc_oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
