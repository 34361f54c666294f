/*
 * lwc_oracle.c -- TEST INFRASTRUCTURE ONLY (see lwc_oracle.h).  Plain C, gcc -O3 -fopenmp.
 */
#include "lwc_oracle.h"

#include <math.h>
#include <omp.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "dec96.h"

int lwc_oracle_num_threads(void) { return omp_get_max_threads(); }

/* ------------------------------------------------------------------ decimal plumbing */

int lwc_oracle_dec_op(const char *a, const char *b, int op, char *out, size_t outlen) {
    dec96 x, y, z;
    int rc = dec96_from_str(a, &x);
    if (rc) return rc;
    rc = dec96_from_str(b, &y);
    if (rc) return rc;
    switch (op) {
        case 0: rc = dec96_add(&x, &y, &z); break;
        case 1: rc = dec96_sub(&x, &y, &z); break;
        case 2: rc = dec96_mul(&x, &y, &z); break;
        case 3: rc = dec96_div(&x, &y, &z); break;
        case 4: snprintf(out, outlen, "%d", dec96_cmp(&x, &y)); return DEC96_OK;
        default: return DEC96_PARSE;
    }
    if (rc) return rc;
    dec96_to_str_normalized(&z, out, outlen);
    return DEC96_OK;
}

double lwc_oracle_dec_to_double(const char *a) {
    dec96 x;
    if (dec96_from_str(a, &x)) return NAN;
    return dec96_to_double(&x);
}

int lwc_oracle_dec_from_f32(float x, char *out, size_t outlen) {
    dec96 d;
    int rc = dec96_from_f32(x, &d);
    if (rc) return rc;
    dec96_to_str_normalized(&d, out, outlen);
    return DEC96_OK;
}

/* ------------------------------------------------------------------ reference tally */

/*
 * Literal restatement of client.rs:385-449 on already-parsed voters.
 *   vote[m*n + i], has_vote[m], weight[m], has_weight[m]
 *   -> choice_weight[n], confidence[n], voter_conf[m] (valid where has_vote[m]).
 */
static int tally_dec(int n, int M, const dec96 *vote, const uint8_t *has_vote, const dec96 *weight,
                     const uint8_t *has_weight, dec96 *choice_weight, dec96 *confidence, dec96 *voter_conf) {
    int rc;
    /* client.rs:385  let mut choice_weight = vec![Decimal::ZERO; request_choices_len]; */
    for (int i = 0; i < n; i++) choice_weight[i] = dec96_zero();
    /* client.rs:388,410-415  for each voter choice with a vote: choice_weight[i] += v * weight.unwrap_or(ZERO) */
    for (int m = 0; m < M; m++) {
        if (!has_vote[m]) continue;
        dec96 wm = has_weight[m] ? weight[m] : dec96_zero();
        for (int i = 0; i < n; i++) {
            dec96 prod, acc;
            if ((rc = dec96_mul(&vote[(size_t)m * n + i], &wm, &prod))) return rc;
            if ((rc = dec96_add(&choice_weight[i], &prod, &acc))) return rc;
            choice_weight[i] = acc;
        }
    }
    /* client.rs:422-424  choice_weight_sum = choice_weight.iter().sum() (fold from ZERO, in order) */
    dec96 sum = dec96_zero();
    for (int i = 0; i < n; i++) {
        dec96 acc;
        if ((rc = dec96_add(&sum, &choice_weight[i], &acc))) return rc;
        sum = acc;
    }
    dec96 zero = dec96_zero();
    int positive = dec96_cmp(&sum, &zero) > 0;
    /* client.rs:428-437  confidence = if sum > 0 { weight / sum } else { 0 } */
    for (int i = 0; i < n; i++) {
        if (positive) {
            if ((rc = dec96_div(&choice_weight[i], &sum, &confidence[i]))) return rc;
        } else {
            confidence[i] = zero;
        }
    }
    /* client.rs:438-449  voter confidence = sum_i ( (cw_i / sum) * v_i ), the quotient recomputed per term */
    for (int m = 0; m < M; m++) {
        voter_conf[m] = zero;
        if (!has_vote[m]) continue;
        for (int i = 0; i < n; i++) {
            dec96 q = zero, term, acc;
            if (positive && (rc = dec96_div(&choice_weight[i], &sum, &q))) return rc;
            if ((rc = dec96_mul(&q, &vote[(size_t)m * n + i], &term))) return rc;
            if ((rc = dec96_add(&voter_conf[m], &term, &acc))) return rc;
            voter_conf[m] = acc;
        }
    }
    return DEC96_OK;
}

static const char *next_token(const char *p, char sep, char *tok, size_t toklen) {
    size_t k = 0;
    while (*p && *p != sep) {
        if (k + 1 < toklen) tok[k++] = *p;
        p++;
    }
    tok[k] = 0;
    return *p ? p + 1 : p;
}

static void csv_append(char *out, size_t outlen, size_t *pos, const char *s, int first) {
    size_t need = strlen(s) + (first ? 0 : 1);
    if (*pos + need + 1 > outlen) return;
    if (!first) out[(*pos)++] = ',';
    memcpy(out + *pos, s, strlen(s));
    *pos += strlen(s);
    out[*pos] = 0;
}

int lwc_oracle_tally_decimal(int n, int M, const char *votes_csv, const char *weights_csv, char *weight_out,
                             char *confidence_out, char *voter_conf_out, size_t outlen) {
    dec96 *vote = calloc((size_t)(M > 0 ? M : 1) * n, sizeof(dec96));
    dec96 *weight = calloc((size_t)(M > 0 ? M : 1), sizeof(dec96));
    uint8_t *has_vote = calloc((size_t)(M > 0 ? M : 1), 1);
    uint8_t *has_weight = calloc((size_t)(M > 0 ? M : 1), 1);
    dec96 *cw = calloc((size_t)n, sizeof(dec96));
    dec96 *conf = calloc((size_t)n, sizeof(dec96));
    dec96 *vconf = calloc((size_t)(M > 0 ? M : 1), sizeof(dec96));
    int rc = DEC96_OK;
    char row[1 << 16], tok[96];
    const char *vp = votes_csv, *wp = weights_csv;
    for (int m = 0; m < M && !rc; m++) {
        vp = next_token(vp, ';', row, sizeof row);
        if (strcmp(row, "-") != 0) {
            has_vote[m] = 1;
            const char *rp = row;
            for (int i = 0; i < n && !rc; i++) {
                rp = next_token(rp, ',', tok, sizeof tok);
                rc = dec96_from_str(tok, &vote[(size_t)m * n + i]);
            }
        }
        wp = next_token(wp, ',', tok, sizeof tok);
        if (!rc && strcmp(tok, "-") != 0) {
            has_weight[m] = 1;
            rc = dec96_from_str(tok, &weight[m]);
        }
    }
    if (!rc) rc = tally_dec(n, M, vote, has_vote, weight, has_weight, cw, conf, vconf);
    if (!rc) {
        size_t p0 = 0, p1 = 0, p2 = 0;
        weight_out[0] = confidence_out[0] = voter_conf_out[0] = 0;
        for (int i = 0; i < n; i++) {
            dec96_to_str_normalized(&cw[i], tok, sizeof tok);
            csv_append(weight_out, outlen, &p0, tok, i == 0);
            dec96_to_str_normalized(&conf[i], tok, sizeof tok);
            csv_append(confidence_out, outlen, &p1, tok, i == 0);
        }
        for (int m = 0; m < M; m++) {
            if (has_vote[m])
                dec96_to_str_normalized(&vconf[m], tok, sizeof tok);
            else
                strcpy(tok, "-");
            csv_append(voter_conf_out, outlen, &p2, tok, m == 0);
        }
    }
    free(vote), free(weight), free(has_vote), free(has_weight), free(cw), free(conf), free(vconf);
    return rc;
}

int lwc_oracle_merge_error_codes(const uint16_t *codes, int M) {
    /* client.rs:386-409: walk voters while all have errored; first code is kept, an unequal
     * later code collapses to 400 if both are 4xx, else 500. client.rs:459-463 + error.rs:41:
     * AllVotesFailed(Some(c)) -> c, AllVotesFailed(None) -> 500. */
    int all_error = 1, have_code = 0;
    uint16_t code = 0;
    for (int m = 0; m < M; m++) {
        if (!all_error) break;
        uint16_t e = codes[m];
        if (e == 0) {
            all_error = 0;
        } else if (!have_code) {
            code = e;
            have_code = 1;
        } else if (e != code) {
            if (e >= 400 && e < 500 && code >= 400 && code < 500)
                code = 400;
            else
                code = 500;
        }
    }
    if (!all_error) return 0;
    return have_code ? code : 500;
}

/* ------------------------------------------------------------------ f64 stages */

static inline double dot_f64(const double *a, const double *b, int D) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0, s6 = 0, s7 = 0;
    int k = 0;
    for (; k + 8 <= D; k += 8) {
        s0 += a[k] * b[k];
        s1 += a[k + 1] * b[k + 1];
        s2 += a[k + 2] * b[k + 2];
        s3 += a[k + 3] * b[k + 3];
        s4 += a[k + 4] * b[k + 4];
        s5 += a[k + 5] * b[k + 5];
        s6 += a[k + 6] * b[k + 6];
        s7 += a[k + 7] * b[k + 7];
    }
    for (; k < D; k++) s0 += a[k] * b[k];
    return ((s0 + s1) + (s2 + s3)) + ((s4 + s5) + (s6 + s7));
}

/* steps 3b-6 given t[] (f64) and, for cosine, the f64 copy of E and the norms */
static void finish_request(int N, int D, int M, int similarity, const double *Ed, const double *t,
                           const float *Vr, double *scratch_cw, double *conf, double *cw, double *agree,
                           int32_t *winner, int32_t *rank) {
    double *c = scratch_cw;
    if (similarity == LWC_ORACLE_SIM_COSINE) {
        double *nrm = c + N; /* scratch has 2N doubles */
        for (int i = 0; i < N; i++) nrm[i] = sqrt(dot_f64(Ed + (size_t)i * D, Ed + (size_t)i * D, D));
        for (int j = 0; j < N; j++) c[j] = 0.0;
        for (int i = 0; i < N; i++) {
            for (int j = i; j < N; j++) {
                double s = 0.0;
                if (nrm[i] > 0.0 && nrm[j] > 0.0)
                    s = dot_f64(Ed + (size_t)i * D, Ed + (size_t)j * D, D) / (nrm[i] * nrm[j]);
                c[j] += t[i] * s;
                if (j != i) c[i] += t[j] * s;
            }
        }
    } else {
        for (int j = 0; j < N; j++) c[j] = t[j];
    }
    double sum = 0.0;
    for (int j = 0; j < N; j++) sum += c[j];
    int positive = sum > 0.0;
    double *cf = c + 2 * (size_t)N; /* third N-slice of scratch: confidences */
    for (int j = 0; j < N; j++) cf[j] = positive ? c[j] / sum : 0.0;
    if (cw) memcpy(cw, c, sizeof(double) * N);
    if (conf) memcpy(conf, cf, sizeof(double) * N);
    if (agree) {
        for (int m = 0; m < M; m++) {
            const float *v = Vr + (size_t)m * N;
            if (isnan(v[0])) {
                agree[m] = NAN;
                continue;
            }
            double a = 0.0;
            for (int i = 0; i < N; i++) a += cf[i] * (double)v[i];
            agree[m] = a;
        }
    }
    if (winner) {
        int best = -1;
        if (positive) {
            best = 0;
            for (int j = 1; j < N; j++)
                if (cf[j] > cf[best]) best = j;
        }
        *winner = best;
    }
    if (rank) { /* stable insertion sort, confidence descending */
        for (int j = 0; j < N; j++) rank[j] = j;
        for (int a = 1; a < N; a++) {
            int32_t x = rank[a];
            int b = a - 1;
            while (b >= 0 && cf[rank[b]] < cf[x]) {
                rank[b + 1] = rank[b];
                b--;
            }
            rank[b + 1] = x;
        }
    }
}

int lwc_oracle_score_f64(int R, int N, int D, int M, int similarity, const float *E, const float *V,
                         const float *w, int w_per_request, double *conf, double *cw, double *agree,
                         int32_t *winner, int32_t *rank, int nthreads) {
    if (R < 0 || N < 1 || D < 0 || M < 0) return 1;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    int cosine = similarity == LWC_ORACLE_SIM_COSINE;
#pragma omp parallel num_threads(nthreads)
    {
        double *Ed = cosine ? malloc(sizeof(double) * (size_t)N * (D > 0 ? D : 1)) : NULL;
        double *t = malloc(sizeof(double) * (size_t)N);
        double *scratch = malloc(sizeof(double) * 3 * (size_t)N);
#pragma omp for schedule(dynamic, 4)
        for (int r = 0; r < R; r++) {
            const float *Vr = V + (size_t)r * M * N;
            const float *wr = w + (w_per_request ? (size_t)r * M : 0);
            for (int i = 0; i < N; i++) t[i] = 0.0;
            for (int m = 0; m < M; m++) {
                const float *v = Vr + (size_t)m * N;
                if (isnan(v[0])) continue;
                double wm = isnan(wr[m]) ? 0.0 : (double)wr[m];
                for (int i = 0; i < N; i++) t[i] += (double)v[i] * wm;
            }
            if (cosine) {
                const float *Er = E + (size_t)r * N * D;
                for (size_t k = 0; k < (size_t)N * D; k++) Ed[k] = (double)Er[k];
            }
            finish_request(N, D, M, similarity, Ed, t, Vr, scratch, conf ? conf + (size_t)r * N : NULL,
                           cw ? cw + (size_t)r * N : NULL, agree ? agree + (size_t)r * M : NULL,
                           winner ? winner + r : NULL, rank ? rank + (size_t)r * N : NULL);
        }
        free(Ed), free(t), free(scratch);
    }
    return 0;
}

/* The "port": Decimal tally on Decimal votes/weights, f64 embeddings. The fp32 -> Decimal
 * conversion of the synthetic inputs is done up front and is not what a caller should time;
 * see lwc_oracle_port_last_seconds(). */
static double g_port_seconds = 0.0;
double lwc_oracle_port_last_seconds(void) { return g_port_seconds; }

int lwc_oracle_score_port(int R, int N, int D, int M, int similarity, const float *E, const float *V,
                          const float *w, int w_per_request, double *conf, double *cw, double *agree,
                          int32_t *winner, int32_t *rank, int nthreads) {
    if (R < 0 || N < 1 || D < 0 || M < 0) return 1;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    int cosine = similarity == LWC_ORACLE_SIM_COSINE;
    size_t nv = (size_t)R * M * N, nw = w_per_request ? (size_t)R * M : (size_t)M;
    dec96 *Vd = malloc(sizeof(dec96) * (nv ? nv : 1));
    dec96 *Wd = malloc(sizeof(dec96) * (nw ? nw : 1));
    uint8_t *hasV = malloc((size_t)R * M + 1), *hasW = malloc(nw + 1);
    int bad = 0;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(| : bad)
    for (long rm = 0; rm < (long)R * M; rm++) {
        const float *v = V + (size_t)rm * N;
        hasV[rm] = !isnan(v[0]);
        for (int i = 0; i < N; i++) {
            if (hasV[rm])
                bad |= dec96_from_f32(v[i], &Vd[(size_t)rm * N + i]);
            else
                Vd[(size_t)rm * N + i] = dec96_zero();
        }
    }
    for (size_t k = 0; k < nw; k++) {
        hasW[k] = !isnan(w[k]);
        Wd[k] = dec96_zero();
        if (hasW[k]) bad |= dec96_from_f32(w[k], &Wd[k]);
    }
    if (bad) {
        free(Vd), free(Wd), free(hasV), free(hasW);
        return 2;
    }
    int fail = 0;
    double t0 = omp_get_wtime();
#pragma omp parallel num_threads(nthreads) reduction(| : fail)
    {
        double *Ed = cosine ? malloc(sizeof(double) * (size_t)N * (D > 0 ? D : 1)) : NULL;
        double *t = malloc(sizeof(double) * (size_t)N);
        double *scratch = malloc(sizeof(double) * 3 * (size_t)N);
        dec96 *cwd = malloc(sizeof(dec96) * (size_t)N), *cfd = malloc(sizeof(dec96) * (size_t)N);
        dec96 *vcd = malloc(sizeof(dec96) * (size_t)(M > 0 ? M : 1));
#pragma omp for schedule(dynamic, 4)
        for (int r = 0; r < R; r++) {
            const dec96 *vr = Vd + (size_t)r * M * N;
            const dec96 *wr = Wd + (w_per_request ? (size_t)r * M : 0);
            const uint8_t *hv = hasV + (size_t)r * M;
            const uint8_t *hw = hasW + (w_per_request ? (size_t)r * M : 0);
            fail |= tally_dec(N, M, vr, hv, wr, hw, cwd, cfd, vcd);
            double *conf_r = conf ? conf + (size_t)r * N : NULL;
            double *cw_r = cw ? cw + (size_t)r * N : NULL;
            double *ag_r = agree ? agree + (size_t)r * M : NULL;
            if (!cosine) { /* pure reference semantics: every number below is a Decimal */
                dec96 zero = dec96_zero();
                int best = -1;
                for (int i = 0; i < N; i++) {
                    scratch[i] = dec96_to_double(&cfd[i]);
                    if (cw_r) cw_r[i] = dec96_to_double(&cwd[i]);
                    if (conf_r) conf_r[i] = scratch[i];
                    if (dec96_cmp(&cfd[i], &zero) > 0 && (best < 0 || dec96_cmp(&cfd[i], &cfd[best]) > 0)) best = i;
                }
                if (ag_r)
                    for (int m = 0; m < M; m++) ag_r[m] = hv[m] ? dec96_to_double(&vcd[m]) : NAN;
                if (winner) winner[r] = best;
                if (rank) {
                    int32_t *rk = rank + (size_t)r * N;
                    for (int j = 0; j < N; j++) rk[j] = j;
                    for (int a = 1; a < N; a++) {
                        int32_t x = rk[a];
                        int b = a - 1;
                        while (b >= 0 && dec96_cmp(&cfd[rk[b]], &cfd[x]) < 0) {
                            rk[b + 1] = rk[b];
                            b--;
                        }
                        rk[b + 1] = x;
                    }
                }
            } else {
                for (int i = 0; i < N; i++) t[i] = dec96_to_double(&cwd[i]);
                const float *Er = E + (size_t)r * N * D;
                for (size_t k = 0; k < (size_t)N * D; k++) Ed[k] = (double)Er[k];
                finish_request(N, D, M, similarity, Ed, t, V + (size_t)r * M * N, scratch, conf_r, cw_r, ag_r,
                               winner ? winner + r : NULL, rank ? rank + (size_t)r * N : NULL);
            }
        }
        free(Ed), free(t), free(scratch), free(cwd), free(cfd), free(vcd);
    }
    g_port_seconds = omp_get_wtime() - t0;
    free(Vd), free(Wd), free(hasV), free(hasW);
    return fail ? 3 : 0;
}

/* ------------------------------------------------------------------ one big pool */

int lwc_oracle_pool_f64(int64_t N, int D, int M, int method, const float *E, const float *V, const float *w,
                        double *conf, double *cw, double *agree, int64_t *winner, int nthreads) {
    if (N < 1 || D < 1 || M < 0) return 1;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    double *t = calloc((size_t)N, sizeof(double));
    double *inv = malloc(sizeof(double) * (size_t)N);
    double *c = calloc((size_t)N, sizeof(double));
    for (int m = 0; m < M; m++) {
        const float *v = V + (size_t)m * N;
        if (isnan(v[0])) continue;
        double wm = isnan(w[m]) ? 0.0 : (double)w[m];
        for (int64_t i = 0; i < N; i++) t[i] += (double)v[i] * wm;
    }
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < N; i++) {
        const float *e = E + (size_t)i * D;
        double s = 0.0;
        for (int k = 0; k < D; k++) s += (double)e[k] * (double)e[k];
        inv[i] = s > 0.0 ? 1.0 / sqrt(s) : 0.0;
    }
    if (method == 1) {
        /* centroid c_vec = sum_i t_i E^_i, then cw_j = <E^_j, c_vec> */
        double *cv = calloc((size_t)D, sizeof(double));
#pragma omp parallel num_threads(nthreads)
        {
            double *loc = calloc((size_t)D, sizeof(double));
#pragma omp for schedule(static)
            for (int64_t i = 0; i < N; i++) {
                double f = t[i] * inv[i];
                if (f == 0.0) continue;
                const float *e = E + (size_t)i * D;
                for (int k = 0; k < D; k++) loc[k] += f * (double)e[k];
            }
#pragma omp critical
            for (int k = 0; k < D; k++) cv[k] += loc[k];
            free(loc);
        }
#pragma omp parallel for num_threads(nthreads) schedule(static)
        for (int64_t j = 0; j < N; j++) {
            const float *e = E + (size_t)j * D;
            double s = 0.0;
            for (int k = 0; k < D; k++) s += (double)e[k] * cv[k];
            c[j] = s * inv[j];
        }
        free(cv);
    } else {
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 8)
        for (int64_t j = 0; j < N; j++) {
            const float *ej = E + (size_t)j * D;
            double acc = 0.0;
            for (int64_t i = 0; i < N; i++) {
                if (t[i] == 0.0) continue;
                const float *ei = E + (size_t)i * D;
                double s = 0.0;
                for (int k = 0; k < D; k++) s += (double)ei[k] * (double)ej[k];
                acc += t[i] * (s * inv[i] * inv[j]);
            }
            c[j] = acc;
        }
    }
    double sum = 0.0;
    for (int64_t j = 0; j < N; j++) sum += c[j];
    int positive = sum > 0.0;
    int64_t best = -1;
    double bestv = 0.0;
    for (int64_t j = 0; j < N; j++) {
        double cf = positive ? c[j] / sum : 0.0;
        if (cw) cw[j] = c[j];
        if (positive && (best < 0 || cf > bestv)) {
            best = j;
            bestv = cf;
        }
        c[j] = cf;
    }
    if (conf) memcpy(conf, c, sizeof(double) * (size_t)N);
    if (winner) *winner = best;
    if (agree) {
        for (int m = 0; m < M; m++) {
            const float *v = V + (size_t)m * N;
            if (isnan(v[0])) {
                agree[m] = NAN;
                continue;
            }
            double a = 0.0;
            for (int64_t i = 0; i < N; i++) a += c[i] * (double)v[i];
            agree[m] = a;
        }
    }
    free(t), free(inv), free(c);
    return 0;
}

/* ------------------------------------------------------------------ get_vote numeric tail */

int lwc_oracle_vote_from_logprobs(int n, int n_top, const double *logprob, const int32_t *leaf, double *vote) {
    /* client.rs:1719 vote = [0; n]; 1765-1784 p = exp(logprob), vote[leaf] += p, sum += p;
     * 1789-1791 vote /= sum.  (The reference's Decimal::exp is a truncated series; f64 exp is
     * the exact-limit restatement and the tolerance is stated in the tests.) */
    for (int i = 0; i < n; i++) vote[i] = 0.0;
    double sum = 0.0;
    for (int k = 0; k < n_top; k++) {
        if (leaf[k] < 0 || leaf[k] >= n || isnan(logprob[k])) continue;
        double p = exp(logprob[k]);
        vote[leaf[k]] += p;
        sum += p;
    }
    if (sum == 0.0) return 1; /* client.rs:1786-1788 unreachable!() in the reference */
    for (int i = 0; i < n; i++) vote[i] /= sum;
    return 0;
}

/* ------------------------------------------------------------------ training-table weights (8f2) */

int lwc_oracle_table_weights(int64_t rows, int D, int M, const float *T, const float *Q, int n_prompts,
                             const float *q, int top, const float *base, const float *minw, const float *maxw,
                             double *weights, int32_t *top_idx, double *top_sim) {
    /* The reference has no arithmetic here (weight.rs:99-117 is `unimplemented!()`): this restates
     * the BUILD's definition (include/lwc_b200.h, lwc_table_weights) in f64 so the CUDA path has
     * an independent checker.  Limits follow llm/mod.rs:671-687, `top >= 1` model/mod.rs:292-297. */
    if (rows < 1 || top < 1) return 1;
    for (int m = 0; m < M; m++)
        if (!(minw[m] > 0.f) || !(base[m] >= minw[m]) || !(base[m] <= maxw[m])) return 1;
    double *sc = (double *)malloc(sizeof(double) * (size_t)rows);
    double *rn = (double *)malloc(sizeof(double) * (size_t)rows);
    double *rho = (double *)malloc(sizeof(double) * (size_t)M);
    for (int64_t t = 0; t < rows; t++) {
        double ss = 0.0;
        for (int k = 0; k < D; k++) ss += (double)T[t * D + k] * (double)T[t * D + k];
        rn[t] = ss > 0.0 ? 1.0 / sqrt(ss) : 0.0;
    }
    for (int p = 0; p < n_prompts; p++) {
        const float *qp = q + (size_t)p * D;
        double qq = 0.0;
        for (int k = 0; k < D; k++) qq += (double)qp[k] * (double)qp[k];
        double qn = qq > 0.0 ? 1.0 / sqrt(qq) : 0.0;
        for (int64_t t = 0; t < rows; t++) {
            double d = 0.0;
            for (int k = 0; k < D; k++) d += (double)T[t * D + k] * (double)qp[k];
            sc[t] = d * rn[t] * qn;
        }
        double ssum = 0.0;
        for (int m = 0; m < M; m++) rho[m] = 0.0;
        for (int k = 0; k < top; k++) {
            int64_t arg = -1;
            double best = -INFINITY;
            for (int64_t t = 0; t < rows; t++)
                if (sc[t] > best) best = sc[t], arg = t;
            top_idx[(size_t)p * top + k] = (int32_t)arg;
            top_sim[(size_t)p * top + k] = arg >= 0 ? best : 0.0;
            if (arg < 0) continue;
            sc[arg] = -INFINITY;
            double s = best > 0.0 ? best : 0.0;
            ssum += s;
            for (int m = 0; m < M; m++) rho[m] += s * (double)Q[arg * M + m];
        }
        double mean = 0.0;
        for (int m = 0; m < M; m++) {
            if (ssum > 0.0) rho[m] /= ssum;
            mean += rho[m];
        }
        mean /= (double)M;
        for (int m = 0; m < M; m++) {
            double w = base[m];
            if (ssum > 0.0 && mean > 0.0) w = (double)base[m] * (rho[m] / mean);
            if (w < minw[m]) w = minw[m];
            if (w > maxw[m]) w = maxw[m];
            weights[(size_t)p * M + m] = w;
        }
    }
    free(sc), free(rn), free(rho);
    return 0;
}
