/*
 * CPU oracle for the CassMantle guess scorer, plain C -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load the
 * library built from this file (oracle/Makefile -> oracle/_build/liboracle_scorer.so).
 * The product library (cassmantle_b200/libcassmantle_b200.so) never links or calls it.
 *
 * PARITY UNPINNED: the reference ships no tests or golden vectors and its arithmetic sits
 * in gensim (requirements.txt:5, unpinned, absent from /root/reference).  This file
 * restates
 *   - src/backend.py:303-310  compute_score: equal strings -> 1.0, KeyError -> min_score,
 *                             max(min_score, score) keeps the cosine only if it is greater;
 *   - gensim KeyedVectors.similarity (call site src/backend.py:307):
 *         dot(unitvec(row_a), unitvec(row_b))
 *   - gensim matutils.unitvec: norm and scaling through the DOUBLE precision BLAS nrm2 /
 *     scal, result cast back to fp32; a zero vector is returned unchanged;
 *   - the labelled N>1 extension = gensim classic n_similarity: fp32 in-order mean of the
 *     rows (numpy mean(axis=0): running fp32 sum, then one fp32 divide), unitvec, dot.
 * It is checked against oracle/scorer_oracle.py (which calls the real scipy BLAS routines
 * gensim binds) and against tests/golden/scorer_kat.json.
 *
 * The fp32 dot is accumulated left to right in fp32; BLAS sdot's order is implementation
 * defined, so agreement with the numpy oracle is to ~1e-7, not bit-exact.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* reference-BLAS dnrm2 on an fp32 vector promoted to fp64 (scaled sum of squares). */
static double nrm2_promoted(const float *v, int n) {
    double scale = 0.0, ssq = 1.0;
    for (int i = 0; i < n; ++i) {
        double x = (double)v[i];
        if (x != 0.0) {
            double ax = fabs(x);
            if (scale < ax) {
                ssq = 1.0 + ssq * (scale / ax) * (scale / ax);
                scale = ax;
            } else {
                ssq += (ax / scale) * (ax / scale);
            }
        }
    }
    return scale * sqrt(ssq);
}

/* gensim matutils.unitvec: out = (double)v * (1/|v|) rounded to fp32; unchanged if |v| == 0 */
void oracle_unitvec(const float *v, int n, float *out) {
    double len = nrm2_promoted(v, n);
    if (len > 0.0) {
        double inv = 1.0 / len;
        for (int i = 0; i < n; ++i) out[i] = (float)((double)v[i] * inv);
    } else {
        memcpy(out, v, (size_t)n * sizeof(float));
    }
}

static float dot_f32(const float *a, const float *b, int n) {
    float acc = 0.0f;
    for (int i = 0; i < n; ++i) acc += a[i] * b[i];
    return acc;
}

float oracle_similarity(const float *table, int dim, int64_t i, int64_t j) {
    float ua[1024], ub[1024];
    oracle_unitvec(table + i * dim, dim, ua);
    oracle_unitvec(table + j * dim, dim, ub);
    return dot_f32(ua, ub, dim);
}

typedef struct {
    const float *table; int dim;
    const int32_t *guess_idx; int N; const int32_t *guess_len;
    const int32_t *answer_rows; const int32_t *answer_slot; const uint8_t *exact;
    float min_score; float *out; int lo, hi;
} job_t;

static void *score_range(void *p) {
    job_t *j = (job_t *)p;
    const int D = j->dim;
    float *mean = (float *)malloc(sizeof(float) * 3 * (size_t)D);
    float *ug = mean + D, *ua = mean + 2 * D;
    for (int b = j->lo; b < j->hi; ++b) {
        if (j->exact && j->exact[b]) { j->out[b] = 1.0f; continue; }
        int n = j->guess_len[b];
        int32_t a = j->answer_rows[j->answer_slot[b]];
        int oov = (n <= 0) || (a < 0);
        for (int k = 0; k < n && !oov; ++k) oov = j->guess_idx[(size_t)b * j->N + k] < 0;
        if (oov) { j->out[b] = j->min_score; continue; }            /* KeyError path */
        const float *r0 = j->table + (size_t)j->guess_idx[(size_t)b * j->N] * D;
        if (n == 1) {
            oracle_unitvec(r0, D, ug);
        } else {
            memcpy(mean, r0, sizeof(float) * D);
            for (int k = 1; k < n; ++k) {
                const float *r = j->table + (size_t)j->guess_idx[(size_t)b * j->N + k] * D;
                for (int d = 0; d < D; ++d) mean[d] += r[d];
            }
            float fn = (float)n;
            for (int d = 0; d < D; ++d) mean[d] /= fn;
            oracle_unitvec(mean, D, ug);
        }
        oracle_unitvec(j->table + (size_t)a * D, D, ua);
        float s = dot_f32(ug, ua, D);
        j->out[b] = (s > j->min_score) ? s : j->min_score;          /* max(min_score, s) */
    }
    free(mean);
    return NULL;
}

/* Index-form batch scorer; returns 0, or -1 on bad arguments. */
int oracle_score_batch(const float *table, int dim,
                       const int32_t *guess_idx, int B, int N, const int32_t *guess_len,
                       const int32_t *answer_rows, const int32_t *answer_slot,
                       const uint8_t *exact, float min_score, float *out, int threads) {
    if (!table || !guess_idx || !guess_len || !answer_rows || !answer_slot || !out) return -1;
    if (dim <= 0 || dim > 1024 || B < 0 || N <= 0) return -1;
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if (threads > B) threads = B > 0 ? B : 1;
    pthread_t tid[256];
    job_t jobs[256];
    int per = (B + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        job_t j = {table, dim, guess_idx, N, guess_len, answer_rows, answer_slot, exact,
                   min_score, out, t * per, (t + 1) * per < B ? (t + 1) * per : B};
        jobs[t] = j;
        if (threads == 1) score_range(&jobs[t]);
        else pthread_create(&tid[t], NULL, score_range, &jobs[t]);
    }
    if (threads > 1) for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
    return 0;
}

/* ------------------------------------------------------------------------------------ *
 * Seeded synthetic table, host side (same definition as cassmantle_b200/synth.py; the
 * numpy version is the specification, this one only exists because filling 3.6 GB with
 * numpy takes half a minute).  Used by bench.py's CPU legs and checked bit-for-bit against
 * synth.py in tests/test_oracle.py.
 * ------------------------------------------------------------------------------------ */
static uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
static int bytesum(uint32_t h) {
    return (int)((h & 0xffu) + ((h >> 8) & 0xffu) + ((h >> 16) & 0xffu) + (h >> 24)) - 510;
}

typedef struct { float *table; int64_t lo, hi; int dim; uint32_t seed; } fill_job_t;

static void *fill_range(void *p) {
    fill_job_t *j = (fill_job_t *)p;
    const uint32_t k_noise = mix32(j->seed ^ 0x9e3779b9u);
    const uint32_t k_clu = mix32(j->seed ^ 0x85ebca6bu);
    const uint32_t k_cen = mix32(j->seed ^ 0xc2b2ae35u);
    const float inv_unit = (float)(1.0 / 3620.0);
    const uint32_t D = (uint32_t)j->dim;
    for (int64_t r = j->lo; r < j->hi; ++r) {
        const uint32_t row = (uint32_t)r;
        const float u = (float)(mix32(row * 0x9e3779b1u + j->seed) >> 8) * 5.9604644775390625e-08f;
        const float t = 4.5f * u;
        const float rs = 0.5f + t;
        const float coef = rs * inv_unit;
        const uint32_t clu = mix32(row ^ k_clu) & 1023u;
        float *dst = j->table + (size_t)r * D;
        for (uint32_t c = 0; c < D; ++c) {
            const int s = bytesum(mix32((row * D + c) ^ k_noise)) + bytesum(mix32((clu * D + c) ^ k_cen));
            dst[c] = (float)s * coef;
        }
    }
    return NULL;
}

int oracle_synth_fill(float *table, int64_t n_rows, int dim, uint32_t seed, int threads) {
    if (!table || n_rows <= 0 || dim <= 0 || (uint64_t)n_rows * (uint64_t)dim > 0xffffffffull) return -1;
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    fill_job_t jobs[256];
    int64_t per = (n_rows + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        int64_t lo = t * per, hi = (t + 1) * per < n_rows ? (t + 1) * per : n_rows;
        if (lo > hi) lo = hi;
        fill_job_t j = {table, lo, hi, dim, seed};
        jobs[t] = j;
        pthread_create(&tid[t], NULL, fill_range, &jobs[t]);
    }
    for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
    return 0;
}
