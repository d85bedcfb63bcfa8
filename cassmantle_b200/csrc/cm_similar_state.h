// Shared between cm_similar.cu (one query) and cm_similar_batch.cu (many queries per table pass).  Not installed.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "cm_internal.h"

struct cm_similar_batch;                 // cm_similar_batch.cu
struct cm_similar_select;                // SelectState / PickState live in cm_similar.cu

struct cm_similar_state {
    float *d_norms = nullptr;
    float *d_mean = nullptr;
    float *d_dists = nullptr;
    void *d_sel = nullptr;               // SelectState
    uint32_t *d_cand_key = nullptr;
    int32_t *d_cand_row = nullptr;
    int32_t *d_rows_out = nullptr;
    float *d_sims_out = nullptr;
    int *d_n_out = nullptr;
    void *d_pick = nullptr;              // PickState
    unsigned char *h_result = nullptr;   // page-locked: rows[kMaxTopK] | sims[kMaxTopK] | n -- the kernels write it directly
    unsigned long long *d_cand = nullptr;
    bool norms_ready = false;
    cm_similar_batch *batch = nullptr;   // scratch of cm_most_similar_batch, allocated on first use
};

namespace cm {

constexpr int kSimilarMaxTopK = 1024;

// order-preserving map float -> uint32 (ascending); NaN -> 0 (smallest)
__device__ __forceinline__ uint32_t similar_key_of(float d) {
    if (d != d) return 0u;
    const uint32_t b = __float_as_uint(d);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Which scan arithmetic cm_most_similar uses for a table of dim4 float4 per row (the 16-lane kernel for
// 300-d tables, the generic 32-lane one otherwise): the batched path re-scores its candidates the same way.
bool similar_uses_scan16(int dim4);

// With e->mu held and e->device current: the scratch exists and norms[r] = ||row r|| are computed (enqueued on e->stream).
int similar_prepare(cm_engine *e, cm_similar_state **out);

// cm_most_similar without taking e->mu (the caller holds it).
int most_similar_locked(cm_engine *e, int32_t row, int k, int32_t *rows_out, float *sims_out, int *n_out);

void similar_batch_free(cm_similar_state *s);      // cm_similar_batch.cu

}  // namespace cm
