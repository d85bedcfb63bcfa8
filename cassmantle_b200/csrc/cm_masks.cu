// cassmantle_b200 -- the numeric half of mask selection (include/...: cm_semantic_distances).
//
// SURVEY.md §8f rank 4.  Reference: semantic_distance, src/utils.py:73-79, called for every
// filtered word of the round's prompt by select_descriptive_words (src/utils.py:91):
//     mean = np.mean([model[w] for w in word_list if w in model], axis=0)   fp32 running sum / n
//     d(w) = np.linalg.norm(model[w] - mean)      if w in model else 0       sqrt(dot(x, x)), fp32
// Once per round and a few dozen rows: nothing here is bandwidth- or latency-critical.  It is on
// the device because the table is (the host keeps no copy of the 3.6 GB), and it is one launch:
// a single CTA sums the known rows in list order (as numpy's reduction over axis 0 does), keeps
// the mean in shared memory, then one warp per word forms ||row - mean||.

#include <cuda_runtime.h>

#include <cstring>

#include "cm_common.cuh"
#include "cm_internal.h"

namespace {

using cm::fail;

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
cm_semantic_distance_kernel(const float *__restrict__ table, int64_t n_rows, int dim,
                            const int32_t *__restrict__ rows, int n, int n_mean, float *__restrict__ out) {
    extern __shared__ float mean[];                          // [dim]
    __shared__ int s_known;
    if (threadIdx.x == 0) {
        int known = 0;
        for (int i = 0; i < n_mean; ++i) known += (rows[i] >= 0 && (int64_t)rows[i] < n_rows) ? 1 : 0;
        s_known = known;
    }
    __syncthreads();
    const int known = s_known;
    for (int d = threadIdx.x; d < dim; d += kThreads) {
        float s = 0.f;
        for (int i = 0; i < n_mean; ++i) {                   // list order, like add.reduce over axis 0
            const int32_t r = rows[i];
            if (r >= 0 && (int64_t)r < n_rows) s += table[(size_t)r * dim + d];
        }
        mean[d] = known > 0 ? __fdiv_rn(s, (float)known) : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = warp; i < n; i += kThreads / 32) {
        const int32_t r = rows[i];
        float ss = 0.f;
        if (r >= 0 && (int64_t)r < n_rows) {
            for (int d = lane; d < dim; d += 32) {
                const float x = __fsub_rn(table[(size_t)r * dim + d], mean[d]);
                ss = fmaf(x, x, ss);
            }
        }
        ss = cm::warp_sum(ss);
        if (lane == 0) out[i] = (r >= 0 && (int64_t)r < n_rows) ? sqrtf(ss) : 0.f;
    }
}

}  // namespace

extern "C" int cm_semantic_distances(cm_engine *e, const int32_t *rows, int n, int n_mean, float *out) {
    if (int rc = cm::check_engine(e)) return rc;
    if (n < 0 || n > CM_MAX_SEQ) return fail(CM_ERR_INVALID, "n %d outside [0, %d]", n, CM_MAX_SEQ);
    if (n_mean < 0 || n_mean > n) return fail(CM_ERR_INVALID, "n_mean %d outside [0, n = %d]", n_mean, n);
    if (n == 0) return CM_OK;
    if (rows == nullptr || out == nullptr) return fail(CM_ERR_INVALID, "rows or out is NULL");
    std::lock_guard<std::mutex> lock(e->mu);
    cm::DeviceGuard guard(e->device);
    // rows in, distances out through the engine's page-locked block (a few KB either way)
    const size_t off_out = cm::align_up((size_t)n * sizeof(int32_t), 16);
    if (int rc = cm::ensure_stage(e, off_out + (size_t)n * sizeof(float))) return rc;
    memcpy(e->h_stage, rows, (size_t)n * sizeof(int32_t));
    cm_semantic_distance_kernel<<<1, kThreads, (size_t)e->dim * sizeof(float), e->stream>>>(
        e->d_table, e->n_rows, e->dim, reinterpret_cast<const int32_t *>(e->h_stage), n, n_mean,
        reinterpret_cast<float *>(e->h_stage + off_out));
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    CM_CUDA(cudaStreamSynchronize(e->stream));
    memcpy(out, e->h_stage + off_out, (size_t)n * sizeof(float));
    return CM_OK;
}
