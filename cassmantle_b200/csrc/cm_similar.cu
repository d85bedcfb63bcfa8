// cassmantle_b200 -- most_similar on the device (include/...: cm_most_similar).
//
// SURVEY.md §8f rank 2.  Reference: Backend.most_similar, src/backend.py:297-301 =
// gensim KeyedVectors.most_similar(positive=[word], topn):
//     norms = ||vectors[i]||_2  (fp32, once)                       fill_norms()
//     mean  = unitvec(vectors[w] / norms[w])                        one positive word, weight 1
//     dists = dot(vectors, mean) / norms                            3M x 300 mat-vec
//     best  = argsort(-dists)[: topn + 1], the query word itself dropped
// One query is a single pass over the 3.6 GB table -- HBM-bound like the scorer, no reuse, so
// no tensor-core work either (a BATCH of queries would make it a skinny GEMM; not built).
//
//   cm_row_norm_kernel   once per engine: norms[V]
//   cm_dist_kernel       dists[r] = dot(row r, mean) / norms[r]; 4 rows per warp in flight
//   cm_hist_kernel x4    radix select of the k-th largest key, 8 bits per pass, over the
//                        12 MB of dists (L2-resident); each pass re-derives the prefix chosen
//                        by the previous one from its histogram, so there is no host round trip
//   cm_collect_kernel    everything above the threshold + ties until k candidates
//   cm_sort_kernel       bitonic sort of the <= 1024 candidates, (dist desc, row asc)
//
// NaN distances (NaN rows, zero norms) rank last, as numpy's argsort places them.

#include <cuda_runtime.h>

#include <new>

#include "cm_common.cuh"
#include "cm_internal.h"

namespace {

using cm::fail;

constexpr int kMaxTopK = 1024;
constexpr int kRowsPerWarp = 4;

// order-preserving map float -> uint32 (ascending); NaN -> 0 (smallest)
__device__ __forceinline__ uint32_t key_of(float d) {
    if (d != d) return 0u;
    const uint32_t b = __float_as_uint(d);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void __launch_bounds__(256)
cm_row_norm_kernel(const float *__restrict__ table, int64_t n_rows, int dim4, float *__restrict__ norms) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const float4 *t4 = reinterpret_cast<const float4 *>(table);
    for (int64_t r = warp; r < n_rows; r += n_warps) {
        float s = 0.f;
        for (int j = lane; j < dim4; j += 32) {
            const float4 v = cm::ld_stream_f4(t4 + r * dim4 + j);
            s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        }
        s = cm::warp_sum(s);
        if (lane == 0) norms[r] = sqrtf(s);
    }
}

// mean = unitvec(vectors[w] / norms[w]) -- one warp
__global__ void cm_query_kernel(const float *__restrict__ table, int dim, int32_t row,
                                const float *__restrict__ norms, float *__restrict__ mean) {
    const int lane = threadIdx.x;
    const float nrm = norms[row];
    double ss = 0.0;
    for (int j = lane; j < dim; j += 32) {
        const float q = table[(size_t)row * dim + j] / nrm;          // get_vector(key, norm=True)
        mean[j] = q;
        ss = fma((double)q, (double)q, ss);
    }
    ss = cm::warp_sum(ss);
    __syncwarp();
    if (ss > 0.0) {                                                  // matutils.unitvec
        const double inv = 1.0 / sqrt(ss);
        for (int j = lane; j < dim; j += 32) mean[j] = (float)((double)mean[j] * inv);
    }
}

template <int VPL>
__global__ void __launch_bounds__(256)
cm_dist_kernel(const float *__restrict__ table, int64_t n_rows, int dim4, const float *__restrict__ mean,
               const float *__restrict__ norms, float *__restrict__ dists) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const float4 *t4 = reinterpret_cast<const float4 *>(table);
    float4 q[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        const int j = lane + 32 * v;
        q[v] = j < dim4 ? reinterpret_cast<const float4 *>(mean)[j] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int64_t r0 = warp * kRowsPerWarp; r0 < n_rows; r0 += n_warps * kRowsPerWarp) {
        float4 buf[kRowsPerWarp][VPL];
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
            const int64_t r = r0 + i;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                const int j = lane + 32 * v;
                buf[i][v] = (r < n_rows && j < dim4) ? cm::ld_stream_f4(t4 + r * dim4 + j)
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        float p[kRowsPerWarp];
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
            float s = 0.f;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                s = fmaf(buf[i][v].x, q[v].x, s); s = fmaf(buf[i][v].y, q[v].y, s);
                s = fmaf(buf[i][v].z, q[v].z, s); s = fmaf(buf[i][v].w, q[v].w, s);
            }
            p[i] = s;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int i = 0; i < kRowsPerWarp; ++i) p[i] += __shfl_xor_sync(cm::kFullMask, p[i], o);
        }
        if (lane < kRowsPerWarp && r0 + lane < n_rows) {
            float mine = p[0];
#pragma unroll
            for (int i = 1; i < kRowsPerWarp; ++i) mine = (lane == i) ? p[i] : mine;
            dists[r0 + lane] = mine / norms[r0 + lane];
        }
    }
}

// Radix-select state: hist[pass][256]; the prefix chosen so far is re-derived by every CTA.
struct SelectState {
    unsigned hist[4][256];
    unsigned n_above;       // candidates strictly above the threshold (exactly k - ties needed)
    unsigned n_tie;         // candidates equal to it that were admitted
};

// prefix/`remaining` after `pass` completed passes (pass in 0..4), scanning bins from the top
__device__ void derive_prefix(const SelectState *st, int pass, unsigned k, uint32_t *prefix, unsigned *remaining) {
    uint32_t pre = 0;
    unsigned rem = k;
    for (int p = 0; p < pass; ++p) {
        unsigned acc = 0;
        int b = 255;
        for (; b > 0; --b) {
            const unsigned c = st->hist[p][b];
            if (acc + c >= rem) break;
            acc += c;
        }
        rem -= acc;
        pre |= (uint32_t)b << (24 - 8 * p);
    }
    *prefix = pre;
    *remaining = rem;
}

__global__ void __launch_bounds__(256)
cm_hist_kernel(const float *__restrict__ dists, int64_t n, SelectState *st, int pass, unsigned k) {
    __shared__ unsigned sh[256];
    __shared__ uint32_t s_prefix;
    sh[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        unsigned rem;
        uint32_t pre;
        derive_prefix(st, pass, k, &pre, &rem);
        s_prefix = pre;
    }
    __syncthreads();
    const uint32_t prefix = s_prefix;
    const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (32 - 8 * pass));
    const int shift = 24 - 8 * pass;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t key = key_of(dists[i]);
        if ((key & mask) == prefix) atomicAdd(&sh[(key >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&st->hist[pass][threadIdx.x], sh[threadIdx.x]);
}

__global__ void __launch_bounds__(256)
cm_collect_kernel(const float *__restrict__ dists, int64_t n, SelectState *st, unsigned k,
                  uint32_t *__restrict__ cand_key, int32_t *__restrict__ cand_row) {
    __shared__ uint32_t s_thr;
    __shared__ unsigned s_ties;
    if (threadIdx.x == 0) {
        unsigned rem;
        uint32_t thr;
        derive_prefix(st, 4, k, &thr, &rem);
        s_thr = thr;                                 // key of the k-th largest distance
        s_ties = rem;                                // how many of the entries EQUAL to it belong to the top k
    }
    __syncthreads();
    const uint32_t thr = s_thr;
    const unsigned ties = s_ties, above = k - ties;  // exactly `above` keys are strictly larger
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t key = key_of(dists[i]);
        if (key > thr) {
            const unsigned slot = atomicAdd(&st->n_above, 1u);
            cand_key[slot] = key; cand_row[slot] = (int32_t)i;
        } else if (key == thr) {                     // which of several exact ties get in is arbitrary
            const unsigned t = atomicAdd(&st->n_tie, 1u);
            if (t < ties) { cand_key[above + t] = key; cand_row[above + t] = (int32_t)i; }
        }
    }
}

// one CTA: sort candidates by (key desc, row asc), write the first k
__global__ void __launch_bounds__(1024)
cm_sort_kernel(const SelectState *st, const uint32_t *__restrict__ cand_key, const int32_t *__restrict__ cand_row,
               const float *__restrict__ dists, unsigned k, int32_t *__restrict__ rows_out,
               float *__restrict__ sims_out, int *__restrict__ n_out) {
    __shared__ unsigned long long s[kMaxTopK];
    const unsigned n = min(k, (unsigned)kMaxTopK);      // the collect pass admitted exactly k
    const unsigned t = threadIdx.x;
    // ascending sort of (~key, row) == descending key, ascending row
    s[t] = t < n ? (((unsigned long long)(~cand_key[t])) << 32) | (uint32_t)cand_row[t] : ~0ull;
    __syncthreads();
    for (unsigned size = 2; size <= (unsigned)kMaxTopK; size <<= 1) {
        for (unsigned stride = size >> 1; stride > 0; stride >>= 1) {
            const unsigned partner = t ^ stride;
            if (partner > t) {
                const bool up = (t & size) == 0;
                const unsigned long long a = s[t], b = s[partner];
                if ((a > b) == up) { s[t] = b; s[partner] = a; }
            }
            __syncthreads();
        }
    }
    const unsigned m = min(n, k);
    if (t < m) {
        const int32_t r = (int32_t)(s[t] & 0xffffffffull);
        rows_out[t] = r;
        sims_out[t] = dists[r];
    }
    if (t == 0) *n_out = (int)m;
}

}  // namespace

struct cm_similar_state {
    float *d_norms = nullptr;
    float *d_mean = nullptr;
    float *d_dists = nullptr;
    SelectState *d_sel = nullptr;
    uint32_t *d_cand_key = nullptr;
    int32_t *d_cand_row = nullptr;
    int32_t *d_rows_out = nullptr;
    float *d_sims_out = nullptr;
    int *d_n_out = nullptr;
    bool norms_ready = false;
};

namespace cm {

void similar_state_free(cm_engine *e) {
    cm_similar_state *s = e->similar;
    if (s == nullptr) return;
    void *ptrs[] = {s->d_norms, s->d_mean, s->d_dists, s->d_sel, s->d_cand_key, s->d_cand_row,
                    s->d_rows_out, s->d_sims_out, s->d_n_out};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete s;
    e->similar = nullptr;
}

}  // namespace cm

static int similar_state(cm_engine *e, cm_similar_state **out) {
    if (e->similar == nullptr) {
        cm_similar_state *s = new (std::nothrow) cm_similar_state();
        if (s == nullptr) return fail(CM_ERR_NOMEM, "most_similar state allocation failed");
        e->similar = s;
        cudaError_t err = cudaMalloc(&s->d_norms, (size_t)e->n_rows * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_dists, (size_t)e->n_rows * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_mean, (size_t)e->dim * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_sel, sizeof(SelectState));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_cand_key, kMaxTopK * sizeof(uint32_t));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_cand_row, kMaxTopK * sizeof(int32_t));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_rows_out, kMaxTopK * sizeof(int32_t));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_sims_out, kMaxTopK * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_n_out, sizeof(int));
        if (err != cudaSuccess) {
            cudaGetLastError();
            cm::similar_state_free(e);
            return fail(CM_ERR_NOMEM, "most_similar scratch (%zu bytes) failed: %s",
                        (size_t)e->n_rows * 8, cudaGetErrorString(err));
        }
    }
    *out = e->similar;
    return CM_OK;
}

extern "C" int cm_most_similar(cm_engine *e, int32_t row, int k, int32_t *rows_out, float *sims_out, int *n_out) {
    if (int rc = cm::check_engine(e)) return rc;
    if (row < 0 || (int64_t)row >= e->n_rows)
        return fail(CM_ERR_INVALID, "row %d outside the %lld-row table", row, (long long)e->n_rows);
    if (k < 1 || k > kMaxTopK) return fail(CM_ERR_INVALID, "k %d outside [1, %d]", k, kMaxTopK);
    if (!rows_out || !sims_out || !n_out) return fail(CM_ERR_INVALID, "an output pointer is NULL");
    std::lock_guard<std::mutex> lock(e->mu);
    cm::DeviceGuard guard(e->device);
    cm_similar_state *s = nullptr;
    if (int rc = similar_state(e, &s)) return rc;
    cudaStream_t st = e->stream;
    const int dim4 = e->dim / 4;
    const int grid = 148 * 8;
    if (!s->norms_ready) {                                           // gensim fill_norms(), once
        cm_row_norm_kernel<<<grid, 256, 0, st>>>(e->d_table, e->n_rows, dim4, s->d_norms);
        cm::count_launch();
        s->norms_ready = true;
    }
    cm_query_kernel<<<1, 32, 0, st>>>(e->d_table, e->dim, row, s->d_norms, s->d_mean);
    cm::count_launch();
    switch ((dim4 + 31) / 32) {
#define CM_CASE(V) case V: cm_dist_kernel<V><<<grid, 256, 0, st>>>(e->d_table, e->n_rows, dim4, s->d_mean, s->d_norms, s->d_dists); break;
        CM_CASE(1) CM_CASE(2) CM_CASE(3) CM_CASE(4) CM_CASE(5) CM_CASE(6) CM_CASE(7)
        default: cm_dist_kernel<8><<<grid, 256, 0, st>>>(e->d_table, e->n_rows, dim4, s->d_mean, s->d_norms, s->d_dists); break;
#undef CM_CASE
    }
    cm::count_launch();
    CM_CUDA(cudaMemsetAsync(s->d_sel, 0, sizeof(SelectState), st));
    const unsigned kk = (unsigned)((int64_t)k < e->n_rows ? k : e->n_rows);
    for (int pass = 0; pass < 4; ++pass) {
        cm_hist_kernel<<<148 * 4, 256, 0, st>>>(s->d_dists, e->n_rows, s->d_sel, pass, kk);
        cm::count_launch();
    }
    cm_collect_kernel<<<148 * 4, 256, 0, st>>>(s->d_dists, e->n_rows, s->d_sel, kk, s->d_cand_key, s->d_cand_row);
    cm::count_launch();
    cm_sort_kernel<<<1, 1024, 0, st>>>(s->d_sel, s->d_cand_key, s->d_cand_row, s->d_dists, kk, s->d_rows_out,
                                       s->d_sims_out, s->d_n_out);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    CM_CUDA(cudaMemcpyAsync(rows_out, s->d_rows_out, kk * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CM_CUDA(cudaMemcpyAsync(sims_out, s->d_sims_out, kk * sizeof(float), cudaMemcpyDeviceToHost, st));
    CM_CUDA(cudaMemcpyAsync(n_out, s->d_n_out, sizeof(int), cudaMemcpyDeviceToHost, st));
    CM_CUDA(cudaStreamSynchronize(st));
    return CM_OK;
}

// The table changed (upload / synthetic fill): the cached row norms are stale.
namespace cm {
void similar_invalidate(cm_engine *e) {
    if (e->similar) e->similar->norms_ready = false;
}
}  // namespace cm
