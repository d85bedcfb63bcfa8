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
//   cm_query_kernel      mean (one warp)
//   cm_scan_kernel       dists[r] = dot(row r, mean) / norms[r], and counts every distance into
//                        a 16-bit-prefix histogram on the way
//   cm_pick_kernel       finds the bin of the k-th largest distance from the histograms, collects
//                        that bin and everything above it; the last CTA sorts (dist desc, row asc)
// A query whose k-th distance shares its bin with thousands of others (duplicated rows) is
// finished by the exact radix select instead: cm_hist_kernel x4 (8 bits per pass over the 12 MB
// of dists, each pass re-deriving the previous one's prefix on the device), cm_collect_kernel,
// cm_sort_kernel.
//
// NaN distances (NaN rows, zero norms) rank last, as numpy's argsort places them.

#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>

#include "cm_common.cuh"
#include "cm_internal.h"
#include "cm_similar_state.h"

namespace {

using cm::fail;

constexpr int kMaxTopK = cm::kSimilarMaxTopK;
constexpr int kRowsPerWarp = 4;

__device__ __forceinline__ uint32_t key_of(float d) { return cm::similar_key_of(d); }

__global__ void __launch_bounds__(256)
cm_row_norm_kernel(const float *__restrict__ table, int64_t n_rows, int dim4, float *__restrict__ norms) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const float4 *t4 = reinterpret_cast<const float4 *>(table);
    for (int64_t r = warp; r < n_rows; r += n_warps) {
        float s = 0.f;
        for (int j = lane; j < dim4; j += 32) {
            const float4 v = cm::ld_stream_seq_f4(t4 + r * dim4 + j);
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

// ---- fast path: scan and select in three launches --------------------------------------
// The scan already holds every distance in a register once, so it also counts them: a
// 65536-bin histogram of the top 16 key bits (one fire-and-forget global add per row; the bins
// around zero are the busiest and still see only one add per ~60 ns) and, through shared
// memory, the 256-bin histogram of the top 8 bits.  cm_pick_kernel then finds the 16-bit bin
// holding the k-th largest key from 2 x 256 counters, takes everything above it plus that one
// bin's population (a few dozen rows: the bin is 1/128 of its value wide), and the last CTA to
// finish sorts them.  If that bin is too crowded to sort in one CTA (thousands of equal
// distances: duplicated rows) the call falls back to the 4-pass radix select below.
constexpr int kBucketCap = 3072;                       // + kMaxTopK = 4096 sortable entries

struct PickState {
    unsigned hist8[256];
    unsigned n_above, n_bucket, ticket, pad;
    unsigned hist16[65536];
};

// dist = dot / norm for one row: store it and count its key
__device__ __forceinline__ void emit_dist(int64_t row, float dot, float norm, float *__restrict__ dists,
                                          unsigned *__restrict__ hist16, unsigned *sh_hist8) {
    const float d = dot / norm;
    dists[row] = d;
    const uint32_t key = key_of(d);
    atomicAdd(&hist16[key >> 16], 1u);
    atomicAdd(&sh_hist8[key >> 24], 1u);
}

// The scan: dists[r] = dot(row r, mean) / norms[r] for every row, one row per warp-wide load
// (lane + 32 v), kRowsPerWarp consecutive rows per warp per round, rounds interleaved across the
// whole grid.  Loads and FMAs are written row by row on purpose: the compiler then requests a
// row, reduces it while the next rows' requests go out, and never holds more than a few
// registers per row.  Measured alternatives (B200, 3M x 300, whole query):
// all rows of the round buffered first 685-720 us; 16 lanes per row with two register buffers
// in flight 702 us; one row per round at full occupancy 750-764 us; contiguous slices per CTA
// handed out dynamically 654 us; this form 620-635 us.
template <int VPL, int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS)
cm_scan_kernel(const float *__restrict__ table, int64_t n_rows, int dim4, const float *__restrict__ mean,
               const float *__restrict__ norms, float *__restrict__ dists, PickState *__restrict__ ps) {
    __shared__ unsigned sh_hist8[256];
    sh_hist8[threadIdx.x] = 0;
    __syncthreads();
    cm::pdl_launch_dependents();
    cm::pdl_wait();                                     // the query kernel wrote `mean`
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
        float p[kRowsPerWarp];
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
            const int64_t r = r0 + i;
            float s = 0.f;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                const int j = lane + 32 * v;
                if (r < n_rows && j < dim4) {
                    const float4 x = cm::ld_stream_seq_f4(t4 + r * dim4 + j);
                    s = fmaf(x.x, q[v].x, s); s = fmaf(x.y, q[v].y, s);
                    s = fmaf(x.z, q[v].z, s); s = fmaf(x.w, q[v].w, s);
                }
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
            emit_dist(r0 + lane, mine, norms[r0 + lane], dists, ps->hist16, sh_hist8);
        }
    }
    __syncthreads();
    if (sh_hist8[threadIdx.x]) atomicAdd(&ps->hist8[threadIdx.x], sh_hist8[threadIdx.x]);
}

// The scan for 300-d tables, laid out like the scoring kernel's gather (rows of 65..80 float4): 16 lanes
// x 5 vectors per row, two rows per load instruction, one half-warp walks 16 consecutive rows
// keeping one partial dot product per row, and a 15-shuffle transpose-reduce at the end leaves
// row r0 + sub's distance in lane sub.  Many short CTAs (128 rows each, 7 per SM) instead of a
// persistent grid.  Each row is staged in an explicit register buffer: written with the load
// inside the FMA loop, ptxas leaves every load alone between its four FFMAs (one load in flight
// per lane, 948 us per query); staged, it keeps about 13 loads in flight per lane.
template <int VPL>
__global__ void __launch_bounds__(128, 7)
cm_scan16_kernel(const float *__restrict__ table, int64_t n_rows, int dim4, const float *__restrict__ mean,
                 const float *__restrict__ norms, float *__restrict__ dists, PickState *__restrict__ ps) {
    constexpr int LPG = 16, ROWS = 16;
    __shared__ unsigned sh_hist8[256];
    sh_hist8[threadIdx.x] = 0;
    sh_hist8[threadIdx.x + 128] = 0;
    __syncthreads();
    cm::pdl_launch_dependents();
    cm::pdl_wait();                                     // the query kernel wrote `mean`
    const int lane = threadIdx.x & 31;
    const int sub = lane & (LPG - 1);
    const int64_t half = (int64_t)blockIdx.x * (128 / LPG) + (threadIdx.x / LPG);
    const int64_t r0 = half * ROWS;
    const float4 *t4 = reinterpret_cast<const float4 *>(table);
    if (r0 - (int64_t)(lane / LPG) * ROWS < n_rows) {   // warp-uniform: the warp's first half-warp has rows
        float4 q[VPL];
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const int j = sub + LPG * v;
            q[v] = j < dim4 ? reinterpret_cast<const float4 *>(mean)[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const int64_t my_row = r0 + sub;
        const float nrm = my_row < n_rows ? norms[my_row] : 1.f;
        float p[ROWS];
#pragma unroll
        for (int k = 0; k < ROWS; ++k) {
            const int64_t r = r0 + k;
            float4 buf[VPL];                            // the whole row is requested before any of it is used
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                const int j = sub + LPG * v;
                buf[v] = (r < n_rows && j < dim4) ? cm::ld_stream_seq_f4(t4 + r * dim4 + j)
                                                  : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            float s = 0.f;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                s = fmaf(buf[v].x, q[v].x, s); s = fmaf(buf[v].y, q[v].y, s);
                s = fmaf(buf[v].z, q[v].z, s); s = fmaf(buf[v].w, q[v].w, s);
            }
            p[k] = s;
        }
        // transpose-reduce: after the stage with offset o a lane keeps the rows whose bit o matches its own
#pragma unroll
        for (int o = ROWS / 2; o > 0; o >>= 1) {
            const bool upper = (sub & o) != 0;
#pragma unroll
            for (int j = 0; j < o; ++j) {
                const float send = upper ? p[j] : p[j + o];
                const float keep = upper ? p[j + o] : p[j];
                p[j] = keep + __shfl_xor_sync(cm::kFullMask, send, o);
            }
        }
        if (my_row < n_rows) emit_dist(my_row, p[0], nrm, dists, ps->hist16, sh_hist8);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += 128)
        if (sh_hist8[i]) atomicAdd(&ps->hist8[i], sh_hist8[i]);
}

// Which of 256 bins, scanned from bin 255 downwards, holds the k-th entry; *above = entries in
// the bins before it.  Called by one full warp; every lane returns the same answer.
__device__ int find_bin_from_top(const unsigned *h, unsigned k, unsigned *above) {
    const int lane = threadIdx.x & 31;
    const int first = 255 - 8 * lane;                   // this lane's bins: first, first-1, ..., first-7
    unsigned mine = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) mine += h[first - i];
    unsigned incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(cm::kFullMask, incl, o);
        if (lane >= o) incl += up;
    }
    const unsigned reached = __ballot_sync(cm::kFullMask, incl >= k);
    const int owner = reached ? __ffs(reached) - 1 : 31;
    unsigned acc = incl - mine;
    int bin = first - 7;
    for (int i = 0; i < 8; ++i) {
        const unsigned c = h[first - i];
        if (acc + c >= k) { bin = first - i; break; }
        if (i < 7) acc += c;
    }
    *above = __shfl_sync(cm::kFullMask, acc, owner);
    return __shfl_sync(cm::kFullMask, bin, owner);
}

// Collect everything in or above the 16-bit bin of the k-th largest key; the last CTA to finish
// sorts the candidates by (distance desc, row asc), writes the first k and clears the counters
// for the next query.  *n_out = k, or -1 when the bin holds more than bucket_cap (<= kBucketCap) rows.
__global__ void __launch_bounds__(256)
cm_pick_kernel(const float *__restrict__ dists, int64_t n, PickState *ps, unsigned k, unsigned bucket_cap,
               unsigned long long *cand, int32_t *__restrict__ rows_out, float *__restrict__ sims_out,
               int *__restrict__ n_out) {
    __shared__ unsigned long long s[kMaxTopK + kBucketCap];
    __shared__ unsigned sh[256];
    __shared__ unsigned s_bin, s_above, s_pop;
    __shared__ bool s_last;
    const unsigned t = threadIdx.x;
    cm::pdl_wait();                                     // the scan has finished and its counts are visible
    sh[t] = __ldcg(&ps->hist8[t]);
    __syncthreads();
    if (t < 32) {
        unsigned above;
        const int c = find_bin_from_top(sh, k, &above);
        if (t == 0) { s_bin = (unsigned)c; s_above = above; }
    }
    __syncthreads();
    const unsigned coarse = s_bin, above_coarse = s_above;
    __syncthreads();
    sh[t] = __ldcg(&ps->hist16[coarse * 256 + t]);
    __syncthreads();
    if (t < 32) {
        unsigned above;
        const int f = find_bin_from_top(sh, k - above_coarse, &above);
        if (t == 0) { s_bin = coarse * 256 + (unsigned)f; s_above = above_coarse + above; s_pop = sh[f]; }
    }
    __syncthreads();
    const uint32_t bin = s_bin;                         // top 16 key bits of the k-th largest distance
    const unsigned total = s_above + s_pop;             // rows in or above that bin: >= k
    const bool crowded = s_pop > bucket_cap;
    if (!crowded) {
        const int64_t n4 = n >> 2;
        const float4 *d4 = reinterpret_cast<const float4 *>(dists);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + t; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
            const float4 v = d4[i];
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint32_t key = key_of(e[c]);
                if ((key >> 16) >= bin) {
                    const unsigned slot = atomicAdd(&ps->n_above, 1u);
                    cand[slot] = ((unsigned long long)(~key) << 32) | (uint32_t)(4 * i + c);
                }
            }
        }
        if (blockIdx.x == 0 && t < (unsigned)(n & 3)) {
            const int64_t i = (n4 << 2) + t;
            const uint32_t key = key_of(dists[i]);
            if ((key >> 16) >= bin) {
                const unsigned slot = atomicAdd(&ps->n_above, 1u);
                cand[slot] = ((unsigned long long)(~key) << 32) | (uint32_t)i;
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (t == 0) s_last = atomicAdd(&ps->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (crowded) {
        if (t == 0) *n_out = -1;
    } else {
        unsigned size = 64;
        while (size < total) size <<= 1;
        for (unsigned i = t; i < size; i += blockDim.x) s[i] = i < total ? __ldcg(&cand[i]) : ~0ull;
        __syncthreads();
        for (unsigned span = 2; span <= size; span <<= 1) {         // bitonic, ascending = best first
            for (unsigned stride = span >> 1; stride > 0; stride >>= 1) {
                for (unsigned c = t; c < size / 2; c += blockDim.x) {
                    const unsigned lo = 2 * c - (c & (stride - 1));      // c-th pair of this stride
                    const unsigned hi = lo + stride;
                    const bool up = (lo & span) == 0;
                    const unsigned long long a = s[lo], b = s[hi];
                    if ((a > b) == up) { s[lo] = b; s[hi] = a; }
                }
                __syncthreads();
            }
        }
        for (unsigned i = t; i < k; i += blockDim.x) {
            const int32_t r = (int32_t)(s[i] & 0xffffffffull);
            rows_out[i] = r;
            sims_out[i] = dists[r];
        }
        if (t == 0) *n_out = (int)k;
    }
    // leave the counters clear for the next query
    unsigned *w = reinterpret_cast<unsigned *>(ps);
    for (unsigned i = t; i < sizeof(PickState) / sizeof(unsigned); i += blockDim.x) w[i] = 0u;
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

namespace cm {

void similar_state_free(cm_engine *e) {
    cm_similar_state *s = e->similar;
    if (s == nullptr) return;
    void *ptrs[] = {s->d_norms, s->d_mean, s->d_dists, s->d_sel, s->d_cand_key, s->d_cand_row,
                    s->d_rows_out, s->d_sims_out, s->d_n_out, s->d_pick, s->d_cand};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (s->h_result) cudaFreeHost(s->h_result);
    similar_batch_free(s);
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
        if (err == cudaSuccess) err = cudaMalloc(&s->d_pick, sizeof(PickState));
        if (err == cudaSuccess) err = cudaMalloc(&s->d_cand, (kMaxTopK + kBucketCap) * sizeof(unsigned long long));
        if (err == cudaSuccess) err = cudaMemset(s->d_pick, 0, sizeof(PickState));
        if (err == cudaSuccess) err = cudaHostAlloc(&s->h_result, kMaxTopK * 8 + 16, cudaHostAllocDefault);
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

static bool generic_scan_forced() {
    // Test hook: CM_SIMILAR_GENERIC_SCAN=1 forces the generic kernel so the two can be compared.
    static const bool v = [] { const char *c = getenv("CM_SIMILAR_GENERIC_SCAN"); return c != nullptr && *c == '1'; }();
    return v;
}

// Called by cm_engine_create right after the table is allocated.  Where the 12 MB of distances and norms land matters:
// allocated late in a process's life (after pipes, pinned blocks, other libraries' pools) the same scan has been measured
// at 650-870 us per query instead of 575 (tools/probe_similar_alloc_order.py, bench.py before this call existed).
// Failure is not an error here: cm_most_similar allocates on first use as before.
void cm::similar_reserve(cm_engine *e) {
    cm_similar_state *s = nullptr;
    if (similar_state(e, &s) != CM_OK) cudaGetLastError();
}

bool cm::similar_uses_scan16(int dim4) { return !generic_scan_forced() && dim4 > 64 && dim4 <= 80; }

int cm::similar_prepare(cm_engine *e, cm_similar_state **out) {
    cm_similar_state *s = nullptr;
    if (int rc = similar_state(e, &s)) return rc;
    if (!s->norms_ready) {                                           // gensim fill_norms(), once
        cm_row_norm_kernel<<<148 * 8, 256, 0, e->stream>>>(e->d_table, e->n_rows, e->dim / 4, s->d_norms);
        cm::count_launch();
        s->norms_ready = true;
    }
    *out = s;
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
    return cm::most_similar_locked(e, row, k, rows_out, sims_out, n_out);
}

int cm::most_similar_locked(cm_engine *e, int32_t row, int k, int32_t *rows_out, float *sims_out, int *n_out) {
    cm_similar_state *s = nullptr;
    if (int rc = cm::similar_prepare(e, &s)) return rc;
    cudaStream_t st = e->stream;
    const int dim4 = e->dim / 4;
    const unsigned kk = (unsigned)((int64_t)k < e->n_rows ? k : e->n_rows);
    cm_query_kernel<<<1, 32, 0, st>>>(e->d_table, e->dim, row, s->d_norms, s->d_mean);
    cm::count_launch();
    // scan + pick ride on programmatic dependent launches: each starts its prologue while the
    // previous kernel drains and waits (griddepcontrol.wait) before it touches that kernel's output
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // test hook: a smaller cap (0 = always) sends queries down the radix-select fallback
    static const unsigned bucket_cap = [] {
        const char *v = getenv("CM_SIMILAR_BUCKET_CAP");
        const long c = v != nullptr && *v != 0 ? atol(v) : (long)kBucketCap;
        return (unsigned)(c < 0 ? 0 : (c > kBucketCap ? kBucketCap : c));
    }();
    const float *table = e->d_table, *mean = s->d_mean, *norms = s->d_norms;
    // 300-d tables (65..80 float4 per row) take the 16-lane scan: 580 against 594 us per query at
    // 3M rows, A/B on one box, two runs each (profiles/r2_similar_ab.txt).
    if (cm::similar_uses_scan16(dim4)) {
        cfg.blockDim = dim3(128);
        cfg.gridDim = dim3((unsigned)((e->n_rows + 127) / 128));
        cudaLaunchKernelEx(&cfg, cm_scan16_kernel<5>, table, e->n_rows, dim4, mean, norms, s->d_dists, (PickState *)s->d_pick);
    } else {
    cfg.blockDim = dim3(256);
#define CM_CASE(V, MB)                                                                             \
    case V:                                                                                        \
        cfg.gridDim = dim3(148 * MB);                                                              \
        cudaLaunchKernelEx(&cfg, cm_scan_kernel<V, MB>, table, e->n_rows, dim4, mean, norms, s->d_dists, (PickState *)s->d_pick); \
        break;
    switch ((dim4 + 31) / 32) {
        CM_CASE(1, 3) CM_CASE(2, 3) CM_CASE(3, 3) CM_CASE(4, 2) CM_CASE(5, 2) CM_CASE(6, 2) CM_CASE(7, 2)
        default:
            cfg.gridDim = dim3(148 * 2);
            cudaLaunchKernelEx(&cfg, cm_scan_kernel<8, 2>, table, e->n_rows, dim4, mean, norms, s->d_dists, (PickState *)s->d_pick);
            break;
    }
#undef CM_CASE
    }
    cm::count_launch();
    cfg.gridDim = dim3(148 * 4);
    cfg.blockDim = dim3(256);
    // the pick kernel writes the result straight into page-locked host memory: no copies to enqueue
    int32_t *h_rows = reinterpret_cast<int32_t *>(s->h_result);
    float *h_sims = reinterpret_cast<float *>(s->h_result + kMaxTopK * 4);
    int *h_n = reinterpret_cast<int *>(s->h_result + kMaxTopK * 8);
    cudaLaunchKernelEx(&cfg, cm_pick_kernel, (const float *)s->d_dists, e->n_rows, (PickState *)s->d_pick,
                       kk, bucket_cap, s->d_cand, h_rows, h_sims, h_n);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    CM_CUDA(cudaStreamSynchronize(st));
    if (*h_n >= 0) {
        memcpy(rows_out, h_rows, kk * sizeof(int32_t));
        memcpy(sims_out, h_sims, kk * sizeof(float));
        *n_out = *h_n;
        return CM_OK;
    }
    // The k-th distance sits in a 16-bit bin with thousands of rows (duplicates): exact 4-pass
    // radix select over the distances the scan left behind.
    CM_CUDA(cudaMemsetAsync(s->d_sel, 0, sizeof(SelectState), st));
    for (int pass = 0; pass < 4; ++pass) {
        cm_hist_kernel<<<148 * 4, 256, 0, st>>>(s->d_dists, e->n_rows, (SelectState *)s->d_sel, pass, kk);
        cm::count_launch();
    }
    cm_collect_kernel<<<148 * 4, 256, 0, st>>>(s->d_dists, e->n_rows, (SelectState *)s->d_sel, kk, s->d_cand_key, s->d_cand_row);
    cm::count_launch();
    cm_sort_kernel<<<1, 1024, 0, st>>>((const SelectState *)s->d_sel, s->d_cand_key, s->d_cand_row, s->d_dists, kk, s->d_rows_out,
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
