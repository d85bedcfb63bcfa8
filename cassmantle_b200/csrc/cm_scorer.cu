// cassmantle_b200 -- guess-scoring kernels and their C ABI (include/cassmantle_b200.h).
//
// Hot path replaced: Backend.compute_score / compute_scores, src/backend.py:303-317, i.e.
// gensim KeyedVectors.similarity (src/backend.py:307) = dot(unitvec(row_a), unitvec(row_b)).
// Everything here is an HBM-bound gather: one 1200-byte table row per guessed word, summed
// in fp32 in word order (numpy mean(axis=0)), normalised through an fp64 norm (gensim binds
// the double-precision BLAS nrm2/scal), dotted in fp32 with the answer's unit vector,
// floored at min_score.  There is no dense contraction on this path, so no tensor-core work.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a (see cassmantle_b200/build.py).

#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "cm_common.cuh"
#include "cm_internal.h"

namespace cm {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

}  // namespace cm

namespace {

using cm::DeviceGuard;
using cm::fail;

constexpr int kWarpsPerCta = 8;

// --------------------------------------------------------------------------------------
// Synthetic table fill -- bit-identical to cassmantle_b200/synth.py: table_rows().
// --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
cm_synth_fill_kernel(float *__restrict__ table, int64_t n_rows, int dim, uint32_t seed) {
    const uint32_t k_noise = cm::mix32(seed ^ 0x9e3779b9u);
    const uint32_t k_clu = cm::mix32(seed ^ 0x85ebca6bu);
    const uint32_t k_cen = cm::mix32(seed ^ 0xc2b2ae35u);
    const float inv_unit = (float)(1.0 / 3620.0);
    const int dim4 = dim >> 2;
    const int64_t total = n_rows * dim4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t row = (uint32_t)(i / dim4);
        const uint32_t c0 = (uint32_t)(i % dim4) * 4u;
        const uint32_t hs = cm::mix32(row * 0x9e3779b1u + seed);
        const float u = __fmul_rn(__uint2float_rn(hs >> 8), 5.9604644775390625e-08f);  // 2^-24
        const float coef = __fmul_rn(__fadd_rn(0.5f, __fmul_rn(4.5f, u)), inv_unit);
        const uint32_t clu = cm::mix32(row ^ k_clu) & 1023u;
        float v[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint32_t col = c0 + t;
            const int s = cm::bytesum(cm::mix32((row * (uint32_t)dim + col) ^ k_noise)) +
                          cm::bytesum(cm::mix32((clu * (uint32_t)dim + col) ^ k_cen));
            v[t] = __fmul_rn(__int2float_rn(s), coef);
        }
        reinterpret_cast<float4 *>(table)[i] = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// --------------------------------------------------------------------------------------
// unitvec() in registers.  A group of LPG lanes (LPG = 32, 16 or 8) holds one vector: lane
// `sub` of the group owns float4 elements sub, sub+LPG, ...  gensim matutils.unitvec:
// veclen = dnrm2(v); v * (1/veclen) in fp64, cast back to fp32; a zero vector is unchanged.
// --------------------------------------------------------------------------------------
template <int LPG, typename T>
__device__ __forceinline__ T group_sum(T v) {
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) v += __shfl_xor_sync(cm::kFullMask, v, o);
    return v;
}

// A lane's share of the sum of squares, in fp64 and in element order.
template <int VPL>
__device__ __forceinline__ double lane_sumsq(const float4 (&m)[VPL]) {
    double ss = 0.0;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        ss = fma((double)m[v].x, (double)m[v].x, ss);
        ss = fma((double)m[v].y, (double)m[v].y, ss);
        ss = fma((double)m[v].z, (double)m[v].z, ss);
        ss = fma((double)m[v].w, (double)m[v].w, ss);
    }
    return ss;
}

template <int VPL, int LPG>
__device__ __forceinline__ void group_scale(float4 (&m)[VPL], double ss);

template <int VPL, int LPG>
__device__ __forceinline__ void group_unitvec(float4 (&m)[VPL]) {
    group_scale<VPL, LPG>(m, lane_sumsq<VPL>(m));
}

// ss: the lane's share (lane_sumsq); the group's lanes add theirs and scale the vector by 1/norm.
template <int VPL, int LPG>
__device__ __forceinline__ void group_scale(float4 (&m)[VPL], double ss) {
    ss = group_sum<LPG>(ss);
    if (ss > 0.0) {
        // (float)((double)x * inv) without going through the conversion pipe for every element:
        // inv = ih + il to ~2^-48, so fma(x, ih, x*il) rounds the same exact product once.
        const double inv = 1.0 / sqrt(ss);
        const float ih = (float)inv;
        if (ih > 1e-30f && ih < 1e30f) {
            const float il = (float)(inv - (double)ih);
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                m[v].x = fmaf(m[v].x, ih, __fmul_rn(m[v].x, il));
                m[v].y = fmaf(m[v].y, ih, __fmul_rn(m[v].y, il));
                m[v].z = fmaf(m[v].z, ih, __fmul_rn(m[v].z, il));
                m[v].w = fmaf(m[v].w, ih, __fmul_rn(m[v].w, il));
            }
        } else {                                   // norms near the ends of the fp32 range
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                float4 t = m[v];
                // launder: keeps the compiler from holding 4*VPL doubles live for this cold path
                asm volatile("" : "+f"(t.x), "+f"(t.y), "+f"(t.z), "+f"(t.w));
                m[v].x = (float)((double)t.x * inv);
                m[v].y = (float)((double)t.y * inv);
                m[v].z = (float)((double)t.z * inv);
                m[v].w = (float)((double)t.w * inv);
            }
        }
    }
}

// fp32 dot of a lane's share of the (unit) guess with its share of the answer's unit vector: the one
// definition every scoring kernel uses, so that they agree bit for bit.
struct DotAcc {
    float p = 0.f;
    __device__ __forceinline__ void add(const float4 &a, const float4 &w) {
        p = fmaf(a.x, w.x, p); p = fmaf(a.y, w.y, p); p = fmaf(a.z, w.z, p); p = fmaf(a.w, w.w, p);
    }
    __device__ __forceinline__ float total() const { return p; }
};

// Measurement aid (cm_engine_flush_l2): a pure READ of the first `n16` 16-byte pieces of the table, which
// leaves L2 full of clean lines (a write flush leaves dirty ones for the next kernels to evict).
__global__ void __launch_bounds__(256)
cm_flush_read_kernel(const uint4 *__restrict__ p, size_t n16, unsigned *__restrict__ sink) {
    unsigned acc = 0;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) {
        const uint4 v = __ldcg(p + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x5eed5eedu && sink != nullptr) *sink = acc;      // keeps the loads; practically never taken
}

// One warp per registered answer: ans_unit[a] = unitvec(table[answer_rows[a]]).
template <int VPL>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
cm_answer_unit_kernel(const float *__restrict__ table, int64_t n_rows, int dim4,
                      const int32_t *__restrict__ answer_rows, int n_answers,
                      float *__restrict__ ans_unit) {
    const int lane = threadIdx.x & 31;
    const int a = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (a >= n_answers) return;
    const int32_t r = answer_rows[a];
    const bool in_vocab = r >= 0 && (int64_t)r < n_rows;
    float4 m[VPL];
    const float4 *src = reinterpret_cast<const float4 *>(table) + (size_t)(in_vocab ? r : 0) * dim4;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        const int j = lane + 32 * v;
        m[v] = (in_vocab && j < dim4) ? __ldg(src + j) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    group_unitvec<VPL, 32>(m);
    float4 *dst = reinterpret_cast<float4 *>(ans_unit) + (size_t)a * dim4;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        const int j = lane + 32 * v;
        if (j < dim4) dst[j] = m[v];
    }
}

// --------------------------------------------------------------------------------------
// The scoring kernel.  A group of LPG lanes owns one guess, so a warp scores 32/LPG guesses
// side by side (LPG = 32 for phrases, 8 for the reference's one-word guesses, where the
// per-guess latency chain, not bandwidth, is what has to be hidden).
//   phase 0  every independent metadata load at once: exact flag, length, answer slot,
//            the first LPG word indices (speculatively, bounded by n_slots)
//   phase 1  gather: R rows requested per round with 16-byte loads, then added in word
//            order in fp32 (numpy mean(axis=0) is a running fp32 sum)
//   phase 2  / N, unitvec through an fp64 norm, fp32 dot with the answer's unit vector,
//            floor at min_score, optional hundredths of a percent
// --------------------------------------------------------------------------------------
// SYS = true is the resident kernel (below): the batch is a block of per-guess records that the CTA
// has staged in shared memory, and the answers -- rewritten by cm_set_answers while the kernel is
// resident -- are read with L2-only loads, issued before the gather (a micro-batch is a latency
// chain, not bandwidth).  SYS = false: the launched kernels, packed arrays.
// IDX24: guess_idx is really an array of 3-byte little-endian row numbers (0xFFFFFF = not in the
// vocabulary), the packed form the pipelined host path can take: a quarter fewer input bytes per
// guess across PCIe, which is what bounds eight GPUs behind one host.
template <bool IDX24>
__device__ __forceinline__ int32_t ld_idx(const int32_t *base, size_t k) {
    if constexpr (IDX24) {
        const uint8_t *p = reinterpret_cast<const uint8_t *>(base) + 3 * k;
        const uint32_t v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
        return v == 0xffffffu ? -1 : (int32_t)v;
    } else {
        return base[k];
    }
}

// Index c + sub of the guess whose first index is element g0, for every lane of an LPG-lane group
// (-1 beyond n_slots).  IDX24 with n_slots a multiple of 4: the chunk's 3 * LPG bytes start on a word
// boundary, so the group fetches them as LPG * 3 / 4 aligned words with ONE load instruction and every
// lane cuts its three bytes out of two of them -- when the batch sits in page-locked host memory a
// byte load is a PCIe request of its own.
template <int LPG, bool IDX24>
__device__ __forceinline__ int32_t ld_idx_chunk(const int32_t *base, size_t g0, int c, int n_slots, int sub, int grp_base) {
    if constexpr (IDX24) {
        if ((n_slots & 3) == 0) {                            // uniform over the launch
            const uint32_t *w32 = reinterpret_cast<const uint32_t *>(reinterpret_cast<const uint8_t *>(base) + 3 * (g0 + c));
            const int n_idx = min(LPG, n_slots - c);         // indices of this chunk (a multiple of 4)
            const uint32_t word = sub < (3 * n_idx) / 4 ? w32[sub] : 0u;
            const int w = (3 * sub) >> 2, sh = ((3 * sub) & 3) * 8;
            const uint32_t lo = __shfl_sync(cm::kFullMask, word, grp_base + (w & (LPG - 1)));
            const uint32_t hi = __shfl_sync(cm::kFullMask, word, grp_base + ((w + 1) & (LPG - 1)));
            const uint32_t v = __funnelshift_r(lo, hi, sh) & 0xffffffu;
            return (sub < n_idx && v != 0xffffffu) ? (int32_t)v : -1;
        }
    }
    return (c + sub < n_slots) ? ld_idx<IDX24>(base, g0 + c + sub) : -1;
}

template <int VPL, int LPG, int R, bool STREAM, int THREADS, bool SINGLE, bool SYS = false, bool IDX24 = false>
__device__ __forceinline__ void
score_groups(int warp, const float *__restrict__ table, int64_t n_rows, int dim4,
             const int32_t *__restrict__ guess_idx, const int32_t *__restrict__ guess_len,
             int n_guesses, int n_slots,
             const float *__restrict__ ans_unit, const int32_t *__restrict__ ans_rows,
             int n_answers, const int32_t *__restrict__ answer_slot,
             const uint8_t *__restrict__ exact, float min_score,
             float *__restrict__ scores, int32_t *__restrict__ centi, uint32_t flags, int rec_stride = 0) {
    constexpr int G = 32 / LPG;
    // Programmatic dependent launch: let the next launch on the stream start scheduling now,
    // and unless the caller vouched for the inputs (CM_SCORE_OVERLAP_PREVIOUS) wait for the
    // previous kernel on the stream before reading anything.
    cm::pdl_launch_dependents();
    const bool overlap = (flags & CM_SCORE_OVERLAP_PREVIOUS) != 0u;
    if (!overlap) cm::pdl_wait();
    const int lane = threadIdx.x & 31;
    const int sub = lane & (LPG - 1);
    const int grp_base = lane & ~(LPG - 1);
    const int b_raw = warp * G + lane / LPG;
    if (warp * G >= n_guesses) return;                      // whole warp out of range
    const bool have = b_raw < n_guesses;
    const int b = have ? b_raw : n_guesses - 1;             // keep every lane's loads in bounds

    // ---- phase 0: independent loads, one DRAM round trip ---------------------------------
    // SYS: guess b is one record of rec_stride ints, {len, slot, exact, index 0, index 1, ...}, which
    // guess_len points at (guess_idx points at its fourth int); shared memory, broadcast reads.
    const size_t g0 = (size_t)b * (SYS ? rec_stride : n_slots);       // first index of guess b
    bool is_exact;
    int n, slot;
    if constexpr (SYS) {
        const int32_t *rec = guess_len + (size_t)b * rec_stride;
        n = rec[0];
        slot = rec[1];
        is_exact = rec[2] != 0;
    } else {
        is_exact = (exact != nullptr) && (exact[b] != 0);
        n = guess_len[b];
        slot = answer_slot[b];
    }
    // SINGLE (n_slots == 1, the reference's own shape): every lane of the group reads the one index
    int32_t mine = SINGLE ? ld_idx<IDX24>(guess_idx, g0) : ld_idx_chunk<LPG, IDX24>(guess_idx, g0, 0, n_slots, sub, grp_base);
    n = n < 0 ? 0 : (n > n_slots ? n_slots : n);
    const bool slot_ok = slot >= 0 && slot < n_answers;

    // KeyError scan: a used slot outside the table, an empty guess, or an unknown answer.
    bool bad;
    if constexpr (SINGLE) {
        bad = (n > 0) && (mine < 0 || (int64_t)mine >= n_rows);
    } else {
        bad = (sub < n) && (mine < 0 || (int64_t)mine >= n_rows);
        for (int c = LPG + sub; c < n; c += LPG) {
            const int32_t r = ld_idx<IDX24>(guess_idx, g0 + c);
            bad |= (r < 0) || ((int64_t)r >= n_rows);
        }
        const unsigned gmask = (LPG == 32) ? cm::kFullMask : (((1u << LPG) - 1u) << grp_base);
        bad = (__ballot_sync(cm::kFullMask, bad) & gmask) != 0u;
    }
    bad = bad || n == 0 || !slot_ok;
    const bool active = have && !is_exact && !bad;
    const int n_live = active ? n : 0;
    const int n_max = (G == 1) ? n_live : __reduce_max_sync(cm::kFullMask, n_live);

    // SYS: the answer's row and unit vector depend only on the slot -- ask for them before the gather
    int32_t ar_pre = -1;
    float4 w_pre[SYS ? VPL : 1];
    if constexpr (SYS) {
        ar_pre = slot_ok ? __ldcg(ans_rows + slot) : -1;
        const float4 *u = reinterpret_cast<const float4 *>(ans_unit) + (size_t)(slot_ok ? slot : 0) * dim4;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const int j = sub + LPG * v;
            w_pre[v] = j < dim4 ? __ldcg(u + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    float4 acc[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 *table4 = reinterpret_cast<const float4 *>(table);

    // ---- phase 1: gather ------------------------------------------------------------------
    if constexpr (SINGLE) {                                  // one row, straight into acc
        const float4 *src = table4 + (size_t)(n_live > 0 ? mine : 0) * dim4;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const int j = sub + LPG * v;
            if (n_live > 0 && j < dim4) acc[v] = STREAM ? cm::ld_stream_f4(src + j) : __ldg(src + j);
        }
    } else
    for (int c = 0; c < n_max; c += LPG) {
        if (c > 0) {                                         // every lane takes part (shuffles); dead slots become row 0, unused
            mine = ld_idx_chunk<LPG, IDX24>(guess_idx, g0, c, n_slots, sub, grp_base);
            if (c + sub >= n_live) mine = 0;
        }
        const int m = min(LPG, n_max - c);
        for (int k = 0; k < m; k += R) {
            float4 buf[R][VPL];
#pragma unroll
            for (int i = 0; i < R; ++i) {
                const int32_t r = __shfl_sync(cm::kFullMask, mine, grp_base + ((k + i) & (LPG - 1)));
                const bool live = (k + i) < m && (c + k + i) < n_live;
                const float4 *src = table4 + (size_t)(live ? r : 0) * dim4;
#pragma unroll
                for (int v = 0; v < VPL; ++v) {
                    const int j = sub + LPG * v;
                    if (live && j < dim4) buf[i][v] = STREAM ? cm::ld_stream_f4(src + j) : __ldg(src + j);
                    else buf[i][v] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#pragma unroll
            for (int i = 0; i < R; ++i) {
#pragma unroll
                for (int v = 0; v < VPL; ++v) {              // word order, like mean(axis=0)
                    acc[v].x += buf[i][v].x;
                    acc[v].y += buf[i][v].y;
                    acc[v].z += buf[i][v].z;
                    acc[v].w += buf[i][v].w;
                }
            }
        }
    }

    // ---- phase 2: mean, unitvec, dot, floor -----------------------------------------------
    // The answers (ans_rows / ans_unit) and the outputs are the only things an overlapped
    // launch shares with its predecessor: from here on the previous kernel must be complete.
    if (overlap) cm::pdl_wait();
    {
        const int32_t ar = SYS ? ar_pre : (slot_ok ? ans_rows[slot] : -1);     // unknown answer = KeyError too
        bad = bad || ar < 0 || (int64_t)ar >= n_rows;
    }
    if (!SINGLE && n_max > 1) {                              // warp-uniform; x / 1 is exact anyway
        const float fn = (float)(n_live > 1 ? n_live : 1);
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            acc[v].x = __fdiv_rn(acc[v].x, fn);
            acc[v].y = __fdiv_rn(acc[v].y, fn);
            acc[v].z = __fdiv_rn(acc[v].z, fn);
            acc[v].w = __fdiv_rn(acc[v].w, fn);
        }
    }
    float p = 0.f;
    if (n_max > 0) {                                         // warp-uniform
        group_unitvec<VPL, LPG>(acc);
        const float4 *u = reinterpret_cast<const float4 *>(ans_unit) + (size_t)(active ? slot : 0) * dim4;
        DotAcc dot;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const int j = sub + LPG * v;
            if (j < dim4) {
                const float4 w = SYS ? w_pre[SYS ? v : 0] : u[j];   // not .nc: written by the answer kernel, maybe just now
                dot.add(acc[v], w);
            }
        }
        p = group_sum<LPG>(dot.total());
    }
    float result;
    if (is_exact) result = 1.0f;                             // src/backend.py:305
    else if (bad) result = min_score;                        // src/backend.py:308-310
    else result = (p > min_score) ? p : min_score;           // max(min_score, score): NaN -> min
    if (have && sub == 0) {
        scores[b] = result;
        if (centi != nullptr) centi[b] = (int32_t)floor((double)result * 10000.0 + 0.5);
    }
}

template <int VPL, int LPG, int R, bool STREAM, int THREADS, int MIN_BLOCKS, bool SINGLE, bool IDX24>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
cm_score_kernel(const float *__restrict__ table, int64_t n_rows, int dim4,
                const int32_t *__restrict__ guess_idx, const int32_t *__restrict__ guess_len,
                int n_guesses, int n_slots,
                const float *__restrict__ ans_unit, const int32_t *__restrict__ ans_rows,
                int n_answers, const int32_t *__restrict__ answer_slot,
                const uint8_t *__restrict__ exact, float min_score,
                float *__restrict__ scores, int32_t *__restrict__ centi, uint32_t flags,
                int n_ship_ctas, const cm::ShipArgs ship) {
    // Several GPUs (cm_gather.cu): the first `n_ship_ctas` CTAs do not score -- they ship the
    // PREVIOUS step's scores to the other ranks while the rest of the launch gathers its rows.
    // (First, because CTAs are dispatched in index order and the ship should start at once.)
    if ((int)blockIdx.x < n_ship_ctas) {
        cm::pdl_launch_dependents();
        cm::ship_duty(ship, (int)blockIdx.x, n_ship_ctas);
        return;
    }
    score_groups<VPL, LPG, R, STREAM, THREADS, SINGLE, false, IDX24>(
        ((int)blockIdx.x - n_ship_ctas) * (THREADS / 32) + (int)(threadIdx.x >> 5),
        table, n_rows, dim4, guess_idx, guess_len, n_guesses, n_slots, ans_unit, ans_rows, n_answers,
        answer_slot, exact, min_score, scores, centi, flags);
}

// --------------------------------------------------------------------------------------
// One-word guesses, many of them (more than one wave): a persistent, software-pipelined form of
// the same arithmetic.  Each warp walks over groups of 32/LPG guesses; while the rows of the
// current group are in flight it already loads the metadata (index, length, slot, exact flag) of
// its next group, so the per-guess chain is  row latency + math  instead of
// metadata latency + row latency + math  (r1_v3 profile: 58 % of the stall samples were
// long-scoreboard waits at 43 % occupancy).  Bit-identical results to cm_score_kernel<SINGLE>.
// --------------------------------------------------------------------------------------
template <int VPL, int LPG, bool STREAM, int THREADS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
cm_score1_kernel(const float *__restrict__ table, int64_t n_rows, int dim4,
                 const int32_t *__restrict__ guess_idx, const int32_t *__restrict__ guess_len,
                 int n_guesses,
                 const float *__restrict__ ans_unit, const int32_t *__restrict__ ans_rows,
                 int n_answers, const int32_t *__restrict__ answer_slot,
                 const uint8_t *__restrict__ exact, float min_score,
                 float *__restrict__ scores, int32_t *__restrict__ centi, uint32_t flags,
                 int n_ship_ctas, const cm::ShipArgs ship) {
    constexpr int G = 32 / LPG;
    if ((int)blockIdx.x < n_ship_ctas) {
        cm::pdl_launch_dependents();
        cm::ship_duty(ship, (int)blockIdx.x, n_ship_ctas);
        return;
    }
    cm::pdl_launch_dependents();
    const bool overlap = (flags & CM_SCORE_OVERLAP_PREVIOUS) != 0u;
    if (!overlap) cm::pdl_wait();
    bool waited = !overlap;
    const int lane = threadIdx.x & 31;
    const int sub = lane & (LPG - 1);
    const int grp = lane / LPG;
    const int warps_total = ((int)gridDim.x - n_ship_ctas) * (THREADS / 32);
    const float4 *table4 = reinterpret_cast<const float4 *>(table);

    int w = ((int)blockIdx.x - n_ship_ctas) * (THREADS / 32) + (threadIdx.x >> 5);
    // metadata of the current group (loaded one iteration ahead)
    int b = min(w * G + grp, n_guesses - 1);
    int32_t m_row = 0; int m_len = 0, m_slot = 0; uint8_t m_exact = 0;
    if (w * G < n_guesses) {
        m_row = guess_idx[b]; m_len = guess_len[b]; m_slot = answer_slot[b];
        m_exact = exact != nullptr ? exact[b] : (uint8_t)0;
    }
    while (w * G < n_guesses) {
        const bool have = w * G + grp < n_guesses;
        const bool is_exact = m_exact != 0;
        const int slot = m_slot;
        const bool slot_ok = slot >= 0 && slot < n_answers;
        const int n = m_len < 0 ? 0 : (m_len > 1 ? 1 : m_len);
        bool bad = (n > 0 && (m_row < 0 || (int64_t)m_row >= n_rows)) || n == 0 || !slot_ok;
        const bool active = have && !is_exact && !bad;
        // rows of this group
        float4 acc[VPL];
        const float4 *src = table4 + (size_t)(active ? m_row : 0) * dim4;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const int j = sub + LPG * v;
            if (active && j < dim4) acc[v] = STREAM ? cm::ld_stream_f4(src + j) : __ldg(src + j);
            else acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // metadata of the next group, in flight while this one is reduced
        const int b_cur = b;
        const int w_next = w + warps_total;
        if (w_next * G < n_guesses) {
            b = min(w_next * G + grp, n_guesses - 1);
            m_row = guess_idx[b]; m_len = guess_len[b]; m_slot = answer_slot[b];
            m_exact = exact != nullptr ? exact[b] : (uint8_t)0;
        }
        if (!waited) { cm::pdl_wait(); waited = true; }           // answers / outputs: predecessor done
        {
            const int32_t ar = slot_ok ? ans_rows[slot] : -1;
            bad = bad || ar < 0 || (int64_t)ar >= n_rows;
        }
        float p = 0.f;
        if (__any_sync(cm::kFullMask, active)) {
            group_unitvec<VPL, LPG>(acc);
            const float4 *u = reinterpret_cast<const float4 *>(ans_unit) + (size_t)(active ? slot : 0) * dim4;
            DotAcc dot;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
                const int j = sub + LPG * v;
                if (j < dim4) dot.add(acc[v], u[j]);
            }
            p = group_sum<LPG>(dot.total());
        }
        float result;
        if (is_exact) result = 1.0f;
        else if (bad) result = min_score;
        else result = (p > min_score) ? p : min_score;
        if (have && sub == 0) {
            scores[b_cur] = result;
            if (centi != nullptr) centi[b_cur] = (int32_t)floor((double)result * 10000.0 + 0.5);
        }
        w = w_next;
    }
    if (!waited) cm::pdl_wait();
}


// --------------------------------------------------------------------------------------
// Resident kernel for micro-batches (cm_engine_arm_resident).  One blocking host call for a few
// hundred guesses is dominated by the launch and the stream synchronisation around a 4 us
// kernel (configs[1]).  Armed, the engine keeps this kernel resident -- one CTA per SM, spinning at
// low duty -- and a call costs two PCIe crossings and no launch:
//   host   packs the batch into the page-locked mailbox, one record per guess, pre-fills the results
//          with a sentinel, then stores the 64-bit doorbell (sequence number, batch shape)
//   CTA 0  polls doorbell + parameters (one 16-byte read across PCIe) and republishes them in device memory
//   all    spin on that device word; each CTA fetches the records of ITS consecutive guesses once, into
//          shared memory, scores them with the same code as the launched kernels (score_groups) and
//          writes the results back into the mailbox; the host watches the sentinels disappear (plain
//          posted stores, no fence: the host touches nothing of a request before all its results landed)
// Measured on B200: when several warps read the SAME 32-byte sector of host memory the reads are
// served one PCIe round trip (0.8 us) after the other -- 592 warps reading one parameter word took
// 470 us.  Hence one reader for the header, a batch layout of one record per guess in sectors of its
// own (resident_rec_stride) instead of the packed arrays of the launch path, and every sector asked
// for exactly once.
// The kernel leaves by itself after `idle_ns` without a doorbell (so a device-wide synchronisation
// elsewhere in the process -- cudaFree, cudaDeviceSynchronize -- never waits longer than that) and on
// a stop doorbell; the host re-launches it on demand.
// --------------------------------------------------------------------------------------
struct ResidentMailbox {                 // page-locked host memory; the first 16 bytes are read in one load
    unsigned long long doorbell;         // seq << 32 | (n_slots - 1) << 20 | (n_guesses - 1) << 8 | stop << 1 | centi
    unsigned long long params;           // n_answers << 32 | bits of min_score
    unsigned alive;                      // written 0 by the kernel when it leaves
    unsigned pad[11];
};
constexpr unsigned long long kDoorStop = 2ull, kDoorIdle = 4ull;      // the host never sets kDoorIdle
constexpr int kResidentThreads = 128, kResidentCtas = 148;
// Measured on B200 (tools/probe_resident.py, blocking call p50, launch path / resident): 2x1 16.3 / 9.9 us,
// 64x1 18.2 / 11.2, 512x8 26.6 / 19.6, 1024x1 19.1 / 16.6, 1024x4 24.4 / 23.1, 2048x1 20.8 / 22.2,
// 2048x8 34.0 / 50.1 -- beyond ~1000 guesses or ~4000 rows the PCIe reads of the records and the four
// warps per SM cost more than the launch they save.
constexpr int kResidentMaxGuesses = 1024, kResidentMaxRows = 4096;
constexpr int kRecHeaderInts = 3;        // record: len, slot, exact, then the indices
// a CTA stages ceil(B / 148) records: at most 7 x 8 ints (1024 x 4) or 1 x 4104 (1 x 4096)
constexpr int kResidentSmemInts = 4224;

// ints per record, rounded up to whole 32-byte sectors
__host__ __device__ inline int resident_rec_stride(int n_slots) { return (kRecHeaderInts + n_slots + 7) & ~7; }

struct ResidentArgs {
    const float *table;
    int64_t n_rows;
    int dim4;
    const float *ans_unit;
    const int32_t *ans_rows;
    ResidentMailbox *box;                // host
    const int32_t *records;              // host: the batch, one record per guess
    float *scores;                       // host: the results
    int32_t *centi;
    unsigned long long *flag;            // device: [0] the doorbell as CTA 0 last saw it, [1] its params
    unsigned long long last;             // doorbell value the host considers done
    unsigned long long idle_ns;
    unsigned long long *trace;           // developer aid (CM_RESIDENT_TRACE=1): device timestamps in the mailbox, or NULL
};

template <int VPL>
__global__ void __launch_bounds__(kResidentThreads)
cm_resident_kernel(const ResidentArgs a) {
    __shared__ unsigned long long s_door, s_params;
    __shared__ int32_t s_rec[kResidentSmemInts];
    unsigned long long last = a.last;
    for (;;) {
        if (threadIdx.x == 0) {
            unsigned long long d, prm = 0;
            if (blockIdx.x == 0) {
                // the ONLY reader of the mailbox header: doorbell and params in one 16-byte PCIe read
                const unsigned long long t0 = cm::global_ns();
                for (;;) {
                    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(d), "=l"(prm) : "l"(&a.box->doorbell) : "memory");
                    if (d != last) break;
                    if (cm::global_ns() - t0 > a.idle_ns) { d = last | kDoorStop | kDoorIdle; break; }    // idle: leave
                }
                // No system-scope fence here (measured: +1.5 us per request).  The records are read only
                // after this value has travelled through the device flag to the other CTAs, by PCIe reads
                // issued later than the one that returned it, and the host wrote them before the doorbell.
                a.flag[1] = prm;
                asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(a.flag), "l"(d) : "memory");
            } else {
                for (;;) {
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(d) : "l"(a.flag) : "memory");
                    if (d != last) break;
                    __nanosleep(64);
                }
                prm = __ldcg(a.flag + 1);
            }
            s_door = d;
            s_params = prm;
        }
        __syncthreads();
        const unsigned long long door = s_door, params = s_params;
        __syncthreads();
        if (a.trace != nullptr && threadIdx.x == 0 && blockIdx.x < 2) a.trace[blockIdx.x] = cm::global_ns();   // 0: doorbell seen, 1: flag seen
        if (door & kDoorStop) break;
        last = door;
        const int n_guesses = (int)((door >> 8) & 0xfffu) + 1;
        const int n_slots = (int)((door >> 20) & 0xfffu) + 1;
        const float min_score = __uint_as_float((unsigned)params);
        const int n_answers = (int)(params >> 32);
        int32_t *centi = (door & 1ull) ? a.centi : nullptr;
        // This CTA's share: `per` consecutive guesses.  Their records are contiguous in the mailbox and are
        // fetched once, by the whole CTA, with coalesced volatile loads (never served from a cache line of
        // an earlier batch; every 32-byte sector asked for exactly once) into shared memory.
        const int rec = resident_rec_stride(n_slots);
        const int per = (n_guesses + (int)gridDim.x - 1) / (int)gridDim.x;
        const int g0 = (int)blockIdx.x * per;
        const int cnt = min(per, n_guesses - g0);
        if (cnt > 0) {
            const int32_t *src = a.records + (size_t)g0 * rec;
            for (int i = threadIdx.x; i < cnt * rec; i += kResidentThreads)
                asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(s_rec[i]) : "l"(src + i) : "memory");
        }
        __syncthreads();
        if (a.trace != nullptr && threadIdx.x == 0 && blockIdx.x == 1) a.trace[3] = cm::global_ns();           // 3: records staged
        if (cnt > 0) {
            const int lw = (int)(threadIdx.x >> 5);
            float *scores = a.scores + g0;
            int32_t *cen = centi != nullptr ? centi + g0 : nullptr;
            bool done = false;
            if constexpr (VPL == 3) {    // 300-d tables, one word per guess: the 8-lane kernel's arithmetic
                if (n_slots == 1 && a.dim4 == 75) {
                    for (int w = lw; w * 4 < cnt; w += kResidentThreads / 32)
                        score_groups<10, 8, 1, true, kResidentThreads, true, true>(
                            w, a.table, a.n_rows, a.dim4, s_rec + kRecHeaderInts, s_rec, cnt, 1, a.ans_unit, a.ans_rows,
                            n_answers, nullptr, nullptr, min_score, scores, cen, 0u, rec);
                    done = true;
                }
            }
            if (!done) {
                // R = 8 rows per round where registers allow: a resident micro-batch is latency, not bytes in flight
                for (int w = lw; w < cnt; w += kResidentThreads / 32)
                    score_groups<VPL, 32, (VPL <= 3 ? 8 : 2), true, kResidentThreads, false, true>(
                        w, a.table, a.n_rows, a.dim4, s_rec + kRecHeaderInts, s_rec, cnt, n_slots, a.ans_unit, a.ans_rows,
                        n_answers, nullptr, nullptr, min_score, scores, cen, 0u, rec);
            }
        }
        if (a.trace != nullptr && threadIdx.x == 0 && blockIdx.x == 1) a.trace[2] = cm::global_ns();           // 2: scored
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(&a.box->alive), "r"(0u) : "memory");
    }
}

}  // namespace

namespace {

// The four scoring kernels that ship for 300-d tables; each is the default for one class of shapes:
//   wide       : 32 lanes per guess, 2 rows per round   -- small (latency-bound) phrase batches
//   paired     : 16 lanes per guess, 2 guesses per warp -- bandwidth-bound phrase batches
//   single     : 8 lanes per guess, 4 guesses per warp  -- one-word guesses (the reference's shape)
//   persistent : the same with the next group's metadata prefetched -- very large one-word batches
// Other table widths take the generic 32-lane kernel.  `CM_SCORE_KERNEL=wide|paired|single|persistent`
// (environment, read once) overrides only the shape heuristic, for A/B timing of these four.
enum ScoreKernel { kAuto = 0, kWide, kPaired, kSingle, kPersistent };

template <int VPL, int LPG, int R, bool STREAM, int THREADS, int MIN_BLOCKS, bool SINGLE = false, bool IDX24 = false>
void launch_variant(const cm_engine *e, const int32_t *gi, const int32_t *gl, int B, int N,
                    const int32_t *slot, const uint8_t *exact, float min_score, float *scores,
                    int32_t *centi, uint32_t flags, cudaStream_t st, const cm::ShipArgs *ship) {
    constexpr int per_cta = (THREADS / 32) * (32 / LPG);
    cudaLaunchConfig_t cfg = {};
    const int score_ctas = (B + per_cta - 1) / per_cta;
    const bool shipping = ship != nullptr && ship->world > 0;
    const int n_ship_ctas = shipping ? cm::ship_ctas(ship->n) : 0;
    cfg.gridDim = dim3((unsigned)(score_ctas + n_ship_ctas));
    cfg.blockDim = dim3(THREADS);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, cm_score_kernel<VPL, LPG, R, STREAM, THREADS, MIN_BLOCKS, SINGLE, IDX24>,
                       (const float *)e->d_table, e->n_rows, e->dim / 4, gi, gl, B, N,
                       (const float *)e->d_ans_unit, (const int32_t *)e->d_ans_rows, e->n_answers,
                       slot, exact, min_score, scores, centi, flags, n_ship_ctas,
                       shipping ? *ship : cm::ShipArgs{});
    cm::count_launch();
}

template <int VPL, int LPG, bool STREAM, int THREADS, int MIN_BLOCKS>
void launch_single_persistent(const cm_engine *e, const int32_t *gi, const int32_t *gl, int B,
                              const int32_t *slot, const uint8_t *exact, float min_score, float *scores,
                              int32_t *centi, uint32_t flags, cudaStream_t st, const cm::ShipArgs *ship) {
    constexpr int per_cta = (THREADS / 32) * (32 / LPG);
    const int needed = (B + per_cta - 1) / per_cta;
    const int resident = 148 * MIN_BLOCKS;
    const int score_ctas = needed < resident ? needed : resident;
    const bool shipping = ship != nullptr && ship->world > 0;
    const int n_ship_ctas = shipping ? cm::ship_ctas(ship->n) : 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(score_ctas + n_ship_ctas));
    cfg.blockDim = dim3(THREADS);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, cm_score1_kernel<VPL, LPG, STREAM, THREADS, MIN_BLOCKS>,
                       (const float *)e->d_table, e->n_rows, e->dim / 4, gi, gl, B,
                       (const float *)e->d_ans_unit, (const int32_t *)e->d_ans_rows, e->n_answers,
                       slot, exact, min_score, scores, centi, flags, n_ship_ctas,
                       shipping ? *ship : cm::ShipArgs{});
    cm::count_launch();
}

ScoreKernel forced_kernel() {
    static const ScoreKernel v = [] {
        const char *s = getenv("CM_SCORE_KERNEL");
        if (s == nullptr) return kAuto;
        if (strcmp(s, "wide") == 0) return kWide;
        if (strcmp(s, "paired") == 0) return kPaired;
        if (strcmp(s, "single") == 0) return kSingle;
        if (strcmp(s, "persistent") == 0) return kPersistent;
        return kAuto;
    }();
    return v;
}

template <int VPL, bool IDX24 = false>
void launch_score(const cm_engine *e, const int32_t *gi, const int32_t *gl, int B, int N,
                  const int32_t *slot, const uint8_t *exact, float min_score, float *scores,
                  int32_t *centi, uint32_t flags, cudaStream_t st, const cm::ShipArgs *ship = nullptr) {
#define CM_V(LPG, R, STREAM, THREADS, MINB, VP) \
    launch_variant<VP, LPG, R, STREAM, THREADS, MINB, false, IDX24>(e, gi, gl, B, N, slot, exact, min_score, scores, centi, flags, st, ship)
#define CM_S(LPG, STREAM, THREADS, MINB, VP) \
    launch_variant<VP, LPG, 1, STREAM, THREADS, MINB, true, IDX24>(e, gi, gl, B, N, slot, exact, min_score, scores, centi, flags, st, ship)
    if constexpr (VPL == 3) {
        if (e->dim == 300) {
            // Measured on B200 (tools/probe_kernel.py, us per launch, CM_SCORE_OVERLAP_PREVIOUS):
            //   4096x16 : wide 12.9 | paired 11.6;  65536x16 200 | 181;  16384x8 30.4 | 24.1;  512x8 3.8 | 6.0
            //   65536x1 : single 15.2 | persistent 15.5;  131072x1 : 28.2 | 27.0;  262144x1 : 55.9 | 52.7;
            //   10^6 x1 : 214 | 202
            ScoreKernel k = forced_kernel();
            if ((k == kSingle || k == kPersistent) && N != 1) k = kAuto;
            if (k == kAuto)
                k = (N == 1) ? (B >= (1 << 17) ? kPersistent : kSingle)
                             : ((int64_t)B * N >= 32768 ? kPaired : kWide);
            if (IDX24 && k == kPersistent) k = kSingle;          // the persistent form takes 4-byte indices only
            switch (k) {
                case kPersistent:
                    if constexpr (!IDX24)
                        launch_single_persistent<10, 8, true, 128, 6>(e, gi, gl, B, slot, exact, min_score, scores, centi, flags, st, ship);
                    return;
                case kSingle: CM_S(8, true, 128, 8, 10); return;
                case kPaired: CM_V(16, 1, true, 128, 7, 5); return;
                default: CM_V(32, 2, true, 128, 8, 3); return;
            }
        }
    }
    CM_V(32, 4, true, 256, 1, VPL);
#undef CM_V
#undef CM_S
}

template <int VPL>
void launch_answers(const cm_engine *e, const int32_t *d_rows, int A, cudaStream_t st) {
    const int grid = (A + kWarpsPerCta - 1) / kWarpsPerCta;
    cm_answer_unit_kernel<VPL><<<grid, kWarpsPerCta * 32, 0, st>>>(
        e->d_table, e->n_rows, e->dim / 4, d_rows, A, e->d_ans_unit);
    cm::count_launch();
}

#define CM_DISPATCH_VPL(dim, CALL)                          \
    switch (((dim) / 4 + 31) / 32) {                        \
        case 1: CALL(1); break;                             \
        case 2: CALL(2); break;                             \
        case 3: CALL(3); break;                             \
        case 4: CALL(4); break;                             \
        case 5: CALL(5); break;                             \
        case 6: CALL(6); break;                             \
        case 7: CALL(7); break;                             \
        default: CALL(8); break;                            \
    }

}  // namespace

namespace cm {

int check_engine(const cm_engine *e) {
    if (e == nullptr) return fail(CM_ERR_INVALID, "engine is NULL");
    if (e->d_table == nullptr) return fail(CM_ERR_STATE, "engine has no resident table");
    return CM_OK;
}

bool is_pinned_host(const void *p) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
}

int enqueue_score(cm_engine *e, const int32_t *guess_idx, const int32_t *guess_len, int n_guesses,
                  int n_slots, const int32_t *answer_slot, const uint8_t *exact, float min_score,
                  float *scores, int32_t *centi, uint32_t flags, cudaStream_t st, const ShipArgs *ship,
                  bool idx24) {
    if (int rc = check_engine(e)) return rc;
    if (idx24 && ((uintptr_t)guess_idx & 3u))
        return fail(CM_ERR_INVALID, "the block of 3-byte indices must start on a 4-byte boundary");
    if (idx24 && e->n_rows > 0xffffffLL)
        return fail(CM_ERR_INVALID, "3-byte indices need a table of at most 16,777,215 rows, this one has %lld",
                    (long long)e->n_rows);
    if (n_guesses < 0) return fail(CM_ERR_INVALID, "n_guesses %d is negative", n_guesses);
    if (n_slots < 1 || n_slots > CM_MAX_SEQ)
        return fail(CM_ERR_INVALID, "n_slots %d outside [1, %d]", n_slots, CM_MAX_SEQ);
    if (flags & ~(uint32_t)CM_SCORE_OVERLAP_PREVIOUS)
        return fail(CM_ERR_INVALID, "unknown flag bits 0x%x", flags);
    if (n_guesses == 0) return CM_OK;
    if (!guess_idx || !guess_len || !answer_slot || !scores)
        return fail(CM_ERR_INVALID, "a required buffer pointer is NULL");
    if (e->n_answers < 1) return fail(CM_ERR_STATE, "cm_set_answers has not been called");
    DeviceGuard guard(e->device);
    if (idx24) {
#define CM_CALL(V) launch_score<V, true>(e, guess_idx, guess_len, n_guesses, n_slots, answer_slot, exact, \
                                         min_score, scores, centi, flags, st, ship)
        CM_DISPATCH_VPL(e->dim, CM_CALL)
#undef CM_CALL
    } else {
#define CM_CALL(V) launch_score<V>(e, guess_idx, guess_len, n_guesses, n_slots, answer_slot, exact, \
                                   min_score, scores, centi, flags, st, ship)
        CM_DISPATCH_VPL(e->dim, CM_CALL)
#undef CM_CALL
    }
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

// memcpy split over a few threads: one core moves ~2-3 GB/s out of pageable (page-cache / mmap)
// memory, far below what the PCIe link takes.
static void parallel_memcpy(unsigned char *dst, const unsigned char *src, size_t len, int threads) {
    if (threads < 2 || len < ((size_t)4 << 20)) { memcpy(dst, src, len); return; }
    const size_t part = ((len + threads - 1) / threads + 4095) & ~(size_t)4095;
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) {
        const size_t off = part * t;
        if (off >= len) break;
        pool.emplace_back([=] { memcpy(dst + off, src + off, len - off < part ? len - off : part); });
    }
    memcpy(dst, src, len < part ? len : part);
    for (auto &th : pool) th.join();
}

// Pinned staging block of the host entry point (only used when the caller's memory is pageable).
int ensure_stage(cm_engine *e, size_t bytes) {
    if (bytes <= e->stage_bytes) return CM_OK;
    size_t want = e->stage_bytes ? e->stage_bytes : (size_t)1 << 20;
    while (want < bytes) want *= 2;
    if (e->h_stage) cudaFreeHost(e->h_stage);
    e->h_stage = nullptr;
    e->stage_bytes = 0;
    if (cudaHostAlloc(&e->h_stage, want, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return fail(CM_ERR_NOMEM, "cudaHostAlloc of %zu staging bytes failed", want);
    }
    e->stage_bytes = want;
    return CM_OK;
}

}  // namespace cm

using cm::align_up;
using cm::check_engine;
using cm::DeviceGuard;
using cm::fail;
using cm::is_pinned_host;

// ======================================================================================
// C ABI
// ======================================================================================
extern "C" {

const char *cm_last_error(void) { return cm::g_err; }
const char *cm_version(void) { return "0.1.0"; }
int64_t cm_kernel_launch_count(void) { return cm::g_launches.load(std::memory_order_relaxed); }

int cm_engine_create(cm_engine **out, int device, int64_t n_rows, int dim) {
    if (out == nullptr) return fail(CM_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (n_rows <= 0 || n_rows > 0x7fffffffLL)
        return fail(CM_ERR_INVALID, "n_rows %lld outside [1, 2^31)", (long long)n_rows);
    if (dim < 4 || dim > CM_MAX_DIM || dim % 4 != 0)
        return fail(CM_ERR_INVALID, "dim %d must be a multiple of 4 in [4, %d]", dim, CM_MAX_DIM);
    int n_dev = 0;
    CM_CUDA(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev)
        return fail(CM_ERR_INVALID, "device %d not in [0, %d)", device, n_dev);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CM_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    cudaDeviceProp prop;
    CM_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(CM_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only",
                    device, prop.major, prop.minor);
    cm_engine *e = new (std::nothrow) cm_engine();
    if (e == nullptr) return fail(CM_ERR_NOMEM, "engine allocation failed");
    e->device = device; e->n_rows = n_rows; e->dim = dim;
    const size_t table_bytes = (size_t)n_rows * dim * sizeof(float);
    cudaError_t err = cudaMalloc(&e->d_table, table_bytes);
    if (err == cudaSuccess) err = cudaMalloc(&e->d_ans_unit, (size_t)CM_MAX_ANSWERS * dim * sizeof(float));
    if (err == cudaSuccess) err = cudaMalloc(&e->d_ans_rows, (size_t)CM_MAX_ANSWERS * sizeof(int32_t));
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&e->answers_done, cudaEventDisableTiming);
    if (err != cudaSuccess) {
        cudaGetLastError();
        const int code = fail(err == cudaErrorMemoryAllocation ? CM_ERR_NOMEM : CM_ERR_CUDA,
                              "engine setup (%zu table bytes) failed: %s", table_bytes,
                              cudaGetErrorString(err));
        cm_engine_destroy(e);
        return code;
    }
    cm::similar_reserve(e);
    *out = e;
    return CM_OK;
}

int cm_engine_destroy(cm_engine *e) {
    if (e == nullptr) return CM_OK;
    {
        DeviceGuard guard(e->device);
        cm::resident_free(e);
        if (e->stream) { cudaStreamSynchronize(e->stream); cudaStreamDestroy(e->stream); }
        if (e->answers_done) cudaEventDestroy(e->answers_done);
        cm::similar_state_free(e);
        if (e->d_table) cudaFree(e->d_table);
        if (e->d_ans_unit) cudaFree(e->d_ans_unit);
        if (e->d_ans_rows) cudaFree(e->d_ans_rows);
        if (e->h_stage) cudaFreeHost(e->h_stage);
    }
    delete e;
    return CM_OK;
}

int cm_engine_upload_rows(cm_engine *e, const float *rows, int64_t row0, int64_t n) {
    if (int rc = check_engine(e)) return rc;
    if (rows == nullptr) return fail(CM_ERR_INVALID, "rows is NULL");
    if (row0 < 0 || n < 0 || row0 + n > e->n_rows)
        return fail(CM_ERR_INVALID, "rows [%lld, %lld) outside the %lld-row table",
                    (long long)row0, (long long)(row0 + n), (long long)e->n_rows);
    DeviceGuard guard(e->device);
    cm::resident_stop(e);                    // it reads the table through the non-coherent path
    const size_t bytes = (size_t)n * e->dim * sizeof(float);
    unsigned char *dst = reinterpret_cast<unsigned char *>(e->d_table + (size_t)row0 * e->dim);
    const unsigned char *src = reinterpret_cast<const unsigned char *>(rows);
    constexpr size_t kChunk = (size_t)32 << 20;
    if (bytes <= kChunk || is_pinned_host(rows)) {
        CM_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, e->stream));
        CM_CUDA(cudaStreamSynchronize(e->stream));
    } else {
        // pageable source (gensim's mmap): stage through two pinned buffers so the host copy of
        // chunk i+1 overlaps the DMA of chunk i, the host copy itself split over up to 8 threads
        unsigned char *pin[2] = {nullptr, nullptr};
        cudaEvent_t done[2] = {nullptr, nullptr};
        static const int copy_threads = [] {
            const char *v = getenv("CM_UPLOAD_THREADS");
            const unsigned hw = std::thread::hardware_concurrency();
            const int n = v != nullptr && *v != 0 ? atoi(v) : (int)(hw > 8 ? 8 : hw);
            return n < 1 ? 1 : n;
        }();
        cudaError_t err = cudaHostAlloc(&pin[0], kChunk, cudaHostAllocDefault);
        if (err == cudaSuccess) err = cudaHostAlloc(&pin[1], kChunk, cudaHostAllocDefault);
        if (err == cudaSuccess) err = cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming);
        if (err == cudaSuccess) err = cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming);
        for (size_t off = 0, i = 0; err == cudaSuccess && off < bytes; off += kChunk, ++i) {
            const size_t len = bytes - off < kChunk ? bytes - off : kChunk;
            if (i >= 2) err = cudaEventSynchronize(done[i & 1]);
            if (err != cudaSuccess) break;
            cm::parallel_memcpy(pin[i & 1], src + off, len, copy_threads);
            err = cudaMemcpyAsync(dst + off, pin[i & 1], len, cudaMemcpyHostToDevice, e->stream);
            if (err == cudaSuccess) err = cudaEventRecord(done[i & 1], e->stream);
        }
        if (err == cudaSuccess) err = cudaStreamSynchronize(e->stream);
        for (int i = 0; i < 2; ++i) {
            if (done[i]) cudaEventDestroy(done[i]);
            if (pin[i]) cudaFreeHost(pin[i]);
        }
        if (err != cudaSuccess) {
            cudaGetLastError();
            return fail(CM_ERR_CUDA, "staged upload failed: %s", cudaGetErrorString(err));
        }
    }
    cm::similar_invalidate(e);
    return CM_OK;
}

int cm_engine_download_rows(cm_engine *e, float *rows, int64_t row0, int64_t n) {
    if (int rc = check_engine(e)) return rc;
    if (rows == nullptr) return fail(CM_ERR_INVALID, "rows is NULL");
    if (row0 < 0 || n < 0 || row0 + n > e->n_rows)
        return fail(CM_ERR_INVALID, "rows [%lld, %lld) outside the %lld-row table",
                    (long long)row0, (long long)(row0 + n), (long long)e->n_rows);
    DeviceGuard guard(e->device);
    CM_CUDA(cudaMemcpyAsync(rows, e->d_table + (size_t)row0 * e->dim,
                            (size_t)n * e->dim * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CM_CUDA(cudaStreamSynchronize(e->stream));
    return CM_OK;
}

int cm_engine_fill_synthetic(cm_engine *e, uint32_t seed) {
    if (int rc = check_engine(e)) return rc;
    if ((uint64_t)e->n_rows * (uint64_t)e->dim > 0xffffffffull)
        return fail(CM_ERR_INVALID, "synthetic fill needs n_rows*dim < 2^32");
    DeviceGuard guard(e->device);
    cm::resident_stop(e);
    cm_synth_fill_kernel<<<148 * 8, 256, 0, e->stream>>>(e->d_table, e->n_rows, e->dim, seed);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    CM_CUDA(cudaStreamSynchronize(e->stream));
    cm::similar_invalidate(e);
    return CM_OK;
}

int cm_engine_flush_l2(cm_engine *e, void *stream) {
    if (int rc = check_engine(e)) return rc;
    DeviceGuard guard(e->device);
    const size_t table_bytes = (size_t)e->n_rows * e->dim * sizeof(float);
    const size_t bytes = table_bytes < ((size_t)512 << 20) ? table_bytes : ((size_t)512 << 20);
    cm_flush_read_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4 *>(e->d_table),
                                                                    bytes / 16, nullptr);
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

int cm_engine_table(cm_engine *e, void **d_table, int64_t *n_rows, int *dim) {
    if (int rc = check_engine(e)) return rc;
    if (d_table) *d_table = e->d_table;
    if (n_rows) *n_rows = e->n_rows;
    if (dim) *dim = e->dim;
    return CM_OK;
}

int cm_engine_stream(cm_engine *e, void **stream) {
    if (e == nullptr || stream == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    *stream = e->stream;
    return CM_OK;
}

int cm_engine_synchronize(cm_engine *e) {
    if (e == nullptr) return fail(CM_ERR_INVALID, "engine is NULL");
    DeviceGuard guard(e->device);
    CM_CUDA(cudaStreamSynchronize(e->stream));
    return CM_OK;
}

int cm_set_answers_device(cm_engine *e, const int32_t *d_answer_rows, int n_answers, void *stream) {
    if (int rc = check_engine(e)) return rc;
    if (d_answer_rows == nullptr) return fail(CM_ERR_INVALID, "answer_rows is NULL");
    if (n_answers < 1 || n_answers > CM_MAX_ANSWERS)
        return fail(CM_ERR_INVALID, "n_answers %d outside [1, %d]", n_answers, CM_MAX_ANSWERS);
    DeviceGuard guard(e->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (d_answer_rows != e->d_ans_rows)
        CM_CUDA(cudaMemcpyAsync(e->d_ans_rows, d_answer_rows, (size_t)n_answers * sizeof(int32_t),
                                cudaMemcpyDeviceToDevice, st));
    e->n_answers = n_answers;
#define CM_CALL(V) launch_answers<V>(e, e->d_ans_rows, n_answers, st)
    CM_DISPATCH_VPL(e->dim, CM_CALL)
#undef CM_CALL
    CM_CUDA(cudaGetLastError());
    if (e->resident != nullptr) {            // the resident kernel is not ordered behind `st` by the stream
        CM_CUDA(cudaEventRecord(e->answers_done, st));
        e->answers_pending = true;
    }
    return CM_OK;
}

int cm_set_answers(cm_engine *e, const int32_t *answer_rows, int n_answers, void *stream) {
    if (int rc = check_engine(e)) return rc;
    if (answer_rows == nullptr) return fail(CM_ERR_INVALID, "answer_rows is NULL");
    if (n_answers < 1 || n_answers > CM_MAX_ANSWERS)
        return fail(CM_ERR_INVALID, "n_answers %d outside [1, %d]", n_answers, CM_MAX_ANSWERS);
    DeviceGuard guard(e->device);
    cudaStream_t st = (cudaStream_t)stream;
    // pageable source: the runtime stages it before returning, so the caller's buffer is free
    CM_CUDA(cudaMemcpyAsync(e->d_ans_rows, answer_rows, (size_t)n_answers * sizeof(int32_t),
                            cudaMemcpyHostToDevice, st));
    return cm_set_answers_device(e, e->d_ans_rows, n_answers, st);
}

int cm_score_batch_device(cm_engine *e, const int32_t *d_guess_idx, const int32_t *d_guess_len,
                          int n_guesses, int n_slots, const int32_t *d_answer_slot,
                          const uint8_t *d_exact, float min_score, float *d_scores,
                          int32_t *d_centi, uint32_t flags, void *stream) {
    return cm::enqueue_score(e, d_guess_idx, d_guess_len, n_guesses, n_slots, d_answer_slot, d_exact,
                             min_score, d_scores, d_centi, flags, (cudaStream_t)stream);
}

}  // extern "C"

// ---- resident micro-batch kernel: host side ---------------------------------------------------
struct cm_resident {
    unsigned char *h_block = nullptr;        // page-locked: mailbox header, records, results
    ResidentMailbox *box = nullptr;
    int32_t *records = nullptr, *centi = nullptr;
    float *scores = nullptr;
    unsigned long long *h_last = nullptr;    // page-locked source of the flag reset
    unsigned long long *d_flag = nullptr;
    cudaStream_t stream = nullptr;
    unsigned long long last = 0;             // doorbell value of the last completed request
    uint32_t seq = 0;
    unsigned long long idle_ns = 0;
    bool armed = false;
    double host_trace[3] = {0, 0, 0};        // us: entry -> doorbell, doorbell -> first result, doorbell -> all results
};

namespace {

int resident_launch(cm_engine *e) {
    cm_resident *r = e->resident;
    ResidentArgs a;
    a.table = e->d_table; a.n_rows = e->n_rows; a.dim4 = e->dim / 4;
    a.ans_unit = e->d_ans_unit; a.ans_rows = e->d_ans_rows;
    a.box = r->box; a.records = r->records;
    a.scores = r->scores; a.centi = r->centi; a.flag = r->d_flag;
    a.last = r->last; a.idle_ns = r->idle_ns;
    a.trace = getenv("CM_RESIDENT_TRACE") != nullptr ? reinterpret_cast<unsigned long long *>(r->h_block + 192) : nullptr;
    *r->h_last = r->last;                    // workers compare the device flag with `last`: start them equal
    CM_CUDA(cudaMemcpyAsync(r->d_flag, r->h_last, sizeof(unsigned long long), cudaMemcpyHostToDevice, r->stream));
    r->box->alive = 1u;
    std::atomic_thread_fence(std::memory_order_seq_cst);
#define CM_CALL(V) cm_resident_kernel<V><<<kResidentCtas, kResidentThreads, 0, r->stream>>>(a)
    CM_DISPATCH_VPL(e->dim, CM_CALL)
#undef CM_CALL
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

inline bool resident_alive(const cm_resident *r) {
    return *reinterpret_cast<const volatile unsigned *>(&r->box->alive) != 0u;
}

inline void ring(cm_resident *r, unsigned long long door) {
    std::atomic_thread_fence(std::memory_order_release);
    *reinterpret_cast<volatile unsigned long long *>(&r->box->doorbell) = door;
}

// One micro-batch through the resident kernel (e->mu held, device current).
int resident_score(cm_engine *e, const int32_t *guess_idx, const int32_t *guess_len, int n_guesses,
                   int n_slots, const int32_t *answer_slot, const uint8_t *exact, float min_score,
                   float *scores, int32_t *centi) {
    cm_resident *r = e->resident;
    const auto t_enter = std::chrono::steady_clock::now();
    const size_t B = (size_t)n_guesses, N = (size_t)n_slots;
    if (e->answers_pending) {                // cm_set_answers ran on some stream: its unit vectors must be complete
        CM_CUDA(cudaEventSynchronize(e->answers_done));
        e->answers_pending = false;
    }
    {   // one record per guess: {len, slot, exact, indices...}, whole 32-byte sectors
        const int rec = resident_rec_stride(n_slots);
        int32_t *out = r->records;
        const int32_t *src = guess_idx;
        for (size_t b = 0; b < B; ++b, out += rec, src += N) {
            out[0] = guess_len[b];
            out[1] = answer_slot[b];
            out[2] = exact ? (int32_t)exact[b] : 0;
            if (N <= 16) {                   // a call to memcpy per guess costs more than the copy
                for (size_t k = 0; k < N; ++k) out[kRecHeaderInts + k] = src[k];
            } else {
                memcpy(out + kRecHeaderInts, src, N * 4);
            }
        }
    }
    memset(r->scores, 0xff, B * 4);          // sentinel: a NaN no result can be, and not a valid hundredths value
    if (centi) memset(r->centi, 0xff, B * 4);
    uint32_t ms_bits;
    memcpy(&ms_bits, &min_score, 4);
    *reinterpret_cast<volatile unsigned long long *>(&r->box->params) =
        ((unsigned long long)(uint32_t)e->n_answers << 32) | ms_bits;
    if (!resident_alive(r))
        if (int rc = resident_launch(e)) return rc;
    const unsigned long long door = ((unsigned long long)(++r->seq) << 32) | ((unsigned long long)(N - 1) << 20) |
                                    ((unsigned long long)(B - 1) << 8) | (centi ? 1ull : 0ull);
    const auto t_ring = std::chrono::steady_clock::now();
    r->host_trace[0] = std::chrono::duration<double, std::micro>(t_ring - t_enter).count();
    ring(r, door);
    bool first_seen = false;
    const volatile uint32_t *ws = reinterpret_cast<const volatile uint32_t *>(r->scores);
    const volatile uint32_t *wc = centi ? reinterpret_cast<const volatile uint32_t *>(r->centi) : nullptr;
    size_t i = 0, j = 0;
    for (long spins = 0;; ++spins) {
        while (i < B && ws[i] != 0xffffffffu) ++i;
        while (wc != nullptr && j < B && wc[j] != 0xffffffffu) ++j;
        if (!first_seen && i > 0) {
            first_seen = true;
            r->host_trace[1] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_ring).count();
        }
        if (i == B && (wc == nullptr || j == B)) break;
        if ((spins & 0x3ff) != 0x3ff) continue;
        if (std::chrono::steady_clock::now() - t_ring > std::chrono::seconds(5)) {
            cm::resident_stop(e);
            r->armed = false;
            return fail(CM_ERR_CUDA, "resident kernel did not answer within 5 s; disarmed");
        }
        if (!resident_alive(r)) {
            // it left (idle limit) just as the doorbell rang, or it died: a fresh instance picks the doorbell up
            cudaError_t err = cudaStreamSynchronize(r->stream);
            if (err != cudaSuccess) return fail(CM_ERR_CUDA, "resident kernel failed: %s", cudaGetErrorString(err));
            if (int rc = resident_launch(e)) return rc;
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    r->host_trace[2] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_ring).count();
    r->last = door;
    memcpy(scores, r->scores, B * 4);
    if (centi) memcpy(centi, r->centi, B * 4);
    return CM_OK;
}

}  // namespace

namespace cm {

void resident_stop(cm_engine *e) {
    cm_resident *r = e->resident;
    if (r == nullptr || r->stream == nullptr) return;
    if (resident_alive(r)) {
        const unsigned long long door = ((unsigned long long)(++r->seq) << 32) | kDoorStop;
        ring(r, door);
        r->last = door;
    }
    cudaStreamSynchronize(r->stream);
}

void resident_free(cm_engine *e) {
    cm_resident *r = e->resident;
    if (r == nullptr) return;
    resident_stop(e);
    if (r->stream) cudaStreamDestroy(r->stream);
    if (r->d_flag) cudaFree(r->d_flag);
    if (r->h_block) cudaFreeHost(r->h_block);
    delete r;
    e->resident = nullptr;
}

}  // namespace cm

extern "C" {

int cm_engine_arm_resident(cm_engine *e, int idle_ms) {
    if (int rc = check_engine(e)) return rc;
    if (idle_ms < 0 || idle_ms > 10000) return fail(CM_ERR_INVALID, "idle_ms %d outside [0, 10000]", idle_ms);
    std::lock_guard<std::mutex> lock(e->mu);
    DeviceGuard guard(e->device);
    if (e->resident == nullptr) {
        cm_resident *r = new (std::nothrow) cm_resident();
        if (r == nullptr) return fail(CM_ERR_NOMEM, "resident state allocation failed");
        e->resident = r;
        // records: B * resident_rec_stride(N) ints with B <= kResidentMaxGuesses and B * N <= kResidentMaxRows
        const size_t off_rec = 256, rec_bytes = ((size_t)kResidentMaxGuesses * (kRecHeaderInts + 7) + kResidentMaxRows + 8) * 4;
        const size_t off_scores = off_rec + rec_bytes, off_centi = off_scores + (size_t)kResidentMaxGuesses * 4;
        const size_t bytes = off_centi + (size_t)kResidentMaxGuesses * 4;
        cudaError_t err = cudaHostAlloc(&r->h_block, bytes, cudaHostAllocDefault);
        if (err == cudaSuccess) err = cudaMalloc(&r->d_flag, 2 * sizeof(unsigned long long));
        if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
        if (err != cudaSuccess) {
            cudaGetLastError();
            cm::resident_free(e);
            return fail(CM_ERR_NOMEM, "resident mailbox setup failed: %s", cudaGetErrorString(err));
        }
        memset(r->h_block, 0, bytes);
        r->box = reinterpret_cast<ResidentMailbox *>(r->h_block);
        r->h_last = reinterpret_cast<unsigned long long *>(r->h_block + 128);
        r->records = reinterpret_cast<int32_t *>(r->h_block + off_rec);
        r->scores = reinterpret_cast<float *>(r->h_block + off_scores);
        r->centi = reinterpret_cast<int32_t *>(r->h_block + off_centi);
        CM_CUDA(cudaDeviceSynchronize());    // answers registered before arming are complete from here on
        e->answers_pending = false;
    }
    cm_resident *r = e->resident;
    r->idle_ns = (unsigned long long)(idle_ms == 0 ? 20 : idle_ms) * 1000000ull;
    r->armed = true;
    if (!resident_alive(r)) {
        cudaStreamSynchronize(r->stream);
        if (int rc = resident_launch(e)) return rc;
    }
    return CM_OK;
}

int cm_engine_disarm_resident(cm_engine *e) {
    if (int rc = check_engine(e)) return rc;
    std::lock_guard<std::mutex> lock(e->mu);
    DeviceGuard guard(e->device);
    if (e->resident != nullptr) {
        e->resident->armed = false;
        cm::resident_stop(e);
    }
    return CM_OK;
}

// Developer aid (CM_RESIDENT_TRACE=1 in the environment before arming): timestamps of the last
// resident request.  out[0..3] device globaltimer ns: doorbell seen by CTA 0, flag seen by CTA 1, CTA 1
// scored, CTA 1 has its records staged; out[4..6]: host us from entry to the doorbell, from the doorbell
// to the first result, from the doorbell to all results.
int cm_engine_resident_trace(cm_engine *e, double *out7) {
    if (e == nullptr || out7 == nullptr || e->resident == nullptr) return fail(CM_ERR_INVALID, "no resident state");
    const unsigned long long *t = reinterpret_cast<const unsigned long long *>(e->resident->h_block + 192);
    for (int i = 0; i < 4; ++i) out7[i] = (double)t[i];
    for (int i = 0; i < 3; ++i) out7[4 + i] = e->resident->host_trace[i];
    return CM_OK;
}

int cm_score_batch_device_idx24(cm_engine *e, const uint8_t *d_guess_idx24, const int32_t *d_guess_len,
                                int n_guesses, int n_slots, const int32_t *d_answer_slot,
                                const uint8_t *d_exact, float min_score, float *d_scores,
                                int32_t *d_centi, uint32_t flags, void *stream) {
    return cm::enqueue_score(e, reinterpret_cast<const int32_t *>(d_guess_idx24), d_guess_len, n_guesses, n_slots,
                             d_answer_slot, d_exact, min_score, d_scores, d_centi, flags, (cudaStream_t)stream,
                             nullptr, true);
}

int cm_score_batch_host(cm_engine *e, const int32_t *guess_idx, const int32_t *guess_len,
                        int n_guesses, int n_slots, const int32_t *answer_slot,
                        const uint8_t *exact, float min_score, float *scores, int32_t *centi) {
    if (int rc = check_engine(e)) return rc;
    if (n_guesses < 0) return fail(CM_ERR_INVALID, "n_guesses %d is negative", n_guesses);
    if (n_slots < 1 || n_slots > CM_MAX_SEQ)
        return fail(CM_ERR_INVALID, "n_slots %d outside [1, %d]", n_slots, CM_MAX_SEQ);
    if (n_guesses == 0) return CM_OK;
    if (!guess_idx || !guess_len || !answer_slot || !scores)
        return fail(CM_ERR_INVALID, "a required host pointer is NULL");
    if (e->n_answers < 1) return fail(CM_ERR_STATE, "cm_set_answers has not been called");

    // One synchronous call = one kernel launch and no DMA: the kernel reads the batch from
    // page-locked host memory and writes the scores back into it across PCIe.  Pageable
    // buffers are first copied into the engine's own pinned block.
    std::lock_guard<std::mutex> lock(e->mu);     // one staging block: callers take turns
    DeviceGuard guard(e->device);
    // armed and small enough: no launch at all, the resident kernel takes the batch from its mailbox
    if (e->resident != nullptr && e->resident->armed && n_guesses <= kResidentMaxGuesses &&
        (int64_t)n_guesses * n_slots <= kResidentMaxRows && min_score == min_score &&
        ((n_guesses + kResidentCtas - 1) / kResidentCtas) * resident_rec_stride(n_slots) <= kResidentSmemInts)
        return resident_score(e, guess_idx, guess_len, n_guesses, n_slots, answer_slot, exact, min_score, scores, centi);
    const size_t B = (size_t)n_guesses, N = (size_t)n_slots;
    const bool in_pinned = is_pinned_host(guess_idx) && is_pinned_host(guess_len) &&
                           is_pinned_host(answer_slot) && (!exact || is_pinned_host(exact));
    const bool out_pinned = is_pinned_host(scores) && (!centi || is_pinned_host(centi));
    const size_t off_len = align_up(B * N * 4, 16);
    const size_t off_slot = off_len + B * 4;
    const size_t off_exact = off_slot + B * 4;
    const size_t off_scores = align_up(off_exact + B, 16);
    const size_t off_centi = off_scores + B * 4;
    if (!in_pinned || !out_pinned)
        if (int rc = cm::ensure_stage(e, off_centi + B * 4)) return rc;

    const int32_t *k_idx = guess_idx, *k_len = guess_len, *k_slot = answer_slot;
    const uint8_t *k_exact = exact;
    float *k_scores = scores;
    int32_t *k_centi = centi;
    if (!in_pinned) {
        memcpy(e->h_stage, guess_idx, B * N * 4);
        memcpy(e->h_stage + off_len, guess_len, B * 4);
        memcpy(e->h_stage + off_slot, answer_slot, B * 4);
        if (exact) memcpy(e->h_stage + off_exact, exact, B);
        k_idx = (const int32_t *)e->h_stage;
        k_len = (const int32_t *)(e->h_stage + off_len);
        k_slot = (const int32_t *)(e->h_stage + off_slot);
        k_exact = exact ? e->h_stage + off_exact : nullptr;
    }
    if (!out_pinned) {
        k_scores = (float *)(e->h_stage + off_scores);
        k_centi = centi ? (int32_t *)(e->h_stage + off_centi) : nullptr;
    }
    // Completion without a driver round trip: the result buffers are host memory the kernel
    // writes across PCIe, so they are pre-filled with a bit pattern no result can be (0xffffffff:
    // a NaN, and not a valid hundredths value) and the host simply watches them fill up.  Small
    // batches only -- for large ones the scan costs more than cudaStreamSynchronize saves.
    static const bool poll_enabled = getenv("CM_HOST_NO_POLL") == nullptr;
    const bool poll = poll_enabled && B <= 2048 && min_score == min_score;
    if (poll) {
        memset(k_scores, 0xff, B * 4);
        if (k_centi) memset(k_centi, 0xff, B * 4);
    }
    if (int rc = cm::enqueue_score(e, k_idx, k_len, n_guesses, n_slots, k_slot, k_exact, min_score,
                                   k_scores, k_centi, 0u, e->stream))
        return rc;
    if (poll) {
        const volatile uint32_t *ws = reinterpret_cast<const volatile uint32_t *>(k_scores);
        const volatile uint32_t *wc = reinterpret_cast<const volatile uint32_t *>(k_centi);
        size_t i = 0, j = 0;
        for (long spins = 0;; ++spins) {
            while (i < B && ws[i] != 0xffffffffu) ++i;
            while (wc != nullptr && j < B && wc[j] != 0xffffffffu) ++j;
            if (i == B && (wc == nullptr || j == B)) break;
            if ((spins & 0xfff) == 0xfff && cudaStreamQuery(e->stream) != cudaErrorNotReady) {
                CM_CUDA(cudaStreamSynchronize(e->stream));       // finished (or failed): take the slow path
                break;
            }
        }
        std::atomic_thread_fence(std::memory_order_acquire);
    } else {
        CM_CUDA(cudaStreamSynchronize(e->stream));
    }
    if (!out_pinned) {
        memcpy(scores, k_scores, B * 4);
        if (centi) memcpy(centi, k_centi, B * 4);
    }
    return CM_OK;
}

}  // extern "C"

// ======================================================================================
// cm_score_batch_noalloc -- scoring for callers that are not allowed to allocate.
//
// Who needs it: a host program that has given every byte of GPU memory to an allocator of
// its own (a serving framework's arena, a memory plan fixed at start-up, a tensor library
// that owns the table), or one that must keep cudaMalloc's device-wide synchronisation out
// of its request path.  Such a caller cannot create a cm_engine, because the engine
// allocates the table replica and the answers' unit vectors.  Here the caller brings the
// table, the batch and the output buffers; the only memory the library adds is the fixed
// array below, which is part of the loaded module on every GPU until the process exits.
//
// The scratch holds unitvec(table[answer_rows[a]]) for a < n_answers -- exactly what
// cm_set_answers keeps in engine memory -- so its size bounds the call:
//   n_answers <= CM_NOALLOC_MAX_ANSWERS (64) and dim <= CM_MAX_DIM (1024): 256 KiB.
// Each call first rewrites rows [0, n_answers) in full (an unknown answer gets zeros) and
// reads nothing else, so nothing has to start at zero and stale rows from an earlier call
// are never seen.  All calls share the one array: two calls whose kernels can overlap on the
// GPU (different streams, no ordering between them) race on it -- keep them on one stream or
// order them with events.
//
// In host code the name `g_noalloc_scratch` is only a handle for the symbol, not a device
// address.  The address is looked up with cudaGetSymbolAddress once per device, with that
// device current, the first time the entry point is used on it; make that first call before
// any stream capture begins.  That first call may also be the moment the CUDA runtime loads
// this library's module onto the device (lazy loading), which allocates and synchronises inside
// the driver: it belongs to start-up, not to the request path.  The address is cached for the
// life of the process; cudaDeviceReset() invalidates it (and everything else the process holds
// on that device), so a program that resets a device must not call this entry point on it again.
// ======================================================================================
static __device__ __align__(256) float g_noalloc_scratch[CM_NOALLOC_MAX_ANSWERS * CM_MAX_DIM];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float *g_noalloc_ptr[kMaxDevices] = {};

bool g_noalloc_checked[kMaxDevices] = {};

// Once per device: it exists and is an sm_100 part -- the checks cm_engine_create makes -- so
// that a wrong ordinal is reported as such and not as a cudaSetDevice or launch error.
int noalloc_check_device(int device) {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_checked[device]) return CM_OK;
    int n_dev = 0;
    CM_CUDA(cudaGetDeviceCount(&n_dev));
    if (device >= n_dev) return fail(CM_ERR_INVALID, "device %d not in [0, %d)", device, n_dev);
    int major = 0, minor = 0;
    CM_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    CM_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10)
        return fail(CM_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only",
                    device, major, minor);
    g_noalloc_checked[device] = true;
    return CM_OK;
}

// Device address of the scratch array on `device` (which must be current).
int noalloc_scratch_address(int device, float **out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[device] == nullptr) {
        float *p = nullptr;
        cudaError_t err = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (err != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            return fail(CM_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d failed: %s", device,
                        cudaGetErrorString(err));
        }
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return CM_OK;
}

}  // namespace

extern "C" int cm_score_batch_noalloc(int device, const float *d_table, int64_t n_rows, int dim,
                                      const int32_t *d_answer_rows, int n_answers,
                                      const int32_t *d_guess_idx, const int32_t *d_guess_len,
                                      int n_guesses, int n_slots, const int32_t *d_answer_slot,
                                      const uint8_t *d_exact, float min_score, float *d_scores,
                                      int32_t *d_centi, void *stream) {
    if (device < 0 || device >= kMaxDevices)
        return fail(CM_ERR_INVALID, "device %d outside [0, %d)", device, kMaxDevices);
    if (n_rows <= 0 || n_rows > 0x7fffffffLL)
        return fail(CM_ERR_INVALID, "n_rows %lld outside [1, 2^31)", (long long)n_rows);
    if (dim < 4 || dim > CM_MAX_DIM || dim % 4 != 0)
        return fail(CM_ERR_INVALID, "dim %d must be a multiple of 4 in [4, %d]", dim, CM_MAX_DIM);
    if (n_answers < 1 || n_answers > CM_NOALLOC_MAX_ANSWERS)
        return fail(CM_ERR_INVALID,
                    "n_answers %d outside [1, %d]: the fixed scratch array holds at most %d answer "
                    "vectors; use a cm_engine for more", n_answers, CM_NOALLOC_MAX_ANSWERS,
                    CM_NOALLOC_MAX_ANSWERS);
    if (n_guesses < 0) return fail(CM_ERR_INVALID, "n_guesses %d is negative", n_guesses);
    if (n_slots < 1 || n_slots > CM_MAX_SEQ)
        return fail(CM_ERR_INVALID, "n_slots %d outside [1, %d]", n_slots, CM_MAX_SEQ);
    if (!d_table || !d_answer_rows) return fail(CM_ERR_INVALID, "table or answer_rows is NULL");
    if (n_guesses == 0) return CM_OK;
    if (!d_guess_idx || !d_guess_len || !d_answer_slot || !d_scores)
        return fail(CM_ERR_INVALID, "a required buffer pointer is NULL");

    if (int rc = noalloc_check_device(device)) return rc;
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CM_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    float *scratch = nullptr;
    if (int rc = noalloc_scratch_address(device, &scratch)) return rc;

    // A borrowed, stack-only view in the shape the launchers expect: the caller's table and
    // answer rows, the static array for the unit vectors.  Nothing here owns memory.
    cm_engine view;
    view.device = device;
    view.n_rows = n_rows;
    view.dim = dim;
    view.d_table = const_cast<float *>(d_table);
    view.d_ans_unit = scratch;
    view.d_ans_rows = const_cast<int32_t *>(d_answer_rows);
    view.n_answers = n_answers;
    cudaStream_t st = (cudaStream_t)stream;
#define CM_CALL(V) launch_answers<V>(&view, d_answer_rows, n_answers, st)
    CM_DISPATCH_VPL(dim, CM_CALL)
#undef CM_CALL
    CM_CUDA(cudaGetLastError());
#define CM_CALL(V) launch_score<V>(&view, d_guess_idx, d_guess_len, n_guesses, n_slots, d_answer_slot, \
                                   d_exact, min_score, d_scores, d_centi, 0u, st)
    CM_DISPATCH_VPL(dim, CM_CALL)
#undef CM_CALL
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}
