// cassmantle_b200 -- most_similar for a BATCH of query words in one pass over the table
// (include/...: cm_most_similar_batch).
//
// Reference: Backend.most_similar, src/backend.py:297-301, called once per word; the micro-batcher
// (cassmantle_b200/server.py) can hold many such calls.  One query is a 3M x 300 mat-vec bound by the
// 3.6 GB table read (cm_similar.cu, 580 us).  Q queries are  [3M x 300] . [300 x Q]  -- the one
// GEMM-shaped piece of work on this path -- and the table is read once for all of them:
//
//   cm_query_batch_kernel  mean[q] = unitvec(vectors[w_q] / norms[w_q])              (as cm_query_kernel)
//   cm_simscan_tc_kernel   tcgen05.mma kind::tf32, M = 128 table rows per tile, N = Q (padded to 16),
//                          K = 300 in ten 128-byte chunks fed by TMA (SWIZZLE_128B, out-of-range columns
//                          zero-filled), fp32 accumulators in tensor memory, double-buffered so the epilogue
//                          of one tile overlaps the MMAs of the next.  One persistent CTA per SM:
//                          warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-9 = epilogue.  Above 64 queries
//                          per pass the two CTAs of an SM pair share 256-row tiles (cta_group::2), each
//                          holding half of the query block.
//                          The tf32 cosines are only a FILTER (|error| < 2^-9, see kBand):
//                            mode 0  sample pass over every S-th tile: approximate cosines written out,
//                                    cm_kth_kernel takes the k-th largest per query as threshold tau_q
//                            mode 1  full pass: every (row, q) with cosine >= tau_q - band is a candidate, appended by
//                                    the epilogue warp to its own list (cm_bucket_kernel sorts the lists by query)
//   cm_finish_kernel       per query: candidates cut down to those within the error band of the k-th largest tf32
//                          cosine, re-scored in fp32 with exactly the arithmetic of the one-query scan (same lanes,
//                          same FMA order, dot / norm), sorted (cosine desc, row asc), top k out.
// The k-th largest of a sample is a lower bound of the k-th largest overall, and the band covers the tf32
// error on both sides, so every true top-k row is among the candidates: rows and cosines are those of Q
// calls of cm_most_similar, bit for bit (ties at the k-th place may resolve differently, as they may there).
// A query with more candidates than kCandCap (or fewer than k) is finished by the one-query path.

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>

#include "cm_common.cuh"
#include "cm_internal.h"
#include "cm_similar_state.h"

namespace {

using cm::fail;

constexpr int kTileRows = 128;          // UMMA M
constexpr int kChunkFloats = 32;        // one 128-byte swizzle row of K
constexpr int kMaxChunks = 10;          // dim <= 320
constexpr int kMaxStages = 8;           // A ring: up to 8 x 16 KB (what the query block leaves free)
constexpr int kStageBytes = kTileRows * kChunkFloats * 4;
constexpr int kMaxNQ = 256;             // queries per pass (UMMA N): 128 on one CTA per SM, 256 on CTA pairs
constexpr int kScanThreads = 320;        // TMA warp, MMA warp, eight epilogue warps
constexpr unsigned kCandCap = 16384;    // candidates per query the full pass may collect
constexpr int kEpilogueWarps = 8;
constexpr unsigned kEmitCap = 4096;     // (tf32 key, row, query) entries one epilogue warp may emit per pass (64 KB)
// tf32 keeps 10 mantissa bits (truncation): each product is off by < (2^-9 + 2^-20) relative, so a cosine (unit
// query, division by the row norm) by < 1.954e-3 absolute, plus the fp32 accumulation of 300 terms (< 2e-5).
// Threshold and candidate are both approximate: twice that = 3.95e-3, and a margin (tests/test_tf32_band.py).
constexpr float kBand = 4.4e-3f;
constexpr unsigned long long kWaitCycles = 4000000000ull;   // ~2 s: a wait that long is a bug, trap instead of hanging

// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded: records `code` and traps when the barrier does not flip within kWaitCycles.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned *err, unsigned code) {
    if (mbar_try_wait(bar, parity)) return;
    const unsigned long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > kWaitCycles) {
            atomicExch(err, code);
            __threadfence_system();
            __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, tf32 inputs, fp32 accumulate (ACC = false overwrites D)
template <bool ACC>
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "n"(ACC ? 1 : 0) : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major operand tile in the SWIZZLE_128B layout TMA writes: rows of 128 bytes, 8-row groups 1024 bytes apart.
// (cute::UMMA::SmemDescriptor: start >> 4 | LBO 1 << 16 | SBO 64 << 32 | version 1 << 46 | SWIZZLE_128B 2 << 61)
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
    const uint32_t lo = ((smem_addr >> 4) & 0x3fffu) | (1u << 16);
    const uint32_t hi = 64u | (1u << 14) | (2u << 29);
    return ((uint64_t)hi << 32) | lo;
}

enum : unsigned { kErrNone = 0, kErrEmpty = 1, kErrFull = 2, kErrQueries = 3, kErrTmemEmpty = 4, kErrTmemFull = 5 };

struct ScanArgs {
    const float *norms;
    int64_t n_rows;
    int n_visits;            // tiles (of 128 rows per CTA of the unit) this launch visits: tile = visit * tile_stride
    int tile_stride;
    int nq;                  // UMMA N: queries padded to a multiple of 16, <= kMaxNQ
    int dim;
    int mode;                // 0 = sample (write cosines), 1 = filter (collect candidates)
    int stages;              // A-ring depth (2 .. kMaxStages): what fits beside the query block
    float *sample_out;       // [nq][sample_ld]
    int64_t sample_ld;
    const float *tau;        // [nq] thresholds, band already subtracted
    uint4 *emit;             // [gridDim.x * kEpilogueWarps][kEmitCap]: (key of the tf32 cosine, row, query, 0), per epilogue warp
    unsigned *emit_cnt;      // [gridDim.x * kEpilogueWarps] entries written (> kEmitCap: the warp overflowed)
    unsigned *err;
};

// --------------------------------------------------------------------------------------------------
// CTAS = 1: one CTA per SM, M = 128.  CTAS = 2: the two CTAs of a cluster (one SM pair) work on one tile of 256 rows
// with tcgen05.mma.cta_group::2 -- each loads its 128 table rows and HALF of the query block (the hardware reads the
// other half from the partner's shared memory), so the query block costs half the shared memory per SM and the ring
// is twice as deep at the same number of queries.  Rank 0 issues the MMAs for both; every TMA load of the pair
// completes on rank 0's barrier (no multicast: each CTA loads its own bytes, all from this GPU's memory), and
// tcgen05.commit arrives on the same barrier in both CTAs.
constexpr uint32_t kLeaderMask = 0xFEFFFFFFu;            // shared::cluster address of the same offset in CTA rank 0

template <int CTAS>
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    if constexpr (CTAS == 1) {
        tma_load_2d(dst, map, c0, c1, bar);
    } else {
        asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     :: "r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & kLeaderMask), "r"(c0), "r"(c1) : "memory");
    }
}
template <int CTAS>
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
    if constexpr (CTAS == 1) {
        tc_commit(bar);
    } else {
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     :: "r"(bar), "h"((uint16_t)3) : "memory");
    }
}
template <int CTAS, bool ACC>
__device__ __forceinline__ void tc_mma_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc) {
    if constexpr (CTAS == 1) {
        tc_mma_tf32<ACC>(d_tmem, a_desc, b_desc, idesc);
    } else {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "setp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                     :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "n"(ACC ? 1 : 0) : "memory");
    }
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int CTAS>
__global__ void __launch_bounds__(kScanThreads, 1)
cm_simscan_tc_kernel(const __grid_constant__ CUtensorMap map_table, const __grid_constant__ CUtensorMap map_queries,
                     const ScanArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // carve: [A ring][this CTA's share of the queries: n_chunks x nq / CTAS x 128 B][tau][barriers][tmem ptr]
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int n_chunks = (a.dim + kChunkFloats - 1) / kChunkFloats;
    const int nq_local = a.nq / CTAS;
    const uint32_t q_chunk_bytes = (uint32_t)nq_local * 128u;
    unsigned char *s_a = smem;
    unsigned char *s_q = s_a + a.stages * kStageBytes;
    float *s_tau = reinterpret_cast<float *>(s_q + (size_t)n_chunks * q_chunk_bytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_tau + kMaxNQ);
    uint64_t *bar_full = bars, *bar_empty = bars + kMaxStages, *bar_q = bars + 2 * kMaxStages;
    uint64_t *bar_tfull = bars + 2 * kMaxStages + 1, *bar_tempty = bars + 2 * kMaxStages + 3;
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 5);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank = 0;
    if constexpr (CTAS == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int unit = (int)blockIdx.x / CTAS, n_units = (int)gridDim.x / CTAS;     // a unit = the CTA(s) sharing tiles
    const int my_visits = a.n_visits > unit ? (a.n_visits - unit + n_units - 1) / n_units : 0;
    constexpr int kUnitRows = kTileRows * CTAS;
    uint32_t tmem_cols = 32;
    while (tmem_cols < 2u * (uint32_t)a.nq) tmem_cols <<= 1;                      // two accumulators of nq columns

    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kMaxStages; ++i) { mbar_init(smem_u32(bar_full + i), 1); mbar_init(smem_u32(bar_empty + i), 1); }
        mbar_init(smem_u32(bar_q), 1);
        for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(bar_tfull + i), 1); mbar_init(smem_u32(bar_tempty + i), 8 * CTAS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                                      // one warp (of each CTA) allocates tensor memory, and frees it at the end
        if constexpr (CTAS == 1) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(s_tmem)), "r"(tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(s_tmem)), "r"(tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    for (int i = threadIdx.x; i < kMaxNQ; i += kScanThreads) s_tau[i] = (a.mode == 1 && i < a.nq) ? a.tau[i] : 0.f;
    tc_fence_before();
    __syncthreads();
    if constexpr (CTAS == 2) cluster_sync_all();          // the partner's barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == 0) {
        // ===== TMA producer (both CTAs of a pair; rank 0 posts the byte counts for both) =====
        if (lane == 0 && my_visits > 0) {
            if (rank == 0) mbar_expect_tx(smem_u32(bar_q), (uint32_t)n_chunks * q_chunk_bytes * CTAS);
            for (int c = 0; c < n_chunks; ++c)
                tma_load_2d_pair<CTAS>(smem_u32(s_q) + (uint32_t)c * q_chunk_bytes, &map_queries, c * kChunkFloats, (int)rank * nq_local, smem_u32(bar_q));
            uint32_t s = 0, ph = 0;
            for (int v = 0; v < my_visits; ++v) {
                const int row0 = (unit + v * n_units) * a.tile_stride * kUnitRows + (int)rank * kTileRows;
                for (int c = 0; c < n_chunks; ++c) {
                    mbar_wait(smem_u32(bar_empty + s), ph ^ 1u, a.err, kErrEmpty);
                    if (rank == 0) mbar_expect_tx(smem_u32(bar_full + s), kStageBytes * CTAS);
                    tma_load_2d_pair<CTAS>(smem_u32(s_a) + s * kStageBytes, &map_table, c * kChunkFloats, row0, smem_u32(bar_full + s));
                    if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (rank 0 only) =====
        if (lane == 0 && my_visits > 0 && rank == 0) {
            // instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate, tf32 x tf32, both K-major, N, M = 128 per CTA
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(a.nq >> 3) << 17) | ((uint32_t)(kUnitRows >> 4) << 24);
            mbar_wait(smem_u32(bar_q), 0, a.err, kErrQueries);
            // The issuing thread is the serial resource of this kernel (one instruction per 8 columns of K): descriptors
            // are running sums, the ring position a counter, the four MMAs of a chunk straight-line code.
            const uint64_t da0 = umma_desc_k_sw128(smem_u32(s_a));
            const uint64_t db0 = umma_desc_k_sw128(smem_u32(s_q));
            const uint64_t db_step = q_chunk_bytes >> 4, da_step = kStageBytes >> 4;
            const int last_steps = (a.dim - (n_chunks - 1) * kChunkFloats + 7) / 8;           // tf32: K = 8 per instruction
            uint32_t s = 0, ph = 0;
            uint64_t da = da0;
            for (int v = 0; v < my_visits; ++v) {
                const uint32_t acc = (uint32_t)v & 1u, tph = ((uint32_t)v >> 1) & 1u;
                mbar_wait(smem_u32(bar_tempty + acc), tph ^ 1u, a.err, kErrTmemEmpty);     // the epilogues have drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * (uint32_t)a.nq;
                uint64_t db = db0;
                for (int c = 0; c < n_chunks; ++c) {
                    mbar_wait(smem_u32(bar_full + s), ph, a.err, kErrFull);
                    tc_fence_after();
                    if (c == 0) tc_mma_pair<CTAS, false>(d_tmem, da, db, idesc);              // +32 bytes of K = +2 in the address field
                    else tc_mma_pair<CTAS, true>(d_tmem, da, db, idesc);
                    if (c + 1 < n_chunks || last_steps == 4) {
                        tc_mma_pair<CTAS, true>(d_tmem, da + 2, db + 2, idesc);
                        tc_mma_pair<CTAS, true>(d_tmem, da + 4, db + 4, idesc);
                        tc_mma_pair<CTAS, true>(d_tmem, da + 6, db + 6, idesc);
                    } else {
                        for (int k = 1; k < last_steps; ++k) tc_mma_pair<CTAS, true>(d_tmem, da + 2 * k, db + 2 * k, idesc);
                    }
                    tc_commit_pair<CTAS>(smem_u32(bar_empty + s));                            // stage free (in both CTAs) once these MMAs have read it
                    db += db_step;
                    da += da_step;
                    if (++s == (uint32_t)a.stages) { s = 0; ph ^= 1u; da = da0; }
                }
                tc_commit_pair<CTAS>(smem_u32(bar_tfull + acc));                              // accumulator complete (in both CTAs)
            }
        }
    } else {
        // ===== epilogue: warps 2..9.  A warp may only read TMEM lanes 32 * (warp % 4) .. + 32 (= this CTA's tile rows); the
        // two warps of a quarter split the query columns in blocks of 16 =====
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row_in_unit = (int)rank * kTileRows + quarter * 32 + lane;
        // this row's norm, fetched two tiles ahead: under the scan's own traffic a load takes about as long as a tile
        auto norm_of = [&](int v) -> float {
            if (v >= my_visits) return 1.f;
            const int64_t r = (int64_t)(unit + v * n_units) * a.tile_stride * kUnitRows + row_in_unit;
            return r < a.n_rows ? a.norms[r] : __int_as_float(0x7fc00000);          // rows past the end: NaN, never a candidate
        };
        float nrm0 = norm_of(0), nrm1 = norm_of(1);
        const size_t emit_id = (size_t)blockIdx.x * kEpilogueWarps + (warp - 2);
        uint4 *my_emit = a.emit + emit_id * kEmitCap;
        unsigned cursor = 0;
        for (int v = 0; v < my_visits; ++v) {
            const int visit = unit + v * n_units;
            const uint32_t acc = (uint32_t)v & 1u, tph = ((uint32_t)v >> 1) & 1u;
            const int64_t row = (int64_t)visit * a.tile_stride * kUnitRows + row_in_unit;
            const float nrm2 = norm_of(v + 2);
            const float inv = 1.0f / nrm0;
            nrm0 = nrm1;
            nrm1 = nrm2;
            mbar_wait(smem_u32(bar_tfull + acc), tph, a.err, kErrTmemFull);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * (uint32_t)a.nq;
            for (int c0 = 16 * half; c0 < a.nq; c0 += 32) {
                float d[16];
                __syncwarp();                                    // tcgen05.ld is warp-wide: rejoin after the emit branch
                tc_ld16(taddr + (uint32_t)c0, d);
                if (a.mode == 0) {
                    float *out = a.sample_out + (int64_t)c0 * a.sample_ld + (int64_t)visit * kUnitRows + row_in_unit;
#pragma unroll
                    for (int i = 0; i < 16; ++i) out[(int64_t)i * a.sample_ld] = d[i] * inv;
                } else {
                    // straight-line compare of 16 cosines against their thresholds (broadcast reads), one rarely taken branch
                    const float4 *t4 = reinterpret_cast<const float4 *>(s_tau + c0);
                    unsigned hits = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 t = t4[i];
                        d[4 * i + 0] *= inv; d[4 * i + 1] *= inv; d[4 * i + 2] *= inv; d[4 * i + 3] *= inv;
                        hits |= (d[4 * i + 0] >= t.x ? 1u : 0u) << (4 * i);
                        hits |= (d[4 * i + 1] >= t.y ? 1u : 0u) << (4 * i + 1);
                        hits |= (d[4 * i + 2] >= t.z ? 1u : 0u) << (4 * i + 2);
                        hits |= (d[4 * i + 3] >= t.w ? 1u : 0u) << (4 * i + 3);
                    }
                    // hits are rare (about k * S per query and pass): the warp appends them to its own list, the cursor
                    // a register -- an atomicAdd whose result is needed would cost a memory round trip per tile
                    while (__any_sync(cm::kFullMask, hits != 0u)) {
                        const bool has = hits != 0u;
                        const unsigned vote = __ballot_sync(cm::kFullMask, has);
                        const int i = has ? __ffs(hits) - 1 : 0;
                        hits &= hits - 1u;
                        const unsigned pos = cursor + (unsigned)__popc(vote & ((1u << lane) - 1u));
                        if (has && pos < kEmitCap)
                            my_emit[pos] = make_uint4(cm::similar_key_of(d[i]), (uint32_t)row, (uint32_t)(c0 + i), 0u);
                        cursor += (unsigned)__popc(vote);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {                                     // the issuer (rank 0) counts the epilogue warps of both CTAs
                if constexpr (CTAS == 1) mbar_arrive(smem_u32(bar_tempty + acc));
                // relaxed: what the arrival publishes are TMEM reads, ordered by the tcgen05 fence above -- a release at
                // cluster scope would also wait for this warp's global stores (18 % of the stall samples of the first version)
                else asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" :: "r"(smem_u32(bar_tempty + acc) & kLeaderMask) : "memory");
            }
        }
        if (a.mode == 1 && lane == 0) a.emit_cnt[emit_id] = cursor;
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (CTAS == 2) cluster_sync_all();          // neither CTA leaves while the other may still signal it
    if (warp == 2) {
        tc_fence_after();
        if constexpr (CTAS == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(tmem_cols) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}


// mean[q] = unitvec(vectors[w_q] / norms[w_q]) -- one warp per query, the arithmetic of cm_query_kernel; rows of
// `mean` past n_queries (padding up to the UMMA N) are zeroed.
__global__ void cm_query_batch_kernel(const float *__restrict__ table, int dim, const int32_t *__restrict__ rows,
                                      int n_queries, const float *__restrict__ norms, float *__restrict__ mean) {
    const int lane = threadIdx.x, q = blockIdx.x;
    float *m = mean + (size_t)q * dim;
    if (q >= n_queries) {
        for (int j = lane; j < dim; j += 32) m[j] = 0.f;
        return;
    }
    const int32_t row = rows[q];
    const float nrm = norms[row];
    double ss = 0.0;
    for (int j = lane; j < dim; j += 32) {
        const float x = table[(size_t)row * dim + j] / nrm;          // get_vector(key, norm=True)
        m[j] = x;
        ss = fma((double)x, (double)x, ss);
    }
    ss = cm::warp_sum(ss);
    __syncwarp();
    if (ss > 0.0) {                                                  // matutils.unitvec
        const double inv = 1.0 / sqrt(ss);
        for (int j = lane; j < dim; j += 32) m[j] = (float)((double)m[j] * inv);
    }
}

__device__ __forceinline__ float float_of_key(uint32_t key) {        // inverse of similar_key_of (NaN was 0)
    if (key == 0u) return __int_as_float(0x7fc00000);
    return __uint_as_float((key & 0x80000000u) ? (key & 0x7fffffffu) : ~key);
}

// The top 8 * PASSES bits of the k-th largest of key_at(0 .. n), the rest zero (so: a lower bound of that key that at
// least k keys reach).  Radix select from the top, 8 bits per pass; called by every thread of the CTA.
// hist: 256 x 32 counters in shared memory -- [bin][lane], so the 32 adds of a warp never share a bank.
template <int PASSES, class KeyAt>
__device__ uint32_t block_kth_prefix(KeyAt key_at, int64_t n, unsigned k, unsigned *hist, uint32_t *s_prefix, unsigned *s_rem) {
    if (threadIdx.x == 0) { *s_prefix = 0u; *s_rem = k; }
    for (int pass = 0; pass < PASSES; ++pass) {
        for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x) hist[i] = 0u;
        __syncthreads();
        const uint32_t prefix = *s_prefix;
        const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (32 - 8 * pass));
        const int shift = 24 - 8 * pass;
        for (int64_t i0 = threadIdx.x; i0 < n; i0 += 8 * (int64_t)blockDim.x) {     // eight loads in flight per thread
            uint32_t key[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int64_t i = i0 + u * (int64_t)blockDim.x;
                key[u] = i < n ? key_at(i) : 0u;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (i0 + u * (int64_t)blockDim.x < n && (key[u] & mask) == prefix)
                    atomicAdd(&hist[((key[u] >> shift) & 0xffu) * 32 + (threadIdx.x & 31)], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 256) {                                      // fold the 32 copies into copy 0
            unsigned sum = 0;
            for (int l = 0; l < 32; ++l) sum += hist[threadIdx.x * 32 + ((l + threadIdx.x) & 31)];
            __syncwarp();
            hist[threadIdx.x * 32] = sum;
        }
        __syncthreads();
        if (threadIdx.x < 32) {                                       // which bin, from the top, holds the rem-th entry
            const int lane = threadIdx.x;
            const int first = 255 - 8 * lane;                         // this lane's bins: first, first - 1, ..., first - 7
            const unsigned rem = *s_rem;
            unsigned cnt8[8], mine = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) { cnt8[i] = hist[(first - i) * 32]; mine += cnt8[i]; }
            unsigned incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned up = __shfl_up_sync(cm::kFullMask, incl, o);
                if (lane >= o) incl += up;
            }
            const unsigned reached = __ballot_sync(cm::kFullMask, incl >= rem);
            const int owner = reached ? __ffs(reached) - 1 : 31;
            if (lane == owner) {
                unsigned acc = incl - mine;
                int b = first - 7;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (acc + cnt8[i] >= rem) { b = first - i; break; }
                    if (i < 7) acc += cnt8[i];
                }
                *s_rem = rem - acc;
                *s_prefix = prefix | ((uint32_t)b << shift);
            }
        }
        __syncthreads();
    }
    return *s_prefix;
}

// tau[q] = (a lower bound, good to 16 key bits = 1/128 of its value, of the k-th largest of the n sampled cosines of
// query q) - band; one CTA per query.
__global__ void __launch_bounds__(1024)
cm_kth_kernel(const float *__restrict__ sample, int64_t ld, int64_t n, unsigned k, float band, float *__restrict__ tau, int n_real) {
    __shared__ unsigned hist[256 * 32];
    __shared__ uint32_t s_prefix;
    __shared__ unsigned s_rem;
    const float *d = sample + (int64_t)blockIdx.x * ld;
    if ((int)blockIdx.x >= n_real) {                                  // padding up to the UMMA N: nothing is a candidate
        if (threadIdx.x == 0) tau[blockIdx.x] = __int_as_float(0x7f800000);
        return;
    }
    const uint32_t prefix = block_kth_prefix<2>([d](int64_t i) { return cm::similar_key_of(__ldcg(d + i)); }, n, k, hist, &s_prefix, &s_rem);
    if (threadIdx.x == 0) {
        const float kth = float_of_key(prefix);
        tau[blockIdx.x] = (kth == kth) ? kth - band : -__int_as_float(0x7f800000);    // fewer than k valid cosines: everything passes
    }
}

// Exact cosine of `row` against the query held in q[] -- the lane layout, FMA order and reduction tree of the one-query
// scans (cm_scan16_kernel: 16 lanes x 5 float4, xor tree 8,4,2,1; cm_scan_kernel: 32 lanes, xor tree 16..1).
template <int LPG, int VPL>
__device__ __forceinline__ float exact_cosine(const float4 *__restrict__ t4, int dim4, int64_t row, const float4 (&q)[VPL],
                                              int sub, float norm) {
    float4 buf[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        const int j = sub + LPG * v;
        buf[v] = j < dim4 ? __ldg(t4 + row * dim4 + j) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float s = 0.f;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        if (LPG == 32 && sub + LPG * v >= dim4) continue;            // the 32-lane scan skips absent pieces, the 16-lane one adds zeros
        s = fmaf(buf[v].x, q[v].x, s); s = fmaf(buf[v].y, q[v].y, s);
        s = fmaf(buf[v].z, q[v].z, s); s = fmaf(buf[v].w, q[v].w, s);
    }
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) s += __shfl_xor_sync(cm::kFullMask, s, o);
    return s / norm;
}

// The epilogue warps' lists sorted into per-query candidate lists: one CTA per list.  Counts per query in shared memory,
// one global atomicAdd per (list, query) reserves the range, a second walk over the list scatters.
// A list that overflowed poisons every query (cnt = 2^31): the host then runs them one by one.
__global__ void __launch_bounds__(256)
cm_bucket_kernel(const uint4 *__restrict__ emit, const unsigned *__restrict__ emit_cnt, int nq,
                 unsigned long long *__restrict__ cand, unsigned *__restrict__ cnt) {
    __shared__ unsigned hist[kMaxNQ], cursor[kMaxNQ];
    const unsigned n_raw = emit_cnt[blockIdx.x];
    if (n_raw > kEmitCap) {
        for (int q = threadIdx.x; q < nq; q += blockDim.x) atomicOr(cnt + q, 0x80000000u);
        return;
    }
    if (n_raw == 0) return;
    const uint4 *list = emit + (size_t)blockIdx.x * kEmitCap;
    for (int q = threadIdx.x; q < kMaxNQ; q += blockDim.x) hist[q] = 0u;
    __syncthreads();
    for (unsigned i = threadIdx.x; i < n_raw; i += blockDim.x) atomicAdd(&hist[list[i].z], 1u);
    __syncthreads();
    for (int q = threadIdx.x; q < nq; q += blockDim.x) cursor[q] = hist[q] ? atomicAdd(cnt + q, hist[q]) : 0u;
    __syncthreads();
    for (unsigned i = threadIdx.x; i < n_raw; i += blockDim.x) {
        const uint4 e = list[i];
        const unsigned pos = atomicAdd(&cursor[e.z], 1u);
        if (pos < kCandCap) cand[(size_t)e.z * kCandCap + pos] = ((unsigned long long)e.x << 32) | e.y;
    }
}

// One CTA per query finishes it from its candidates (tf32 cosine, row):
//   1. a_k = the k-th largest tf32 cosine (to 16 key bits, rounded down).  A true top-k row has an exact cosine of at
//      least a_k - eps (k rows have tf32 cosines >= a_k), so a tf32 cosine of at least a_k - 2 eps: keep those only
//   2. the survivors (about k of them) re-scored exactly, sorted (cosine desc, row asc), the best k written out.
// n_out[q] = k, or -1 when the candidates overflowed kCandCap, fewer than k were collected, or more than kFinishCap
// survive step 1 (the host then runs the one-query path for q).
constexpr unsigned kFinishCap = 2048;

template <int LPG, int VPL>
__global__ void __launch_bounds__(1024)
cm_finish_kernel(const float *__restrict__ table, int dim4, const float *__restrict__ mean, const float *__restrict__ norms,
                 const unsigned long long *__restrict__ cand, const unsigned *__restrict__ cnt, unsigned k, float band,
                 int32_t *__restrict__ rows_out, float *__restrict__ sims_out, int *__restrict__ n_out) {
    __shared__ __align__(16) unsigned hist[256 * 32];                 // later the sort buffer
    __shared__ int32_t s_rows[kFinishCap];
    __shared__ uint32_t s_prefix;
    __shared__ unsigned s_rem, s_m;
    unsigned long long *s_keys = reinterpret_cast<unsigned long long *>(hist);
    static_assert(sizeof(hist) >= kFinishCap * sizeof(unsigned long long), "sort buffer");
    const int q_id = blockIdx.x;
    const unsigned n = cnt[q_id];
    if (n > kCandCap || n < k) {
        if (threadIdx.x == 0) n_out[q_id] = -1;
        return;
    }
    const unsigned long long *c = cand + (size_t)q_id * kCandCap;
    if (threadIdx.x == 0) s_m = 0u;
    const uint32_t prefix = block_kth_prefix<2>([c](int64_t i) { return (uint32_t)(__ldcg(c + i) >> 32); }, n, k, hist, &s_prefix, &s_rem);
    const uint32_t keep = cm::similar_key_of(float_of_key(prefix) - band);
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned long long e = __ldcg(c + i);
        if ((uint32_t)(e >> 32) >= keep) {
            const unsigned slot = atomicAdd(&s_m, 1u);
            if (slot < kFinishCap) s_rows[slot] = (int32_t)(e & 0xffffffffull);
        }
    }
    __syncthreads();
    const unsigned m = s_m;
    if (m > kFinishCap) {                                            // uniform over the CTA
        if (threadIdx.x == 0) n_out[q_id] = -1;
        return;
    }
    const int lane = threadIdx.x & 31, sub = lane & (LPG - 1);
    const float4 *t4 = reinterpret_cast<const float4 *>(table);
    float4 q[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        const int j = sub + LPG * v;
        q[v] = j < dim4 ? reinterpret_cast<const float4 *>(mean + (size_t)q_id * dim4 * 4)[j] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    constexpr unsigned kGroups = 1024 / LPG;
    for (unsigned base = (threadIdx.x >> 5) * (32 / LPG); base < m; base += kGroups) {   // whole warps (the reduction shuffles)
        const unsigned i = base + (unsigned)(lane / LPG);
        const bool have = i < m;
        const int32_t row = have ? s_rows[i] : 0;
        const float d = exact_cosine<LPG, VPL>(t4, dim4, row, q, sub, norms[row]);
        if (have && sub == 0) s_keys[i] = ((unsigned long long)(~cm::similar_key_of(d)) << 32) | (uint32_t)row;
    }
    unsigned size = 64;
    while (size < m) size <<= 1;
    for (unsigned i = m + threadIdx.x; i < size; i += blockDim.x) s_keys[i] = ~0ull;
    __syncthreads();
    for (unsigned span = 2; span <= size; span <<= 1) {              // bitonic, ascending = best first
        for (unsigned stride = span >> 1; stride > 0; stride >>= 1) {
            for (unsigned p = threadIdx.x; p < size / 2; p += blockDim.x) {
                const unsigned lo = 2 * p - (p & (stride - 1));
                const unsigned hi = lo + stride;
                const bool up = (lo & span) == 0;
                const unsigned long long x = s_keys[lo], y = s_keys[hi];
                if ((x > y) == up) { s_keys[lo] = y; s_keys[hi] = x; }
            }
            __syncthreads();
        }
    }
    for (unsigned i = threadIdx.x; i < k; i += blockDim.x) {
        const unsigned long long e = s_keys[i];
        rows_out[(size_t)q_id * cm::kSimilarMaxTopK + i] = (int32_t)(e & 0xffffffffull);
        sims_out[(size_t)q_id * cm::kSimilarMaxTopK + i] = float_of_key(~(uint32_t)(e >> 32));
    }
    if (threadIdx.x == 0) n_out[q_id] = (int)k;
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------
struct cm_similar_batch {
    float *d_mean = nullptr;             // [kMaxNQ][dim]
    int32_t *d_rows = nullptr;           // [kMaxNQ]
    float *d_tau = nullptr;              // [kMaxNQ]
    unsigned *d_cnt = nullptr;           // [kMaxNQ]
    unsigned *d_err = nullptr;
    unsigned long long *d_cand = nullptr; // [kMaxNQ][kCandCap] (key of the tf32 cosine, row)
    uint4 *d_emit = nullptr;             // [sm_count * kEpilogueWarps][kEmitCap] the epilogue warps' own lists
    unsigned *d_emit_cnt = nullptr;      // [sm_count * kEpilogueWarps]
    float *d_sample = nullptr;           // [kMaxNQ][sample_ld]
    size_t sample_floats = 0;
    unsigned char *h_result = nullptr;   // page-locked: rows [kMaxNQ][1024] | sims [kMaxNQ][1024] | n [kMaxNQ] | err
    CUtensorMap map_table;
    const float *map_table_ptr = nullptr;
    int64_t map_table_rows = 0;
    int sm_count = 148;
    bool attrs_set = false;
};

namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            p = nullptr;
        }
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// [n_rows][dim] fp32, row-major; boxes of box_rows x 32 floats in the 128-byte swizzle, columns past dim read as zero
int make_map(CUtensorMap *map, const float *base, int64_t n_rows, int dim, int box_rows) {
    EncodeTiledFn fn = encode_tiled();
    if (fn == nullptr) return fail(CM_ERR_CUDA, "the driver does not export cuTensorMapEncodeTiled");
    const cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)n_rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)dim * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)kChunkFloats, (cuuint32_t)box_rows};
    const cuuint32_t estride[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), gdim, gstride, box, estride,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(CM_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return CM_OK;
}

constexpr size_t kSmemLimit = 232448;                  // 227 KB, the most one CTA may have
constexpr size_t kScanSmemFixed = 1024 + kMaxNQ * 4 + 32 * 8;

// A-ring depth that fits beside the query block
int scan_stages(int nq, int n_chunks) {
    const size_t left = kSmemLimit - kScanSmemFixed - (size_t)n_chunks * nq * 128;
    const int st = (int)(left / kStageBytes);
    return st > kMaxStages ? kMaxStages : st;
}
size_t scan_smem(int nq, int n_chunks, int stages) { return kScanSmemFixed + (size_t)n_chunks * nq * 128 + (size_t)stages * kStageBytes; }
constexpr size_t kResultRows = (size_t)kMaxNQ * cm::kSimilarMaxTopK;

int batch_state(cm_engine *e, cm_similar_state *s, cm_similar_batch **out) {
    if (s->batch == nullptr) {
        cm_similar_batch *b = new (std::nothrow) cm_similar_batch();
        if (b == nullptr) return fail(CM_ERR_NOMEM, "most_similar batch state allocation failed");
        s->batch = b;
        cudaError_t err = cudaMalloc(&b->d_mean, (size_t)kMaxNQ * e->dim * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_rows, kMaxNQ * sizeof(int32_t));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_tau, kMaxNQ * sizeof(float));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_cnt, kMaxNQ * sizeof(unsigned));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_err, sizeof(unsigned));
        if (err == cudaSuccess) err = cudaMemset(b->d_err, 0, sizeof(unsigned));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_cand, (size_t)kMaxNQ * kCandCap * sizeof(unsigned long long));
        if (err == cudaSuccess) err = cudaHostAlloc(&b->h_result, kResultRows * 8 + kMaxNQ * sizeof(int), cudaHostAllocDefault);
        if (err == cudaSuccess) {
            int sms = 0;
            if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device) == cudaSuccess && sms > 0) b->sm_count = sms;
        }
        if (err == cudaSuccess) err = cudaMalloc(&b->d_emit, (size_t)b->sm_count * kEpilogueWarps * kEmitCap * sizeof(uint4));
        if (err == cudaSuccess) err = cudaMalloc(&b->d_emit_cnt, (size_t)b->sm_count * kEpilogueWarps * sizeof(unsigned));
        if (err != cudaSuccess) {
            cudaGetLastError();
            cm::similar_batch_free(s);
            return fail(CM_ERR_NOMEM, "most_similar batch scratch failed: %s", cudaGetErrorString(err));
        }
    }
    *out = s->batch;
    return CM_OK;
}

template <int LPG, int VPL>
void launch_finish(cm_engine *e, cm_similar_state *s, cm_similar_batch *b, int n_real, unsigned kk, int32_t *h_rows, float *h_sims, int *h_n) {
    cm_finish_kernel<LPG, VPL><<<n_real, 1024, 0, e->stream>>>(e->d_table, e->dim / 4, b->d_mean, s->d_norms, b->d_cand, b->d_cnt, kk, kBand,
                                                               h_rows, h_sims, h_n);
    cm::count_launch();
}

}  // namespace

void cm::similar_batch_free(cm_similar_state *s) {
    cm_similar_batch *b = s->batch;
    if (b == nullptr) return;
    void *ptrs[] = {b->d_mean, b->d_rows, b->d_tau, b->d_cnt, b->d_err, b->d_cand, b->d_emit, b->d_emit_cnt, b->d_sample};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (b->h_result) cudaFreeHost(b->h_result);
    delete b;
    s->batch = nullptr;
}

extern "C" int cm_most_similar_batch(cm_engine *e, const int32_t *rows, int n_queries, int k,
                                     int32_t *rows_out, float *sims_out, int *n_out) {
    if (int rc = cm::check_engine(e)) return rc;
    if (n_queries < 0) return fail(CM_ERR_INVALID, "n_queries %d is negative", n_queries);
    if (k < 1 || k > cm::kSimilarMaxTopK) return fail(CM_ERR_INVALID, "k %d outside [1, %d]", k, cm::kSimilarMaxTopK);
    if (n_queries == 0) return CM_OK;
    if (!rows || !rows_out || !sims_out || !n_out) return fail(CM_ERR_INVALID, "a buffer pointer is NULL");
    for (int q = 0; q < n_queries; ++q)
        if (rows[q] < 0 || (int64_t)rows[q] >= e->n_rows)
            return fail(CM_ERR_INVALID, "rows[%d] = %d outside the %lld-row table", q, rows[q], (long long)e->n_rows);
    std::lock_guard<std::mutex> lock(e->mu);
    cm::DeviceGuard guard(e->device);
    cm_similar_state *s = nullptr;
    if (int rc = cm::similar_prepare(e, &s)) return rc;
    const unsigned kk = (unsigned)((int64_t)k < e->n_rows ? k : e->n_rows);
    // Test hook / escape hatch: CM_SIMILAR_BATCH=0 runs the queries one by one.
    static const bool batch_off = [] { const char *v = getenv("CM_SIMILAR_BATCH"); return v != nullptr && *v == '0'; }();
    const bool tensor_path = !batch_off && n_queries >= 2 && e->dim % 4 == 0 && e->dim >= kChunkFloats && e->dim <= kMaxChunks * kChunkFloats &&
                             e->n_rows >= 2 * kTileRows;              // a TMA box is not larger than the table
    if (!tensor_path) {
        for (int q = 0; q < n_queries; ++q)
            if (int rc = cm::most_similar_locked(e, rows[q], k, rows_out + (size_t)q * k, sims_out + (size_t)q * k, n_out + q)) return rc;
        return CM_OK;
    }
    cm_similar_batch *b = nullptr;
    if (int rc = batch_state(e, s, &b)) return rc;
    // The scan wants all of an SM's shared memory: beside a CTA of the resident micro-batch kernel (cm_engine_arm_resident)
    // it would never be scheduled.  That kernel is stopped here and comes back with the next blocking score call.
    cm::resident_stop(e);
    cudaStream_t st = e->stream;
    if (!b->attrs_set) {
        CM_CUDA(cudaFuncSetAttribute(cm_simscan_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLimit));
        CM_CUDA(cudaFuncSetAttribute(cm_simscan_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLimit));
        b->attrs_set = true;
    }
    // CTA pairs (tcgen05 cta_group::2) for passes of more than 64 queries: half the query block per SM, so the ring stays
    // eight deep up to 128 queries (684 against 744 us per pass) and 256 queries fit at all.  CM_SIMILAR_PAIR=0|1 forces it.
    static const int pair_forced = [] { const char *v = getenv("CM_SIMILAR_PAIR"); return v != nullptr && *v != 0 ? atoi(v) : -1; }();
    if (b->map_table_ptr != e->d_table || b->map_table_rows != e->n_rows) {
        if (int rc = make_map(&b->map_table, e->d_table, e->n_rows, e->dim, kTileRows)) return rc;
        b->map_table_ptr = e->d_table;
        b->map_table_rows = e->n_rows;
    }
    int32_t *h_rows = reinterpret_cast<int32_t *>(b->h_result);
    float *h_sims = reinterpret_cast<float *>(b->h_result + kResultRows * 4);
    int *h_n = reinterpret_cast<int *>(b->h_result + kResultRows * 8);
    const int dim4 = e->dim / 4;
    // measured per query at 160 / 192 / 224 / 256 queries per pass: 6.3 / 6.1 / 5.9 / 5.6 us -- the largest pass wins
    const int pass_max = pair_forced == 0 ? kMaxNQ / 2 : kMaxNQ;
    for (int q0 = 0; q0 < n_queries; q0 += pass_max) {
        const int n_real = n_queries - q0 < pass_max ? n_queries - q0 : pass_max;
        const int nq = (n_real + 15) / 16 * 16;
        const bool pair = pair_forced >= 0 ? pair_forced != 0 : nq > 64;
        const int ctas = pair ? 2 : 1;
        // tiles of 128 rows per CTA of a unit (a CTA, or a CTA pair); the sample is every S-th tile, S chosen so that
        // about 33,000 rows are sampled and the k-th largest of the sample admits about k * S candidates
        const int unit_rows = kTileRows * ctas;
        const int n_tiles = (int)((e->n_rows + unit_rows - 1) / unit_rows);
        int stride = n_tiles / (256 / ctas);
        if (stride > (int)(kCandCap / 4) / (int)kk) stride = (int)(kCandCap / 4) / (int)kk;
        if (stride > 128) stride = 128;
        if (stride < 1) stride = 1;
        const int n_sample_visits = (n_tiles + stride - 1) / stride;
        const int64_t sample_ld = (int64_t)n_sample_visits * unit_rows;
        if (b->sample_floats < (size_t)sample_ld * kMaxNQ) {
            CM_CUDA(cudaStreamSynchronize(st));
            if (b->d_sample) cudaFree(b->d_sample);
            b->d_sample = nullptr;
            b->sample_floats = 0;
            CM_CUDA(cudaMalloc(&b->d_sample, (size_t)sample_ld * kMaxNQ * sizeof(float)));
            b->sample_floats = (size_t)sample_ld * kMaxNQ;
        }
        CM_CUDA(cudaMemcpyAsync(b->d_rows, rows + q0, n_real * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        CM_CUDA(cudaMemsetAsync(b->d_cnt, 0, kMaxNQ * sizeof(unsigned), st));
        cm_query_batch_kernel<<<nq, 32, 0, st>>>(e->d_table, e->dim, b->d_rows, n_real, s->d_norms, b->d_mean);
        cm::count_launch();
        CUtensorMap map_q;
        if (int rc = make_map(&map_q, b->d_mean, nq, e->dim, nq / ctas)) return rc;      // a CTA of a pair loads its half of the block
        ScanArgs a = {};
        a.norms = s->d_norms;
        a.n_rows = e->n_rows;
        a.nq = nq;
        a.dim = e->dim;
        a.sample_out = b->d_sample;
        a.sample_ld = sample_ld;
        a.tau = b->d_tau;
        a.emit = b->d_emit;
        a.emit_cnt = b->d_emit_cnt;
        a.err = b->d_err;
        const int n_chunks = (e->dim + kChunkFloats - 1) / kChunkFloats;
        a.stages = scan_stages(nq / ctas, n_chunks);
        const size_t smem = scan_smem(nq / ctas, n_chunks, a.stages);
        auto launch_scan = [&](int n_units_to_visit) {
            int grid = b->sm_count / ctas < n_units_to_visit ? b->sm_count / ctas : n_units_to_visit;      // units: CTAs or CTA pairs
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(grid * ctas));
            cfg.blockDim = dim3(kScanThreads);
            cfg.dynamicSmemBytes = smem;
            cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned)ctas;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            if (pair) cudaLaunchKernelEx(&cfg, cm_simscan_tc_kernel<2>, b->map_table, map_q, a);
            else cudaLaunchKernelEx(&cfg, cm_simscan_tc_kernel<1>, b->map_table, map_q, a);
            cm::count_launch();
            return grid * ctas;
        };
        // sample pass
        a.mode = 0;
        a.n_visits = n_sample_visits;
        a.tile_stride = stride;
        launch_scan(a.n_visits);
        cm_kth_kernel<<<nq, 1024, 0, st>>>(b->d_sample, sample_ld, sample_ld, kk, kBand, b->d_tau, n_real);
        cm::count_launch();
        // full pass
        a.mode = 1;
        a.n_visits = n_tiles;
        a.tile_stride = 1;
        const int scan_ctas = launch_scan(a.n_visits);
        cm_bucket_kernel<<<scan_ctas * kEpilogueWarps, 256, 0, st>>>(b->d_emit, b->d_emit_cnt, nq, b->d_cand, b->d_cnt);
        cm::count_launch();
        if (cm::similar_uses_scan16(dim4)) launch_finish<16, 5>(e, s, b, n_real, kk, h_rows, h_sims, h_n);
        else if (dim4 <= 32) launch_finish<32, 1>(e, s, b, n_real, kk, h_rows, h_sims, h_n);
        else if (dim4 <= 64) launch_finish<32, 2>(e, s, b, n_real, kk, h_rows, h_sims, h_n);
        else launch_finish<32, 3>(e, s, b, n_real, kk, h_rows, h_sims, h_n);
        CM_CUDA(cudaGetLastError());
        const cudaError_t sync = cudaStreamSynchronize(st);
        if (sync != cudaSuccess)
            return fail(CM_ERR_CUDA, "most_similar batch scan failed: %s (a pipeline wait inside the kernel may have timed out)",
                        cudaGetErrorString(sync));
        for (int q = 0; q < n_real; ++q) {
            int32_t *ro = rows_out + (size_t)(q0 + q) * k;
            float *so = sims_out + (size_t)(q0 + q) * k;
            if (h_n[q] >= 0) {
                memcpy(ro, h_rows + (size_t)q * cm::kSimilarMaxTopK, kk * sizeof(int32_t));
                memcpy(so, h_sims + (size_t)q * cm::kSimilarMaxTopK, kk * sizeof(float));
                n_out[q0 + q] = h_n[q];
            } else if (int rc = cm::most_similar_locked(e, rows[q0 + q], k, ro, so, n_out + q0 + q)) {
                return rc;
            }
        }
    }
    return CM_OK;
}
