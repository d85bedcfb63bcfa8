// Device side of the score all-gather (see cm_gather.cu for the protocol and its rationale).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cm {

constexpr int kMaxPeers = 8;
constexpr uint32_t kShipSentinel = 0xffffffffu;   // a NaN pattern no score can be (NaN -> min_score)
constexpr int kShipChunk = 4096;                  // scores handled by one ship CTA

// Work order for the CTAs that ship ONE finished step of scores to every rank.  world == 0: none.
struct ShipArgs {
    float *dst[kMaxPeers];        // rank r's receive buffer of that step: [world][stride] floats
    uint32_t *flag[kMaxPeers];    // rank r's flag array [world]; flag[r][me] = last step I shipped
    const uint32_t *my_flags;     // my own flag array
    uint32_t *stage;              // local staging slot of that step (scores or sentinel)
    float *copy_to;               // optional: a caller buffer that also receives my scores
    unsigned *counter;            // CTA arrival counter (only touched when several CTAs ship)
    int *err;                     // 0, or who timed out
    unsigned long long *trace;    // developer trace (CM_GATHER_TRACE=1): 8 timestamps per step, or NULL
    long long timeout_cycles;
    uint32_t step;
    int world, rank, stride, n;
    int spin;                     // 0: do not wait for the peers; k > 0: wait until they have shipped step-(k-1)
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__host__ __device__ inline int ship_ctas(int n) { return n > 0 ? (n + kShipChunk - 1) / kShipChunk : 1; }

__device__ __forceinline__ void spin_on_flags(const uint32_t *flags, int world, uint32_t step,
                                              long long timeout_cycles, int *err) {
    const int r = threadIdx.x;
    if (r >= world) return;
    const long long t0 = clock64();
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
        if ((int32_t)(v - step) >= 0) break;
        if (clock64() - t0 > timeout_cycles) {          // a peer died or never launched: do not hang
            atomicExch(err, r + 1);
            break;
        }
        __nanosleep(100);
    }
}

// Executed by `n_ctas` whole CTAs (index `cta`), any block size that is a multiple of 32.
//  0. wait until this rank has published step-1 (ships are strictly ordered)
//  1. poll the staging slot until no sentinel is left in my chunk: the data is the flag, so the
//     scoring launch needs no fence, atomic or tail of its own
//  2. copy the scores into every rank's receive buffer (16-byte stores through peer mappings),
//     optionally into the caller's buffer, and reset the slot to the sentinel
//  3. fence.sys; the last CTA publishes the step in my entry of every peer's flag array
//  4. unless deferred, wait (bounded) until every rank's flag in MY array has reached the step
//     minus a lag: ranks may drift a few steps apart (ring depth) but never further
__device__ __forceinline__ void ship_duty(const ShipArgs &a, int cta, int n_ctas) {
    __shared__ int s_last;
    const int lo = cta * kShipChunk;
    const int hi = min(a.n, lo + kShipChunk);
    const long long t0 = clock64();
    unsigned long long *tr = (a.trace != nullptr && cta == 0 && threadIdx.x == 0) ? a.trace + (a.step % 64u) * 8u : nullptr;
    if (tr) { tr[0] = a.step; tr[1] = global_ns(); }
    // 0. ships run strictly in step order.  A ship CTA does not honour the grid dependency of
    //    its launch, so with short kernels the ship of step s+depth can be resident before step
    //    s has been shipped -- and both use the same staging slot.  My own entry of my flag
    //    array says which step I published last: wait until it is step-1.
    if (threadIdx.x == 0) {
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.my_flags + a.rank) : "memory");
            if ((int32_t)(v - (a.step - 1u)) >= 0) break;
            if (clock64() - t0 > a.timeout_cycles) {
                atomicExch(a.err, -(int)a.step);
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    if (tr) tr[2] = global_ns();
    // Up to kBatch 16-byte pieces per thread are asked for at once (independent loads: one L2 round trip,
    // not kBatch of them -- the ship of one step used to spend 5.5 us here, and with two fences behind it
    // took longer than the 11 us scoring step it rides in); a piece that still holds a sentinel is polled.
    constexpr int kBatch = 8;
    const int stride = 4 * (int)blockDim.x;
    for (int base = lo + 4 * (int)threadIdx.x; base < hi; base += stride * kBatch) {
        uint4 v[kBatch];
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const int i = base + j * stride;
            if (i < hi)
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(a.stage + i) : "memory");
        }
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const int i = base + j * stride;
            if (i >= hi) continue;
            for (;;) {
                const bool pending = v[j].x == kShipSentinel || (i + 1 < hi && v[j].y == kShipSentinel) ||
                                     (i + 2 < hi && v[j].z == kShipSentinel) || (i + 3 < hi && v[j].w == kShipSentinel);
                if (!pending) break;
                if (clock64() - t0 > a.timeout_cycles) {
                    atomicExch(a.err, -(int)a.step);        // the scoring launch of that step never delivered
                    break;
                }
                __nanosleep(100);
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(a.stage + i) : "memory");
            }
#pragma unroll
            for (int r = 0; r < kMaxPeers; ++r)
                if (r < a.world)
                    *reinterpret_cast<uint4 *>(a.dst[r] + (size_t)a.rank * a.stride + i) = v[j];
            if (a.copy_to != nullptr) {
                a.copy_to[i] = __uint_as_float(v[j].x);
                if (i + 1 < hi) a.copy_to[i + 1] = __uint_as_float(v[j].y);
                if (i + 2 < hi) a.copy_to[i + 2] = __uint_as_float(v[j].z);
                if (i + 3 < hi) a.copy_to[i + 3] = __uint_as_float(v[j].w);
            }
            *reinterpret_cast<uint4 *>(a.stage + i) =
                make_uint4(kShipSentinel, kShipSentinel, kShipSentinel, kShipSentinel);
        }
    }
    if (tr) tr[3] = global_ns();
    // One system-scope fence, by the threads that publish, AFTER the barrier: fences are cumulative, so it
    // covers every thread's peer stores that the barrier ordered before it.  (A fence per thread before the
    // barrier and another before the flags cost 2.5-3.5 us each at 8 GPUs: measured, cm_gather_trace.)
    __syncthreads();
    if (tr) tr[4] = global_ns();
    if (threadIdx.x == 0) {
        if (n_ctas == 1) {
            s_last = 1;
        } else {
            __threadfence_system();                  // this CTA's stores, before the other CTAs may publish for it
            s_last = atomicAdd(a.counter, 1u) == (unsigned)n_ctas - 1u;
            if (s_last) *a.counter = 0u;
        }
    }
    __syncthreads();
    // one thread per peer publishes: eight release stores from one thread cost ~2 us EACH
    // (each waits for the previous remote store), eight threads overlap them
    if (s_last && (int)threadIdx.x < a.world) {
        __threadfence_system();
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(a.flag[threadIdx.x] + a.rank), "r"(a.step)
                     : "memory");
    }
    __syncthreads();
    if (tr) tr[5] = global_ns();
    if (s_last && a.spin > 0 && a.step >= (uint32_t)a.spin)
        spin_on_flags(a.my_flags, a.world, a.step - (uint32_t)(a.spin - 1), a.timeout_cycles, a.err);
    if (tr) tr[6] = global_ns();
}

}  // namespace cm
