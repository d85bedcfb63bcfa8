// cassmantle_b200 -- score all-gather over NVLink peer memory (include/...: cm_gather_*).
//
// SURVEY.md §8e / north_star: every GPU holds a table replica and scores a slice of the batch;
// the per-guess scores are then needed on every rank.  That exchange is 4 bytes per guess, so
// a library collective is pure latency (measured on 2 B200: a 16 KiB NCCL all-gather costs
// ~18 us per step next to a 13 us scoring kernel).  Here the exchange rides inside the
// scoring launches themselves, one step behind:
//
//   launch s      scores step s into a local staging slot that was pre-filled with a sentinel
//                 bit pattern (0xffffffff, a NaN no score can be: the floor maps NaN to
//                 min_score).  The scoring CTAs carry no fence, atomic or tail for the gather.
//   launch s+1    has one extra CTA per 4096 scores (cm_ship.cuh: ship_duty).  While the other
//                 CTAs gather rows for step s+1 it polls the slot of step s until no sentinel is
//                 left (the data is the flag), copies the scores into every rank's receive
//                 buffer with 16-byte peer stores, resets the slot, fence.sys, publishes the
//                 step in its entry of every peer's flag array (st.release.sys) and waits,
//                 bounded, until every rank's flag in ITS array has reached the step.
//   cm_gather_wait  ships the last step with a stand-alone kernel when nothing follows it.
//
// So a stream of gathered steps costs ONE launch per step and the exchange never sits on the
// critical path.  Ordering: launch k+1's scoring CTAs store only after griddepcontrol.wait,
// i.e. after launch k has finished including its ship CTA, so a staging slot (ring of `depth`
// >= 2) is reset before it is written again.  A ship waits for the peers with a lag of depth/2
// steps (a wait for the very same step couples all ranks every launch: 25.5 us/step on 8 GPUs),
// so ranks drift at most depth/2 steps apart and a receive slot is safe for depth/2 launches.
//
// Designs measured and dropped (2 B200, 4096x16, us per step; one GPU alone: 12.7):
//   NCCL all_gather_into_tensor after every scoring launch                              30.3
//   per-warp peer stores + system fences inside the scoring kernel                      21.4
//   last-CTA tail inside the scoring kernel (an atomic per CTA: +25 % at 16k CTAs)      19.5
//   exchange kernel chained behind the scoring kernel on the same stream                16.7
//     (the next launch parks on its grid dependency holding every SM slot meanwhile)
//   ship kernel on a high-priority side stream, sentinel polling                        17.6
//     (two launches + an event per step: host-bound at 12.6 us per call)
//
// Peer blocks are exchanged as CUDA IPC handles by the host language (torch.distributed is the
// plumbing in dist.py); cm_gather_import_ptr links blocks inside one process.  No reference
// counterpart.

#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>

#include "cm_common.cuh"
#include "cm_internal.h"

struct cm_gather {
    cm_engine *engine = nullptr;
    int rank = 0, world = 0, stride = 0, depth = 0;
    unsigned char *block = nullptr;   // [depth][world][stride] f32 | stage [depth][stride] | flags | counter | err
    size_t block_bytes = 0, off_stage = 0, off_flags = 0, off_counter = 0, off_err = 0, off_trace = 0;
    bool trace = false;
    unsigned char *peer[cm::kMaxPeers] = {};    // mapped blocks of every rank (peer[rank] == block)
    bool ipc_opened[cm::kMaxPeers] = {};
    uint32_t step = 0;                          // last step launched (scored)
    uint32_t shipped = 0;                       // last step whose ship has been enqueued
    int n_of_step = 0;                          // guesses of step `step`
    float *copy_of_step = nullptr;              // caller's d_scores of step `step` (may be NULL)
    bool defer_of_step = false;
};

namespace {

using cm::fail;

using cm::ShipArgs;

// Stand-alone ship: used when no scoring launch follows the step (cm_gather_wait, empty slices).
__global__ void __launch_bounds__(128) cm_gather_ship_kernel(ShipArgs a) {
    cm::ship_duty(a, (int)blockIdx.x, (int)gridDim.x);
}

__global__ void cm_gather_wait_kernel(const uint32_t *flags, int world, uint32_t step,
                                      long long timeout_cycles, int *err) {
    cm::spin_on_flags(flags, world, step, timeout_cycles, err);
}

bool all_linked(const cm_gather *g) {
    for (int r = 0; r < g->world; ++r)
        if (g->peer[r] == nullptr) return false;
    return true;
}

size_t buffer_offset(const cm_gather *g, uint32_t step) {
    return (size_t)(step % (uint32_t)g->depth) * g->world * g->stride * sizeof(float);
}

constexpr long long kTimeoutCycles = 4000000000LL;   // ~2 s at 2 GHz: far beyond a healthy step

}  // namespace

extern "C" {

int cm_gather_destroy(cm_gather *g) {
    if (g == nullptr) return CM_OK;
    {
        cm::DeviceGuard guard(g->engine->device);
        cudaDeviceSynchronize();
        for (int r = 0; r < g->world; ++r)
            if (g->ipc_opened[r] && g->peer[r]) cudaIpcCloseMemHandle(g->peer[r]);
        if (g->block) cudaFree(g->block);
    }
    delete g;
    return CM_OK;
}

int cm_gather_create(cm_gather **out, cm_engine *e, int rank, int world, int max_guesses_per_rank,
                     int depth) {
    if (out == nullptr) return fail(CM_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (int rc = cm::check_engine(e)) return rc;
    if (world < 1 || world > cm::kMaxPeers)
        return fail(CM_ERR_INVALID, "world %d outside [1, %d]", world, cm::kMaxPeers);
    if (rank < 0 || rank >= world) return fail(CM_ERR_INVALID, "rank %d not in [0, %d)", rank, world);
    if (max_guesses_per_rank < 1) return fail(CM_ERR_INVALID, "max_guesses_per_rank < 1");
    if (depth < 2 || depth > 16) return fail(CM_ERR_INVALID, "depth %d outside [2, 16]", depth);
    cm_gather *g = new (std::nothrow) cm_gather();
    if (g == nullptr) return fail(CM_ERR_NOMEM, "gather allocation failed");
    g->engine = e; g->rank = rank; g->world = world; g->depth = depth;
    g->stride = (int)cm::align_up((size_t)max_guesses_per_rank, 64);
    g->off_stage = cm::align_up((size_t)depth * world * g->stride * sizeof(float), 256);
    g->off_flags = g->off_stage + cm::align_up((size_t)depth * g->stride * sizeof(float), 256);
    g->off_counter = g->off_flags + 256;
    g->off_err = g->off_counter + 256;
    g->off_trace = g->off_err + 256;
    g->block_bytes = g->off_trace + 64 * 8 * sizeof(unsigned long long);
    g->trace = getenv("CM_GATHER_TRACE") != nullptr;
    cm::DeviceGuard guard(e->device);
    cudaError_t err = cudaMalloc(&g->block, g->block_bytes);
    if (err == cudaSuccess) err = cudaMemset(g->block, 0, g->block_bytes);
    if (err == cudaSuccess)     // staging slots start out "nothing has landed"
        err = cudaMemset(g->block + g->off_stage, 0xff, g->off_flags - g->off_stage);
    if (err == cudaSuccess) err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        cudaGetLastError();
        const int code = fail(CM_ERR_NOMEM, "gather block (%zu bytes) failed: %s", g->block_bytes,
                              cudaGetErrorString(err));
        cm_gather_destroy(g);
        return code;
    }
    g->peer[rank] = g->block;
    *out = g;
    return CM_OK;
}

int cm_gather_export(cm_gather *g, void *handle64) {
    if (g == nullptr || handle64 == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is exchanged as 64 bytes");
    cm::DeviceGuard guard(g->engine->device);
    cudaIpcMemHandle_t h;
    CM_CUDA(cudaIpcGetMemHandle(&h, g->block));
    memcpy(handle64, &h, 64);
    return CM_OK;
}

int cm_gather_import(cm_gather *g, int peer_rank, const void *handle64) {
    if (g == nullptr || handle64 == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    if (peer_rank < 0 || peer_rank >= g->world || peer_rank == g->rank)
        return fail(CM_ERR_INVALID, "peer_rank %d invalid for rank %d of %d", peer_rank, g->rank, g->world);
    cm::DeviceGuard guard(g->engine->device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *p = nullptr;
    CM_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    g->peer[peer_rank] = (unsigned char *)p;
    g->ipc_opened[peer_rank] = true;
    return CM_OK;
}

int cm_gather_import_ptr(cm_gather *g, int peer_rank, void *d_block) {
    if (g == nullptr || d_block == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    if (peer_rank < 0 || peer_rank >= g->world || peer_rank == g->rank)
        return fail(CM_ERR_INVALID, "peer_rank %d invalid for rank %d of %d", peer_rank, g->rank, g->world);
    g->peer[peer_rank] = (unsigned char *)d_block;
    return CM_OK;
}

int cm_gather_block(cm_gather *g, void **d_block, int64_t *bytes) {
    if (g == nullptr) return fail(CM_ERR_INVALID, "gather is NULL");
    if (d_block) *d_block = g->block;
    if (bytes) *bytes = (int64_t)g->block_bytes;
    return CM_OK;
}

// Work order that ships step `g->step` (the latest scored one).
static ShipArgs make_ship(const cm_gather *g, bool spin) {
    ShipArgs a = {};
    const size_t buf = buffer_offset(g, g->step);
    for (int r = 0; r < g->world; ++r) {
        a.dst[r] = (float *)(g->peer[r] + buf);
        a.flag[r] = (uint32_t *)(g->peer[r] + g->off_flags);
    }
    a.my_flags = (const uint32_t *)(g->block + g->off_flags);
    a.stage = (uint32_t *)(g->block + g->off_stage) + (size_t)(g->step % (uint32_t)g->depth) * g->stride;
    a.copy_to = g->copy_of_step;
    a.counter = (unsigned *)(g->block + g->off_counter);
    a.err = (int *)(g->block + g->off_err);
    a.trace = g->trace ? (unsigned long long *)(g->block + g->off_trace) : nullptr;
    a.timeout_cycles = kTimeoutCycles;
    a.step = g->step;
    a.world = g->world;
    a.rank = g->rank;
    a.stride = g->stride;
    a.n = g->n_of_step;
    // lagged coupling: a rank may run depth/2 steps ahead of the slowest peer, not more, so a
    // receive slot is never overwritten while its owner may still be reading it
    a.spin = spin ? 1 + g->depth / 2 : 0;
    return a;
}

// Ships the latest step with a stand-alone kernel if no launch has taken it along yet.
static int flush_pending(cm_gather *g, cudaStream_t st) {
    if (g->shipped == g->step) return CM_OK;
    const ShipArgs a = make_ship(g, !g->defer_of_step);
    cm_gather_ship_kernel<<<cm::ship_ctas(a.n), 128, 0, st>>>(a);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    g->shipped = g->step;
    return CM_OK;
}

int cm_score_batch_gather(cm_engine *e, cm_gather *g, const int32_t *d_guess_idx,
                          const int32_t *d_guess_len, int n_guesses, int n_slots,
                          const int32_t *d_answer_slot, const uint8_t *d_exact, float min_score,
                          float *d_scores, int32_t *d_centi, uint32_t flags, void *stream) {
    if (g == nullptr) return fail(CM_ERR_INVALID, "gather is NULL");
    if (g->engine != e) return fail(CM_ERR_INVALID, "gather belongs to another engine");
    if (!all_linked(g)) return fail(CM_ERR_STATE, "not every peer block has been imported");
    if (n_guesses < 0 || n_guesses > g->stride)
        return fail(CM_ERR_INVALID, "n_guesses %d outside [0, %d]", n_guesses, g->stride);
    if (flags & ~(uint32_t)(CM_SCORE_OVERLAP_PREVIOUS | CM_GATHER_DEFER_WAIT))
        return fail(CM_ERR_INVALID, "unknown flag bits 0x%x", flags);
    if (!(min_score == min_score)) return fail(CM_ERR_INVALID, "min_score is NaN");
    cudaStream_t st = (cudaStream_t)stream;
    cm::DeviceGuard guard(e->device);
    const uint32_t step = g->step + 1;
    uint32_t *stage = (uint32_t *)(g->block + g->off_stage) + (size_t)(step % (uint32_t)g->depth) * g->stride;
    if (n_guesses == 0) {
        // nothing to launch for this step: get the previous one out, this one ships as "empty"
        if (int rc = flush_pending(g, st)) return rc;
    } else {
        ShipArgs prev = {};
        const bool take_along = g->shipped != g->step;      // the previous step still waits to be shipped
        if (take_along) prev = make_ship(g, !g->defer_of_step);
        if (int rc = cm::enqueue_score(e, d_guess_idx, d_guess_len, n_guesses, n_slots, d_answer_slot,
                                       d_exact, min_score, (float *)stage, d_centi,
                                       flags & CM_SCORE_OVERLAP_PREVIOUS, st, take_along ? &prev : nullptr))
            return rc;
        if (take_along) g->shipped = g->step;
    }
    g->step = step;
    g->n_of_step = n_guesses;
    g->copy_of_step = d_scores;
    g->defer_of_step = (flags & CM_GATHER_DEFER_WAIT) != 0;
    return CM_OK;
}

// Several batches in one boundary crossing (a micro-batcher flushing more than one queue, or a
// producer that has k steps ready): the calls are enqueued back to back, in order, exactly as k
// separate cm_score_batch_device / cm_score_batch_gather calls would be.
int cm_score_batch_list_device(cm_engine *e, cm_gather *g, const cm_batch *batches, int n_batches,
                               float min_score, uint32_t flags, void *stream) {
    if (n_batches < 0) return fail(CM_ERR_INVALID, "n_batches %d is negative", n_batches);
    if (n_batches > 0 && batches == nullptr) return fail(CM_ERR_INVALID, "batches is NULL");
    for (int i = 0; i < n_batches; ++i) {
        const cm_batch &b = batches[i];
        const int rc = g != nullptr
            ? cm_score_batch_gather(e, g, b.guess_idx, b.guess_len, b.n_guesses, b.n_slots, b.answer_slot,
                                    b.exact, min_score, b.scores, b.centi, flags, stream)
            : cm_score_batch_device(e, b.guess_idx, b.guess_len, b.n_guesses, b.n_slots, b.answer_slot,
                                    b.exact, min_score, b.scores, b.centi, flags, stream);
        if (rc != CM_OK) return rc;
    }
    return CM_OK;
}

int cm_gather_result(cm_gather *g, float **d_gathered) {
    if (g == nullptr || d_gathered == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    if (g->step == 0) return fail(CM_ERR_STATE, "no gathered launch yet");
    *d_gathered = (float *)(g->block + buffer_offset(g, g->step));
    return CM_OK;
}

int cm_gather_flush(cm_gather *g, void *stream) {
    if (g == nullptr) return fail(CM_ERR_INVALID, "gather is NULL");
    cm::DeviceGuard guard(g->engine->device);
    return flush_pending(g, (cudaStream_t)stream);
}

int cm_gather_wait(cm_gather *g, float **d_gathered, void *stream) {
    if (g == nullptr) return fail(CM_ERR_INVALID, "gather is NULL");
    if (g->step == 0) return fail(CM_ERR_STATE, "no gathered launch yet");
    cm::DeviceGuard guard(g->engine->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = flush_pending(g, st)) return rc;
    // a deferred ship has not waited for the peers: do it now (a no-op when they are all there)
    cm_gather_wait_kernel<<<1, 32, 0, st>>>((const uint32_t *)(g->block + g->off_flags), g->world,
                                           g->step, kTimeoutCycles, (int *)(g->block + g->off_err));
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    if (d_gathered) *d_gathered = (float *)(g->block + buffer_offset(g, g->step));
    return CM_OK;
}

int cm_gather_status(cm_gather *g, int *timed_out_rank_plus_1) {
    if (g == nullptr || timed_out_rank_plus_1 == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    cm::DeviceGuard guard(g->engine->device);
    CM_CUDA(cudaMemcpy(timed_out_rank_plus_1, g->block + g->off_err, sizeof(int), cudaMemcpyDeviceToHost));
    return CM_OK;
}

int cm_gather_stride(cm_gather *g) { return g ? g->stride : 0; }

// Developer aid (CM_GATHER_TRACE=1): the last 64 ship traces, 8 x u64 each:
// step, start, order-wait done, stores issued, fence done, flags published, peer wait done (ns).
int cm_gather_trace(cm_gather *g, unsigned long long *out512) {
    if (g == nullptr || out512 == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    cm::DeviceGuard guard(g->engine->device);
    CM_CUDA(cudaMemcpy(out512, g->block + g->off_trace, 64 * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return CM_OK;
}

}  // extern "C"
