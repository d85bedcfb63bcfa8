// cassmantle_b200 -- pipelined host path (include/cassmantle_b200.h: cm_pipe_*).
//
// A pipe is a ring of `depth` slots.  Each slot owns a page-locked host block that the caller
// fills in place (no staging copy), a device block of the same layout, a stream and an event.
// cm_pipe_submit() takes one of two routes on the slot's stream:
//   in place     one scoring kernel that reads the batch from the page-locked block across PCIe
//                and writes the scores back into it; the slots' kernels overlap on the GPU (the
//                reads of batch i+1 arrive while batch i is gathering)
//   copy engine  H2D copy of the inputs, kernel, D2H copy of the scores
// chosen by batch size and, for mid-size batches, by timing both (see the constants below).
// Either way the host fills batch i+2 meanwhile.
//
// This is the device-side half of the request micro-batcher that north_star places in
// src/server.py (cassmantle_b200/server.py drives it); the reference has no counterpart.

#include <cuda_runtime.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "cm_internal.h"
#include "cm_route_tuner.h"

struct cm_pipe_slot {
    unsigned char *h = nullptr;     // pinned: [idx | len | slot | exact | scores | centi]
    unsigned char *d = nullptr;     // device: same layout (used by the DMA routes only)
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    int n_guesses = 0, n_slots = 0;
    size_t off_len = 0, off_slot = 0, off_exact = 0, in_bytes = 0, off_scores = 0, off_centi = 0;
    bool in_flight = false;
};

struct cm_pipe {
    cm_engine *engine = nullptr;
    int max_guesses = 0, max_slots = 0;
    size_t slot_bytes = 0;
    std::vector<cm_pipe_slot> slots;
    size_t direct_in_max = 0;       // inputs up to this many bytes are read in place by the kernel
    int direct_out_max = 0;         // up to this many guesses the kernel writes scores in place
    cm::RouteTuner tuner;           // mid-size batches (kSmallIn < inputs <= direct_in_max): cm_route_tuner.h
    bool idx24 = false;             // slots hold 3-byte indices (cm_pipe_set_idx24)
};

namespace {

using cm::align_up;
using cm::fail;

void layout(cm_pipe_slot &s, size_t B, size_t N, bool idx24) {
    s.off_len = align_up(B * N * (idx24 ? 3 : 4), 16);
    s.off_slot = s.off_len + align_up(B * 4, 16);
    s.off_exact = s.off_slot + align_up(B * 4, 16);
    s.in_bytes = align_up(s.off_exact + B, 256);
    s.off_scores = s.in_bytes;
    s.off_centi = s.off_scores + align_up(B * 4, 16);
}

// Route thresholds (measured on B200, tools/probe_pipe.py).  Below them the
// kernel reads the batch from / writes the scores to the page-locked block itself, which saves a
// DMA enqueue per step and lets the PCIe reads of batch i+1 hide under the gather of batch i;
// above them the copy engine moves whole blocks, which small PCIe transactions cannot match.
constexpr size_t kDirectInMax = 640u << 10;   // bytes of packed inputs
constexpr size_t kDirectOutMax = 8192;        // guesses; also caps the direct-input route

// Up to this many input bytes a step is bound by launch/enqueue cost, not by PCIe: always in place.
constexpr size_t kSmallIn = 64u << 10;
// true: this mid-size batch goes in place; false: through the copy engine
bool pick_route(cm_pipe *p) {
    const double now = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    return p->tuner.next(now);
}

size_t env_size(const char *name, size_t fallback) {
    const char *v = getenv(name);
    return v != nullptr && *v != 0 ? (size_t)strtoull(v, nullptr, 10) : fallback;
}

int check_slot(const cm_pipe *p, int slot) {
    if (p == nullptr) return fail(CM_ERR_INVALID, "pipe is NULL");
    if (slot < 0 || slot >= (int)p->slots.size())
        return fail(CM_ERR_INVALID, "slot %d not in [0, %d)", slot, (int)p->slots.size());
    return CM_OK;
}

}  // namespace

extern "C" {

int cm_pipe_destroy(cm_pipe *p) {
    if (p == nullptr) return CM_OK;
    {
        cm::DeviceGuard guard(p->engine->device);
        for (auto &s : p->slots) {
            if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
            if (s.done) cudaEventDestroy(s.done);
            if (s.d) cudaFree(s.d);
            if (s.h) cudaFreeHost(s.h);
        }
    }
    delete p;
    return CM_OK;
}

int cm_pipe_create(cm_pipe **out, cm_engine *e, int depth, int max_guesses, int max_slots) {
    if (out == nullptr) return fail(CM_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (int rc = cm::check_engine(e)) return rc;
    if (depth < 1 || depth > 64) return fail(CM_ERR_INVALID, "depth %d outside [1, 64]", depth);
    if (max_guesses < 1) return fail(CM_ERR_INVALID, "max_guesses %d < 1", max_guesses);
    if (max_slots < 1 || max_slots > CM_MAX_SEQ)
        return fail(CM_ERR_INVALID, "max_slots %d outside [1, %d]", max_slots, CM_MAX_SEQ);
    cm_pipe *p = new (std::nothrow) cm_pipe();
    if (p == nullptr) return fail(CM_ERR_NOMEM, "pipe allocation failed");
    p->engine = e;
    p->max_guesses = max_guesses;
    p->max_slots = max_slots;
    cm_pipe_slot probe;
    layout(probe, (size_t)max_guesses, (size_t)max_slots, false);
    p->slot_bytes = probe.off_centi + align_up((size_t)max_guesses * 4, 256);
    p->slots.resize(depth);
    p->direct_in_max = env_size("CM_PIPE_DIRECT_IN_MAX", kDirectInMax);
    p->direct_out_max = (int)env_size("CM_PIPE_DIRECT_OUT_MAX", kDirectOutMax);
    if (const char *r = getenv("CM_PIPE_ROUTE")) {          // "inplace" | "copy" pin the mid-size route
        if (strcmp(r, "inplace") == 0) p->tuner.pin(cm::RouteTuner::kInPlace);
        else if (strcmp(r, "copy") == 0) p->tuner.pin(cm::RouteTuner::kCopy);
    }
    cm::DeviceGuard guard(e->device);
    for (auto &s : p->slots) {
        cudaError_t err = cudaHostAlloc(&s.h, p->slot_bytes, cudaHostAllocDefault);
        if (err == cudaSuccess) err = cudaMalloc(&s.d, p->slot_bytes);
        if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (err == cudaSuccess) err = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming);
        if (err != cudaSuccess) {
            cudaGetLastError();
            const int code = fail(err == cudaErrorMemoryAllocation ? CM_ERR_NOMEM : CM_ERR_CUDA,
                                  "pipe slot setup (%zu bytes) failed: %s", p->slot_bytes,
                                  cudaGetErrorString(err));
            cm_pipe_destroy(p);
            return code;
        }
    }
    *out = p;
    return CM_OK;
}

int cm_pipe_set_idx24(cm_pipe *p, int on) {
    if (p == nullptr) return fail(CM_ERR_INVALID, "pipe is NULL");
    for (auto &s : p->slots)
        if (s.in_flight) return fail(CM_ERR_STATE, "a slot is in flight");
    if (on && p->engine->n_rows > 0xffffffLL)
        return fail(CM_ERR_INVALID, "3-byte indices need a table of at most 16,777,215 rows");
    p->idx24 = on != 0;
    for (auto &s : p->slots) s.n_guesses = 0;            // layouts change: cm_pipe_buffers must be called again
    return CM_OK;
}

int cm_pipe_buffers(cm_pipe *p, int slot, int n_guesses, int n_slots, int32_t **guess_idx,
                    int32_t **guess_len, int32_t **answer_slot, uint8_t **exact, float **scores,
                    int32_t **centi) {
    if (int rc = check_slot(p, slot)) return rc;
    if (n_guesses < 1 || n_guesses > p->max_guesses)
        return fail(CM_ERR_INVALID, "n_guesses %d outside [1, %d]", n_guesses, p->max_guesses);
    if (n_slots < 1 || n_slots > p->max_slots)
        return fail(CM_ERR_INVALID, "n_slots %d outside [1, %d]", n_slots, p->max_slots);
    cm_pipe_slot &s = p->slots[slot];
    if (s.in_flight) return fail(CM_ERR_STATE, "slot %d is in flight; cm_pipe_wait it first", slot);
    s.n_guesses = n_guesses;
    s.n_slots = n_slots;
    layout(s, (size_t)n_guesses, (size_t)n_slots, p->idx24);
    if (guess_idx) *guess_idx = (int32_t *)s.h;
    if (guess_len) *guess_len = (int32_t *)(s.h + s.off_len);
    if (answer_slot) *answer_slot = (int32_t *)(s.h + s.off_slot);
    if (exact) *exact = s.h + s.off_exact;
    if (scores) *scores = (float *)(s.h + s.off_scores);
    if (centi) *centi = (int32_t *)(s.h + s.off_centi);
    return CM_OK;
}

int cm_pipe_submit(cm_pipe *p, int slot, float min_score, int want_centi) {
    if (int rc = check_slot(p, slot)) return rc;
    cm_pipe_slot &s = p->slots[slot];
    if (s.n_guesses < 1) return fail(CM_ERR_STATE, "slot %d has no batch; call cm_pipe_buffers", slot);
    if (s.in_flight) return fail(CM_ERR_STATE, "slot %d is already in flight", slot);
    cm_engine *e = p->engine;
    cm::DeviceGuard guard(e->device);
    bool direct_out = s.n_guesses <= p->direct_out_max;
    bool direct_in = direct_out && s.in_bytes <= p->direct_in_max;
    if (direct_in && s.in_bytes > kSmallIn && !pick_route(p)) direct_in = direct_out = false;
    const unsigned char *in = direct_in ? s.h : s.d;
    unsigned char *res = direct_out ? s.h : s.d;
    if (!direct_in) CM_CUDA(cudaMemcpyAsync(s.d, s.h, s.in_bytes, cudaMemcpyHostToDevice, s.stream));
    if (int rc = cm::enqueue_score(e, (const int32_t *)in, (const int32_t *)(in + s.off_len),
                                   s.n_guesses, s.n_slots, (const int32_t *)(in + s.off_slot),
                                   in + s.off_exact, min_score, (float *)(res + s.off_scores),
                                   want_centi ? (int32_t *)(res + s.off_centi) : nullptr, 0u, s.stream,
                                   nullptr, p->idx24))
        return rc;
    if (!direct_out) {
        const size_t n = want_centi ? s.off_centi + (size_t)s.n_guesses * 4 - s.off_scores
                                    : (size_t)s.n_guesses * 4;
        CM_CUDA(cudaMemcpyAsync(s.h + s.off_scores, s.d + s.off_scores, n, cudaMemcpyDeviceToHost, s.stream));
    }
    CM_CUDA(cudaEventRecord(s.done, s.stream));
    s.in_flight = true;
    return CM_OK;
}

int cm_pipe_wait(cm_pipe *p, int slot) {
    if (int rc = check_slot(p, slot)) return rc;
    cm_pipe_slot &s = p->slots[slot];
    if (!s.in_flight) return CM_OK;
    cm::DeviceGuard guard(p->engine->device);
    CM_CUDA(cudaEventSynchronize(s.done));
    s.in_flight = false;
    return CM_OK;
}

// Steady-state step of a producer that keeps several batches in flight: wait for an older slot
// (wait_slot < 0: none) and submit a filled one, in a single boundary crossing.
int cm_pipe_cycle(cm_pipe *p, int wait_slot, int submit_slot, float min_score, int want_centi) {
    if (wait_slot >= 0)
        if (int rc = cm_pipe_wait(p, wait_slot)) return rc;
    return cm_pipe_submit(p, submit_slot, min_score, want_centi);
}

// The same for a producer that has `count` batches ready: for i < count, wait for slot
// (wait_first + i) % depth and submit slot (submit_first + i) % depth, so the number of batches in
// flight does not change during the call.  Either index may be negative: no waits / no submits.
int cm_pipe_cycle_many(cm_pipe *p, int wait_first, int submit_first, int count, float min_score,
                       int want_centi) {
    if (p == nullptr) return fail(CM_ERR_INVALID, "pipe is NULL");
    const int depth = (int)p->slots.size();
    if (count < 0 || count > depth) return fail(CM_ERR_INVALID, "count %d outside [0, %d]", count, depth);
    if (wait_first >= depth || submit_first >= depth)
        return fail(CM_ERR_INVALID, "slot index beyond the pipe's %d slots", depth);
    for (int i = 0; i < count; ++i) {
        if (wait_first >= 0)
            if (int rc = cm_pipe_wait(p, (wait_first + i) % depth)) return rc;
        if (submit_first >= 0)
            if (int rc = cm_pipe_submit(p, (submit_first + i) % depth, min_score, want_centi)) return rc;
    }
    return CM_OK;
}

// Developer aid / measurement: the copy-engine route of `steps` steps WITHOUT the kernel -- H2D copy of
// the slot's inputs, D2H copy of its scores, `in_flight` slots busy at a time -- so that a caller can
// tell what the host and the PCIe path alone sustain (bench.py reports it as the ceiling of the
// end-to-end figure when several GPUs share one host).  Slots must have been sized with cm_pipe_buffers.
int cm_pipe_probe_copies(cm_pipe *p, int steps, int in_flight, double *seconds) {
    if (p == nullptr || seconds == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    const int depth = (int)p->slots.size();
    if (in_flight < 1 || in_flight > depth) return fail(CM_ERR_INVALID, "in_flight %d outside [1, %d]", in_flight, depth);
    cm::DeviceGuard guard(p->engine->device);
    for (auto &s : p->slots)
        if (s.in_flight || s.n_guesses < 1) return fail(CM_ERR_STATE, "every slot must be sized and idle");
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; ++i) {
        cm_pipe_slot &s = p->slots[i % depth];
        if (i >= in_flight) CM_CUDA(cudaEventSynchronize(p->slots[(i - in_flight) % depth].done));
        CM_CUDA(cudaMemcpyAsync(s.d, s.h, s.in_bytes, cudaMemcpyHostToDevice, s.stream));
        CM_CUDA(cudaMemcpyAsync(s.h + s.off_scores, s.d + s.off_scores, (size_t)s.n_guesses * 4,
                                cudaMemcpyDeviceToHost, s.stream));
        CM_CUDA(cudaEventRecord(s.done, s.stream));
    }
    for (int i = steps > in_flight ? steps - in_flight : 0; i < steps; ++i)
        CM_CUDA(cudaEventSynchronize(p->slots[i % depth].done));
    *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return CM_OK;
}

int cm_pipe_depth(cm_pipe *p) { return p ? (int)p->slots.size() : 0; }

}  // extern "C"
