// Host-side internals shared by the translation units of libcassmantle_b200.so (not installed).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>

#include "../../include/cassmantle_b200.h"
#include "cm_ship.cuh"

struct cm_similar_state;   // scratch of cm_most_similar (cm_similar.cu), allocated on first use
struct cm_resident;        // the resident micro-batch kernel's mailbox and state (cm_scorer.cu)

struct cm_engine {
    int device = 0;
    int64_t n_rows = 0;
    int dim = 0;
    float *d_table = nullptr;
    float *d_ans_unit = nullptr;     // [CM_MAX_ANSWERS][dim] unit vectors of the round's answers
    int32_t *d_ans_rows = nullptr;   // [CM_MAX_ANSWERS] their table rows (-1 = not in vocabulary)
    int n_answers = 0;
    cudaStream_t stream = nullptr;
    // pinned staging for the host-buffer entry point when the caller's memory is pageable
    unsigned char *h_stage = nullptr;
    size_t stage_bytes = 0;
    cm_similar_state *similar = nullptr;
    cm_resident *resident = nullptr;         // non-NULL once cm_engine_arm_resident has been called
    cudaEvent_t answers_done = nullptr;      // recorded after cm_set_answers launches while `resident` exists
    bool answers_pending = false;            // the resident path must wait for answers_done once
    std::mutex mu;
};

namespace cm {

// Records the message for cm_last_error() on this thread and returns `code`.
int fail(int code, const char *fmt, ...);
void count_launch();

#define CM_CUDA(expr)                                                                       \
    do {                                                                                    \
        cudaError_t err__ = (expr);                                                         \
        if (err__ != cudaSuccess)                                                           \
            return cm::fail(CM_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(err__)); \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// True when `p` is page-locked host memory the GPU can address (cudaHostAlloc / cudaHostRegister).
bool is_pinned_host(const void *p);

int check_engine(const cm_engine *e);

void similar_state_free(cm_engine *e);     // cm_similar.cu
void similar_reserve(cm_engine *e);        // cm_similar.cu: allocate cm_most_similar's scratch now (best effort)
int ensure_stage(cm_engine *e, size_t bytes);   // e->h_stage holds at least `bytes` (call with e->mu held)
void similar_invalidate(cm_engine *e);     // the table changed: cached row norms are stale
void resident_stop(cm_engine *e);          // stop the resident kernel if it runs (it is re-launched on demand)
void resident_free(cm_engine *e);

// Validates the arguments and enqueues the scoring kernel; pointers may be device memory or
// pinned host memory (the kernel then reads / writes it across PCIe itself).  `ship`, if not
// NULL, adds CTAs to the launch that ship an EARLIER step's scores to the other ranks.
int enqueue_score(cm_engine *e, const int32_t *guess_idx, const int32_t *guess_len, int n_guesses,
                  int n_slots, const int32_t *answer_slot, const uint8_t *exact, float min_score,
                  float *scores, int32_t *centi, uint32_t flags, cudaStream_t st,
                  const ShipArgs *ship = nullptr, bool idx24 = false);   // idx24: guess_idx holds 3-byte indices

}  // namespace cm
