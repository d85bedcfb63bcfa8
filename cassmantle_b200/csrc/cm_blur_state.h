// Host-side state of a cm_blur object, shared by cm_blur.cu (the blur) and cm_jpeg.cu (the JPEG
// encode of the blurred images).  Not installed.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>

struct cm_jpeg_state;             // scratch of cm_blur_batch_jpeg (cm_jpeg.cu), allocated on first use

struct cm_blur {
    int device = 0;
    int max_h = 0, max_w = 0, channels = 0, max_batch = 0;
    int h = 0, w = 0;                 // current image
    uint8_t *d_src = nullptr;         // the round's image
    uint8_t *d_a = nullptr;           // ping  [max_batch][h*w*c]
    uint8_t *d_b = nullptr;           // pong
    uint8_t *h_pin = nullptr;         // pinned staging for results
    void *d_params = nullptr;         // BoxParams[max_batch]
    int max_box_radius = 0;           // largest box radius among the uploaded requests (chooses the blur kernel)
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // results of chunk i travel while chunk i+1 is blurred
    cudaEvent_t blurred[2] = {nullptr, nullptr};
    cm_jpeg_state *jpeg = nullptr;
    std::mutex mu;
};

namespace cm {

// Enqueues the blur of the current image for n sessions on b->stream (b->mu held, device current);
// the n blurred images end up back to back in b->d_b.  No copy, no synchronisation.
int blur_enqueue(cm_blur *b, const float *radii, int n);
void jpeg_state_free(cm_blur *b);          // cm_jpeg.cu

}  // namespace cm
