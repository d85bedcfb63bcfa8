// cassmantle_b200 -- per-session image blur on the device (include/...: cm_blur_*).
//
// SURVEY.md §8f rank 3.  Reference: Backend.score_to_blur + Backend.mask_image,
// src/backend.py:319-324, called per request from Server.fetch_masked_image
// (src/server.py:129-133):
//     blur  = min_blur + (1 - score**2) * (max_blur - min_blur)
//     image.filter(ImageFilter.GaussianBlur(blur))
// Pillow's GaussianBlur is NOT a Gaussian convolution: it is three passes of an "extended box
// blur" per axis (libImaging/BoxBlur.c), all in 32-bit integer arithmetic with a rounding to
// uint8 after every pass.  This file restates that arithmetic so the result is bit-identical:
//
//   box radius  r' = l + a   from sigma^2 = radius^2 / 3 (Gwosdek et al., float32 as in Pillow)
//   one pass    out[x] = (acc*ww + (in[x-r-1] + in[x+r+1])*fw + 2^23) >> 24      (uint32 wrap)
//               acc = sum of in[clamp(i)] for i in [x-r, x+r], r = int(r'),
//               ww = (uint32)(2^24 / (2 r' + 1)),  fw = (2^24 - (2r+1) ww) / 2
//   image       3 passes along x, transpose, 3 passes along x again, transpose back (what
//               Pillow does for the y direction; it also keeps every pass coalesced here)
//
// The running accumulator of the C code becomes a per-line prefix sum in shared memory (exact:
// it is an integer sum), so every output pixel of a line is computed independently.  One CTA
// per (line, request); the same source image serves a whole batch of sessions, each with its
// own radius -- that is the batch dimension of this path.

#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <new>
#include <vector>

#include "cm_internal.h"

#include "cm_blur_state.h"

namespace {

using cm::fail;

struct BoxParams {
    int radius;         // int(r')
    uint32_t ww, fw;
    int identity;       // GaussianBlur(0) returns a copy
};

// Pillow _gaussian_blur_radius + ImagingHorizontalBoxBlur's weights, float32 step by step.
BoxParams box_params(float gaussian_radius) {
    BoxParams p = {0, 1u << 24, 0u, 0};
    if (gaussian_radius == 0.0f) { p.identity = 1; return p; }
    volatile float sigma2 = gaussian_radius * gaussian_radius / 3.0f;
    volatile float L = (float)std::sqrt(12.0 * (double)sigma2 + 1.0);
    volatile float l = (float)std::floor(((double)L - 1.0) / 2.0);
    volatile float t1 = 2.0f * l + 1.0f;
    volatile float t2 = l * (l + 1.0f);
    volatile float t3 = 3.0f * sigma2;
    volatile float t4 = t2 - t3;
    volatile float a = t1 * t4;
    volatile float lp1 = l + 1.0f;
    volatile float t5 = lp1 * lp1;
    volatile float t6 = sigma2 - t5;
    volatile float t7 = 6.0f * t6;
    a = a / t7;
    volatile float fr = l + a;
    p.radius = (int)fr;
    volatile float denom = fr * 2.0f + 1.0f;
    volatile float q = (float)(1u << 24) / denom;
    p.ww = (uint32_t)q;
    p.fw = ((1u << 24) - (uint32_t)(p.radius * 2 + 1) * p.ww) / 2u;
    return p;
}

// ---- the same three passes on a PLANAR copy of the line (round 2, late) ---------------------------------
// The interleaved form below spends about ten shared-memory accesses per output byte and pass (byte loads of
// pixels, two prefix sums, two far pixels, two edge pixels).  Here the line is de-interleaved into one plane per
// channel, padded on both sides with kPad copies of the edge pixel -- so Pillow's index clamping and its
// n_left / n_right edge terms become ordinary window sums -- and a thread produces FOUR consecutive bytes of a plane:
//   the window sum of the first of them is (2r + 1) / 4 word loads added up with __vsadu4 (no prefix sums: a
//   block-wide scan costs more instructions per pass than it saves, and two barriers),
//   the two far pixels of the four outputs are two unaligned words (two aligned loads + a funnel shift each),
//   and the window sum slides from one output to the next with the bytes of those same two words.  Integer arithmetic and rounding are those of the
// interleaved form (uint32 wrap included): the output is the same, byte for byte.  Needs n % 4 == 0 and
// radius <= kPad - 2 (the game's blur <= 15 gives a box radius <= 15); other lines take the interleaved form.
constexpr int kPad = 64;

template <int C>
__device__ __forceinline__ void box3_planar(const uint8_t *__restrict__ in, uint8_t *__restrict__ out, int n,
                                            const BoxParams &bp, int passes, uint32_t *smem32) {
    const int t = threadIdx.x;
    const int W = n + 2 * kPad, W4 = W >> 2, n4 = n >> 2;
    uint32_t *cur = smem32;                                // [C][W4] the padded planes
    uint32_t *nxt = cur + C * W4;
    const int bytes = n * C;
    const bool rgb_words = C == 3 && ((((uintptr_t)in) | ((uintptr_t)out)) & 3) == 0;
    if (rgb_words) {   // four RGB pixels = three words in, one word per plane out: six byte permutes
        for (int q = t; q < n4; q += 256) {
            const uint32_t *w = reinterpret_cast<const uint32_t *>(in) + 3 * q;
            const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];         // R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
            const int at = (kPad >> 2) + q;
            cur[0 * W4 + at] = __byte_perm(__byte_perm(w0, w1, 0x0630), w2, 0x5210);
            cur[1 * W4 + at] = __byte_perm(__byte_perm(w0, w1, 0x0741), w2, 0x6210);
            cur[2 * W4 + at] = __byte_perm(__byte_perm(w0, w1, 0x0052), w2, 0x7410);
        }
    } else {   // load and de-interleave byte by byte
        uint8_t *c8 = reinterpret_cast<uint8_t *>(cur);
        int done = 0;
        if (((uintptr_t)in & 15) == 0) {
            const int n16 = bytes >> 4;
            for (int v = t; v < n16; v += 256) {
                const uint4 q = reinterpret_cast<const uint4 *>(in)[v];
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const int idx = 16 * v + k;
                    c8[(idx % C) * W + kPad + idx / C] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
                }
            }
            done = n16 << 4;
        }
        for (int idx = done + t; idx < bytes; idx += 256) c8[(idx % C) * W + kPad + idx / C] = in[idx];
    }
    __syncthreads();
    const int r = bp.radius;
    constexpr int kPad4 = kPad >> 2;
    for (int pass = 0; pass < passes; ++pass) {
        // pads: kPad copies of the first and of the last pixel of every plane
        if (t < 2 * kPad4) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const uint8_t *c8 = reinterpret_cast<const uint8_t *>(cur + c * W4);
                if (t < kPad4) cur[c * W4 + t] = 0x01010101u * c8[kPad];
                else cur[c * W4 + kPad4 + n4 + (t - kPad4)] = 0x01010101u * c8[kPad + n - 1];
            }
        }
        __syncthreads();
        // four consecutive output bytes of one plane per step
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const uint32_t *pw = cur + c * W4;
            for (int q = t; q < n4; q += 256) {
                const int p = kPad + 4 * q;                          // padded position of the first output byte
                const int ihi = p + r + 1, ilo = p - r;              // window [ilo, ihi) of the first output
                const int first = ilo >> 2, last = ihi >> 2;
                const int sh_hi = (ihi & 3) * 8, sh_lo = (ilo & 3) * 8;
                const uint32_t whi = pw[last], wlo = pw[first];
                // window sum = whole words [first, last) + the bytes of word `last` below ihi - the bytes of word `first`
                // below ilo.  A block-wide prefix sum would make this two lookups, but building it costs more
                // instructions per pass than these (2r + 1) / 4 adds (measured: 370 against 250 for the outputs themselves).
                uint32_t acc = __vsadu4(sh_hi ? (whi << (32 - sh_hi)) : 0u, 0u) - __vsadu4(sh_lo ? (wlo << (32 - sh_lo)) : 0u, 0u);
                for (int j = first; j < last; ++j) acc += __vsadu4(pw[j], 0u);
                // bytes [ilo - 1, ilo + 3) and [ihi, ihi + 4): far pixels of the four outputs, and what the window gains / loses
                const int il = ilo - 1;
                const uint32_t left4 = __funnelshift_r(pw[il >> 2], pw[(il >> 2) + 1], (il & 3) * 8);
                const uint32_t right4 = __funnelshift_r(whi, pw[last + 1], sh_hi);
                uint32_t packed = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t fl = (left4 >> (8 * k)) & 0xffu, fr = (right4 >> (8 * k)) & 0xffu;
                    const uint32_t bulk = acc * bp.ww + (fl + fr) * bp.fw;
                    packed |= ((bulk + (1u << 23)) >> 24) << (8 * k);
                    if (k < 3) acc += fr - ((left4 >> (8 * (k + 1))) & 0xffu);
                }
                nxt[c * W4 + (p >> 2)] = packed;
            }
        }
        __syncthreads();
        uint32_t *tmp = cur; cur = nxt; nxt = tmp;
    }
    if (rgb_words) {
        for (int q = t; q < n4; q += 256) {
            const int at = (kPad >> 2) + q;
            const uint32_t R = cur[0 * W4 + at], G = cur[1 * W4 + at], B = cur[2 * W4 + at];
            uint32_t *w = reinterpret_cast<uint32_t *>(out) + 3 * q;
            w[0] = __byte_perm(__byte_perm(R, G, 0x1040), B, 0x3410);
            w[1] = __byte_perm(__byte_perm(G, R, 0x2601), B, 0x3250);
            w[2] = __byte_perm(__byte_perm(R, G, 0x0730), B, 0x7216);
        }
    } else {   // interleave and store
        const uint8_t *c8 = reinterpret_cast<const uint8_t *>(cur);
        int done = 0;
        if (((uintptr_t)out & 15) == 0) {
            const int n16 = bytes >> 4;
            for (int v = t; v < n16; v += 256) {
                uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const int idx = 16 * v + k;
                    w[k >> 2] |= (uint32_t)c8[(idx % C) * W + kPad + idx / C] << (8 * (k & 3));
                }
                reinterpret_cast<uint4 *>(out)[v] = make_uint4(w[0], w[1], w[2], w[3]);
            }
            done = n16 << 4;
        }
        for (int idx = done + t; idx < bytes; idx += 256) out[idx] = c8[(idx % C) * W + kPad + idx / C];
    }
}

template <int C>
__global__ void __launch_bounds__(256, 8)
cm_boxblur3_planar_kernel(const uint8_t *__restrict__ src, size_t src_req_stride, uint8_t *__restrict__ dst,
                          size_t dst_req_stride, int n, size_t line_stride, const BoxParams *__restrict__ params,
                          int passes) {
    extern __shared__ uint32_t smem[];
    const BoxParams bp = params[blockIdx.y];
    const uint8_t *in = src + blockIdx.y * src_req_stride + blockIdx.x * line_stride;
    uint8_t *out = dst + blockIdx.y * dst_req_stride + blockIdx.x * line_stride;
    box3_planar<C>(in, out, n, bp, bp.identity ? 0 : passes, smem);       // GaussianBlur(0): de-interleave, re-interleave
}

// The three box-blur passes of one axis along a "line" of `n` pixels of C channels, in ONE kernel: the
// line is read from global memory once, blurred three times in shared memory (Pillow rounds to uint8
// after every pass, so the intermediate lines are bytes too) and written once.  Pixel x of line `line`
// of request `b` lives at  base + b*req_stride + line*line_stride + x*C  (bytes, contiguous lines).
// Round 1 launched one kernel per pass: six 58 us launches per batch of eight 1024x1024 images,
// 0.9 TB/s each -- per-CTA latency (load, five barriers, store), not bandwidth.
template <int C>
__global__ void __launch_bounds__(256)
cm_boxblur3_kernel(const uint8_t *__restrict__ src, size_t src_req_stride, uint8_t *__restrict__ dst,
                   size_t dst_req_stride, int n, size_t line_stride, const BoxParams *__restrict__ params,
                   int passes) {
    extern __shared__ uint32_t smem[];
    uint32_t *P = smem;                                   // [C][n + 1] exclusive prefix sums
    uint8_t *buf0 = reinterpret_cast<uint8_t *>(smem) + (((size_t)C * (n + 1) * sizeof(uint32_t) + 15) & ~(size_t)15);   // [n][C] the line
    uint8_t *buf1 = buf0 + (((size_t)n * C + 15) & ~(size_t)15);               // the line after a pass
    __shared__ uint32_t warp_tot[8][C];
    const BoxParams bp = params[blockIdx.y];
    const uint8_t *in = src + blockIdx.y * src_req_stride + blockIdx.x * line_stride;
    uint8_t *out = dst + blockIdx.y * dst_req_stride + blockIdx.x * line_stride;
    const int t = threadIdx.x;
    const int bytes = n * C;
    // load: 16 bytes per thread where the line allows it
    if ((((uintptr_t)in | (uintptr_t)buf0) & 15) == 0) {
        const int n16 = bytes >> 4;
        for (int i = t; i < n16; i += 256) reinterpret_cast<uint4 *>(buf0)[i] = reinterpret_cast<const uint4 *>(in)[i];
        for (int i = (n16 << 4) + t; i < bytes; i += 256) buf0[i] = in[i];
    } else {
        for (int i = t; i < bytes; i += 256) buf0[i] = in[i];
    }
    __syncthreads();
    uint8_t *px = buf0, *nx = buf1;
    if (!bp.identity && n > 0) {
        const int chunk = (n + 255) / 256;
        const int x0 = min(t * chunk, n), x1 = min(x0 + chunk, n);
        const int lane = t & 31, wid = t >> 5;
        const int r = bp.radius, last = n - 1;
        for (int pass = 0; pass < passes; ++pass) {
            // block-wide exclusive scan: thread t owns pixels [x0, x1)
            uint32_t local[C];
#pragma unroll
            for (int c = 0; c < C; ++c) local[c] = 0;
            for (int x = x0; x < x1; ++x) {
#pragma unroll
                for (int c = 0; c < C; ++c) local[c] += px[x * C + c];
            }
            uint32_t incl[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                uint32_t v = local[c];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += u;
                }
                incl[c] = v;
                if (lane == 31) warp_tot[wid][c] = v;
            }
            __syncthreads();
            uint32_t run[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                uint32_t base = 0;
                for (int w = 0; w < wid; ++w) base += warp_tot[w][c];
                run[c] = base + incl[c] - local[c];           // exclusive prefix at x0
            }
            for (int x = x0; x < x1; ++x) {
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    P[c * (n + 1) + x] = run[c];
                    run[c] += px[x * C + c];
                }
            }
            if (x1 == n && x0 < n) {
#pragma unroll
                for (int c = 0; c < C; ++c) P[c * (n + 1) + n] = run[c];
            }
            __syncthreads();
            for (int x = t; x < n; x += 256) {
                const int hi = min(x + r, last), lo = max(x - r, 0);
                const uint32_t n_left = (uint32_t)max(r - x, 0), n_right = (uint32_t)max(x + r - last, 0);
                const int far_l = max(x - r - 1, 0), far_r = min(x + r + 1, last);
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const uint32_t acc = P[c * (n + 1) + hi + 1] - P[c * (n + 1) + lo] + n_left * px[c] +
                                         n_right * px[last * C + c];
                    const uint32_t bulk = acc * bp.ww + ((uint32_t)px[far_l * C + c] + (uint32_t)px[far_r * C + c]) * bp.fw;
                    nx[x * C + c] = (uint8_t)((bulk + (1u << 23)) >> 24);
                }
            }
            __syncthreads();
            uint8_t *tmp = px; px = nx; nx = tmp;
        }
    }
    if ((((uintptr_t)out | (uintptr_t)px) & 15) == 0) {
        const int n16 = bytes >> 4;
        for (int i = t; i < n16; i += 256) reinterpret_cast<uint4 *>(out)[i] = reinterpret_cast<const uint4 *>(px)[i];
        for (int i = (n16 << 4) + t; i < bytes; i += 256) out[i] = px[i];
    } else {
        for (int i = t; i < bytes; i += 256) out[i] = px[i];
    }
}

template <int C>
int launch_pass(cm_blur *b, const uint8_t *src, size_t src_req_stride, uint8_t *dst, size_t dst_req_stride,
                int n_lines, int n, size_t line_stride, int batch, const void *params) {
    const size_t line = ((size_t)n * C + 15) & ~(size_t)15;
    const size_t smem = (((size_t)C * (n + 1) * sizeof(uint32_t) + 15) & ~(size_t)15) + 2 * line;
    // Test hook: CM_BLUR_PLANAR=0 keeps every line on the interleaved form so the two can be compared.
    static const bool planar_on = [] { const char *v = getenv("CM_BLUR_PLANAR"); return !(v != nullptr && *v == '0'); }();
    const dim3 grid((unsigned)n_lines, (unsigned)batch);
    if (planar_on && n >= 4 && (n & 3) == 0 && b->max_box_radius <= kPad - 2) {
        const size_t smem_planar = 2 * (size_t)C * ((size_t)n + 2 * kPad);
        if (smem_planar > 48 * 1024)
            CM_CUDA(cudaFuncSetAttribute(cm_boxblur3_planar_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_planar));
        cm_boxblur3_planar_kernel<C><<<grid, 256, smem_planar, b->stream>>>(
            src, src_req_stride, dst, dst_req_stride, n, line_stride, (const BoxParams *)params, 3);
    } else {
        // per device and cheap: set it on every launch rather than cache a process-wide flag
        if (smem > 48 * 1024)
            CM_CUDA(cudaFuncSetAttribute(cm_boxblur3_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cm_boxblur3_kernel<C><<<grid, 256, smem, b->stream>>>(
            src, src_req_stride, dst, dst_req_stride, n, line_stride, (const BoxParams *)params, 3);
    }
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

// Pixel transpose [h][w][C] -> [w][h][C] through a 32x32 shared-memory tile (Pillow's
// ImagingTranspose): the blur along y then runs on contiguous lines like the blur along x.  A
// strided column pass measured 351 us against 56 us for a row pass (1024x1024 RGB, 8 sessions).
template <int C>
__global__ void __launch_bounds__(256)
cm_transpose_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int h, int w, size_t req_stride) {
    __shared__ uint8_t tile[32][32 * C + 4];
    const uint8_t *in = src + blockIdx.z * req_stride;
    uint8_t *out = dst + blockIdx.z * req_stride;
    const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int y = y0 + r, x = x0 + tx;
        if (y < h && x < w) {
#pragma unroll
            for (int c = 0; c < C; ++c) tile[r][tx * C + c] = in[((size_t)y * w + x) * C + c];
        }
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int x = x0 + r, y = y0 + tx;                   // output row x, output column y
        if (x < w && y < h) {
#pragma unroll
            for (int c = 0; c < C; ++c) out[((size_t)x * h + y) * C + c] = tile[tx][r * C + c];
        }
    }
}

template <int C>
int launch_transpose(cm_blur *b, const uint8_t *src, uint8_t *dst, int h, int w, size_t req_stride, int batch) {
    cm_transpose_kernel<C><<<dim3((unsigned)((w + 31) / 32), (unsigned)((h + 31) / 32), (unsigned)batch), 256, 0,
                             b->stream>>>(src, dst, h, w, req_stride);
    cm::count_launch();
    CM_CUDA(cudaGetLastError());
    return CM_OK;
}

// Sessions [first, first + batch): three passes along x, transpose, three along x, transpose back.
// The result is in d_b.
template <int C>
int blur_batch(cm_blur *b, int first, int batch) {
    const size_t img = (size_t)b->h * b->w * C;
    const size_t row = (size_t)b->w * C, col = (size_t)b->h * C;
    uint8_t *a = b->d_a + (size_t)first * img, *o = b->d_b + (size_t)first * img;
    const void *pr = static_cast<const BoxParams *>(b->d_params) + first;
    // along x: lines are rows (three passes in one launch)
    if (int rc = launch_pass<C>(b, b->d_src, 0, a, img, b->h, b->w, row, batch, pr)) return rc;
    // along y: transpose, blur the (now contiguous) columns, transpose back -- as Pillow does
    if (int rc = launch_transpose<C>(b, a, o, b->h, b->w, img, batch)) return rc;
    if (int rc = launch_pass<C>(b, o, img, a, img, b->w, b->h, col, batch, pr)) return rc;
    if (int rc = launch_transpose<C>(b, a, o, b->w, b->h, img, batch)) return rc;
    return CM_OK;
}

}  // namespace

extern "C" {

int cm_blur_destroy(cm_blur *b) {
    if (b == nullptr) return CM_OK;
    {
        cm::DeviceGuard guard(b->device);
        if (b->stream) { cudaStreamSynchronize(b->stream); cudaStreamDestroy(b->stream); }
        if (b->copy_stream) { cudaStreamSynchronize(b->copy_stream); cudaStreamDestroy(b->copy_stream); }
        for (cudaEvent_t ev : b->blurred)
            if (ev) cudaEventDestroy(ev);
        if (b->d_src) cudaFree(b->d_src);
        if (b->d_a) cudaFree(b->d_a);
        if (b->d_b) cudaFree(b->d_b);
        if (b->d_params) cudaFree(b->d_params);
        if (b->h_pin) cudaFreeHost(b->h_pin);
        cm::jpeg_state_free(b);
    }
    delete b;
    return CM_OK;
}

int cm_blur_create(cm_blur **out, int device, int max_height, int max_width, int channels, int max_batch) {
    if (out == nullptr) return fail(CM_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (max_height < 1 || max_width < 1 || max_height > 8192 || max_width > 8192)
        return fail(CM_ERR_INVALID, "image size %dx%d outside [1, 8192]", max_width, max_height);
    if (channels != 1 && channels != 3 && channels != 4)
        return fail(CM_ERR_INVALID, "channels %d must be 1 (L), 3 (RGB) or 4 (RGBA)", channels);
    if (max_batch < 1 || max_batch > 1024) return fail(CM_ERR_INVALID, "max_batch %d outside [1, 1024]", max_batch);
    int n_dev = 0;
    CM_CUDA(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(CM_ERR_INVALID, "device %d not in [0, %d)", device, n_dev);
    cm_blur *b = new (std::nothrow) cm_blur();
    if (b == nullptr) return fail(CM_ERR_NOMEM, "blur allocation failed");
    b->device = device; b->max_h = max_height; b->max_w = max_width; b->channels = channels; b->max_batch = max_batch;
    const size_t img = (size_t)max_height * max_width * channels;
    cm::DeviceGuard guard(device);
    cudaError_t err = cudaMalloc(&b->d_src, img);
    if (err == cudaSuccess) err = cudaMalloc(&b->d_a, img * max_batch);
    if (err == cudaSuccess) err = cudaMalloc(&b->d_b, img * max_batch);
    if (err == cudaSuccess) err = cudaMalloc(&b->d_params, sizeof(BoxParams) * max_batch);
    if (err == cudaSuccess) err = cudaHostAlloc(&b->h_pin, img * max_batch, cudaHostAllocDefault);
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&b->copy_stream, cudaStreamNonBlocking);
    for (cudaEvent_t &ev : b->blurred)
        if (err == cudaSuccess) err = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    if (err != cudaSuccess) {
        cudaGetLastError();
        const int code = fail(CM_ERR_NOMEM, "blur buffers (%zu bytes x %d) failed: %s", img, 2 * max_batch + 1,
                              cudaGetErrorString(err));
        cm_blur_destroy(b);
        return code;
    }
    *out = b;
    return CM_OK;
}

int cm_blur_set_image(cm_blur *b, const uint8_t *pixels, int height, int width) {
    if (b == nullptr || pixels == nullptr) return fail(CM_ERR_INVALID, "NULL argument");
    if (height < 1 || width < 1 || height > b->max_h || width > b->max_w)
        return fail(CM_ERR_INVALID, "image %dx%d exceeds the %dx%d this object was created for", width, height,
                    b->max_w, b->max_h);
    std::lock_guard<std::mutex> lock(b->mu);
    cm::DeviceGuard guard(b->device);
    CM_CUDA(cudaMemcpyAsync(b->d_src, pixels, (size_t)height * width * b->channels, cudaMemcpyHostToDevice, b->stream));
    CM_CUDA(cudaStreamSynchronize(b->stream));
    b->h = height; b->w = width;
    return CM_OK;
}

// Validates the radii and uploads the per-request box parameters (b->mu held, device current).
static int upload_params(cm_blur *b, const float *radii, int n) {
    if (b->h == 0) return fail(CM_ERR_STATE, "cm_blur_set_image has not been called");
    std::vector<BoxParams> params((size_t)n);
    int max_radius = 0;
    for (int i = 0; i < n; ++i) {
        // Pillow squares the radius (_gaussian_blur_radius), so GaussianBlur(-r) == GaussianBlur(r): a
        // cosine that rounds to 1.0000001 makes score_to_blur slightly negative in the reference too
        if (!(fabsf(radii[i]) <= 4096.0f))
            return fail(CM_ERR_INVALID, "radius %g of request %d outside [-4096, 4096]", (double)radii[i], i);
        params[(size_t)i] = box_params(radii[i]);
        if (!params[(size_t)i].identity && params[(size_t)i].radius > max_radius) max_radius = params[(size_t)i].radius;
    }
    b->max_box_radius = max_radius;
    // pageable source: the runtime stages it before returning, so `params` may go out of scope
    CM_CUDA(cudaMemcpyAsync(b->d_params, params.data(), sizeof(BoxParams) * n, cudaMemcpyHostToDevice, b->stream));
    return CM_OK;
}

static int blur_chunk(cm_blur *b, int first, int count) {
    switch (b->channels) {
        case 1: return blur_batch<1>(b, first, count);
        case 3: return blur_batch<3>(b, first, count);
        default: return blur_batch<4>(b, first, count);
    }
}

}  // extern "C"

int cm::blur_enqueue(cm_blur *b, const float *radii, int n) {
    if (int rc = upload_params(b, radii, n)) return rc;
    return blur_chunk(b, 0, n);
}

extern "C" {

// Blurs into the device ping-pong buffers and brings the n images back into b->h_pin (b->mu held).
static int blur_batch_to_pinned(cm_blur *b, const float *radii, int n, size_t *bytes_out) {
    cm::DeviceGuard guard(b->device);
    if (int rc = upload_params(b, radii, n)) return rc;
    // In chunks of about 16 MB of pixels: while one chunk is being blurred the previous one is
    // already on its way to the host (1024x1024 RGB: the copy is the longer of the two).
    const size_t img = (size_t)b->h * b->w * b->channels;
    const int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)n, ((size_t)16 << 20) / img));
    int k = 0;
    for (int first = 0; first < n; first += chunk, ++k) {
        const int count = std::min(chunk, n - first);
        if (int rc = blur_chunk(b, first, count)) return rc;
        CM_CUDA(cudaEventRecord(b->blurred[k & 1], b->stream));
        CM_CUDA(cudaStreamWaitEvent(b->copy_stream, b->blurred[k & 1], 0));
        CM_CUDA(cudaMemcpyAsync(b->h_pin + (size_t)first * img, b->d_b + (size_t)first * img, (size_t)count * img,
                                cudaMemcpyDeviceToHost, b->copy_stream));
    }
    CM_CUDA(cudaStreamSynchronize(b->copy_stream));
    CM_CUDA(cudaStreamSynchronize(b->stream));
    const size_t bytes = img * n;
    *bytes_out = bytes;
    return CM_OK;
}

int cm_blur_batch_host(cm_blur *b, const float *radii, int n, uint8_t *out) {
    if (b == nullptr) return fail(CM_ERR_INVALID, "blur is NULL");
    if (n < 0 || n > b->max_batch) return fail(CM_ERR_INVALID, "n %d outside [0, %d]", n, b->max_batch);
    if (n == 0) return CM_OK;
    if (radii == nullptr || out == nullptr) return fail(CM_ERR_INVALID, "radii or out is NULL");
    std::lock_guard<std::mutex> lock(b->mu);
    size_t bytes = 0;
    if (int rc = blur_batch_to_pinned(b, radii, n, &bytes)) return rc;
    memcpy(out, b->h_pin, bytes);
    return CM_OK;
}

int cm_blur_batch_pinned(cm_blur *b, const float *radii, int n, const uint8_t **out) {
    if (b == nullptr) return fail(CM_ERR_INVALID, "blur is NULL");
    if (n < 1 || n > b->max_batch) return fail(CM_ERR_INVALID, "n %d outside [1, %d]", n, b->max_batch);
    if (radii == nullptr || out == nullptr) return fail(CM_ERR_INVALID, "radii or out is NULL");
    *out = nullptr;
    std::lock_guard<std::mutex> lock(b->mu);
    size_t bytes = 0;
    if (int rc = blur_batch_to_pinned(b, radii, n, &bytes)) return rc;
    *out = b->h_pin;
    return CM_OK;
}

}  // extern "C"
