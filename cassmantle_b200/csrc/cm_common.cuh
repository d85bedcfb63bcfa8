// Shared device helpers for the cassmantle_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cm {

constexpr unsigned kFullMask = 0xffffffffu;

// lowbias32 finaliser -- the same function as cassmantle_b200/synth.py: mix32.
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

__host__ __device__ __forceinline__ int bytesum(uint32_t h) {
    return (int)((h & 0xffu) + ((h >> 8) & 0xffu) + ((h >> 16) & 0xffu) + (h >> 24)) - 510;
}

// Streaming 16-byte load: table rows are touched once per batch, keep them out of L1, and let L2
// fetch no more than 64 bytes around each request.  A row is 1200 bytes at a 16-byte-aligned
// address, so the default (128-byte) L2 fill drags in the neighbours of its first and last line:
// measured on B200, 4096x16: 11.64 us with the default or .L2::128B, 12.07 us with .L2::256B,
// 11.29 us with .L2::64B; 65536x16: 181.0 / 188.2 / 175.4 us (7.2 TB/s).
__device__ __forceinline__ float4 ld_stream_f4(const float4 *p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// The same without the L2 hint, for kernels that walk the table front to back (every line is
// wanted anyway).
__device__ __forceinline__ float4 ld_stream_seq_f4(const float4 *p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Programmatic dependent launch (PDL) controls; both are no-ops for a plain launch.
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

}  // namespace cm
