// liborbgpu internals: error plumbing and small device helpers shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/orbgpu.h"

namespace orbgpu {

void set_error(const char *fmt, ...);

#define ORB_CUDA_TRY(expr)                                                                   \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            orbgpu::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                  \
                              cudaGetErrorString(_e));                                       \
            return ORB_ERR_CUDA;                                                             \
        }                                                                                    \
    } while (0)

#define ORB_REQUIRE(cond, code, ...)                                                         \
    do {                                                                                     \
        if (!(cond)) {                                                                       \
            orbgpu::set_error(__VA_ARGS__);                                                  \
            return (code);                                                                   \
        }                                                                                    \
    } while (0)

// BORDER_REFLECT_101 index (gfedcb|abcdefgh|gfedcba); valid for |overshoot| < n
__host__ __device__ __forceinline__ int reflect101(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return p;
}

// 256-bit Hamming distance of two descriptors held as 8 x u32
__device__ __forceinline__ int hamming256(const uint32_t a[8], const uint32_t b[8]) {
    int d = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) d += __popc(a[i] ^ b[i]);
    return d;
}

// Per-thread call context of the stream-less entry points (orb_descriptor_distance, orb_hamming_top2, orb_top2_merge,
// orb_search_matches_ratio, orb_is_in_frustum, orb_undistort_*, orb_distinctive_descriptors, orb_rot_hist_filter, and
// orb_cvt_color_gray / orb_remap_linear with stream == NULL): they run on `stream` = the stream set with
// orb_set_call_stream(), else the thread's own non-blocking stream -- never on the legacy default stream, which would
// serialise against every other thread of the process.  d_buf / h_buf (pinned) are reusable staging blocks for the
// host-mode forms: no cudaMalloc / cudaFree per call.
struct ThreadCtx {
    int device = -1;
    cudaStream_t own = nullptr, user = nullptr, stream = nullptr;
    bool has_user = false;
    uint8_t *d_buf = nullptr; size_t d_bytes = 0;
    uint8_t *h_buf = nullptr; size_t h_bytes = 0;
};
// context of the calling thread on the current device, with at least d_bytes / h_bytes of staging
int orbgpu_thread_ctx(ThreadCtx **out, size_t d_bytes, size_t h_bytes);

}  // namespace orbgpu
