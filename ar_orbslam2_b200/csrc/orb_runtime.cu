// liborbgpu -- per-thread call context: the stream and the reusable staging blocks of the stream-less entry points.
//
// The reference calls ORBmatcher / MapPoint helpers from three threads at once (Tracking, LocalMapping, LoopClosing:
// ORB_SLAM2/src/System.cc:91-101).  Nothing in this library may therefore run on CUDA's legacy default stream, which
// serialises against every blocking stream of the process and does NOT order against the handles' non-blocking streams.
// Each thread owns one non-blocking stream; orb_set_call_stream() replaces it (e.g. by a handle's stream, to chain a
// helper in front of orb_extract_batch without a host round trip).
#include "orb_common.cuh"

#include <algorithm>

namespace orbgpu {

static thread_local ThreadCtx g_ctx;   // CUDA objects are deliberately not destroyed at thread exit (the runtime may be gone)

int orbgpu_thread_ctx(ThreadCtx **out, size_t d_bytes, size_t h_bytes) {
    ThreadCtx &T = g_ctx;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    int dev = 0;
    ORB_CUDA_TRY(cudaGetDevice(&dev));
    if (T.device != dev) {   // first use on this device: the staging of another device is dropped
        if (T.own) { cudaStreamSynchronize(T.own); cudaStreamDestroy(T.own); }
        if (T.d_buf) cudaFree(T.d_buf);
        if (T.h_buf) cudaFreeHost(T.h_buf);
        T.own = nullptr; T.d_buf = nullptr; T.h_buf = nullptr; T.d_bytes = T.h_bytes = 0;
        ORB_CUDA_TRY(cudaStreamCreateWithFlags(&T.own, cudaStreamNonBlocking));
        T.device = dev;
    }
    T.stream = T.has_user ? T.user : T.own;
    if (T.d_bytes < d_bytes) {
        ORB_CUDA_TRY(cudaStreamSynchronize(T.stream));
        if (T.d_buf) cudaFree(T.d_buf);
        T.d_buf = nullptr; T.d_bytes = 0;
        const size_t want = d_bytes * 3 / 2 + 65536;
        ORB_CUDA_TRY(cudaMalloc(&T.d_buf, want));
        T.d_bytes = want;
    }
    if (T.h_bytes < h_bytes) {
        ORB_CUDA_TRY(cudaStreamSynchronize(T.stream));
        if (T.h_buf) cudaFreeHost(T.h_buf);
        T.h_buf = nullptr; T.h_bytes = 0;
        const size_t want = h_bytes * 3 / 2 + 65536;
        ORB_CUDA_TRY(cudaMallocHost(&T.h_buf, want));
        T.h_bytes = want;
    }
    *out = &T;
    return ORB_OK;
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" {

int orb_set_call_stream(void *cuda_stream, int use_it) {
    g_ctx.has_user = use_it != 0;
    g_ctx.user = use_it ? (cudaStream_t)cuda_stream : nullptr;
    return ORB_OK;
}

void *orb_get_call_stream(void) {
    ThreadCtx *T = nullptr;
    if (orbgpu_thread_ctx(&T, 0, 0) != ORB_OK) return nullptr;
    return (void *)T->stream;
}

int orb_call_stream_sync(void) {
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, 0, 0);
    if (rc != ORB_OK) return rc;
    ORB_CUDA_TRY(cudaStreamSynchronize(T->stream));
    return ORB_OK;
}

}  // extern "C"
