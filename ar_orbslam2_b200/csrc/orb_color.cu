// liborbgpu: colour -> grey conversion in front of the extractor.
//
// ref: the cvtColor calls of Tracking::GrabImageStereo / GrabImageRGBD / GrabImageMonocular
// (ORB_SLAM2/src/Tracking.cc:510-526,545-556,567-577): CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY on 8-bit
// images.  cv::cvtColor is OpenCV code (not vendored); its 8-bit arithmetic is restated and pinned against cv2 4.13.0:
//   grey = (R*9798 + G*19235 + B*3735 + 16384) >> 15          (variant 0, OpenCV >= 3.x/4.x; 0 mismatches on 16.7 M pixels)
//   grey = (R*4899 + G*9617  + B*1868 + 8192)  >> 14          (variant 1, OpenCV 2.4's CV_DESCALE form; unverifiable here)
// Pure streaming, HBM bound: reads 3 or 4 bytes and writes 1 byte per pixel.  A thread converts 16 pixels: 3 (RGB) or 4
// (RGBA) 16-byte loads, one 16-byte store; rows that are not 16-byte aligned and the last width%16 pixels take a
// per-pixel path in the same kernel.
#include "orb_common.cuh"

using namespace orbgpu;

namespace {

struct GrayCoef { int c0, c1, c2, rnd, shift; };   // weights of channel 0,1,2 in memory order

__device__ __forceinline__ unsigned byte_at(const uint32_t *w, int idx) { return (w[idx >> 2] >> ((idx & 3) * 8)) & 0xffu; }

template <int CH>
__global__ void __launch_bounds__(256)
k_cvt_gray(const uint8_t *__restrict__ src, int width, int height, long long src_pitch, long long src_stride,
           uint8_t *__restrict__ dst, long long dst_pitch, long long dst_stride, GrayCoef k, int aligned) {
    const int y = blockIdx.y, b = blockIdx.z;
    const int x0 = (blockIdx.x * 256 + threadIdx.x) * 16;
    if (x0 >= width) return;
    const uint8_t *s = src + b * src_stride + y * src_pitch;
    uint8_t *d = dst + b * dst_stride + y * dst_pitch;
    if (aligned && x0 + 16 <= width) {
        uint32_t w[4 * CH];
        const uint4 *sp = reinterpret_cast<const uint4 *>(s + (long long)x0 * CH);
#pragma unroll
        for (int i = 0; i < CH; i++) { const uint4 v = __ldg(sp + i); w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w; }
        uint32_t o[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t acc = 0;
#pragma unroll
            for (int p = 0; p < 4; p++) {
                const int px = q * 4 + p;
                const unsigned g = (byte_at(w, px * CH) * k.c0 + byte_at(w, px * CH + 1) * k.c1 + byte_at(w, px * CH + 2) * k.c2 + k.rnd) >> k.shift;
                acc |= g << (8 * p);
            }
            o[q] = acc;
        }
        *reinterpret_cast<uint4 *>(d + x0) = make_uint4(o[0], o[1], o[2], o[3]);
    } else {
        const int x1 = min(width, x0 + 16);
        for (int x = x0; x < x1; x++) {
            const uint8_t *p = s + (long long)x * CH;
            d[x] = (uint8_t)((p[0] * k.c0 + p[1] * k.c1 + p[2] * k.c2 + k.rnd) >> k.shift);
        }
    }
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

}  // namespace

extern "C" int orb_cvt_color_gray(const uint8_t *src, int width, int height, int src_pitch, size_t src_stride, int channels,
                                  int rgb_order, int variant, int batch, uint8_t *dst, int dst_pitch, size_t dst_stride, int loc,
                                  void *stream) {
    ORB_REQUIRE(src && dst && width >= 0 && height >= 0 && batch >= 1, ORB_ERR_INVALID, "bad argument");
    ORB_REQUIRE(channels == 3 || channels == 4, ORB_ERR_INVALID, "channels must be 3 or 4 (the reference converts nothing else, Tracking.cc:510-526)");
    ORB_REQUIRE(src_pitch >= width * channels && dst_pitch >= width, ORB_ERR_INVALID, "pitch smaller than a row");
    ORB_REQUIRE(variant == 0 || variant == 1, ORB_ERR_INVALID, "variant must be 0 (OpenCV 4.x) or 1 (OpenCV 2.4)");
    if (width == 0 || height == 0) return ORB_OK;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    GrayCoef k;
    const int wr = variant ? 4899 : 9798, wg = variant ? 9617 : 19235, wb = variant ? 1868 : 3735;
    k.c0 = rgb_order ? wr : wb; k.c1 = wg; k.c2 = rgb_order ? wb : wr;
    k.shift = variant ? 14 : 15; k.rnd = 1 << (k.shift - 1);
    const uint8_t *d_src = src;
    uint8_t *d_dst = dst;
    DevBuf bs, bd;
    cudaStream_t st = (cudaStream_t)stream;
    if (!st) {   // NULL = the thread's call stream (orb_set_call_stream), never the legacy default stream
        ThreadCtx *T = nullptr;
        int rc = orbgpu_thread_ctx(&T, 0, 0);
        if (rc != ORB_OK) return rc;
        st = T->stream;
    }
    const size_t src_bytes = (batch - 1) * src_stride + (size_t)(height - 1) * src_pitch + (size_t)width * channels;
    const size_t dst_bytes = (batch - 1) * dst_stride + (size_t)(height - 1) * dst_pitch + (size_t)width;
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(bs.alloc(src_bytes));
        ORB_CUDA_TRY(bd.alloc(dst_bytes));
        ORB_CUDA_TRY(cudaMemcpyAsync(bs.p, src, src_bytes, cudaMemcpyHostToDevice, st));
        d_src = (const uint8_t *)bs.p; d_dst = (uint8_t *)bd.p;
    }
    const int aligned = ((uintptr_t)d_src % 16 == 0) && ((uintptr_t)d_dst % 16 == 0) && src_pitch % 16 == 0 && dst_pitch % 16 == 0 &&
                        src_stride % 16 == 0 && dst_stride % 16 == 0;
    dim3 grid((unsigned)((width + 16 * 256 - 1) / (16 * 256)), (unsigned)height, (unsigned)batch);
    if (channels == 3) k_cvt_gray<3><<<grid, 256, 0, st>>>(d_src, width, height, src_pitch, (long long)src_stride, d_dst, dst_pitch, (long long)dst_stride, k, aligned);
    else k_cvt_gray<4><<<grid, 256, 0, st>>>(d_src, width, height, src_pitch, (long long)src_stride, d_dst, dst_pitch, (long long)dst_stride, k, aligned);
    ORB_CUDA_TRY(cudaGetLastError());
    if (loc == ORB_HOST) {
        // strided copy back: only the pixels, the caller's padding bytes stay untouched
        if (dst_stride == (size_t)height * dst_pitch || batch == 1)
            ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_pitch, d_dst, dst_pitch, width, (size_t)height * (batch == 1 ? 1 : batch), cudaMemcpyDeviceToHost, st));
        else
            for (int b = 0; b < batch; b++)
                ORB_CUDA_TRY(cudaMemcpy2DAsync(dst + b * dst_stride, dst_pitch, d_dst + b * dst_stride, dst_pitch, width, height, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return ORB_OK;
}
