// liborbgpu: stereo rectification cv::remap(src, dst, map1, map2, INTER_LINEAR) on the device.
//
// ref: ORB_SLAM2/Examples/ROS/ORB_SLAM2/src/ros_stereo.cc:163-164 (maps from cv::initUndistortRectifyMap(..., CV_32F, M1, M2),
// :107-112).  cv::remap is OpenCV code (not vendored); its 8-bit bilinear arithmetic is restated and pinned against cv2 4.13.0
// (tests/golden/cv2_remap.npz, 0 differing pixels): coordinates are rounded to 1/32 pixel (cvRound(map * 32)), the four
// weights are the 15-bit table BilinearTab_i -- (32-fx)(32-fy)*32 etc., with the table's one saturated entry (fx = fy = 0:
// {32767, 0, 0, 1}) kept -- the sum is rounded with (s + 2^14) >> 15, and pixels outside the source read 0 (BORDER_CONSTANT).
// HBM bound: 8 map bytes + ~1 gathered source byte in, 1 byte out per pixel; a thread makes 4 pixels (one 32-bit store).
#include "orb_common.cuh"

using namespace orbgpu;

namespace {

__device__ __forceinline__ uint32_t remap_px(const uint8_t *__restrict__ s, int sw, int sh, long long sp, float mx, float my) {
    const int sx = __float2int_rn(__fmul_rn(mx, 32.0f)), sy = __float2int_rn(__fmul_rn(my, 32.0f));
    const int fx = sx & 31, fy = sy & 31, ix = sx >> 5, iy = sy >> 5;
    int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    if ((fx | fy) == 0) { w00 = 32767; w11 = 1; }   // saturate_cast<short>(32768) and the table's sum correction
    const bool x0 = (unsigned)ix < (unsigned)sw, x1 = (unsigned)(ix + 1) < (unsigned)sw;
    const bool y0 = (unsigned)iy < (unsigned)sh, y1 = (unsigned)(iy + 1) < (unsigned)sh;
    const uint8_t *p = s + (long long)iy * sp + ix;
    const int p00 = (x0 && y0) ? p[0] : 0, p01 = (x1 && y0) ? p[1] : 0, p10 = (x0 && y1) ? p[sp] : 0, p11 = (x1 && y1) ? p[sp + 1] : 0;
    const int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + (1 << 14)) >> 15;
    return (uint32_t)min(max(v, 0), 255);
}

__global__ void __launch_bounds__(128)
k_remap_linear(const uint8_t *__restrict__ src, int sw, int sh, long long sp, long long sstride, const float *__restrict__ mapx,
               const float *__restrict__ mapy, long long mp, uint8_t *__restrict__ dst, int dw, int dh, long long dp,
               long long dstride, int vec) {
    const int x0 = (blockIdx.x * 128 + threadIdx.x) * 4, y = blockIdx.y, b = blockIdx.z;
    if (x0 >= dw) return;
    const uint8_t *s = src + b * sstride;
    uint8_t *d = dst + b * dstride + (long long)y * dp;
    const float *mx = mapx + (long long)y * mp, *my = mapy + (long long)y * mp;
    if (vec && x0 + 4 <= dw) {
        const float4 ax = __ldg(reinterpret_cast<const float4 *>(mx + x0)), ay = __ldg(reinterpret_cast<const float4 *>(my + x0));
        const uint32_t o = remap_px(s, sw, sh, sp, ax.x, ay.x) | (remap_px(s, sw, sh, sp, ax.y, ay.y) << 8) |
                           (remap_px(s, sw, sh, sp, ax.z, ay.z) << 16) | (remap_px(s, sw, sh, sp, ax.w, ay.w) << 24);
        *reinterpret_cast<uint32_t *>(d + x0) = o;
    } else {
        for (int x = x0; x < min(dw, x0 + 4); x++) d[x] = (uint8_t)remap_px(s, sw, sh, sp, mx[x], my[x]);
    }
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

}  // namespace

extern "C" int orb_remap_linear(const uint8_t *src, int src_width, int src_height, int src_pitch, size_t src_stride, const float *map_x,
                                const float *map_y, int map_pitch, uint8_t *dst, int dst_width, int dst_height, int dst_pitch,
                                size_t dst_stride, int batch, int loc, void *stream) {
    ORB_REQUIRE(src && map_x && map_y && dst && batch >= 1, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(src_width > 0 && src_height > 0 && dst_width >= 0 && dst_height >= 0 && src_pitch >= src_width && dst_pitch >= dst_width &&
                    map_pitch >= dst_width, ORB_ERR_INVALID, "bad geometry");
    if (dst_width == 0 || dst_height == 0) return ORB_OK;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    cudaStream_t st = (cudaStream_t)stream;
    if (!st) {   // NULL = the thread's call stream (orb_set_call_stream), never the legacy default stream
        ThreadCtx *T = nullptr;
        int rc = orbgpu_thread_ctx(&T, 0, 0);
        if (rc != ORB_OK) return rc;
        st = T->stream;
    }
    const uint8_t *d_src = src;
    const float *d_mx = map_x, *d_my = map_y;
    uint8_t *d_dst = dst;
    DevBuf bs, bmx, bmy, bd;
    const size_t src_bytes = (batch - 1) * src_stride + (size_t)(src_height - 1) * src_pitch + src_width;
    const size_t dst_bytes = (batch - 1) * dst_stride + (size_t)(dst_height - 1) * dst_pitch + dst_width;
    const size_t map_bytes = ((size_t)(dst_height - 1) * map_pitch + dst_width) * 4;
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(bs.alloc(src_bytes)); ORB_CUDA_TRY(bmx.alloc(map_bytes)); ORB_CUDA_TRY(bmy.alloc(map_bytes)); ORB_CUDA_TRY(bd.alloc(dst_bytes));
        ORB_CUDA_TRY(cudaMemcpyAsync(bs.p, src, src_bytes, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(bmx.p, map_x, map_bytes, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(bmy.p, map_y, map_bytes, cudaMemcpyHostToDevice, st));
        d_src = (const uint8_t *)bs.p; d_mx = (const float *)bmx.p; d_my = (const float *)bmy.p; d_dst = (uint8_t *)bd.p;
    }
    const int vec = ((uintptr_t)d_mx % 16 == 0) && ((uintptr_t)d_my % 16 == 0) && map_pitch % 4 == 0 && ((uintptr_t)d_dst % 4 == 0) &&
                    dst_pitch % 4 == 0 && dst_stride % 4 == 0;
    dim3 grid((unsigned)((dst_width + 511) / 512), (unsigned)dst_height, (unsigned)batch);
    k_remap_linear<<<grid, 128, 0, st>>>(d_src, src_width, src_height, src_pitch, (long long)src_stride, d_mx, d_my, map_pitch, d_dst,
                                         dst_width, dst_height, dst_pitch, (long long)dst_stride, vec);
    ORB_CUDA_TRY(cudaGetLastError());
    if (loc == ORB_HOST) {
        for (int b = 0; b < batch; b++)
            ORB_CUDA_TRY(cudaMemcpy2DAsync(dst + b * dst_stride, dst_pitch, d_dst + b * dst_stride, dst_pitch, dst_width, dst_height,
                                           cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return ORB_OK;
}
