// liborbgpu -- the RGB-D Frame constructor's depth association (ref: Frame::ComputeStereoFromRGBD, src/Frame.cc:645-667,
// called from the RGB-D constructor :151 after ExtractORB and UndistortKeyPoints) with the depth-map conversion in front of it
// folded in (ref: `imDepth.convertTo(imDepth, CV_32F, mDepthMapFactor)`, src/Tracking.cc:553-554): the RGB-D counterpart of
// ComputeStereoMatches -- it fills the same mvuRight / mvDepth.  One thread per key point, one gather each.
#include "orb_common.cuh"

namespace orbgpu {

__global__ void __launch_bounds__(128)
k_depth_from_rgbd(const orb_keypoint *__restrict__ keys, const orb_keypoint *__restrict__ keys_un, int n, const uint8_t *__restrict__ depth_map,
                  int depth_type, int width, int height, int pitch, float factor, int convert, float mbf, float *__restrict__ u_right,
                  float *__restrict__ depth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float ur = -1.f, dp = -1.f;
    const int v = (int)keys[i].y, u = (int)keys[i].x;   // imDepth.at<float>(v, u) with float arguments: converted to int (truncation)
    if (u >= 0 && v >= 0 && u < width && v < height) {
        const uint8_t *row = depth_map + (size_t)v * pitch;
        float d;
        if (depth_type == 1) d = __fmul_rn((float)reinterpret_cast<const uint16_t *>(row)[u], factor);   // 16U -> 32F: (float)src * alpha
        else {
            d = reinterpret_cast<const float *>(row)[u];
            if (convert) d = __fmul_rn(d, factor);
        }
        if (d > 0) {
            dp = d;
            ur = __fsub_rn(keys_un[i].x, __fdiv_rn(mbf, d));   // mvuRight[i] = kpU.pt.x - mbf / d
        }
    }
    u_right[i] = ur;
    depth[i] = dp;
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" int orb_depth_from_rgbd(const orb_keypoint *keys, const orb_keypoint *keys_un, int n, const void *depth_map, int depth_type,
                                   int width, int height, int pitch_bytes, float depth_factor, float mbf, int loc, float *u_right,
                                   float *depth) {
    ORB_REQUIRE(n >= 0 && (n == 0 || (keys && keys_un && u_right && depth)), ORB_ERR_INVALID, "bad argument");
    ORB_REQUIRE(depth_map && width >= 1 && height >= 1 && (depth_type == 0 || depth_type == 1), ORB_ERR_INVALID, "bad depth map");
    ORB_REQUIRE(pitch_bytes >= width * (depth_type == 1 ? 2 : 4), ORB_ERR_INVALID, "pitch too small");
    if (n == 0) return ORB_OK;
    const int convert = depth_factor != 1.0f;   // ref Tracking.cc:553: `mDepthMapFactor != 1 || type != CV_32F`
    auto a256 = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t kb = a256((size_t)n * sizeof(orb_keypoint)), mb = a256((size_t)pitch_bytes * height), ob = a256((size_t)n * 4);
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? 2 * kb + mb + 2 * ob : 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    const orb_keypoint *dk = keys, *du = keys_un;
    const uint8_t *dm = (const uint8_t *)depth_map;
    float *our = u_right, *odp = depth;
    if (loc == ORB_HOST) {
        uint8_t *b = T->d_buf;
        ORB_CUDA_TRY(cudaMemcpyAsync(b, keys, (size_t)n * sizeof(orb_keypoint), cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(b + kb, keys_un, (size_t)n * sizeof(orb_keypoint), cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(b + 2 * kb, depth_map, (size_t)pitch_bytes * height, cudaMemcpyHostToDevice, st));
        dk = (const orb_keypoint *)b; du = (const orb_keypoint *)(b + kb); dm = b + 2 * kb;
        our = (float *)(b + 2 * kb + mb); odp = (float *)(b + 2 * kb + mb + ob);
    }
    k_depth_from_rgbd<<<(n + 127) / 128, 128, 0, st>>>(dk, du, n, dm, depth_type, width, height, pitch_bytes, depth_factor, convert, mbf, our, odp);
    ORB_CUDA_TRY(cudaGetLastError());
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(cudaMemcpyAsync(u_right, our, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(depth, odp, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return ORB_OK;
}
