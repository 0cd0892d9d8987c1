// liborbgpu: Frame::UndistortKeyPoints / ComputeImageBounds on the device.
//
// ref: ORB_SLAM2/src/Frame.cc:409-439 (UndistortKeyPoints) and :441-468 (ComputeImageBounds): cv::undistortPoints(mat, mat, mK,
// mDistCoef, cv::Mat(), mK).  cv::undistortPoints is OpenCV code (not vendored); its arithmetic -- double precision, five
// fixed-point iterations of the inverse Brown model, re-projection with P = K -- is restated and pinned against cv2 4.13.0
// (tests/golden/cv2_undistort.npz: 0 differing floats).  One thread per point; compiled with --fmad=false so that the
// double expressions are evaluated operation by operation like the host build of OpenCV.
#include "orb_common.cuh"
#include "orb_internal.cuh"

#include <cstring>

using namespace orbgpu;

namespace {

struct Cam { double fx, fy, cx, cy, ifx, ify, k[5]; };

__device__ __forceinline__ void undistort_one(const Cam &c, float px, float py, float *ox, float *oy) {
    const double u = px, v = py;
    double x = (u - c.cx) * c.ifx, y = (v - c.cy) * c.ify;
    const double x0 = x, y0 = y;
#pragma unroll 1
    for (int j = 0; j < 5; j++) {   // TermCriteria(MAX_ITER, 5, 0.01) of the cv::undistortPoints overload the reference calls
        const double r2 = x * x + y * y;
        // k[5..7] (rational model) are zero for the reference's 4/5-coefficient vectors: the numerator is 1 + ((0*r2+0)*r2+0)*r2
        const double icdist = (1 + ((0.0 * r2 + 0.0) * r2 + 0.0) * r2) / (1 + ((c.k[4] * r2 + c.k[1]) * r2 + c.k[0]) * r2);
        if (icdist < 0) { x = x0; y = y0; break; }   // OpenCV >= 3.x gives up on a diverging point
        const double deltaX = 2 * c.k[2] * x * y + c.k[3] * (r2 + 2 * x * x) + 0.0 * r2 + 0.0 * r2 * r2;
        const double deltaY = c.k[2] * (r2 + 2 * y * y) + 2 * c.k[3] * x * y + 0.0 * r2 + 0.0 * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    const double xx = c.fx * x + 0.0 * y + c.cx, yy = 0.0 * x + c.fy * y + c.cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
    *ox = (float)(xx * ww);
    *oy = (float)(yy * ww);
}

__global__ void k_undistort_keypoints(const orb_keypoint *__restrict__ in, int n, Cam c, orb_keypoint *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orb_keypoint kp = in[i];
    undistort_one(c, kp.x, kp.y, &kp.x, &kp.y);
    out[i] = kp;
}

__global__ void k_undistort_points(const float2 *__restrict__ in, int n, Cam c, float2 *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 p = in[i];
    float2 o;
    undistort_one(c, p.x, p.y, &o.x, &o.y);
    out[i] = o;
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

int make_cam(float fx, float fy, float cx, float cy, const float *dist, int ndist, Cam *c) {
    ORB_REQUIRE(dist && (ndist == 4 || ndist == 5), ORB_ERR_INVALID, "dist_coef must hold 4 or 5 floats (k1,k2,p1,p2[,k3]; src/Tracking.cc:427-438)");
    c->fx = fx; c->fy = fy; c->cx = cx; c->cy = cy;
    c->ifx = 1. / c->fx; c->ify = 1. / c->fy;
    for (int i = 0; i < 5; i++) c->k[i] = i < ndist ? (double)dist[i] : 0.0;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    return ORB_OK;
}

}  // namespace

int orbgpu_undistort_keypoints_async(const orb_keypoint *keys, int n, float fx, float fy, float cx, float cy, const float *dist_coef,
                                     int ndist, orb_keypoint *keys_un, cudaStream_t stream) {
    Cam c;
    int rc = make_cam(fx, fy, cx, cy, dist_coef, ndist, &c);
    if (rc != ORB_OK) return rc;
    if (n == 0) return ORB_OK;
    k_undistort_keypoints<<<(n + 127) / 128, 128, 0, stream>>>(keys, n, c, keys_un);
    ORB_CUDA_TRY(cudaGetLastError());
    return ORB_OK;
}

extern "C" {

int orb_undistort_keypoints(const orb_keypoint *keys, int n, float fx, float fy, float cx, float cy, const float *dist_coef,
                            int ndist, orb_keypoint *keys_un, int loc) {
    ORB_REQUIRE(n >= 0 && (n == 0 || (keys && keys_un)), ORB_ERR_INVALID, "bad argument");
    Cam c;
    int rc = make_cam(fx, fy, cx, cy, dist_coef, ndist, &c);
    if (rc != ORB_OK) return rc;
    if (n == 0) return ORB_OK;
    const size_t bytes = (size_t)n * sizeof(orb_keypoint);
    ThreadCtx *T = nullptr;
    rc = orbgpu_thread_ctx(&T, 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;   // the thread's call stream (orb_set_call_stream), never the legacy default stream
    if (dist_coef[0] == 0.0f) {   // `mvKeysUn = mvKeys` (Frame.cc:411-415)
        if (keys != keys_un) {
            if (loc == ORB_DEVICE) ORB_CUDA_TRY(cudaMemcpyAsync(keys_un, keys, bytes, cudaMemcpyDeviceToDevice, st));
            else memcpy(keys_un, keys, bytes);
        }
        return ORB_OK;
    }
    if (loc == ORB_DEVICE) {
        k_undistort_keypoints<<<(n + 127) / 128, 128, 0, st>>>(keys, n, c, keys_un);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    DevBuf a, b;
    ORB_CUDA_TRY(a.alloc(bytes));
    ORB_CUDA_TRY(b.alloc(bytes));
    ORB_CUDA_TRY(cudaMemcpyAsync(a.p, keys, bytes, cudaMemcpyHostToDevice, st));
    k_undistort_keypoints<<<(n + 127) / 128, 128, 0, st>>>((const orb_keypoint *)a.p, n, c, (orb_keypoint *)b.p);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(keys_un, b.p, bytes, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_undistort_points(const float *xy, int n, float fx, float fy, float cx, float cy, const float *dist_coef, int ndist,
                         float *xy_un, int loc) {
    ORB_REQUIRE(n >= 0 && (n == 0 || (xy && xy_un)), ORB_ERR_INVALID, "bad argument");
    Cam c;
    int rc = make_cam(fx, fy, cx, cy, dist_coef, ndist, &c);
    if (rc != ORB_OK) return rc;
    if (n == 0) return ORB_OK;
    const size_t bytes = (size_t)n * 8;
    ThreadCtx *T = nullptr;
    rc = orbgpu_thread_ctx(&T, 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    if (loc == ORB_DEVICE) {
        k_undistort_points<<<(n + 127) / 128, 128, 0, st>>>((const float2 *)xy, n, c, (float2 *)xy_un);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    DevBuf a, b;
    ORB_CUDA_TRY(a.alloc(bytes));
    ORB_CUDA_TRY(b.alloc(bytes));
    ORB_CUDA_TRY(cudaMemcpyAsync(a.p, xy, bytes, cudaMemcpyHostToDevice, st));
    k_undistort_points<<<(n + 127) / 128, 128, 0, st>>>((const float2 *)a.p, n, c, (float2 *)b.p);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(xy_un, b.p, bytes, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

}  // extern "C"
