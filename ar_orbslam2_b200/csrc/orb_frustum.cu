// liborbgpu: Frame::isInFrustum for all local map points of a frame.
//
// ref: ORB_SLAM2/src/Frame.cc:274-330, MapPoint::PredictScale (src/MapPoint.cc:357-366), caller Tracking::SearchLocalPoints
// (src/Tracking.cc:1527-1539).  One thread per map point.  The cv::Mat algebra of the reference (3x3 * 3x1 products, L2 norm,
// dot product) accumulates in double and rounds to float, which is written out explicitly here; everything is compiled with
// --fmad=false.  logf restates glibc's algorithm (sysdeps/ieee754/flt-32/e_logf.c: 16-entry table, degree-3 polynomial in
// double): verified bit-identical to the host libm for every positive finite float, with and without FMA contraction.
#include "orb_common.cuh"

using namespace orbgpu;

namespace {

__constant__ double c_logf_tab[16][2] = {
    {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
    {0x1.49539f0f010b0p+0, -0x1.01eae7f513a67p-2}, {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
    {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8ea0p+0, -0x1.1aa2bc79c8100p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
    {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aa0p-1, 0x1.c5e53aa362eb4p-4},
    {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d224770p-3},
    {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};

__device__ float glibc_logf(float x) {
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -__int_as_float(0x7f800000);                                          // log(+-0) = -inf
        if (ix == 0x7f800000u) return x;                                                              // log(inf) = inf
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return __int_as_float(0x7fc00000);           // negative or NaN
        ix = __float_as_uint(__fmul_rn(x, 0x1p23f));                                                  // subnormal: normalise
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) & 15, k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & (0x1ffu << 23));
    const double invc = c_logf_tab[i][0], logc = c_logf_tab[i][1], z = (double)__uint_as_float(iz);
    const double r = z * invc - 1, y0 = logc + (double)k * 0x1.62e42fefa39efp-1, r2 = r * r;
    double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
    y = -0x1.00ea348b88334p-2 * r2 + y;
    y = y * r2 + (y0 + r);
    return (float)y;
}

__global__ void k_is_in_frustum(const orb_frustum_camera cam, int n, const float *__restrict__ wp, const float *__restrict__ nrm,
                                const float *__restrict__ maxd, const float *__restrict__ mind, const uint8_t *__restrict__ consider,
                                float limit, uint8_t *__restrict__ in_view, float *__restrict__ px, float *__restrict__ py,
                                float *__restrict__ pxr, int *__restrict__ level, float *__restrict__ vcos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool ok = !consider || consider[i];
    float u = 0, v = 0, ur = 0, vc = 0;
    int lvl = 0;
    if (ok) {
        const float P[3] = {wp[3 * i], wp[3 * i + 1], wp[3 * i + 2]};
        float Pc[3];
#pragma unroll
        for (int r = 0; r < 3; r++) {   // mRcw * P + mtcw: the product in double (cv::gemm), then a float addition
            double s = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) s += (double)cam.Rcw[3 * r + k] * (double)P[k];
            Pc[r] = __fadd_rn((float)s, cam.tcw[r]);
        }
        ok = !(Pc[2] < 0.0f);                                                         // ref :288-289
        const float invz = __fdiv_rn(1.0f, Pc[2]);
        u = __fadd_rn(__fmul_rn(__fmul_rn(cam.fx, Pc[0]), invz), cam.cx);             // ref :293-294
        v = __fadd_rn(__fmul_rn(__fmul_rn(cam.fy, Pc[1]), invz), cam.cy);
        ok = ok && !(u < cam.min_x || u > cam.max_x) && !(v < cam.min_y || v > cam.max_y);
        const float maxDistance = __fmul_rn(1.2f, maxd[i]), minDistance = __fmul_rn(0.8f, mind[i]);
        const float PO[3] = {__fsub_rn(P[0], cam.Ow[0]), __fsub_rn(P[1], cam.Ow[1]), __fsub_rn(P[2], cam.Ow[2])};
        double s2 = 0, sd = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) { s2 += (double)PO[k] * (double)PO[k]; sd += (double)PO[k] * (double)nrm[3 * i + k]; }
        const float dist = (float)sqrt(s2);                                           // cv::norm(PO), ref :305
        ok = ok && !(dist < minDistance || dist > maxDistance);
        vc = (float)(sd / (double)dist);                                              // PO.dot(Pn)/dist, ref :313
        ok = ok && !(vc < limit);
        const float ratio = __fdiv_rn(maxd[i], dist);                                 // MapPoint::PredictScale
        lvl = (int)ceilf(__fdiv_rn(glibc_logf(ratio), cam.log_scale_factor));
        ur = __fsub_rn(u, __fmul_rn(cam.mbf, invz));                                  // ref :324
    }
    in_view[i] = ok ? 1 : 0;
    px[i] = ok ? u : 0.0f; py[i] = ok ? v : 0.0f; pxr[i] = ok ? ur : 0.0f; vcos[i] = ok ? vc : 0.0f;
    level[i] = ok ? lvl : 0;
}

// The projection pre-loops of the mapping / loop-closing matchers (ref src/ORBmatcher.cc:1491-1531 SearchByProjection(Frame&,
// KeyFrame*), :311-360 SearchByProjection(KeyFrame*, Scw), :846-893 and :1004-1048 Fuse, :1150-1226 SearchBySim3): the same shape
// as isInFrustum with the reference's small differences between the sites kept as rules (how 1/z is rounded, how u is
// associated, which bounds test, which distance, whether the viewing angle is tested).  MapPoint::PredictScale stays with the
// caller: it is a public method of the reference's MapPoint, the raw mfMaxDistance it needs is not.
__global__ void k_project_points(const orb_projection P, int n, const float *__restrict__ wp, const float *__restrict__ nrm,
                                 const float *__restrict__ maxd, const float *__restrict__ mind, const uint8_t *__restrict__ consider,
                                 uint8_t *__restrict__ ok_out, float *__restrict__ pu, float *__restrict__ pv, float *__restrict__ pur,
                                 float *__restrict__ pdist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool ok = !consider || consider[i];
    float u = 0, v = 0, ur = 0, dist = 0;
    if (ok) {
        const float X[3] = {wp[3 * i], wp[3 * i + 1], wp[3 * i + 2]};
        float a[3], c[3];
#pragma unroll
        for (int r = 0; r < 3; r++) {   // R * X + t: the product in double (cv::gemm), then a float addition
            double s = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) s += (double)P.R1[3 * r + k] * (double)X[k];
            a[r] = __fadd_rn((float)s, P.t1[r]);
        }
        if (P.two_stage) {
#pragma unroll
            for (int r = 0; r < 3; r++) {
                double s = 0;
#pragma unroll
                for (int k = 0; k < 3; k++) s += (double)P.R2[3 * r + k] * (double)a[k];
                c[r] = __fadd_rn((float)s, P.t2[r]);
            }
        } else {
            c[0] = a[0]; c[1] = a[1]; c[2] = a[2];
        }
        if (P.reject_negative_z) ok = !(c[2] < 0.0f);
        const float invz = P.invz_rule == 0 ? (float)(1.0 / (double)c[2]) : __fdiv_rn(1.0f, c[2]);
        if (P.uv_rule == 0) {
            u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, c[0]), invz), P.cx);
            v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, c[1]), invz), P.cy);
        } else {
            u = __fadd_rn(__fmul_rn(P.fx, __fmul_rn(c[0], invz)), P.cx);
            v = __fadd_rn(__fmul_rn(P.fy, __fmul_rn(c[1], invz)), P.cy);
        }
        if (P.bounds_rule == 0) ok = ok && !(u < P.min_x || u > P.max_x) && !(v < P.min_y || v > P.max_y);
        else ok = ok && (u >= P.min_x && u < P.max_x && v >= P.min_y && v < P.max_y);
        ur = __fsub_rn(u, __fmul_rn(P.bf, invz));
        double s2 = 0, sd = 0;
        if (P.dist_rule == 0) {
            const float PO[3] = {__fsub_rn(X[0], P.Ow[0]), __fsub_rn(X[1], P.Ow[1]), __fsub_rn(X[2], P.Ow[2])};
#pragma unroll
            for (int k = 0; k < 3; k++) {
                s2 += (double)PO[k] * (double)PO[k];
                if (P.check_view_angle) sd += (double)PO[k] * (double)nrm[3 * i + k];
            }
        } else {
#pragma unroll
            for (int k = 0; k < 3; k++) s2 += (double)c[k] * (double)c[k];
        }
        dist = (float)sqrt(s2);                                                       // cv::norm
        ok = ok && !(dist < mind[i] || dist > maxd[i]);
        if (P.check_view_angle) ok = ok && !(sd < 0.5 * (double)dist);                // PO.dot(Pn) < 0.5 * dist
    }
    ok_out[i] = ok ? 1 : 0;
    pu[i] = ok ? u : 0.0f; pv[i] = ok ? v : 0.0f; pur[i] = ok ? ur : 0.0f; pdist[i] = ok ? dist : 0.0f;
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

}  // namespace

extern "C" int orb_is_in_frustum(const orb_frustum_camera *cam, const orb_frustum_points *pts, float viewing_cos_limit, int loc,
                                 uint8_t *in_view, float *proj_x, float *proj_y, float *proj_xr, int32_t *level, float *view_cos) {
    ORB_REQUIRE(cam && pts && pts->n >= 0, ORB_ERR_INVALID, "bad argument");
    const int n = pts->n;
    if (n == 0) return ORB_OK;
    ORB_REQUIRE(pts->world_pos && pts->normal && pts->max_distance && pts->min_distance && in_view && proj_x && proj_y && proj_xr &&
                    level && view_cos, ORB_ERR_INVALID, "null array");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    const int blocks = (n + 127) / 128;
    ThreadCtx *T = nullptr;
    { int rc = orbgpu_thread_ctx(&T, 0, 0); if (rc != ORB_OK) return rc; }
    cudaStream_t st = T->stream;   // the thread's call stream (orb_set_call_stream), never the legacy default stream
    if (loc == ORB_DEVICE) {
        k_is_in_frustum<<<blocks, 128, 0, st>>>(*cam, n, pts->world_pos, pts->normal, pts->max_distance, pts->min_distance, pts->consider,
                                         viewing_cos_limit, in_view, proj_x, proj_y, proj_xr, level, view_cos);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    // host arrays: one staging buffer in, one out
    const size_t fin = (size_t)n * 8 * 4, bin = fin + (size_t)n, fout = (size_t)n * 5 * 4, bout = fout + (size_t)n;
    DevBuf din, dout;
    ORB_CUDA_TRY(din.alloc(bin));
    ORB_CUDA_TRY(dout.alloc(bout));
    float *d_wp = (float *)din.p, *d_nr = d_wp + 3 * (size_t)n, *d_mx = d_nr + 3 * (size_t)n, *d_mn = d_mx + n;
    uint8_t *d_cons = (uint8_t *)(d_mn + n);
    ORB_CUDA_TRY(cudaMemcpyAsync(d_wp, pts->world_pos, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d_nr, pts->normal, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d_mx, pts->max_distance, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d_mn, pts->min_distance, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if (pts->consider) ORB_CUDA_TRY(cudaMemcpyAsync(d_cons, pts->consider, (size_t)n, cudaMemcpyHostToDevice, st));
    float *o_x = (float *)dout.p, *o_y = o_x + n, *o_xr = o_y + n, *o_vc = o_xr + n;
    int *o_lv = (int *)(o_vc + n);
    uint8_t *o_in = (uint8_t *)(o_lv + n);
    k_is_in_frustum<<<blocks, 128, 0, st>>>(*cam, n, d_wp, d_nr, d_mx, d_mn, pts->consider ? d_cons : nullptr, viewing_cos_limit, o_in, o_x, o_y,
                                     o_xr, o_lv, o_vc);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(proj_x, o_x, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(proj_y, o_y, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(proj_xr, o_xr, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(view_cos, o_vc, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(level, o_lv, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(in_view, o_in, (size_t)n, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

extern "C" int orb_project_points(const orb_projection *proj, const orb_projection_points *pts, int loc, uint8_t *ok, float *u, float *v,
                                  float *u_right, float *dist) {
    ORB_REQUIRE(proj && pts && pts->n >= 0, ORB_ERR_INVALID, "bad argument");
    const int n = pts->n;
    if (n == 0) return ORB_OK;
    ORB_REQUIRE(pts->world_pos && pts->max_distance && pts->min_distance && ok && u && v && u_right && dist, ORB_ERR_INVALID, "null array");
    ORB_REQUIRE(!(proj->check_view_angle && proj->dist_rule == 0) || pts->normal, ORB_ERR_INVALID, "the viewing-angle test needs the normals");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    const int blocks = (n + 127) / 128;
    const size_t N = (size_t)n, a = (N * 4 + 255) / 256 * 256;   // one float array, 256-byte aligned
    const size_t need_d = 3 * a + 3 * a + 2 * a + a + 4 * a + a, need_h = need_d;
    ThreadCtx *T = nullptr;
    { int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? need_d : 0, loc == ORB_HOST ? need_h : 0); if (rc != ORB_OK) return rc; }
    cudaStream_t st = T->stream;
    if (loc == ORB_DEVICE) {
        k_project_points<<<blocks, 128, 0, st>>>(*proj, n, pts->world_pos, pts->normal, pts->max_distance, pts->min_distance, pts->consider, ok,
                                               u, v, u_right, dist);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    // host arrays: packed into the thread's pinned block, one copy in, one copy out
    uint8_t *h = T->h_buf, *d = T->d_buf;
    const size_t o_wp = 0, o_nr = 3 * a, o_mx = 6 * a, o_mn = 7 * a, o_cons = 8 * a, o_out = 9 * a;   // outputs: u, v, ur, dist, ok
    memcpy(h + o_wp, pts->world_pos, N * 12);
    if (pts->normal) memcpy(h + o_nr, pts->normal, N * 12);
    memcpy(h + o_mx, pts->max_distance, N * 4);
    memcpy(h + o_mn, pts->min_distance, N * 4);
    if (pts->consider) memcpy(h + o_cons, pts->consider, N);
    ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_out, cudaMemcpyHostToDevice, st));
    float *du = (float *)(d + o_out), *dv = (float *)(d + o_out + a), *dur = (float *)(d + o_out + 2 * a), *dd = (float *)(d + o_out + 3 * a);
    uint8_t *dok = d + o_out + 4 * a;
    k_project_points<<<blocks, 128, 0, st>>>(*proj, n, (const float *)(d + o_wp), pts->normal ? (const float *)(d + o_nr) : nullptr,
                                           (const float *)(d + o_mx), (const float *)(d + o_mn), pts->consider ? d + o_cons : nullptr, dok, du,
                                           dv, dur, dd);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(h + o_out, d + o_out, 5 * a, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(u, h + o_out, N * 4); memcpy(v, h + o_out + a, N * 4); memcpy(u_right, h + o_out + 2 * a, N * 4);
    memcpy(dist, h + o_out + 3 * a, N * 4); memcpy(ok, h + o_out + 4 * a, N);
    return ORB_OK;
}
