// liborbgpu internals: the fp32 / fp64 device routines whose results must equal the host's bit for bit (cv::fastAtan2,
// glibc's sincosf), shared by the ORB-SLAM2 extractor (orb_extractor.cu) and OpenCV's cv::ORB (orb_cvorb.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbgpu {

__device__ __forceinline__ float dev_fast_atan2(float y, float x) {
    // cv::fastAtan2; every operation rounded separately (no FMA)
    const float p1 = 57.283627f, p3 = -18.667446f, p5 = 8.9140005f, p7 = -2.5397246f;
    const float eps = 2.2204460492503131e-16f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc sincosf (the reference's cosf/sinf, ref :103-104) restated with unfused double
// arithmetic; bit-identical to glibc 2.39 on every fp32 angle in [0, 2*pi).
__device__ __forceinline__ float dev_sincos_poly(double x, double x2, bool alt, int n) {
    if ((n & 1) == 0) {
        const double s1c = -0x1.555545995a603p-3, s2c = 0x1.1107605230bc4p-7, s3c = -0x1.994eb3774cf24p-13;
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(s2c, __dmul_rn(x2, s3c));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, s1c));
        return __double2float_rn(__dadd_rn(s, __dmul_rn(x7, s1)));
    }
    const double sg = alt ? -1.0 : 1.0;
    const double c0 = sg * 0x1p0, c1c = sg * -0x1.ffffffd0c621cp-2, c2c = sg * 0x1.55553e1068f19p-5,
                 c3c = sg * -0x1.6c087e89a359dp-10, c4c = sg * 0x1.99343027bf8c3p-16;
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __dadd_rn(c3c, __dmul_rn(x2, c4c));
    const double c1 = __dadd_rn(c0, __dmul_rn(x2, c1c));
    const double x6 = __dmul_rn(x4, x2);
    const double c = __dadd_rn(c1, __dmul_rn(x4, c2c));
    return __double2float_rn(__dadd_rn(c, __dmul_rn(x6, c2)));
}
__device__ __forceinline__ void dev_sincosf(float y, float *sn, float *cs) {
    double x = (double)y;
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    if (top < 0x3f4) {   // |y| < 0.75
        if (top < 0x398) { *sn = y; *cs = 1.0f; return; }   // |y| < 2^-12
        const double x2 = __dmul_rn(x, x);
        *sn = dev_sincos_poly(x, x2, false, 0);
        *cs = dev_sincos_poly(x, x2, false, 1);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    const bool alt = (n & 2) != 0;
    const double hpi = alt ? -0x1.921FB54442D18p0 : 0x1.921FB54442D18p0;
    // reduce_fast uses table[0].hpi: x - n*hpi
    x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    (void)hpi;
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const double xs = __dmul_rn(x, sgn), x2 = __dmul_rn(x, x);
    *sn = dev_sincos_poly(xs, x2, alt, n);
    *cs = dev_sincos_poly(xs, x2, alt, n ^ 1);
}


}  // namespace orbgpu
