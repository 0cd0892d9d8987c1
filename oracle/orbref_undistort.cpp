// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of cv::undistortPoints(src, dst, K, D, noArray(), K) as Frame::UndistortKeyPoints and
// Frame::ComputeImageBounds call it (ORB_SLAM2/src/Frame.cc:409-468).  OpenCV is an un-vendored dependency; the arithmetic
// (OpenCV >= 3.x: double precision, TermCriteria(MAX_ITER, 5, 0.01), early exit when the radial factor turns negative) is
// pinned against cv2 4.13.0 by tests/golden/cv2_undistort.npz (tools/make_golden_undistort.py).
#include "orbref.h"

extern "C" int orbref_undistort_points(const float *xy, int n, float ffx, float ffy, float fcx, float fcy, const float *dist,
                                       int ndist, float *out) {
    if (ndist != 4 && ndist != 5) return ORB_ERR_INVALID;
    double k[14] = {0};
    for (int i = 0; i < ndist; i++) k[i] = dist[i];
    const double fx = ffx, fy = ffy, cx = fcx, cy = fcy, ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        const double u = xy[2 * i], v = xy[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        out[2 * i] = (float)(xx * ww);
        out[2 * i + 1] = (float)(yy * ww);
    }
    return ORB_OK;
}
