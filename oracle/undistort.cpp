// ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement of Frame::UndistortKeyPoints / UndistortKeyPointsStereo and the
// corner part of Frame::ComputeImageBounds (/root/reference/src/Frame.cc:670-758, 760-887): keypoint coordinates go
// through cv::undistortPoints(mat, mat, K, distCoef, R, P) and come back as float.
// cv::undistortPoints lives in OpenCV (un-vendored, pinned 3.4.1); restated from its published algorithm
// (calib3d undistort: cvUndistortPoints, default criteria = 5 fixed iterations) in double precision, operation for
// operation, and pinned against the cv2 4.13 in this image (tests/test_oracle_cv2.py):
//   x = (u - cx) / fx  [as (u - cx) * (1 / fx)],  y likewise;  x0 = x, y0 = y
//   5 x { r2 = x*x + y*y;  icdist = (1 + ((k6 r2 + k5) r2 + k4) r2) / (1 + ((k3 r2 + k2) r2 + k1) r2);
//         if icdist < 0: x, y = the normalised input, stop;
//         dX = 2 p1 x y + p2 (r2 + 2 x x);  dY = p1 (r2 + 2 y y) + 2 p2 x y;  x = (x0 - dX) icdist;  y = (y0 - dY) icdist }
//   [xx yy ww] = (P * R) [x y 1];  out = (float)(xx / ww) computed as xx * (1 / ww)
#include <cmath>
#include <cstdint>

#include "gforb_oracle.h"

extern "C" void orc_undistort_matrix(const double* P9, const double* R9, double* RR) {
    // RR = P(3x3) * R, row by column, accumulated left to right from zero (cv::gemm on 3x3 doubles)
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += P9[3 * i + k] * R9[3 * k + j];
            RR[3 * i + j] = s;
        }
}

extern "C" void orc_undistort_points(const float* xy_in, int n, const double* K9, const double* dist8, const double* RR,
                                     float* xy_out) {
    const double fx = K9[0], fy = K9[4], cx = K9[2], cy = K9[5], ifx = 1. / fx, ify = 1. / fy;
    const double k1 = dist8[0], k2 = dist8[1], p1 = dist8[2], p2 = dist8[3], k3 = dist8[4], k4 = dist8[5], k5 = dist8[6], k6 = dist8[7];
    for (int i = 0; i < n; ++i) {
        const double u = xy_in[2 * i], v = xy_in[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k6 * r2 + k5) * r2 + k4) * r2) / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
            const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = RR[0] * x + RR[1] * y + RR[2], yy = RR[3] * x + RR[4] * y + RR[5];
        const double ww = 1. / (RR[6] * x + RR[7] * y + RR[8]);
        xy_out[2 * i] = (float)(xx * ww);
        xy_out[2 * i + 1] = (float)(yy * ww);
    }
}
