// Keypoint undistortion: Frame::UndistortKeyPoints / UndistortKeyPointsStereo (reference src/Frame.cc:670-758) and the
// eight border points of Frame::ComputeImageBounds[Stereo] (:760-887).  Both push float coordinates through
// cv::undistortPoints(mat, mat, K, distCoef, R, P) -- OpenCV's fixed 5-iteration inverse of the radial-tangential model
// in double precision -- and store the result as float.  Restated operation for operation (the library is compiled with
// --fmad=false, so no multiply-add is contracted); the oracle of the same sequence is bit-identical to cv2
// (tests/test_oracle_cv2.py), and IEEE double add/mul/div give the same bits on the device.
// One thread per keypoint; everything but pt.x / pt.y is copied through (`cv::KeyPoint kp = mvKeys[i]`, :695-699).
// Latency-bound toy next to the pixel kernels (2,000 points x ~150 double ops): it exists so that the default build's
// path (undistort -> ComputeStereoMatches_Undistorted -> grid) never has to leave the device.
#include "gforb_internal.cuh"

namespace gforb {

struct UndistortArgs {
    double fx, fy, cx, cy, ifx, ify;
    double k1, k2, p1, p2, k3, k4, k5, k6;
    double rr[9];
    int passthrough;            // distCoef[0] == 0: the reference copies the keys unchanged (:672-677, :726-730)
};

__device__ __forceinline__ void undistort_point(const UndistortArgs& a, float uf, float vf, float& xo, float& yo) {
    const double u = uf, v = vf;
    double x = (u - a.cx) * a.ifx, y = (v - a.cy) * a.ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((a.k6 * r2 + a.k5) * r2 + a.k4) * r2) / (1 + ((a.k3 * r2 + a.k2) * r2 + a.k1) * r2);
        if (icdist < 0) { x = (u - a.cx) * a.ifx; y = (v - a.cy) * a.ify; break; }
        const double deltaX = 2 * a.p1 * x * y + a.p2 * (r2 + 2 * x * x);
        const double deltaY = a.p1 * (r2 + 2 * y * y) + 2 * a.p2 * x * y;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    const double xx = a.rr[0] * x + a.rr[1] * y + a.rr[2], yy = a.rr[3] * x + a.rr[4] * y + a.rr[5];
    const double ww = 1. / (a.rr[6] * x + a.rr[7] * y + a.rr[8]);
    xo = (float)(xx * ww);
    yo = (float)(yy * ww);
}

__global__ void __launch_bounds__(128)
undistort_keys_kernel(const UndistortArgs a, const gforb_keypoint* __restrict__ in, const int32_t* __restrict__ counts,
                      int n_fixed, int stride, gforb_keypoint* __restrict__ out) {
    const int img = blockIdx.y;
    const int n = counts ? min(counts[img], stride) : n_fixed;
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    gforb_keypoint kp = in[(size_t)img * stride + i];
    if (!a.passthrough) undistort_point(a, kp.x, kp.y, kp.x, kp.y);
    out[(size_t)img * stride + i] = kp;
}

__global__ void undistort_xy_kernel(const UndistortArgs a, const float* __restrict__ in, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float x = in[2 * i], y = in[2 * i + 1];
    if (!a.passthrough) undistort_point(a, x, y, x, y);
    out[2 * i] = x; out[2 * i + 1] = y;
}

// cam: K[9], dist[8], R[9], P[9] (doubles, row-major); RR = P * R accumulated left to right (cv::gemm on 3x3 doubles)
static UndistortArgs make_args(const double* K, const double* dist, const double* R, const double* P) {
    UndistortArgs a{};
    a.fx = K[0]; a.fy = K[4]; a.cx = K[2]; a.cy = K[5]; a.ifx = 1. / a.fx; a.ify = 1. / a.fy;
    a.k1 = dist[0]; a.k2 = dist[1]; a.p1 = dist[2]; a.p2 = dist[3]; a.k3 = dist[4]; a.k4 = dist[5]; a.k5 = dist[6]; a.k6 = dist[7];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += P[3 * i + k] * R[3 * k + j];
            a.rr[3 * i + j] = s;
        }
    a.passthrough = (float)dist[0] == 0.0f;
    return a;
}

void launch_undistort_keys(const double* K, const double* dist, const double* R, const double* P, const gforb_keypoint* in,
                           const int32_t* counts, int n_fixed, int stride, int nimg, gforb_keypoint* out, cudaStream_t s) {
    const int nmax = counts ? stride : n_fixed;
    if (nimg <= 0 || nmax <= 0) return;
    dim3 grid((nmax + 127) / 128, nimg);
    undistort_keys_kernel<<<grid, 128, 0, s>>>(make_args(K, dist, R, P), in, counts, n_fixed, stride, out);
}

void launch_undistort_xy(const double* K, const double* dist, const double* R, const double* P, const float* in, int n,
                         float* out, cudaStream_t s) {
    if (n <= 0) return;
    undistort_xy_kernel<<<(n + 127) / 128, 128, 0, s>>>(make_args(K, dist, R, P), in, n, out);
}

}  // namespace gforb
