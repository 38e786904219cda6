// ORACLE -- TEST INFRASTRUCTURE ONLY.  The image functions of the cv:: stand-in (cvshim.hpp): each forwards to
// the cv2-pinned restatement in oracle/cvprim.cpp and keeps OpenCV's calling conventions (create() of the
// destination, in-place operation, cleared output vector).
#include "cvshim.hpp"

#include <vector>

#include "../gforb_oracle.h"   // orc_undistort_matrix / orc_undistort_points (oracle/undistort.cpp)

namespace cv {

thread_local FastLog* g_fast_log = nullptr;

// cv::FAST(image, keypoints, threshold, nonmax): TYPE_9_16; keypoints are cleared and refilled row-major as
// KeyPoint(x, y, 7.f, -1, score).  Call sites: src/ORBextractor.cc:811-817.
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    const Mat img = image.getMat();
    assert(img.type() == CV_8UC1);
    std::vector<orc::FastPoint> pts;
    orc::fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmaxSuppression, pts);
    keypoints.clear();
    keypoints.reserve(pts.size());
    for (const orc::FastPoint& p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1.f, (float)p.score));
    if (g_fast_log) {
        g_fast_log->calls.push_back(FastCall{img.data, img.step, g_fast_log->pts.size(), threshold});
        for (const orc::FastPoint& p : pts) g_fast_log->pts.push_back(FastLogEntry{img.data, img.step, p.x, p.y, p.score});
    }
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) on CV_8UC1.  Call site: src/ORBextractor.cc:1189 (dst is the ROI
// view of the padded level buffer; create() leaves it in place).
void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation) {
    const Mat src = src_.getMat();
    assert(src.type() == CV_8UC1 && fx == 0 && fy == 0 && interpolation == INTER_LINEAR);
    (void)fx; (void)fy; (void)interpolation;
    dst_.create(dsize.height, dsize.width, src.type());
    Mat dst = dst_.getMat();
    assert(dst.data != src.data);
    orc::resize_linear_8u(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

// cv::copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101 [+ BORDER_ISOLATED]).  Call sites:
// src/ORBextractor.cc:1191-1197.  Without BORDER_ISOLATED OpenCV would read real pixels around a src that is a
// ROI of a larger image; the level-0 source here is always a whole image (asserted by the caller in ref_capi.cpp),
// so both forms reflect from the source itself.  src may be the interior of dst (level > 0): the interior copy is
// then the identity and only the frame is written.
void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType,
                    const Scalar&) {
    const Mat src = src_.getMat();
    assert(src.type() == CV_8UC1 && (borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    assert(top == bottom && left == right && top == left);
    (void)bottom; (void)right; (void)left; (void)borderType;
    dst_.create(src.rows + 2 * top, src.cols + 2 * top, src.type());
    Mat dst = dst_.getMat();
    orc::copy_make_border_reflect101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top);
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) on CV_8UC1, in place allowed.
// Call site: src/ORBextractor.cc:1155.
void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType) {
    const Mat src = src_.getMat();
    assert(src.type() == CV_8UC1 && ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 &&
           borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    dst_.create(src.rows, src.cols, src.type());
    Mat dst = dst_.getMat();
    orc::gaussian_blur_7x7_s2(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

// cv::undistortPoints(src, dst, K, D, R, P): N x 1 two-channel CV_32F points, every matrix CV_32F as the reference holds them
// (src/Tracking.cc:144-168 converts the yaml's doubles to CV_32F).  OpenCV converts K, D, R, P to double, multiplies P(3x3) * R
// in double and iterates in double: oracle/undistort.cpp, pinned against cv2.undistortPoints by tests/test_oracle_cv2.py.
// Call sites: src/Frame.cc:688,710,745,795,858.
void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray D_, InputArray R_, InputArray P_) {
    const Mat src = src_.getMat(), K = K_.getMat(), D = D_.getMat();
    assert(src.type() == CV_32F && src.channels() == 2 && src.cols == 1);
    assert(K.type() == CV_32F && K.rows == 3 && K.cols == 3 && D.type() == CV_32F);
    const int n = src.rows, nd = D.rows * D.cols;
    assert(nd == 4 || nd == 5 || nd == 8);
    double K9[9], d8[8] = {0, 0, 0, 0, 0, 0, 0, 0}, R9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, P9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, RR[9];
    for (int i = 0; i < 9; ++i) K9[i] = (double)K.at<float>(i / 3, i % 3);
    for (int i = 0; i < nd; ++i) d8[i] = (double)D.at<float>(i);
    if (!R_.empty()) {
        const Mat R = R_.getMat();
        assert(R.type() == CV_32F && R.rows == 3 && R.cols == 3);
        for (int i = 0; i < 9; ++i) R9[i] = (double)R.at<float>(i / 3, i % 3);
    }
    if (!P_.empty()) {
        const Mat P = P_.getMat();
        assert(P.type() == CV_32F && P.rows == 3 && (P.cols == 3 || P.cols == 4));
        for (int i = 0; i < 9; ++i) P9[i] = (double)P.at<float>(i / 3, i % 3);
    }
    orc_undistort_matrix(P9, R9, RR);
    std::vector<float> in((size_t)2 * n + 2), out((size_t)2 * n + 2);
    for (int i = 0; i < n; ++i) { in[2 * i] = src.ptr<float>(i)[0]; in[2 * i + 1] = src.ptr<float>(i)[1]; }
    orc_undistort_points(in.data(), n, K9, d8, RR, out.data());
    Mat* dm = dst_.mat_ptr();
    if (!(dm->data && dm->rows == n && dm->cols == 1 && dm->channels() == 2 && dm->type() == CV_32F))
        dst_.assign(Mat(n, 2, CV_32F).reshape(2));
    Mat dst = dst_.getMat();
    for (int i = 0; i < n; ++i) { dst.ptr<float>(i)[0] = out[2 * i]; dst.ptr<float>(i)[1] = out[2 * i + 1]; }
}

}  // namespace cv
