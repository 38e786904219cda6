// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// CPU restatements of the OpenCV primitives the reference's ORB front-end calls.  OpenCV is an
// un-vendored dependency of the reference (pinned 3.4.1, /root/reference/CMakeLists.txt:116-125);
// each function states the published algorithm it restates and the reference call site it serves.
// Every primitive here is cross-checked bit-for-bit against cv2 4.13.0 in tests/test_oracle_cv2.py.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orc {

// cvRound: round-half-to-even (SSE cvtss2si / lrint semantics).  Call sites:
// src/ORBextractor.cc:80,114,118-119,441,459,1181.
int cv_round(float v);
int cv_round(double v);

// BORDER_REFLECT_101 index map: gfedcb|abcdefgh|gfedcba.  Call sites: src/ORBextractor.cc:1191-1197, 1155.
int reflect101(int p, int len);

// cv::copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101) into a (w+2b) x (h+2b) buffer.
void copy_make_border_reflect101(const uint8_t* src, int w, int h, size_t spitch,
                                 uint8_t* dst, size_t dpitch, int border);

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1: 11-bit fixed-point separable bilinear
// (OpenCV imgproc resize.cpp: resizeGeneric_<HResizeLinear<uchar,int,short,2048>, VResizeLinear<...>>),
// including the exact-2x shortcut to INTER_AREA.  Call site: src/ORBextractor.cc:1189.
void resize_linear_8u(const uint8_t* src, int sw, int sh, size_t spitch,
                      uint8_t* dst, int dw, int dh, size_t dpitch);

// cv::FAST(img, kps, threshold, nonmaxSuppression=true), TYPE_9_16 (OpenCV features2d fast.cpp FAST_t<16>,
// fast_score.cpp cornerScore<16>; in-tree restatement of the same algorithm: src/FAST_NEON.cc:16-66,91-287).
// Call sites: src/ORBextractor.cc:811-817.  Output in emission order (row-major).
struct FastPoint { int x, y, score; };
void fast9_16(const uint8_t* img, int w, int h, size_t pitch, int threshold, bool nonmax,
              std::vector<FastPoint>& out);
// cornerScore<16>: largest threshold for which the pixel is still a FAST-9 corner.
int fast_corner_score(const uint8_t* p, size_t pitch, int threshold);

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 (OpenCV >= 3.4.2 / 4.x
// fixed-point path: Q8 kernel {18,34,48,56,48,34,18}, u16 rows, 32-bit columns, round-half-up).
// Call site: src/ORBextractor.cc:1155.
void gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch);

// cv::fastAtan2(y, x) in degrees [0,360) (OpenCV core mathfuncs_core atan_f32: 7th-order odd polynomial).
// Call site: src/ORBextractor.cc:102.
float fast_atan2(float y, float x);

}  // namespace orc
