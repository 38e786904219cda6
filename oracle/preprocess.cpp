// ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the image pre-processing the reference applies between the
// camera and ORBextractor::operator():
//   * 16-bit -> 8-bit: cv::imread(IMREAD_GRAYSCALE) of the 16-bit TUM-VI PNGs keeps the high byte
//     (/root/reference/test/test_Stereo.cpp:100-106; tools/make_fixtures.py does the same for tests/data);
//   * stereo rectification: cv::remap(src, dst, M1, M2, cv::INTER_LINEAR) with CV_32F maps, BORDER_CONSTANT 0
//     (/root/reference/Examples/ROS/GF_ORB_SLAM2/src/ros_stereo.cc:290-295; test/test_Stereo.cpp:197-210);
//   * colour -> gray: cv::cvtColor(CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY)
//     (/root/reference/src/Tracking.cc:366-382), applied AFTER the caller's remap, as the reference orders them.
// The arithmetic lives in OpenCV (un-vendored, pinned 3.4.1); restated from OpenCV's published algorithm and pinned
// bit-for-bit against the cv2 4.13 in this image (tests/test_oracle_cv2.py):
//   remap 8U bilinear: sx = cvRound(mapx * 32), sy likewise; (ix, iy) = (sx >> 5, sy >> 5) saturated to short,
//   (fx, fy) = low 5 bits; weights = the 32x32 bilinear table in Q15, exact multiples of 32, so
//   dst = ((32-fx)(32-fy) p00 + fx (32-fy) p01 + (32-fx) fy p10 + fx fy p11 + 512) >> 10, taps outside the source = 0;
//   RGB2GRAY 8U: (R * 9798 + G * 19235 + B * 3735 + 16384) >> 15.
#include <algorithm>
#include <cmath>
#include <cstdint>

#include "cvprim.hpp"
#include "gforb_oracle.h"

namespace {
inline int tap(const uint8_t* src, int w, int h, size_t pitch, int fmt, int ch, int x, int y) {
    if (x < 0 || x >= w || y < 0 || y >= h) return 0;
    const uint8_t* row = src + (size_t)y * pitch;
    switch (fmt) {
        case 0: return row[x];
        case 1: return reinterpret_cast<const uint16_t*>(row)[x] >> 8;
        case 2: case 3: return row[3 * x + ch];
        default: return row[4 * x + ch];
    }
}
}  // namespace

extern "C" int orc_preprocess(const void* src_, int sw, int sh, size_t spitch, int fmt, const float* mapx, const float* mapy,
                              int dw, int dh, uint8_t* dst, size_t dpitch) {
    const uint8_t* src = static_cast<const uint8_t*>(src_);
    if (fmt < 0 || fmt > 5) return -1;
    const int nch = fmt <= 1 ? 1 : 3;
    const bool bgr = fmt == 3 || fmt == 5;
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            int v[3];
            if (mapx) {
                const int sx = orc::cv_round(mapx[(size_t)y * dw + x] * 32.0f), sy = orc::cv_round(mapy[(size_t)y * dw + x] * 32.0f);
                const int ix = std::min(std::max(sx >> 5, -32768), 32767), iy = std::min(std::max(sy >> 5, -32768), 32767);
                const int fx = sx & 31, fy = sy & 31;
                for (int c = 0; c < nch; ++c) {
                    const int a = (32 - fx) * (32 - fy) * tap(src, sw, sh, spitch, fmt, c, ix, iy) +
                                  fx * (32 - fy) * tap(src, sw, sh, spitch, fmt, c, ix + 1, iy) +
                                  (32 - fx) * fy * tap(src, sw, sh, spitch, fmt, c, ix, iy + 1) +
                                  fx * fy * tap(src, sw, sh, spitch, fmt, c, ix + 1, iy + 1);
                    v[c] = (a + 512) >> 10;
                }
            } else {
                for (int c = 0; c < nch; ++c) v[c] = tap(src, sw, sh, spitch, fmt, c, x, y);
            }
            int g = v[0];
            if (nch == 3) {
                const int r = bgr ? v[2] : v[0], b = bgr ? v[0] : v[2];
                g = (r * 9798 + v[1] * 19235 + b * 3735 + 16384) >> 15;
            }
            dst[(size_t)y * dpitch + x] = (uint8_t)g;
        }
    return 0;
}
