// ORACLE -- TEST INFRASTRUCTURE ONLY (see cvprim.hpp).  Build with -ffp-contract=off.
#include "cvprim.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace orc {

int cv_round(float v) { return (int)lrintf(v); }   // default FE_TONEAREST => half-to-even
int cv_round(double v) { return (int)lrint(v); }

int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

void copy_make_border_reflect101(const uint8_t* src, int w, int h, size_t spitch,
                                 uint8_t* dst, size_t dpitch, int border) {
    const int W = w + 2 * border, H = h + 2 * border;
    std::vector<int> xmap(W);
    for (int x = 0; x < W; ++x) xmap[x] = reflect101(x - border, w);
    for (int y = 0; y < H; ++y) {
        const uint8_t* s = src + (size_t)reflect101(y - border, h) * spitch;
        uint8_t* d = dst + (size_t)y * dpitch;
        for (int x = 0; x < W; ++x) d[x] = s[xmap[x]];
    }
}

// ---------------------------------------------------------------------------------------------
// resize INTER_LINEAR, 8UC1
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int kCoefBits = 11;
constexpr int kCoefOne = 1 << kCoefBits;  // INTER_RESIZE_COEF_SCALE

inline short sat_short_round(float v) {
    int r = cv_round(v);
    return (short)std::min(std::max(r, -32768), 32767);
}

// per-axis source offset + the two 11-bit weights, with OpenCV's edge handling (weights collapse to
// {2048,0} on a clamped tap)
void linear_axis_table(int slen, int dlen, double scale, std::vector<int>& ofs, std::vector<short>& coef) {
    ofs.resize(dlen);
    coef.resize((size_t)dlen * 2);
    for (int d = 0; d < dlen; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= (float)s;
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= slen - 1) { f = 0.f; s = slen - 1; }
        ofs[d] = s;
        coef[2 * d + 0] = sat_short_round((1.f - f) * (float)kCoefOne);
        coef[2 * d + 1] = sat_short_round(f * (float)kCoefOne);
    }
}
}  // namespace

void resize_linear_8u(const uint8_t* src, int sw, int sh, size_t spitch,
                      uint8_t* dst, int dw, int dh, size_t dpitch) {
    const double inv_x = (double)dw / sw, inv_y = (double)dh / sh;
    const double scale_x = 1. / inv_x, scale_y = 1. / inv_y;

    // cv::resize turns an exact 2x INTER_LINEAR decimation into INTER_AREA (2x2 box, round half up)
    const int isx = cv_round(scale_x), isy = cv_round(scale_y);
    if (std::abs(scale_x - isx) < DBL_EPSILON && std::abs(scale_y - isy) < DBL_EPSILON && isx == 2 && isy == 2) {
        for (int y = 0; y < dh; ++y) {
            const uint8_t* s0 = src + (size_t)(2 * y) * spitch;
            const uint8_t* s1 = s0 + spitch;
            uint8_t* d = dst + (size_t)y * dpitch;
            for (int x = 0; x < dw; ++x)
                d[x] = (uint8_t)((s0[2 * x] + s0[2 * x + 1] + s1[2 * x] + s1[2 * x + 1] + 2) >> 2);
        }
        return;
    }

    std::vector<int> xofs, yofs;
    std::vector<short> alpha, beta;
    linear_axis_table(sw, dw, scale_x, xofs, alpha);
    linear_axis_table(sh, dh, scale_y, yofs, beta);

    std::vector<int> row0(dw), row1(dw);
    auto hresize = [&](int sy, std::vector<int>& out) {
        const uint8_t* s = src + (size_t)std::min(std::max(sy, 0), sh - 1) * spitch;
        for (int x = 0; x < dw; ++x) {
            const int sx = xofs[x];
            const int sx1 = std::min(sx + 1, sw - 1);  // weight is 0 whenever this clamps
            out[x] = s[sx] * alpha[2 * x] + s[sx1] * alpha[2 * x + 1];
        }
    };
    for (int y = 0; y < dh; ++y) {
        const int sy = yofs[y];
        hresize(sy, row0);
        hresize(sy + 1, row1);
        const int b0 = beta[2 * y], b1 = beta[2 * y + 1];
        uint8_t* d = dst + (size_t)y * dpitch;
        for (int x = 0; x < dw; ++x) {
            const int v = (((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2;
            d[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// FAST-9/16
// ---------------------------------------------------------------------------------------------
namespace {
// Bresenham circle r=3, index 0..15 (same table as src/FAST_NEON.cc:3-7)
const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                            {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};
}

int fast_corner_score(const uint8_t* p, size_t pitch, int threshold) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 25; ++k) {
        const int* o = kCircle[k & 15];
        d[k] = v - p[(ptrdiff_t)o[1] * (ptrdiff_t)pitch + o[0]];
    }
    // best "all darker" arc: max over the 16 arcs of 9 of min(d); best "all brighter": min of max(d)
    int a0 = threshold;
    for (int k = 0; k < 16; ++k) {
        int a = d[k];
        for (int i = 1; i < 9; ++i) a = std::min(a, d[k + i]);
        a0 = std::max(a0, a);
    }
    int b0 = -a0;
    for (int k = 0; k < 16; ++k) {
        int b = d[k];
        for (int i = 1; i < 9; ++i) b = std::max(b, d[k + i]);
        b0 = std::min(b0, b);
    }
    return -b0 - 1;
}

void fast9_16(const uint8_t* img, int w, int h, size_t pitch, int threshold, bool nonmax,
              std::vector<FastPoint>& out) {
    out.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return;

    // three rolling score rows (0 = not a corner) and the corner columns found on each
    std::vector<uint8_t> score[3];
    std::vector<int> cols[3];
    for (auto& s : score) s.assign((size_t)w + 1, 0);

    for (int i = 3; i < h - 2; ++i) {
        std::vector<uint8_t>& curr = score[(i - 3) % 3];
        std::vector<int>& ccols = cols[(i - 3) % 3];
        std::fill(curr.begin(), curr.end(), 0);
        ccols.clear();

        if (i < h - 3) {
            const uint8_t* row = img + (size_t)i * pitch;
            for (int j = 3; j < w - 3; ++j) {
                const uint8_t* p = row + j;
                const int v = p[0];
                const int lo = v - threshold, hi = v + threshold;
                // high-speed rejection, as OpenCV does it: a 9-arc contains one pixel of every opposite pair,
                // so each pair must hold a darker (bit 0) / brighter (bit 1) pixel of the same polarity
                auto cls = [&](int k) {
                    const int x = p[(ptrdiff_t)kCircle[k][1] * (ptrdiff_t)pitch + kCircle[k][0]];
                    return (x < lo ? 1 : 0) | (x > hi ? 2 : 0);
                };
                int d = cls(0) | cls(8);
                if (d == 0) continue;
                d &= cls(2) | cls(10);
                d &= cls(4) | cls(12);
                d &= cls(6) | cls(14);
                if (d == 0) continue;
                d &= cls(1) | cls(9);
                d &= cls(3) | cls(11);
                d &= cls(5) | cls(13);
                d &= cls(7) | cls(15);
                if (d == 0) continue;
                // 9 contiguous strictly darker, or 9 contiguous strictly brighter
                int run_d = 0, run_b = 0;
                bool corner = false;
                for (int k = 0; k < 25 && !corner; ++k) {
                    const int* o = kCircle[k & 15];
                    const int x = p[(ptrdiff_t)o[1] * (ptrdiff_t)pitch + o[0]];
                    run_d = (x < lo) ? run_d + 1 : 0;
                    run_b = (x > hi) ? run_b + 1 : 0;
                    corner = run_d > 8 || run_b > 8;
                }
                if (!corner) continue;
                ccols.push_back(j);
                if (nonmax) curr[j] = (uint8_t)fast_corner_score(p, pitch, threshold);
            }
        }
        if (i == 3) continue;

        // emit row i-1 now that its lower neighbour row is known
        const std::vector<uint8_t>& prev = score[(i - 4 + 3) % 3];
        const std::vector<uint8_t>& pprev = score[(i - 5 + 3) % 3];
        for (int j : cols[(i - 4 + 3) % 3]) {
            const int s = prev[j];
            if (!nonmax || (s > prev[j + 1] && s > prev[j - 1] && s > pprev[j - 1] && s > pprev[j] &&
                            s > pprev[j + 1] && s > curr[j - 1] && s > curr[j] && s > curr[j + 1]))
                out.push_back({j, i - 1, s});
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur 7x7 sigma 2, 8UC1, fixed point
// ---------------------------------------------------------------------------------------------
void gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch) {
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};  // Q8, sums to 256
    std::vector<uint16_t> hbuf((size_t)w * h);
    std::vector<int> xm((size_t)w + 6);
    for (int x = -3; x < w + 3; ++x) xm[x + 3] = reflect101(x, w);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * spitch;
        uint16_t* o = &hbuf[(size_t)y * w];
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int k = 0; k < 7; ++k) acc += K[k] * s[xm[x + k]];
            o[x] = (uint16_t)acc;  // <= 255*256
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* r[7];
        for (int k = 0; k < 7; ++k) r[k] = &hbuf[(size_t)reflect101(y + k - 3, h) * w];
        uint8_t* d = dst + (size_t)y * dpitch;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; ++k) acc += (uint32_t)K[k] * r[k][x];
            d[x] = (uint8_t)std::min<uint32_t>((acc + 32768u) >> 16, 255u);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// fastAtan2
// ---------------------------------------------------------------------------------------------
float fast_atan2(float y, float x) {
    static const float s = (float)(180.0 / 3.14159265358979323846);
    static const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s,
                       p5 = 0.1555786518463281f * s, p7 = -0.04432655554792128f * s;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

}  // namespace orc
