// ORACLE -- TEST INFRASTRUCTURE ONLY.  C ABI + CPU baseline driver over the restatement.
#include <climits>
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>

#include "cvprim.hpp"
#include "gforb_oracle.h"
#include "orb_oracle.hpp"

struct orc_extractor {
    orc::Extractor ex;
    orc_extractor(int nf, float sf, int nl, int a, int b) : ex(nf, sf, nl, a, b) {}
};

static_assert(sizeof(orc_keypoint) == 28 && sizeof(orc::KeyPoint) == 28, "cv::KeyPoint layout");

extern "C" {

int orc_cv_round_f(float v) { return orc::cv_round(v); }
int orc_cv_round_d(double v) { return orc::cv_round(v); }
float orc_fast_atan2(float y, float x) { return orc::fast_atan2(y, x); }
void orc_border_reflect101(const uint8_t* s, int w, int h, size_t sp, uint8_t* d, size_t dp, int b) {
    orc::copy_make_border_reflect101(s, w, h, sp, d, dp, b);
}
void orc_resize_linear_8u(const uint8_t* s, int sw, int sh, size_t sp, uint8_t* d, int dw, int dh, size_t dp) {
    orc::resize_linear_8u(s, sw, sh, sp, d, dw, dh, dp);
}
void orc_gaussian_blur_7x7_s2(const uint8_t* s, int w, int h, size_t sp, uint8_t* d, size_t dp) {
    orc::gaussian_blur_7x7_s2(s, w, h, sp, d, dp);
}
int orc_fast9_16(const uint8_t* img, int w, int h, size_t pitch, int th, int nonmax, int* xys, int cap) {
    std::vector<orc::FastPoint> pts;
    orc::fast9_16(img, w, h, pitch, th, nonmax != 0, pts);
    for (int i = 0; i < (int)pts.size() && i < cap; ++i) {
        xys[3 * i] = pts[i].x; xys[3 * i + 1] = pts[i].y; xys[3 * i + 2] = pts[i].score;
    }
    return (int)pts.size();
}
int orc_hamming256(const uint8_t* a, const uint8_t* b) { return orc::hamming256(a, b); }

orc_extractor* orc_extractor_create(int nf, float sf, int nl, int ini, int mn) {
    if (nf < 0 || nl < 1 || nl > 32 || !(sf > 1.0f)) return nullptr;
    return new orc_extractor(nf, sf, nl, ini, mn);
}
void orc_extractor_destroy(orc_extractor* ex) { delete ex; }

void orc_extractor_tables(const orc_extractor* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* quota, int* umax) {
    const orc::Extractor& e = h->ex;
    const size_t n = (size_t)e.nlevels;
    if (scale) std::memcpy(scale, e.scale.data(), n * 4);
    if (inv_scale) std::memcpy(inv_scale, e.inv_scale.data(), n * 4);
    if (sigma2) std::memcpy(sigma2, e.sigma2.data(), n * 4);
    if (inv_sigma2) std::memcpy(inv_sigma2, e.inv_sigma2.data(), n * 4);
    if (quota) std::memcpy(quota, e.quota.data(), n * 4);
    if (umax) std::memcpy(umax, e.umax, 16 * 4);
}
void orc_extractor_pattern(const orc_extractor* h, int8_t* xy) { std::memcpy(xy, h->ex.pattern, 1024); }

int orc_compute_pyramid(orc_extractor* h, const uint8_t* img, int w, int hh, size_t pitch) {
    if (!img || w <= 0 || hh <= 0) return -1;
    h->ex.compute_pyramid(img, w, hh, pitch);
    return 0;
}
int orc_pyramid_level(const orc_extractor* h, int level, const uint8_t** ptr, int* w, int* hh, size_t* pitch) {
    if (level < 0 || level >= h->ex.nlevels || h->ex.pyramid[level].buf.empty()) return -1;
    const orc::Image& im = h->ex.pyramid[level];
    if (ptr) *ptr = im.roi();
    if (w) *w = im.w;
    if (hh) *hh = im.h;
    if (pitch) *pitch = im.pitch;
    return 0;
}
static int put_candidates(const std::vector<orc::Candidate>& c, int* xyr, int cap) {
    for (int i = 0; i < (int)c.size() && i < cap; ++i) {
        xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].response;
    }
    return (int)c.size();
}
int orc_detect_level(const orc_extractor* h, int level, int* xyr, int cap) {
    std::vector<orc::Candidate> c;
    h->ex.detect_level(level, c);
    return put_candidates(c, xyr, cap);
}
int orc_distribute_octree(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* keep, int cap) {
    std::vector<orc::Candidate> c(n);
    for (int i = 0; i < n; ++i) c[i] = {xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
    std::vector<int> k = orc::Extractor::distribute_octree(c, minX, maxX, minY, maxY, N);
    for (int i = 0; i < (int)k.size() && i < cap; ++i) keep[i] = k[i];
    return (int)k.size();
}
int orc_extract(orc_extractor* h, const uint8_t* img, int w, int hh, size_t pitch,
                orc_keypoint* kps, uint8_t* desc, int cap, int* level_counts) {
    std::vector<orc::KeyPoint> k;
    std::vector<uint8_t> d;
    const int n = h->ex.extract(img, w, hh, pitch, k, d);
    const int m = n < cap ? n : cap;
    if (kps && m > 0) std::memcpy(kps, k.data(), (size_t)m * sizeof(orc_keypoint));
    if (desc && m > 0) std::memcpy(desc, d.data(), (size_t)m * 32);
    if (level_counts)
        for (int l = 0; l < h->ex.nlevels; ++l)
            level_counts[l] = l < (int)h->ex.last_level_counts.size() ? h->ex.last_level_counts[l] : 0;
    return n;
}
int orc_last_candidates(const orc_extractor* h, int level, int* xyr, int cap) {
    if (level < 0 || level >= (int)h->ex.last_candidates.size()) return -1;
    return put_candidates(h->ex.last_candidates[level], xyr, cap);
}

int orc_stereo_match(const orc_extractor* exL, const orc_extractor* exR, const orc_stereo_params* p,
                     const orc_keypoint* kL, int nL, const uint8_t* dL,
                     const orc_keypoint* kR, int nR, const uint8_t* dR,
                     const float* min_d, const float* max_d,
                     float* uRight, float* depth, int* dbg_best_dist, int* dbg_best_idx, int* dbg_sad) {
    orc::StereoParams prm{p->bf, p->b, p->mode, p->th_high, p->th_low, p->min_x};
    std::vector<float> u, z;
    orc::StereoDebug dbg;
    const int r = orc::stereo_match(exL->ex, exR->ex, prm, (const orc::KeyPoint*)kL, nL, dL,
                                    (const orc::KeyPoint*)kR, nR, dR, min_d, max_d, u, z, &dbg);
    if (nL > 0) {
        std::memcpy(uRight, u.data(), (size_t)nL * 4);
        std::memcpy(depth, z.data(), (size_t)nL * 4);
        if (dbg_best_dist) std::memcpy(dbg_best_dist, dbg.best_dist.data(), (size_t)nL * 4);
        if (dbg_best_idx) std::memcpy(dbg_best_idx, dbg.best_idx.data(), (size_t)nL * 4);
        if (dbg_sad) std::memcpy(dbg_sad, dbg.sad.data(), (size_t)nL * 4);
    }
    return r;
}

double orc_bench_stereo(const uint8_t* images, int nimg, int w, int h, int nframes, int nthreads, int shape,
                        int nf, float sf, int nl, int ini, int mn, float bf, float b, int do_match,
                        long long* checksum) {
    const size_t isz = (size_t)w * h;
    std::atomic<long long> sum{0};
    const orc::StereoParams prm{bf, b, 0, 100, 50, 0.f};
    auto t0 = std::chrono::steady_clock::now();
    if (shape == 1) {
        // reference-shaped: one frame at a time, one thread per eye, serial stereo match
        orc::Extractor exL(nf, sf, nl, ini, mn), exR(nf, sf, nl, ini, mn);
        std::vector<orc::KeyPoint> kL, kR;
        std::vector<uint8_t> dL, dR;
        std::vector<float> u, z;
        for (int f = 0; f < nframes; ++f) {
            const uint8_t* iL = images + isz * (size_t)((2 * f) % nimg);
            const uint8_t* iR = images + isz * (size_t)((2 * f + 1) % nimg);
            std::thread tl([&] { exL.extract(iL, w, h, (size_t)w, kL, dL); });
            std::thread tr([&] { exR.extract(iR, w, h, (size_t)w, kR, dR); });
            tl.join();
            tr.join();
            long long s = (long long)kL.size() + (long long)kR.size();
            if (do_match)
                s += orc::stereo_match(exL, exR, prm, kL.data(), (int)kL.size(), dL.data(), kR.data(),
                                       (int)kR.size(), dR.data(), nullptr, nullptr, u, z, nullptr);
            sum += s;
        }
    } else {
        std::atomic<int> next{0};
        auto worker = [&] {
            orc::Extractor exL(nf, sf, nl, ini, mn), exR(nf, sf, nl, ini, mn);
            std::vector<orc::KeyPoint> kL, kR;
            std::vector<uint8_t> dL, dR;
            std::vector<float> u, z;
            for (;;) {
                const int f = next.fetch_add(1);
                if (f >= nframes) break;
                const uint8_t* iL = images + isz * (size_t)((2 * f) % nimg);
                const uint8_t* iR = images + isz * (size_t)((2 * f + 1) % nimg);
                exL.extract(iL, w, h, (size_t)w, kL, dL);
                exR.extract(iR, w, h, (size_t)w, kR, dR);
                long long s = (long long)kL.size() + (long long)kR.size();
                if (do_match)
                    s += orc::stereo_match(exL, exR, prm, kL.data(), (int)kL.size(), dL.data(), kR.data(),
                                           (int)kR.size(), dR.data(), nullptr, nullptr, u, z, nullptr);
                sum += s;
            }
        };
        std::vector<std::thread> pool;
        for (int t = 0; t < (nthreads > 0 ? nthreads : 1); ++t) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (checksum) *checksum = sum.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

void orc_match_bruteforce(const uint8_t* a, int nA, const uint8_t* b, int nB, float ratio, int threshold,
                          int* best_idx, int* best_dist, int* second_dist, int* match, int nthreads) {
    auto run = [&](int q0, int q1) {
        for (int i = q0; i < q1; ++i) {
            int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx = -1;   // src/ORBmatcher.cc:544-546
            for (int j = 0; j < nB; ++j) {
                const int dist = orc_hamming256(a + (size_t)i * 32, b + (size_t)j * 32);
                if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx = j; }
                else if (dist < bestDist2) bestDist2 = dist;
            }
            if (best_idx) best_idx[i] = bestIdx;
            if (best_dist) best_dist[i] = bestDist;
            if (second_dist) second_dist[i] = bestDist2;
            if (match) match[i] = (bestIdx >= 0 && bestDist <= threshold && bestDist < (float)bestDist2 * ratio) ? bestIdx : -1;
        }
    };
    if (nthreads <= 1) { run(0, nA); return; }
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; ++t) pool.emplace_back(run, (int)((long long)nA * t / nthreads), (int)((long long)nA * (t + 1) / nthreads));
    for (auto& t : pool) t.join();
}

}  // extern "C"
