// Exercises the C++ mirror (include/gforb/ORBextractor.hpp) the way Frame::Frame does (src/Frame.cc:248-326):
// two extractors, one thread per eye, then ComputeStereoMatches.  Prints counts and FNV hashes so the pytest
// wrapper can compare them with the ctypes path.  usage: shim_test left.pgm right.pgm
#include <cstdio>
#include <fstream>
#include <map>
#include <sstream>
#include <thread>
#include "../../include/gforb/ORBextractor.hpp"

static std::vector<uint8_t> read_pgm(const char* path, int& w, int& h) {
    std::ifstream f(path, std::ios::binary);
    std::string magic; int maxv;
    f >> magic >> w >> h >> maxv;
    f.get();
    std::vector<uint8_t> px((size_t)w * h);
    f.read(reinterpret_cast<char*>(px.data()), px.size());
    return px;
}
static uint64_t fnv(const void* p, size_t n) {
    uint64_t hsh = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) { hsh ^= static_cast<const uint8_t*>(p)[i]; hsh *= 1099511628211ull; }
    return hsh;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    try {
        int w, h, w2, h2;
        std::vector<uint8_t> imL = read_pgm(argv[1], w, h), imR = read_pgm(argv[2], w2, h2);
        gforb::ORBextractor exL(1000, 1.2f, 8, 20, 7), exR(1000, 1.2f, 8, 20, 7);
        std::vector<gforb_keypoint> kL, kR;
        std::vector<uint8_t> dL, dR;
        std::thread tl([&] { exL(imL.data(), w, h, (size_t)w, kL, dL); });      // Frame::ExtractORB(0, imLeft)
        std::thread tr([&] { exR(imR.data(), w2, h2, (size_t)w2, kR, dR); });   // Frame::ExtractORB(1, imRight)
        tl.join(); tr.join();
        std::vector<float> u, z;
        const float bf = 47.90639384423901f, b = bf / 435.2046959714599f;
        const int nm = gforb::ComputeStereoMatches(exL, exR, kL, dL, kR, dR, bf, b, u, z);
        printf("nL=%zu nR=%zu nmatched=%d levels=%d dist=%d\n", kL.size(), kR.size(), nm, exL.GetLevels(),
               gforb::ORBmatcher::DescriptorDistance(dL.data(), dR.data()));
        printf("hash kL=%016llx dL=%016llx kR=%016llx dR=%016llx u=%016llx z=%016llx\n",
               (unsigned long long)fnv(kL.data(), kL.size() * sizeof(gforb_keypoint)), (unsigned long long)fnv(dL.data(), dL.size()),
               (unsigned long long)fnv(kR.data(), kR.size() * sizeof(gforb_keypoint)), (unsigned long long)fnv(dR.data(), dR.size()),
               (unsigned long long)fnv(u.data(), u.size() * 4), (unsigned long long)fnv(z.data(), z.size() * 4));
        // the two eyes as two frames: SearchForInitialization + the feature grid through the C++ mirror
        gforb::ORBmatcher matcher(0.9f, true);
        gforb::FrameView F1{&kL, dL.data(), {0.f, (float)w, 0.f, (float)h}}, F2{&kR, dR.data(), {0.f, (float)w2, 0.f, (float)h2}};
        std::vector<gforb::Point2f> prev(kL.size());
        for (size_t i = 0; i < kL.size(); ++i) prev[i] = {kL[i].x, kL[i].y};
        std::vector<int> m12;
        const int ninit = matcher.SearchForInitialization(F1, F2, prev, m12, 100);
        gforb::FrameGrid grid(matcher, kR, F2.bounds);
        const std::vector<size_t> area = grid.GetFeaturesInArea(w2 / 2.f, h2 / 2.f, 60.f, 0, 2);
        printf("init nmatches=%d m12=%016llx prev=%016llx area=%zu area_hash=%016llx\n", ninit,
               (unsigned long long)fnv(m12.data(), m12.size() * 4), (unsigned long long)fnv(prev.data(), prev.size() * 8),
               area.size(), (unsigned long long)fnv(area.data(), area.size() * sizeof(size_t)));
        // last-frame SearchByProjection (src/ORBmatcher.cc:1440-1593): the right eye's keypoints as the last frame's
        // points projected into the left frame, th = 15, stereo gate against the mvuRight just computed
        std::vector<gforb_proj_query> pq(kR.size());
        for (size_t i = 0; i < kR.size(); ++i) {
            float sc = 1.f;
            for (int l = 0; l < kR[i].octave; ++l) sc *= 1.2f;
            pq[i] = {kR[i].x, kR[i].y, 15.f * sc, kR[i].x - 20.f, kR[i].angle, kR[i].octave - 1, kR[i].octave + 1,
                     GFORB_PQ_ACTIVE | GFORB_PQ_BLOCKS | GFORB_PQ_RIGHT_GATE};
        }
        std::vector<int> kq, kd, qk;
        const int nproj = matcher.SearchByProjection(F1, u.data(), nullptr, pq, dR.data(), gforb::ORBmatcher::TH_HIGH, false,
                                                     matcher.checkOrientation(), kq, &kd, &qk);
        printf("proj nmatches=%d kq=%016llx kd=%016llx qk=%016llx\n", nproj, (unsigned long long)fnv(kq.data(), kq.size() * 4),
               (unsigned long long)fnv(kd.data(), kd.size() * 4), (unsigned long long)fnv(qk.data(), qk.size() * 4));
        // SearchByBoW (src/ORBmatcher.cc:270-404) with the first descriptor byte's high nibble standing in for the node id
        std::map<unsigned, std::vector<unsigned>> fvL, fvR;
        for (size_t i = 0; i < kL.size(); ++i) fvL[dL[32 * i] & 0xF0u].push_back((unsigned)i);
        for (size_t i = 0; i < kR.size(); ++i) fvR[dR[32 * i] & 0xF0u].push_back((unsigned)i);
        gforb::ORBmatcher bow(0.7f, true);
        std::vector<int> mb;
        const int nbow = bow.SearchByBoW(F1, nullptr, fvL, F2, nullptr, fvR, mb);
        printf("bow nmatches=%d mb=%016llx\n", nbow, (unsigned long long)fnv(mb.data(), mb.size() * 4));
        // SearchForTriangulation (src/ORBmatcher.cc:770-935): rectified pair = sideways motion, F12 = [0 0 0; 0 0 -1; 0 1 0]
        gforb_epipolar ep = {};
        ep.F12[5] = -1.f; ep.F12[7] = 1.f; ep.ex = 400.f; ep.ey = 240.f; ep.n_levels = 8;
        for (int l = 0; l < 8; ++l) { ep.scale_factors[l] = l ? ep.scale_factors[l - 1] * 1.2f : 1.f; ep.level_sigma2[l] = ep.scale_factors[l] * ep.scale_factors[l]; }
        std::vector<std::pair<size_t, size_t> > pairs;
        const int ntri = bow.SearchForTriangulation(F1, nullptr, nullptr, fvL, F2, nullptr, nullptr, fvR, ep, pairs);
        printf("tri nmatches=%d pairs=%016llx\n", ntri, (unsigned long long)fnv(pairs.data(), pairs.size() * sizeof(pairs[0])));
        // SearchBySim3 (src/ORBmatcher.cc:1214-1438): every key's own descriptor as its map point's, the two eyes as the two
        // key frames, "projected" 25 px sideways, r = 30 * scale, levels octave - 1 .. octave
        std::vector<gforb_proj_query> q12(kL.size()), q21(kR.size());
        auto fill = [](const std::vector<gforb_keypoint>& k, float shift, std::vector<gforb_proj_query>& q) {
            for (size_t i = 0; i < k.size(); ++i) {
                float sc = 1.f;
                for (int l = 0; l < k[i].octave; ++l) sc *= 1.2f;
                q[i] = {k[i].x + shift, k[i].y, 30.f * sc, 0.f, 0.f, k[i].octave - 1, k[i].octave, GFORB_PQ_ACTIVE};
            }
        };
        fill(kL, -25.f, q12); fill(kR, 25.f, q21);
        std::vector<int> ms;
        const int nsim = matcher.SearchBySim3(F1, F2, q12, dL.data(), q21, dR.data(), ms);
        printf("sim3 nfound=%d ms=%016llx\n", nsim, (unsigned long long)fnv(ms.data(), ms.size() * 4));
        // HASHING::MultiIndexHashing (include/Hashing.h:171-216): insert the left descriptors, query with the right ones
        gforb::MultiIndexHashing mih(256, 32);
        std::vector<int32_t> ids(kL.size());
        for (size_t i = 0; i < ids.size(); ++i) ids[i] = (int32_t)(1000 + i);
        std::vector<uint32_t> hashed;
        const std::vector<uint8_t> okh = mih.insert(dL.data(), ids, hashed);
        std::vector<std::vector<int32_t> > cands;
        mih.query(dR.data(), (int)kR.size(), cands, 8, 20);
        std::vector<int32_t> flat;
        for (const auto& c : cands) { flat.push_back((int32_t)c.size()); flat.insert(flat.end(), c.begin(), c.end()); }
        printf("mih bits=%u inserted=%d okh=%016llx cands=%016llx\n", mih.getSubstringLen(), (int)okh.size(),
               (unsigned long long)fnv(okh.data(), okh.size()), (unsigned long long)fnv(flat.data(), flat.size() * 4));
        // an empty image is the reference's silent no-op
        std::vector<gforb_keypoint> k0; std::vector<uint8_t> d0;
        exL(nullptr, 0, 0, 0, k0, d0);
        printf("empty ok=%d\n", (int)(k0.empty() && d0.empty()));
    } catch (const std::exception& e) {
        printf("EXCEPTION %s\n", e.what());
        return 1;
    }
    return 0;
}
