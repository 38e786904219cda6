// ORACLE -- TEST INFRASTRUCTURE ONLY (see orb_oracle.hpp).  Build with -ffp-contract=off.
// Restates /root/reference/src/Frame.cc:889-1078 (classic) and :1167-1316 + include/Frame.h:230-263
// (_Undistorted), and src/ORBmatcher.cc:1768-1784.
#include <algorithm>
#include <climits>
#include <cmath>

#include "orb_oracle.hpp"

namespace orc {

// src/ORBmatcher.cc:1768-1784 (SWAR popcount over eight 32-bit words)
int hamming256(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t wa, wb;
        __builtin_memcpy(&wa, a + 4 * i, 4);
        __builtin_memcpy(&wb, b + 4 * i, 4);
        uint32_t v = wa ^ wb;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
    }
    return dist;
}

int stereo_match(const Extractor& exL, const Extractor& exR, const StereoParams& prm,
                 const KeyPoint* kL, int nL, const uint8_t* dL, const KeyPoint* kR, int nR, const uint8_t* dR,
                 const float* minD_opt, const float* maxD_opt,
                 std::vector<float>& uRight, std::vector<float>& depth, StereoDebug* dbg) {
    uRight.assign(nL, -1.f);
    depth.assign(nL, -1.f);
    if (dbg) {
        dbg->best_dist.assign(nL, prm.th_high);
        dbg->best_idx.assign(nL, -1);
        dbg->sad.assign(nL, -1);
    }
    const bool classic = prm.mode == 0;
    const int thOrbDist = (prm.th_high + prm.th_low) / 2;
    const int nRows = exL.pyramid[0].h;
    const std::vector<float>& scale = exL.scale;       // Frame copies the LEFT extractor's tables (src/Frame.cc:75-81)
    const std::vector<float>& inv_scale = exL.inv_scale;

    // row table (:899-917 / Frame.h:244-258)
    std::vector<std::vector<int>> rows(nRows);
    for (int iR = 0; iR < nR; ++iR) {
        const float y = kR[iR].y;
        const float r = 2.0f * scale[kR[iR].octave];
        int maxr, minr;
        if (classic) {
            maxr = (int)std::ceil(y + r);
            minr = (int)std::floor(y - r);
            // the reference indexes unclamped; keys sit >= 19*scale px inside so this never triggers
            maxr = std::min(maxr, nRows - 1);
            minr = std::max(minr, 0);
        } else {
            maxr = (int)std::min((float)(nRows - 1), std::ceil(y + r));
            minr = (int)std::max(0.0f, std::floor(y - r));
        }
        for (int yi = minr; yi <= maxr; ++yi) rows[yi].push_back(iR);
    }

    const float minZ = prm.b;
    std::vector<std::pair<int, int>> distIdx;
    int nmatched = 0;

    for (int iL = 0; iL < nL; ++iL) {
        float minD = 0;
        float maxD = prm.bf / minZ;
        const KeyPoint& kp = kL[iL];
        const int levelL = kp.octave;
        const float vL = kp.y, uL = kp.x;

        if (!classic && (vL < 0 || vL > nRows - 1)) continue;
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;  // classic reference would index out of range
        const std::vector<int>& cand = rows[row];
        if (cand.empty()) continue;

        if (!classic && minD_opt && maxD_opt) {  // per-key window from a matched map point (:1220-1231)
            minD = minD_opt[iL];
            maxD = maxD_opt[iL];
        }
        const float minU = uL - maxD, maxU = uL - minD;
        if (classic ? (maxU < 0) : (maxU < prm.min_x)) continue;

        int bestDist = prm.th_high;
        int bestIdxR = 0;
        for (int iR : cand) {
            const KeyPoint& kr = kR[iR];
            if (kr.octave < levelL - 1 || kr.octave > levelL + 1) continue;
            const float uR = kr.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = hamming256(dL + 32 * (size_t)iL, dR + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (dbg && bestDist < prm.th_high) { dbg->best_dist[iL] = bestDist; dbg->best_idx[iL] = bestIdxR; }

        if (!classic) {
            if (bestDist < thOrbDist) {
                float bestuR = kR[bestIdxR].x;
                float disparity = uL - bestuR;
                if (disparity >= minD && disparity < maxD) {
                    if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                    depth[iL] = prm.bf / disparity;
                    uRight[iL] = bestuR;
                    distIdx.push_back({bestDist, iL});
                }
            }
            ++nmatched;
            continue;
        }

        if (bestDist >= thOrbDist) continue;

        // sub-pixel refinement by 11x11 SAD over +-5 px at the LEFT key's octave (:976-1059)
        const float uR0 = kR[bestIdxR].x;
        const float sf = inv_scale[kp.octave];
        const float scaleduL = std::round(kp.x * sf);
        const float scaledvL = std::round(kp.y * sf);
        const float scaleduR0 = std::round(uR0 * sf);
        const int w = 5, L = 5;
        const Image& imL = exL.pyramid[kp.octave];
        const Image& imR = exR.pyramid[kp.octave];

        const float iniu = scaleduR0 + L - w;
        const float endu = scaleduR0 + L + w + 1;
        if (iniu < 0 || endu >= imR.w) continue;

        const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
        const uint8_t* pL = imL.roi();
        const uint8_t* pR = imR.roi();
        const float cL = (float)pL[(ptrdiff_t)cv * imL.pitch + cu];
        int bestSad = INT_MAX, bestInc = 0;
        float dists[2 * L + 1];
        for (int inc = -L; inc <= L; ++inc) {
            const float cR = (float)pR[(ptrdiff_t)cv * imR.pitch + cr + inc];
            double acc = 0;  // cv::norm(NORM_L1) on CV_32F accumulates in double; all terms are integers
            for (int dy = -w; dy <= w; ++dy)
                for (int dx = -w; dx <= w; ++dx) {
                    const float l = (float)pL[(ptrdiff_t)(cv + dy) * imL.pitch + cu + dx] - cL;
                    const float r = (float)pR[(ptrdiff_t)(cv + dy) * imR.pitch + cr + inc + dx] - cR;
                    acc += std::fabs(l - r);
                }
            const float dist = (float)acc;
            if (dist < bestSad) { bestSad = (int)dist; bestInc = inc; }
            dists[L + inc] = dist;
        }
        if (bestInc == -L || bestInc == L) continue;

        const float d1 = dists[L + bestInc - 1], d2 = dists[L + bestInc], d3 = dists[L + bestInc + 1];
        const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
        if (deltaR < -1 || deltaR > 1) continue;

        float bestuR = scale[kp.octave] * ((float)scaleduR0 + (float)bestInc + deltaR);
        float disparity = uL - bestuR;
        if (disparity >= minD && disparity < maxD) {
            if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
            depth[iL] = prm.bf / disparity;
            uRight[iL] = bestuR;
            distIdx.push_back({bestSad, iL});
            if (dbg) dbg->sad[iL] = bestSad;
        }
    }

    // median-based outlier cut (:1064-1077 / :1291-1313); an empty list is undefined in the classic
    // reference and an early return in the undistorted one -- both mean "no cut" here
    if (!distIdx.empty()) {
        std::sort(distIdx.begin(), distIdx.end());
        const float median = (float)distIdx[distIdx.size() / 2].first;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = (int)distIdx.size() - 1; i >= 0; --i) {
            if (distIdx[i].first < thDist) break;
            uRight[distIdx[i].second] = -1;
            depth[distIdx[i].second] = -1;
            --nmatched;
        }
    }
    if (classic) {
        int kept = 0;
        for (int i = 0; i < nL; ++i) kept += depth[i] > 0 ? 1 : 0;
        return kept;
    }
    return nmatched;
}

}  // namespace orc
