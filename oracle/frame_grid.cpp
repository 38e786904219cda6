// ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the Frame feature grid and of
// ORBmatcher::SearchForInitialization, written as the reference writes them (sequential loops, std::vector
// push_back order), so the CUDA path's parallel reformulation can be checked against the literal algorithm.
//
//   Frame::AssignFeaturesToGrid   /root/reference/src/Frame.cc:461-476
//   Frame::PosInGrid              /root/reference/src/Frame.cc:645-656
//   Frame::GetFeaturesInArea      /root/reference/src/Frame.cc:593-643
//   grid constants                /root/reference/include/Frame.h:92-93 (64 x 48), src/Frame.cc:129-130
//   ORBmatcher::SearchForInitialization   /root/reference/src/ORBmatcher.cc:520-633
//   ORBmatcher::ComputeThreeMaxima        /root/reference/src/ORBmatcher.cc:1723-1764
//   HISTO_LENGTH = 30, TH_LOW = 50        /root/reference/src/ORBmatcher.cc:37-39
//
// Parity note: the reference's own tests hold no vectors for these functions (parity unpinned by the reference);
// tests/test_oracle_grid.py checks this file against an independent pure-numpy restatement.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <vector>

#include "gforb_oracle.h"

namespace {
constexpr int kCols = 64, kRows = 48, kHisto = 30;

struct Grid {
    float minX, minY, wInv, hInv;
    std::vector<int> cell[kCols][kRows];
};

// src/Frame.cc:129-130: static_cast<float>(FRAME_GRID_COLS) / (mnMaxX - mnMinX), all float
void grid_scales(const float* bounds, float* wInv, float* hInv) {
    *wInv = static_cast<float>(kCols) / (bounds[1] - bounds[0]);
    *hInv = static_cast<float>(kRows) / (bounds[3] - bounds[2]);
}

// src/Frame.cc:645-656: round() on float (half away from zero), then the range test
bool pos_in_grid(const Grid& g, const orc_keypoint& kp, int* px, int* py) {
    *px = (int)std::round((kp.x - g.minX) * g.wInv);
    *py = (int)std::round((kp.y - g.minY) * g.hInv);
    return !(*px < 0 || *px >= kCols || *py < 0 || *py >= kRows);
}

void build_grid(Grid& g, const orc_keypoint* keys, int n, const float* bounds) {
    g.minX = bounds[0]; g.minY = bounds[2];
    grid_scales(bounds, &g.wInv, &g.hInv);
    for (int i = 0; i < n; ++i) {                      // src/Frame.cc:468-475: ascending i, push_back
        int px, py;
        if (pos_in_grid(g, keys[i], &px, &py)) g.cell[px][py].push_back(i);
    }
}

// src/Frame.cc:593-643, statement for statement
std::vector<int> features_in_area(const Grid& g, const orc_keypoint* keys, float x, float y, float r, int minLevel,
                                  int maxLevel) {
    std::vector<int> out;
    const int nMinCellX = std::max(0, (int)std::floor((x - g.minX - r) * g.wInv));
    if (nMinCellX >= kCols) return out;
    const int nMaxCellX = std::min(kCols - 1, (int)std::ceil((x - g.minX + r) * g.wInv));
    if (nMaxCellX < 0) return out;
    const int nMinCellY = std::max(0, (int)std::floor((y - g.minY - r) * g.hInv));
    if (nMinCellY >= kRows) return out;
    const int nMaxCellY = std::min(kRows - 1, (int)std::ceil((y - g.minY + r) * g.hInv));
    if (nMaxCellY < 0) return out;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (int idx : g.cell[ix][iy]) {
                const orc_keypoint& kp = keys[idx];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float distx = kp.x - x, disty = kp.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) out.push_back(idx);
            }
    return out;
}

// src/ORBmatcher.cc:1723-1764 (sizes instead of vectors)
void three_maxima(const int* size, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = size[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}
}  // namespace

extern "C" {

int orc_assign_features_to_grid(const orc_keypoint* keys, int n, const float* bounds, int* cell_start, int* cell_idx) {
    Grid g;
    build_grid(g, keys, n, bounds);
    int k = 0;
    for (int ix = 0; ix < kCols; ++ix)
        for (int iy = 0; iy < kRows; ++iy) {
            cell_start[ix * kRows + iy] = k;
            for (int idx : g.cell[ix][iy]) cell_idx[k++] = idx;
        }
    cell_start[kCols * kRows] = k;
    return k;
}

int orc_features_in_area(const orc_keypoint* keys, int n, const float* bounds, float x, float y, float r, int minLevel,
                         int maxLevel, int* out, int cap) {
    Grid g;
    build_grid(g, keys, n, bounds);
    std::vector<int> v = features_in_area(g, keys, x, y, r, minLevel, maxLevel);
    for (int i = 0; i < (int)v.size() && i < cap; ++i) out[i] = v[i];
    return (int)v.size();
}

int orc_search_for_initialization(const orc_keypoint* k1, const uint8_t* d1, int n1, const orc_keypoint* k2,
                                  const uint8_t* d2, int n2, const float* bounds, float* prev_xy, int windowSize,
                                  float nnratio, int checkOri, int* matches12) {
    Grid g;
    build_grid(g, k2, n2, bounds);
    int nmatches = 0;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    std::vector<int> rotHist[kHisto];
    const float factor = 1.0f / kHisto;
    std::vector<int> vMatchedDistance(n2, INT_MAX), vnMatches21(n2, -1);
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = k1[i1].octave;
        if (level1 > 0) continue;
        std::vector<int> vIndices2 =
            features_in_area(g, k2, prev_xy[2 * i1], prev_xy[2 * i1 + 1], (float)windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = orc_hamming256(d1 + 32 * (size_t)i1, d2 + 32 * (size_t)i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= 50) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOri) {
                    float rot = k1[i1].angle - k2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == kHisto) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[kHisto];
        for (int i = 0; i < kHisto; ++i) sizes[i] = (int)rotHist[i].size();
        three_maxima(sizes, kHisto, ind1, ind2, ind3);
        for (int i = 0; i < kHisto; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (matches12[i1] >= 0) { prev_xy[2 * i1] = k2[matches12[i1]].x; prev_xy[2 * i1 + 1] = k2[matches12[i1]].y; }
    return nmatches;
}


// ORBmatcher::SearchByProjection, the three per-frame forms of the reference folded into one loop:
//   (Frame&, vector<MapPoint*>, th)                       /root/reference/src/ORBmatcher.cc:155-241  (ratio test :230)
//   (CurrentFrame, LastFrame, th, bMono)                  /root/reference/src/ORBmatcher.cc:1440-1593 (rotation check)
//   (CurrentFrame, KeyFrame*, sAlreadyFound, th, ORBdist) /root/reference/src/ORBmatcher.cc:1595-1721 (rotation check)
// What differs between them is data, not control flow, and comes in through the query record:
//   x, y, r, min_level, max_level  the GetFeaturesInArea call (:179, :1501-1506, :1651)
//   ACTIVE      the guards in front of the window (mbTrackInView / isBad :164-168; pMP, outlier, invzc, bounds :1469-1489)
//   BLOCKS      pMP->Observations() > 0: a key this query takes is skipped by later queries (:197-199, :1519-1521);
//               the third form skips ANY assigned key (:1662-1663) = every query blocks, every assigned key occupied
//   RIGHT_GATE  the mvuRight test (:201-206, :1523-1529); its bound equals r in both forms (r * scale; radius)
//   REPROJ_GATE (8) the candidate search of ORBmatcher::Fuse (:937-1087; the Sim3 form :1089-1212 has no such gate): a
//               candidate's reprojection error e2 * mvInvLevelSigma2[level] must stay below 7.8 (stereo key, with the
//               right coordinate) or 5.99 (:1022-1046); queries neither block nor see occupancy, query_key[q] = bestIdx
// Pose projection, PredictScale and RadiusByViewingCos stay with the caller (they read MapPoint / pose state).
// key_query[k]: query holding key k when the call returns (-1 untouched, -2 set to NULL by the rotation check);
// key_dist[k]: mvpMatchScore (:235); query_key[q]: key taken by query q (-1 none / removed).  Returns nmatches.
int orc_search_by_projection(const orc_keypoint* keys, const uint8_t* desc, const float* u_right,
                             const uint8_t* occupied, int n, const float* bounds, const orc_proj_query* qs,
                             const uint8_t* qdesc, int nq, int th_dist, float nnratio, int use_ratio, int checkOri,
                             const float* inv_level_sigma2, int* key_query, int* key_dist, int* query_key) {
    if (th_dist > 255) return -1;      // bestDist starts at 256 with bestIdx = -1 (:183, :1510): the reference would index [-1]
    Grid g;
    build_grid(g, keys, n, bounds);
    int nmatches = 0;
    std::vector<uint8_t> blocked(n, 0);
    for (int k = 0; k < n; ++k) { key_query[k] = -1; key_dist[k] = 0; blocked[k] = occupied ? (occupied[k] != 0) : 0; }
    for (int q = 0; q < nq; ++q) query_key[q] = -1;
    std::vector<int> rotHist[kHisto], rotQuery[kHisto];
    const float factor = 1.0f / kHisto;
    for (int iq = 0; iq < nq; ++iq) {
        const orc_proj_query& q = qs[iq];
        if (!(q.flags & 1)) continue;
        std::vector<int> vIndices = features_in_area(g, keys, q.x, q.y, q.r, q.min_level, q.max_level);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (blocked[idx]) continue;
            if ((q.flags & 4) && u_right && u_right[idx] > 0) {
                const float er = std::fabs(q.xr - u_right[idx]);
                if (er > q.r) continue;
            }
            if ((q.flags & 8) && inv_level_sigma2) {                  // ORBmatcher::Fuse, :1022-1046: reprojection error gate
                const orc_keypoint& kp = keys[idx];
                if (u_right && u_right[idx] >= 0) {
                    const float ex = q.x - kp.x, ey = q.y - kp.y, er = q.xr - u_right[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_level_sigma2[kp.octave] > 7.8) continue;
                } else {
                    const float ex = q.x - kp.x, ey = q.y - kp.y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_level_sigma2[kp.octave] > 5.99) continue;
                }
            }
            const int dist = orc_hamming256(qdesc + 32 * (size_t)iq, desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = keys[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = keys[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist <= th_dist) {
            if (use_ratio && bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) continue;
            key_query[bestIdx] = iq;
            key_dist[bestIdx] = bestDist;
            query_key[iq] = bestIdx;
            if (q.flags & 2) blocked[bestIdx] = 1;
            nmatches++;
            if (checkOri) {
                float rot = q.angle - keys[bestIdx].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == kHisto) bin = 0;
                rotHist[bin].push_back(bestIdx);
                rotQuery[bin].push_back(iq);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[kHisto];
        for (int i = 0; i < kHisto; ++i) sizes[i] = (int)rotHist[i].size();
        three_maxima(sizes, kHisto, ind1, ind2, ind3);
        for (int i = 0; i < kHisto; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); ++j) {
                key_query[rotHist[i][j]] = -2;
                query_key[rotQuery[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}


// ORBmatcher::SearchBySim3 (/root/reference/src/ORBmatcher.cc:1214-1438) over already-projected queries: the two
// search loops (:1259-1337 KF1's map points in KF2, :1339-1417 KF2's in KF1) and the agreement test (:1420-1435),
// literally.  What precedes GetFeaturesInArea in each iteration (pose / Sim3 projection, IsInImage, distance
// invariance, PredictScale) needs MapPoint / KeyFrame state and arrives as q.x, q.y, q.r, q.max_level = nPredictedLevel;
// flags & 1 == 0 where the reference `continue`s before the window.
static void sim3_one_way(const Grid& g, const orc_keypoint* keys, const uint8_t* desc, const orc_proj_query* qs,
                         const uint8_t* qdesc, int nq, int th_dist, int* vnMatch) {
    for (int i = 0; i < nq; ++i) {
        vnMatch[i] = -1;
        const orc_proj_query& q = qs[i];
        if (!(q.flags & 1)) continue;
        const int nPredictedLevel = q.max_level;
        const std::vector<int> vIndices = features_in_area(g, keys, q.x, q.y, q.r, -1, -1);       // (u, v, radius): all levels
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (int idx : vIndices) {
            const orc_keypoint& kp = keys[idx];
            if (kp.octave < nPredictedLevel - 1 || kp.octave > nPredictedLevel) continue;
            const int dist = orc_hamming256(qdesc + 32 * (size_t)i, desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= th_dist) vnMatch[i] = bestIdx;
    }
}

int orc_search_by_sim3(const orc_keypoint* k1, const uint8_t* d1, int n1, const orc_keypoint* k2, const uint8_t* d2, int n2,
                       const float* bounds, const orc_proj_query* q12, const uint8_t* qd12, const orc_proj_query* q21,
                       const uint8_t* qd21, int th_dist, int* vnMatch1, int* vnMatch2, int* match12) {
    Grid g1, g2;
    build_grid(g1, k1, n1, bounds);
    build_grid(g2, k2, n2, bounds);
    sim3_one_way(g2, k2, d2, q12, qd12, n1, th_dist, vnMatch1);
    sim3_one_way(g1, k1, d1, q21, qd21, n2, th_dist, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < n1; ++i1) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

// ORBmatcher::SearchByBoW, both forms:
//   (KeyFrame* pKF, Frame& F, vpMapPointMatches)        /root/reference/src/ORBmatcher.cc:270-404   bestDist1 <= TH_LOW
//   (KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12)       /root/reference/src/ORBmatcher.cc:635-768   bestDist1 <  TH_LOW
// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>, Thirdparty/DBoW2/DBoW2/FeatureVector.h) arrives
// flattened: node ids ascending, CSR offsets, feature indices in the vector's order.  valid1[i] = the query's MapPoint
// exists and is not bad (:300-306, :679-684); valid2[j] = the same for the second key frame's candidates (:700-706;
// NULL for the Frame form, which has no such test).  A candidate that has been matched is skipped from then on
// (vpMapPointMatches[realIdxF] :321, vbMatched2 :701).  Both forms reduce to matches12[idx1] = idx2: the Frame form's
// vpMapPointMatches[idx2] = pMP(idx1) is its inverse (a candidate is matched once), and its rotation histogram, which
// holds idx2, removes the same pairs.  th_dist is inclusive: TH_LOW for the first form, TH_LOW - 1 for the second.
int orc_search_by_bow(const orc_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, int n1, const int* node_id1,
                      const int* node_start1, const int* node_idx1, int nn1, const orc_keypoint* k2, const uint8_t* d2,
                      const uint8_t* valid2, int n2, const int* node_id2, const int* node_start2, const int* node_idx2,
                      int nn2, int th_dist, float nnratio, int checkOri, int* matches12) {
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    std::vector<uint8_t> matched2(n2 > 0 ? n2 : 1, 0);
    std::vector<int> rotHist[kHisto];
    const float factor = 1.0f / kHisto;
    int nmatches = 0, it1 = 0, it2 = 0;
    while (it1 != nn1 && it2 != nn2) {
        if (node_id1[it1] == node_id2[it2]) {
            for (int a = node_start1[it1]; a < node_start1[it1 + 1]; ++a) {
                const int idx1 = node_idx1[a];
                if (valid1 && !valid1[idx1]) continue;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int b = node_start2[it2]; b < node_start2[it2 + 1]; ++b) {
                    const int idx2 = node_idx2[b];
                    if (matched2[idx2] || (valid2 && !valid2[idx2])) continue;
                    const int dist = orc_hamming256(d1 + 32 * (size_t)idx1, d2 + 32 * (size_t)idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= th_dist) {
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {
                        matches12[idx1] = bestIdx2;
                        matched2[bestIdx2] = 1;
                        if (checkOri) {
                            float rot = k1[idx1].angle - k2[bestIdx2].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == kHisto) bin = 0;
                            rotHist[bin].push_back(idx1);
                        }
                        nmatches++;
                    }
                }
            }
            it1++; it2++;
        } else if (node_id1[it1] < node_id2[it2]) {
            it1 = (int)(std::lower_bound(node_id1 + it1, node_id1 + nn1, node_id2[it2]) - node_id1);
        } else {
            it2 = (int)(std::lower_bound(node_id2 + it2, node_id2 + nn2, node_id1[it1]) - node_id2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[kHisto];
        for (int i = 0; i < kHisto; ++i) sizes[i] = (int)rotHist[i].size();
        three_maxima(sizes, kHisto, ind1, ind2, ind3);
        for (int i = 0; i < kHisto; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { matches12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}


// ORBmatcher::SearchForTriangulation (/root/reference/src/ORBmatcher.cc:770-935) with CheckDistEpipolarLine (:251-268),
// statement for statement.  valid1[i] = feature i of key frame 1 has no MapPoint (and is stereo if bOnlyStereo), :824-831;
// valid2[j] the same for key frame 2 (:844-851; vbMatched2 is never set in the reference, so it filters nothing);
// stereoX[i] = mvuRight[i] >= 0.  F12 row-major 3x3 float, (ex, ey) the epipole in image 2 (:776-783, computed by the
// caller from the poses), scale_factors = pKF2->mvScaleFactors, level_sigma2 = pKF2->mvLevelSigma2.
int orc_search_for_triangulation(const orc_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, const uint8_t* stereo1,
                                 int n1, const int* node_id1, const int* node_start1, const int* node_idx1, int nn1,
                                 const orc_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, const uint8_t* stereo2,
                                 int n2, const int* node_id2, const int* node_start2, const int* node_idx2, int nn2,
                                 const float* F12, float ex, float ey, const float* scale_factors,
                                 const float* level_sigma2, int th_low, int checkOri, int* matches12) {
    (void)n2;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    std::vector<int> rotHist[kHisto];
    const float factor = 1.0f / kHisto;
    int nmatches = 0, it1 = 0, it2 = 0;
    while (it1 != nn1 && it2 != nn2) {
        if (node_id1[it1] == node_id2[it2]) {
            for (int a = node_start1[it1]; a < node_start1[it1 + 1]; ++a) {
                const int idx1 = node_idx1[a];
                if (valid1 && !valid1[idx1]) continue;
                const bool bStereo1 = stereo1 && stereo1[idx1];
                const orc_keypoint& kp1 = k1[idx1];
                int bestDist = th_low, bestIdx2 = -1;
                for (int b = node_start2[it2]; b < node_start2[it2 + 1]; ++b) {
                    const int idx2 = node_idx2[b];
                    if (valid2 && !valid2[idx2]) continue;
                    const bool bStereo2 = stereo2 && stereo2[idx2];
                    const int dist = orc_hamming256(d1 + 32 * (size_t)idx1, d2 + 32 * (size_t)idx2);
                    if (dist > th_low || dist > bestDist) continue;
                    const orc_keypoint& kp2 = k2[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * scale_factors[kp2.octave]) continue;
                    }
                    // CheckDistEpipolarLine, :251-268
                    const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                    const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                    const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                    const float num = la * kp2.x + lb * kp2.y + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) continue;
                    const float dsqr = num * num / den;
                    if (dsqr < 3.84 * level_sigma2[kp2.octave]) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    matches12[idx1] = bestIdx2;
                    nmatches++;
                    if (checkOri) {
                        float rot = kp1.angle - k2[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == kHisto) bin = 0;
                        rotHist[bin].push_back(idx1);
                    }
                }
            }
            it1++; it2++;
        } else if (node_id1[it1] < node_id2[it2]) {
            it1 = (int)(std::lower_bound(node_id1 + it1, node_id1 + nn1, node_id2[it2]) - node_id1);
        } else {
            it2 = (int)(std::lower_bound(node_id2 + it2, node_id2 + nn2, node_id1[it1]) - node_id2);
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[kHisto];
        for (int i = 0; i < kHisto; ++i) sizes[i] = (int)rotHist[i].size();
        three_maxima(sizes, kHisto, ind1, ind2, ind3);
        for (int i = 0; i < kHisto; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { matches12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

}  // extern "C"

// MapPoint::ComputeDistinctiveDescriptors (/root/reference/src/MapPoint.cc:397-458), the descriptor part: for each
// map point, given the descriptors of its observations (set p = rows set_start[p] .. set_start[p+1]), all pairwise
// distances, per row the median = sorted[int(0.5 * (N - 1))], and the FIRST row with the least median.
#include <algorithm>
extern "C" void orc_distinctive_descriptors(const uint8_t* desc, const int* set_start, int npoints, int* best_idx,
                                            int* best_median) {
    for (int p = 0; p < npoints; ++p) {
        const uint8_t* d = desc + 32 * (size_t)set_start[p];
        const int N = set_start[p + 1] - set_start[p];
        best_idx[p] = N > 0 ? 0 : -1;
        best_median[p] = INT_MAX;
        if (N <= 0) continue;
        std::vector<float> D((size_t)N * N);
        for (int i = 0; i < N; i++) {
            D[(size_t)i * N + i] = 0;
            for (int j = i + 1; j < N; j++) {
                const int distij = orc_hamming256(d + 32 * (size_t)i, d + 32 * (size_t)j);
                D[(size_t)i * N + j] = (float)distij;
                D[(size_t)j * N + i] = (float)distij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (int i = 0; i < N; i++) {
            std::vector<int> vDists(D.begin() + (size_t)i * N, D.begin() + (size_t)(i + 1) * N);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best_idx[p] = BestIdx;
        best_median[p] = BestMedian;
    }
}
