// ORACLE -- TEST INFRASTRUCTURE ONLY (see orb_oracle.hpp).  Build with -ffp-contract=off.
// Restates /root/reference/src/ORBextractor.cc; line numbers in comments refer to that file.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>

#include "cvprim.hpp"
#include "orb_oracle.hpp"

namespace orc {

namespace {
constexpr int kPatch = 31, kHalfPatch = 15, kEdge = 19;  // :72-74
const int8_t kPattern[512 * 2] = {
#include "orb_pattern_31.inc"
};
}  // namespace

// :409-469
Extractor::Extractor(int nf, float sf, int nl, int ini, int mn)
    : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf) {
    scale.assign(nl, 1.f);
    sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        scale[i] = (float)(scale[i - 1] * scaleFactor);  // float * double member, stored to float
        sigma2[i] = scale[i] * scale[i];
    }
    inv_scale.resize(nl);
    inv_sigma2.resize(nl);
    for (int i = 0; i < nl; ++i) {
        inv_scale[i] = 1.0f / scale[i];
        inv_sigma2[i] = 1.0f / sigma2[i];
    }
    pyramid.resize(nl);

    quota.assign(nl, 0);
    const float factor = (float)(1.0f / scaleFactor);
    float desired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        quota[l] = cv_round(desired);
        sum += quota[l];
        desired *= factor;
    }
    quota[nl - 1] = std::max(nfeatures - sum, 0);

    std::memcpy(pattern, kPattern, sizeof(pattern));

    // circular patch row half-widths (:453-468)
    std::memset(umax, 0, sizeof(umax));
    const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// :1176-1201
void Extractor::compute_pyramid(const uint8_t* img, int w, int h, size_t pitch) {
    for (int l = 0; l < nlevels; ++l) {
        const float s = inv_scale[l];
        Image& im = pyramid[l];
        im.w = cv_round((float)w * s);
        im.h = cv_round((float)h * s);
        im.pitch = (size_t)im.w + 2 * kEdge;
        im.buf.assign(im.pitch * (size_t)(im.h + 2 * kEdge), 0);
        if (l == 0) {
            copy_make_border_reflect101(img, w, h, pitch, im.buf.data(), im.pitch, kEdge);
        } else {
            const Image& up = pyramid[l - 1];
            std::vector<uint8_t> tmp((size_t)im.w * im.h);
            resize_linear_8u(up.roi(), up.w, up.h, up.pitch, tmp.data(), im.w, im.h, (size_t)im.w);
            copy_make_border_reflect101(tmp.data(), im.w, im.h, (size_t)im.w, im.buf.data(), im.pitch, kEdge);
        }
    }
}

// :767-831 (one level)
void Extractor::detect_level(int level, std::vector<Candidate>& out) const {
    out.clear();
    const Image& im = pyramid[level];
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = im.w - kEdge + 3, maxBY = im.h - kEdge + 3;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const float W = 30;
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    if (nCols <= 0 || nRows <= 0) return;  // reference divides by zero here; defined as "no keypoints"
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);

    std::vector<FastPoint> cell;
    for (int i = 0; i < nRows; ++i) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; ++j) {
            const float iniX = (float)(minBX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = (float)maxBX;
            const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
            const uint8_t* sub = im.roi() + (size_t)y0 * im.pitch + x0;
            fast9_16(sub, cw, ch, im.pitch, iniTh, true, cell);
            if (cell.empty()) fast9_16(sub, cw, ch, im.pitch, minTh, true, cell);
            for (const FastPoint& p : cell) out.push_back({p.x + j * wCell, p.y + i * hCell, p.score});
        }
    }
}

// ---------------------------------------------------------------------------------------------
// quadtree culling (:481-763)
// ---------------------------------------------------------------------------------------------
namespace {
struct QNode {
    int ulx, uly, urx, ury, blx, bly, brx, bry;  // the four corner points, kept individually like the reference
    std::vector<int> pts;                        // candidate indices, always in original order
    bool no_more = false;
    long seq = 0;                                // creation sequence number (stands in for the heap address)
    std::list<QNode>::iterator self;
};

// :481-537
void divide(const QNode& n, const std::vector<Candidate>& c, QNode out[4]) {
    const int halfX = (int)std::ceil((float)(n.urx - n.ulx) / 2);
    const int halfY = (int)std::ceil((float)(n.bry - n.uly) / 2);
    QNode &n1 = out[0], &n2 = out[1], &n3 = out[2], &n4 = out[3];
    n1.ulx = n.ulx;         n1.uly = n.uly;
    n1.urx = n.ulx + halfX; n1.ury = n.uly;
    n1.blx = n.ulx;         n1.bly = n.uly + halfY;
    n1.brx = n.ulx + halfX; n1.bry = n.uly + halfY;

    n2.ulx = n1.urx; n2.uly = n1.ury;
    n2.urx = n.urx;  n2.ury = n.ury;
    n2.blx = n1.brx; n2.bly = n1.bry;
    n2.brx = n.urx;  n2.bry = n.uly + halfY;

    n3.ulx = n1.blx; n3.uly = n1.bly;
    n3.urx = n1.brx; n3.ury = n1.bry;
    n3.blx = n.blx;  n3.bly = n.bly;
    n3.brx = n1.brx; n3.bry = n.bly;

    n4.ulx = n3.urx; n4.uly = n3.ury;
    n4.urx = n2.brx; n4.ury = n2.bry;
    n4.blx = n3.brx; n4.bly = n3.bry;
    n4.brx = n.brx;  n4.bry = n.bry;

    for (int idx : n.pts) {
        const float px = (float)c[idx].x, py = (float)c[idx].y;
        if (px < n1.urx) {
            if (py < n1.bry) n1.pts.push_back(idx);
            else n3.pts.push_back(idx);
        } else if (py < n1.bry) n2.pts.push_back(idx);
        else n4.pts.push_back(idx);
    }
    for (int k = 0; k < 4; ++k) out[k].no_more = out[k].pts.size() == 1;
}
}  // namespace

std::vector<int> Extractor::distribute_octree(const std::vector<Candidate>& c, int minX, int maxX,
                                              int minY, int maxY, int N) {
    std::vector<int> result;
    const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
    if (nIni <= 0 || c.empty()) return result;  // reference: division by zero / empty list; defined as "none"
    const float hX = (float)(maxX - minX) / nIni;

    std::list<QNode> nodes;
    long seq = 0;
    std::vector<QNode*> ini(nIni);
    for (int i = 0; i < nIni; ++i) {
        QNode n;
        n.ulx = (int)(hX * (float)i);       n.uly = 0;
        n.urx = (int)(hX * (float)(i + 1)); n.ury = 0;
        n.blx = n.ulx; n.bly = maxY - minY;
        n.brx = n.urx; n.bry = maxY - minY;
        n.seq = seq++;
        nodes.push_back(n);
        ini[i] = &nodes.back();
    }
    for (int i = 0; i < (int)c.size(); ++i) {
        int slot = (int)((float)c[i].x / hX);
        ini[slot]->pts.push_back(i);
    }
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->pts.size() == 1) { it->no_more = true; ++it; }
        else if (it->pts.empty()) it = nodes.erase(it);
        else ++it;
    }

    using SizeNode = std::pair<int, QNode*>;
    // ties on size: the reference orders by heap address (:591,684); frozen here as creation order
    auto by_size_then_age = [](const SizeNode& a, const SizeNode& b) {
        return a.first != b.first ? a.first < b.first : a.second->seq < b.second->seq;
    };
    auto add_children = [&](QNode kids[4], std::vector<SizeNode>& expandable, int* nToExpand) {
        for (int k = 0; k < 4; ++k) {
            if (kids[k].pts.empty()) continue;
            kids[k].seq = seq++;
            nodes.push_front(std::move(kids[k]));
            nodes.front().self = nodes.begin();
            if (nodes.front().pts.size() > 1) {
                if (nToExpand) ++*nToExpand;
                expandable.push_back({(int)nodes.front().pts.size(), &nodes.front()});
            }
        }
    };

    bool finish = false;
    std::vector<SizeNode> expandable;
    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        expandable.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->no_more) { ++it; continue; }
            QNode kids[4];
            divide(*it, c, kids);
            add_children(kids, expandable, &nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<SizeNode> prev = expandable;
                expandable.clear();
                std::sort(prev.begin(), prev.end(), by_size_then_age);
                for (int j = (int)prev.size() - 1; j >= 0; --j) {
                    QNode kids[4];
                    divide(*prev[j].second, c, kids);
                    add_children(kids, expandable, nullptr);
                    nodes.erase(prev[j].second->self);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }

    // best response per node, first one wins ties (:744-760)
    result.reserve(nodes.size());
    for (const QNode& n : nodes) {
        int best = n.pts[0];
        for (size_t k = 1; k < n.pts.size(); ++k)
            if ((float)c[n.pts[k]].response > (float)c[best].response) best = n.pts[k];
        result.push_back(best);
    }
    return result;
}

// :76-103
float Extractor::ic_angle(const Image& im, int x, int y) const {
    int m01 = 0, m10 = 0;
    const uint8_t* ctr = im.roi() + (ptrdiff_t)y * (ptrdiff_t)im.pitch + x;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * ctr[u];
    const ptrdiff_t step = (ptrdiff_t)im.pitch;
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int lo = ctr[u + v * step], up = ctr[u - v * step];
            vsum += lo - up;
            m10 += u * (lo + up);
        }
        m01 += v * vsum;
    }
    return fast_atan2((float)m01, (float)m10);
}

// :107-146
void Extractor::orb_descriptor(const uint8_t* img, int pitch, int x, int y, float angle_deg, uint8_t* out) const {
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    const float angle = angle_deg * factorPI;
    const float a = cosf(angle), b = sinf(angle);
    const uint8_t* ctr = img + (ptrdiff_t)y * pitch + x;
    auto tap = [&](int idx) -> int {
        const float px = (float)pattern[idx][0], py = (float)pattern[idx][1];
        const int r = cv_round(px * b + py * a);
        const int cc = cv_round(px * a - py * b);
        return ctr[(ptrdiff_t)r * pitch + cc];
    };
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            const int t0 = tap(16 * i + 2 * k), t1 = tap(16 * i + 2 * k + 1);
            val |= (t0 < t1) << k;
        }
        out[i] = (uint8_t)val;
    }
}

// :1112-1174 (+ ComputeKeyPointsOctTree :767-855)
int Extractor::extract(const uint8_t* img, int w, int h, size_t pitch,
                       std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc) {
    kps.clear();
    desc.clear();
    last_candidates.assign(nlevels, {});
    last_level_counts.assign(nlevels, 0);
    if (img == nullptr || w <= 0 || h <= 0) return 0;

    compute_pyramid(img, w, h, pitch);

    std::vector<std::vector<KeyPoint>> all(nlevels);
    for (int l = 0; l < nlevels; ++l) {
        const Image& im = pyramid[l];
        const int minBX = kEdge - 3, minBY = minBX, maxBX = im.w - kEdge + 3, maxBY = im.h - kEdge + 3;
        std::vector<Candidate>& cands = last_candidates[l];
        detect_level(l, cands);
        const std::vector<int> keep = distribute_octree(cands, minBX, maxBX, minBY, maxBY, quota[l]);
        const int patch = (int)(kPatch * scale[l]);
        for (int idx : keep) {
            KeyPoint k;
            k.x = (float)cands[idx].x + (float)minBX;
            k.y = (float)cands[idx].y + (float)minBY;
            k.size = (float)patch;
            k.angle = -1.f;
            k.response = (float)cands[idx].response;
            k.octave = l;
            k.class_id = -1;
            all[l].push_back(k);
        }
        last_level_counts[l] = (int)all[l].size();
    }
    for (int l = 0; l < nlevels; ++l)
        for (KeyPoint& k : all[l]) k.angle = ic_angle(pyramid[l], cv_round(k.x), cv_round(k.y));

    std::vector<uint8_t> blurred;
    for (int l = 0; l < nlevels; ++l) {
        if (all[l].empty()) continue;
        const Image& im = pyramid[l];
        blurred.resize((size_t)im.w * im.h);
        gaussian_blur_7x7_s2(im.roi(), im.w, im.h, im.pitch, blurred.data(), (size_t)im.w);
        for (KeyPoint& k : all[l]) {
            const size_t at = desc.size();
            desc.resize(at + 32);
            orb_descriptor(blurred.data(), im.w, cv_round(k.x), cv_round(k.y), k.angle, &desc[at]);
            if (l != 0) { k.x *= scale[l]; k.y *= scale[l]; }
            kps.push_back(k);
        }
    }
    return (int)kps.size();
}

}  // namespace orc
