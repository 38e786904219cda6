// ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's ORB front-end
// (/root/reference/src/ORBextractor.cc, src/Frame.cc:889-1078,1167-1316, src/ORBmatcher.cc:1768-1784).
// The reference itself cannot be compiled here (needs OpenCV 3.4.1 C++), so parity is
// pinned by (1) primitive-level bit-exact checks against cv2 4.13.0 and (2) golden vectors produced by
// a second, independent composition of the same algorithm over real cv2 primitives
// (tools/gen_golden_cv2.py -> tests/golden/).  See oracle/README.md.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orc {

// layout-identical to cv::KeyPoint (28 bytes); fields as left by ORBextractor::operator()
// (src/ORBextractor.cc:839-849,1164-1170)
struct KeyPoint {
    float x, y, size, angle, response;
    int octave, class_id;
};

struct Image {            // one padded pyramid level; roi() points at pixel (0,0) of the level
    int w = 0, h = 0;
    size_t pitch = 0;     // of the padded buffer
    std::vector<uint8_t> buf;
    uint8_t* roi() { return buf.data() + 19 * pitch + 19; }
    const uint8_t* roi() const { return buf.data() + 19 * pitch + 19; }
};

// one FAST survivor handed to the quadtree; coordinates relative to (minBorderX, minBorderY) = (16,16)
// of its level, in the reference's emission order (src/ORBextractor.cc:791-829)
struct Candidate { int x, y, response; };

class Extractor {
public:
    // src/ORBextractor.cc:409-469
    Extractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

    // src/ORBextractor.cc:1112-1174.  Returns number of keypoints; desc is n x 32.
    int extract(const uint8_t* img, int w, int h, size_t pitch,
                std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);

    // src/ORBextractor.cc:1176-1201
    void compute_pyramid(const uint8_t* img, int w, int h, size_t pitch);

    // cell loop of src/ORBextractor.cc:767-831 for one level (pyramid must be computed)
    void detect_level(int level, std::vector<Candidate>& out) const;

    // src/ORBextractor.cc:539-763 + DivideNode :481-537.  Returns indices into `cands` in result order.
    static std::vector<int> distribute_octree(const std::vector<Candidate>& cands, int minX, int maxX,
                                              int minY, int maxY, int N);

    // src/ORBextractor.cc:76-103 on a level image; (x,y) integer level coordinates
    float ic_angle(const Image& im, int x, int y) const;

    // src/ORBextractor.cc:107-146 on a blurred level image (plain w x h, pitch w)
    void orb_descriptor(const uint8_t* blurred, int pitch, int x, int y, float angle_deg, uint8_t* out32) const;

    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;                       // stored as double like the reference member
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;                   // mnFeaturesPerLevel
    int umax[16];
    int8_t pattern[512][2];
    std::vector<Image> pyramid;               // mvImagePyramid
    std::vector<std::vector<Candidate>> last_candidates;  // per level, kept for stage-level parity tests
    std::vector<int> last_level_counts;       // kept keypoints per level
};

// src/ORBmatcher.cc:1768-1784
int hamming256(const uint8_t* a, const uint8_t* b);

struct StereoParams {
    float bf;        // mbf
    float b;         // mb (the reference reads it before initialising it, src/Frame.cc:920 vs :322; explicit here)
    int mode;        // 0 = classic (SAD + parabola, src/Frame.cc:889-1078), 1 = _Undistorted (src/Frame.cc:1167-1316)
    int th_high;     // ORBmatcher::TH_HIGH = 100
    int th_low;      // ORBmatcher::TH_LOW  = 50
    float min_x;     // mnMinX, mode 1 only
};

struct StereoDebug {                 // optional per-left-key intermediates for stage-level parity
    std::vector<int> best_dist;      // best Hamming (th_high if none)
    std::vector<int> best_idx;       // right index (-1 if none)
    std::vector<int> sad;            // best SAD (-1 if the refinement did not run / was rejected)
};

// Classic + undistorted stereo association.  uRight/depth are resized to nL and filled (-1 = none).
// Returns the number of left keys that keep a depth (classic) / the reference's nmatched (undistorted).
int stereo_match(const Extractor& exL, const Extractor& exR, const StereoParams& prm,
                 const KeyPoint* kL, int nL, const uint8_t* dL, const KeyPoint* kR, int nR, const uint8_t* dR,
                 const float* minD_opt, const float* maxD_opt,
                 std::vector<float>& uRight, std::vector<float>& depth, StereoDebug* dbg);

}  // namespace orc
