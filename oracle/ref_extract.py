#!/usr/bin/env python3
"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Lifts whole member-function definitions, by line range and verbatim, out of the reference's src/Frame.cc,
include/Frame.h, src/ORBmatcher.cc, src/MapPoint.cc and src/KeyFrame.cc into include files under oracle/_ref/gen/ (git-ignored: reference TEXT never
enters the repository).  oracle/ref_capi.cpp compiles them against the stand-in class declarations of
oracle/ref_stub.hpp.  Those two translation-unit-sized files cannot be compiled whole here: they pull in DBoW2,
Eigen, Armadillo, g2o and ROS headers through Tracking.h / KeyFrame.h, none of which this image has.

Every range is anchored: its first line must contain the expected signature and its last line must be the closing
brace at the expected indentation, so a reference checkout whose lines have moved fails loudly instead of lifting
the wrong text.

usage: ref_extract.py <reference root> <output dir>
"""
import sys
from pathlib import Path

# (output file, source file, first line, last line, text the first line must contain, exact last line)
RANGES = [
    # --- class-body inlines of include/Frame.h (included inside the stand-in class Frame)
    ("frame_inline.inc", "include/Frame.h", 179, 190, "inline bool WorldToCameraPoint(const cv::Mat Pw, cv::Mat & Pc){", "    }"),
    ("frame_inline.inc", "include/Frame.h", 230, 263, "inline void PrepareStereoCandidates() {", "    }"),
    # --- member functions of src/Frame.cc
    ("frame_methods.inc", "src/Frame.cc", 461, 476, "void Frame::AssignFeaturesToGrid()", "}"),
    ("frame_methods.inc", "src/Frame.cc", 593, 646, "vector<size_t> Frame::GetFeaturesInArea(", "}"),
    ("frame_methods.inc", "src/Frame.cc", 648, 658, "bool Frame::PosInGrid(", "}"),
    ("frame_methods.inc", "src/Frame.cc", 670, 722, "void Frame::UndistortKeyPointsStereo()", "}"),
    ("frame_methods.inc", "src/Frame.cc", 724, 758, "void Frame::UndistortKeyPoints()", "}"),
    ("frame_methods.inc", "src/Frame.cc", 760, 828, "void Frame::ComputeImageBounds(const cv::Mat &imLeft)", "}"),
    ("frame_methods.inc", "src/Frame.cc", 830, 887, "void Frame::ComputeImageBoundsStereo(const cv::Mat &imLeft) {", "}"),
    ("frame_methods.inc", "src/Frame.cc", 889, 1078, "void Frame::ComputeStereoMatches()", "}"),
    ("frame_methods.inc", "src/Frame.cc", 1167, 1316, "int Frame::ComputeStereoMatches_Undistorted(bool isOnline)", "}"),
    # --- src/ORBmatcher.cc
    ("matcher_methods.inc", "src/ORBmatcher.cc", 37, 43, "const int ORBmatcher::TH_HIGH = 100;", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 155, 241, "int ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint *> &vpMapPoints, const float th)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 243, 249, "float ORBmatcher::RadiusByViewingCos(", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 520, 633, "int ORBmatcher::SearchForInitialization(", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 251, 268, "bool ORBmatcher::CheckDistEpipolarLine(", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 270, 404, "int ORBmatcher::SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint *> &vpMapPointMatches)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 406, 518, "int ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint *> &vpPoints, vector<MapPoint *> &vpMatched, int th)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 635, 768, "int ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint *> &vpMatches12)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 770, 935, "int ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12,", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 937, 1087, "int ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint *> &vpMapPoints, const float th)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 1214, 1438, "int ORBmatcher::SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint *> &vpMatches12,", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 1440, 1593, "int ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame,", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 1595, 1721, "int ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint *> &sAlreadyFound, const float th, const int ORBdist)", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 1723, 1764, "void ORBmatcher::ComputeThreeMaxima(", "}"),
    ("matcher_methods.inc", "src/ORBmatcher.cc", 1768, 1784, "int ORBmatcher::DescriptorDistance(", "}"),
    # --- src/MapPoint.cc, src/KeyFrame.cc: the small accessors the lifted searches call
    ("mappoint_methods.inc", "src/MapPoint.cc", 397, 462, "void MapPoint::ComputeDistinctiveDescriptors()", "}"),
    ("mappoint_methods.inc", "src/MapPoint.cc", 530, 534, "float MapPoint::GetMinDistanceInvariance()", "}"),
    ("mappoint_methods.inc", "src/MapPoint.cc", 536, 540, "float MapPoint::GetMaxDistanceInvariance()", "}"),
    ("mappoint_methods.inc", "src/MapPoint.cc", 542, 557, "int MapPoint::PredictScale(const float &currentDist, KeyFrame* pKF)", "}"),
    ("mappoint_methods.inc", "src/MapPoint.cc", 559, 574, "int MapPoint::PredictScale(const float &currentDist, Frame* pF)", "}"),
    ("keyframe_methods.inc", "src/KeyFrame.cc", 877, 916, "vector<size_t> KeyFrame::GetFeaturesInArea(const float &x, const float &y, const float &r) const", "}"),
    ("keyframe_methods.inc", "src/KeyFrame.cc", 918, 921, "bool KeyFrame::IsInImage(const float &x, const float &y) const", "}"),
]


def main(ref_root, out_dir):
    ref_root, out_dir = Path(ref_root), Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    cache, outputs = {}, {}
    for out, src, a, b, head, tail in RANGES:
        if src not in cache:
            cache[src] = (ref_root / src).read_text(encoding="utf-8", errors="replace").split("\n")
        lines = cache[src]
        first, last = lines[a - 1], lines[b - 1].rstrip("\r")
        if head not in first:
            sys.exit(f"ref_extract: {src}:{a} does not hold {head!r} (found {first!r})")
        if last != tail:
            sys.exit(f"ref_extract: {src}:{b} is {last!r}, expected the closing {tail!r}")
        chunk = outputs.setdefault(out, [])
        chunk.append(f"// ---- verbatim: {src}:{a}-{b}")
        chunk.append(f'#line {a} "{ref_root / src}"')
        chunk.extend(l.rstrip("\r") for l in lines[a - 1:b])
        chunk.append("")
    for out, chunk in outputs.items():
        (out_dir / out).write_text("\n".join(chunk) + "\n")
    (out_dir / "RANGES.txt").write_text(
        "".join(f"{src}:{a}-{b} -> {out}\n" for out, src, a, b, _, _ in RANGES))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
