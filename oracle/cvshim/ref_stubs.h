// oracle/cvshim/ref_stubs.h — TEST INFRASTRUCTURE ONLY (see ../orb_oracle.h).
//
// Force-included (-include) when /root/reference/src/Frame.cc is compiled for oracle/_ref. Frame.h pulls in MapPoint.h and
// KeyFrame.h, and Frame.cc pulls in Converter.h (Eigen, g2o) and ORBmatcher.h — none of which can be compiled here.
// (ORBVocabulary.h is NOT stubbed: the vendored DBoW2 compiles against the OpenCV stand-in, so Frame::ComputeBoW runs the
// real TemplatedVocabulary::transform.) Their include guards are defined up front so that those headers
// read as empty, and the few names Frame.cc needs from them are declared below. Of these only
// ORBmatcher::TH_LOW/TH_HIGH/DescriptorDistance are ever used by the code that RUNS in the tests (ComputeStereoMatches,
// src/Frame.cc:522-700): they are defined in ref_wrap.cpp from src/ORBmatcher.cc:37-38 and :1672-1688. Everything else
// only has to link; those stubs abort.
#pragma once
#define MAPPOINT_H
#define KEYFRAME_H
#define CONVERTER_H
#define ORBMATCHER_H

#include <climits>
#include <cmath>
#include <vector>

#include <opencv2/cvshim.hpp>

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

using namespace std;  // the reference's headers get this from Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:36 (Frame.h uses bare `vector`)

namespace ORB_SLAM2 {

class Frame;

class MapPoint {
public:
    cv::Mat GetWorldPos();
    cv::Mat GetNormal();
    float GetMaxDistanceInvariance();
    float GetMinDistanceInvariance();
    int PredictScale(const float& currentDist, Frame* pF);
    bool mbTrackInView;
    float mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos;
    int mnTrackScaleLevel;
};

class KeyFrame {};

class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};

class ORBmatcher {
public:
    static const int TH_LOW;
    static const int TH_HIGH;
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
};

}  // namespace ORB_SLAM2
