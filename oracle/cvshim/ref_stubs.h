// oracle/cvshim/ref_stubs.h — TEST INFRASTRUCTURE ONLY (see ../orb_oracle.h).
//
// Force-included (-include) when /root/reference/src/Frame.cc is compiled for oracle/_ref. Frame.h pulls in MapPoint.h and
// KeyFrame.h, and Frame.cc pulls in Converter.h (Eigen, g2o) and ORBmatcher.h — none of which can be compiled here.
// (ORBVocabulary.h is NOT stubbed: the vendored DBoW2 compiles against the OpenCV stand-in, so Frame::ComputeBoW runs the
// real TemplatedVocabulary::transform.) Their include guards are defined up front so that those headers
// read as empty, and the names Frame.cc and ORBmatcher.cc need from them are declared below: MapPoint as a plain record
// behind the reference's accessor names, KeyFrame as names only (its users are compiled but never run), Converter's one
// helper. ORBmatcher.h / ORBmatcher.cc are the reference's own (not stubbed).
#pragma once
#define MAPPOINT_H
#define KEYFRAME_H
#define CONVERTER_H

#include <climits>
#include <cmath>
#include <map>
#include <set>
#include <vector>

#include <opencv2/cvshim.hpp>

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

using namespace std;  // the reference's headers get this from Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:36 (Frame.h uses bare `vector`)

namespace ORB_SLAM2 {

class Frame;

class KeyFrame;

// MapPoint: a plain record with the reference's accessor names (include/MapPoint.h:52-114). The two SearchByProjection
// overloads that the tests run only read data through these (world position, descriptor, observation count, the
// mTrack* fields Frame::isInFrustum fills); the mutating calls used by Fuse / SearchBySim3 abort.
class MapPoint {
public:
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetNormal();
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    int Observations() { return nObs; }
    bool isBad() { return mbBad; }
    float GetMaxDistanceInvariance();
    float GetMinDistanceInvariance();
    int PredictScale(const float& currentDist, Frame* pF);
    int PredictScale(const float& currentDist, KeyFrame* pKF);
    void AddObservation(KeyFrame* pKF, size_t idx);
    int GetIndexInKeyFrame(KeyFrame* pKF);
    bool IsInKeyFrame(KeyFrame* pKF);
    void Replace(MapPoint* pMP);
    bool mbTrackInView = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 0;
    int mnTrackScaleLevel = 0;
    long unsigned int mnLastFrameSeen = 0, mnFuseCandidateForKF = 0, mnTrackReferenceForFrame = 0;
    // storage behind the accessors
    cv::Mat mWorldPos, mDescriptor;
    int nObs = 0;
    bool mbBad = false;
};

// KeyFrame: only names (include/KeyFrame.h); every function of ORBmatcher.cc that takes a KeyFrame is compiled but never
// called by the tests, so the methods abort (ref_wrap.cpp).
class KeyFrame {
public:
    cv::Mat GetCameraCenter();
    cv::Mat GetRotation();
    cv::Mat GetTranslation();
    void AddMapPoint(MapPoint* pMP, const size_t& idx);
    std::set<MapPoint*> GetMapPoints();
    std::vector<MapPoint*> GetMapPointMatches();
    MapPoint* GetMapPoint(const size_t& idx);
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;
    bool IsInImage(const float& x, const float& y) const;
    float fx = 0, fy = 0, cx = 0, cy = 0, invfx = 0, invfy = 0, mbf = 0, mb = 0, mThDepth = 0;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors;
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    float mfScaleFactor = 0, mfLogScaleFactor = 0;
    int mnScaleLevels = 0, mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    long unsigned int mnId = 0;
};

class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};

}  // namespace ORB_SLAM2
