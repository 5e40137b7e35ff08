// oracle/ref_stubs_impl.cpp — TEST INFRASTRUCTURE ONLY (see orb_oracle.h).
// Link-time bodies for the names cvshim/ref_stubs.h declares in place of headers that cannot be compiled here (MapPoint,
// KeyFrame, Converter) and for cv::undistortPoints. Compiled into both oracle/_ref libraries (the CPU reference and the
// reference's Frame.cc on top of the GPU extractor).
#include <cstdlib>

#include <opencv2/core/core.hpp>

#include "Frame.h"  // with cvshim/ref_stubs.h force-included (see the Makefile)

namespace cv {
void undistortPoints(const Mat&, Mat&, const Mat&, const Mat&, const Mat&, const Mat&) { std::abort(); }  // zero distortion only
}

namespace ORB_SLAM2 {

// never reached by the functions the tests call; they only have to link
cv::Mat MapPoint::GetNormal() { std::abort(); }
float MapPoint::GetMaxDistanceInvariance() { std::abort(); }
float MapPoint::GetMinDistanceInvariance() { std::abort(); }
int MapPoint::PredictScale(const float&, Frame*) { std::abort(); }
int MapPoint::PredictScale(const float&, KeyFrame*) { std::abort(); }
void MapPoint::AddObservation(KeyFrame*, size_t) { std::abort(); }
int MapPoint::GetIndexInKeyFrame(KeyFrame*) { std::abort(); }
bool MapPoint::IsInKeyFrame(KeyFrame*) { std::abort(); }
void MapPoint::Replace(MapPoint*) { std::abort(); }
cv::Mat KeyFrame::GetCameraCenter() { std::abort(); }
cv::Mat KeyFrame::GetRotation() { std::abort(); }
cv::Mat KeyFrame::GetTranslation() { std::abort(); }
void KeyFrame::AddMapPoint(MapPoint*, const size_t&) { std::abort(); }
std::set<MapPoint*> KeyFrame::GetMapPoints() { std::abort(); }
std::vector<MapPoint*> KeyFrame::GetMapPointMatches() { std::abort(); }
MapPoint* KeyFrame::GetMapPoint(const size_t&) { std::abort(); }
std::vector<size_t> KeyFrame::GetFeaturesInArea(const float&, const float&, const float&) const { std::abort(); }
bool KeyFrame::IsInImage(const float&, const float&) const { std::abort(); }
// Converter::toDescriptorVector, src/Converter.cc:27-35: one 1x32 row header per descriptor
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors)
{
    std::vector<cv::Mat> vDesc;
    vDesc.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
    return vDesc;
}

}  // namespace ORB_SLAM2

