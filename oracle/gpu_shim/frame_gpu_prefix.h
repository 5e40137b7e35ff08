// oracle/gpu_shim/frame_gpu_prefix.h — TEST INFRASTRUCTURE ONLY.
// Force-included when the reference's src/Frame.cc is compiled on top of the GPU extractor (oracle/Makefile, target
// _ref/libref_frame_gpu.so). It applies step 2 of INTEGRATION.md without touching the reference's files: the reference's
// include/ORBextractor.h reads as empty (its include guard is defined here) and ORB_SLAM2::ORBextractor names the shim class
// of include/ORBextractorGPU.h, compiled with ORBX_WITH_OPENCV — the overload with the reference's exact
// operator()(cv::InputArray, cv::InputArray, std::vector<cv::KeyPoint>&, cv::OutputArray) signature — against the OpenCV
// stand-in of oracle/cvshim.
#pragma once
#include "../cvshim/ref_stubs.h"

#define ORBEXTRACTOR_H
#define ORBX_WITH_OPENCV
#include "ORBextractorGPU.h"
namespace ORB_SLAM2 {
using ORBextractor = ORBextractorGPU;
}
