// oracle/ref_gpu_wrap.cpp — TEST INFRASTRUCTURE ONLY (see orb_oracle.h).
//
// The drop-in proof: C entry points around the REFERENCE'S OWN Frame (/root/reference/src/Frame.cc, compiled where it lies
// with the patch of INTEGRATION.md and nothing else changed) running on top of liborbx.so:
//   step 2  ORB_SLAM2::ORBextractor = ORB_SLAM2::ORBextractorGPU (gpu_shim/frame_gpu_prefix.h), whose ORBX_WITH_OPENCV
//           overload has the reference's exact operator() signature (include/ORBextractor.h:75-77 of the reference);
//   step 3  the body of Frame::ComputeStereoMatches() is gpu_shim/ComputeStereoMatches.body.inc.
// The stereo constructor (src/Frame.cc:80-152) then runs as written: scale tables from the extractor's getters, two
// std::threads calling ExtractORB -> operator(), UndistortKeyPoints, ComputeStereoMatches, ComputeImageBounds,
// AssignFeaturesToGrid. tests/test_frame_dropin.py compares what it leaves in the Frame with the same constructor on the
// reference's CPU extractors (ref_wrap.cpp: ref_frame_stereo). Built into oracle/_ref/libref_frame_gpu.so, which links
// liborbx.so; nothing in the product links or loads this file.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>

#include <opencv2/core/core.hpp>

#include "Frame.h"  // the reference's, with gpu_shim/frame_gpu_prefix.h force-included
#include "orb_oracle.h"

#include "ref_frame_entry.inc"

extern "C" {

void* refgpu_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    try {
        return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);  // = ORBextractorGPU
    } catch (const std::exception& e) {
        std::fprintf(stderr, "refgpu_create: %s\n", e.what());
        return nullptr;
    }
}

void refgpu_destroy(void* h) { delete (ORB_SLAM2::ORBextractor*)h; }

// ORBextractorGPU::operator()(cv::InputArray, cv::InputArray, std::vector<cv::KeyPoint>&, cv::OutputArray) alone: the
// overload with the reference's signature, called the way Frame::ExtractORB calls it (src/Frame.cc:296-302)
int refgpu_extract(void* h, const uint8_t* img, int rows, int cols, int stride, orc_keypoint* kps, uint8_t* desc, int capacity)
{
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)h;
    cv::Mat image(rows, cols, CV_8UC1, (void*)img, (size_t)stride), descriptors;
    std::vector<cv::KeyPoint> keypoints;
    try {
        (*e)(image, cv::Mat(), keypoints, descriptors);
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "refgpu_extract: %s\n", ex.what());
        return -3;
    }
    const int n = (int)keypoints.size();
    if (n > capacity) return -1;
    if (n) std::memcpy(kps, keypoints.data(), (size_t)n * sizeof(orc_keypoint));
    for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
    return n;
}

int refgpu_frame_stereo(void* hL, void* hR, const uint8_t* imL, const uint8_t* imR, int rows, int cols, int stride, const float* K4,
                        float bf, float thDepth, float mbPreset, int capacity, orc_keypoint* keys, orc_keypoint* keysRight,
                        orc_keypoint* keysUn, uint8_t* desc, uint8_t* descRight, float* uRight, float* depth, int* gridStart,
                        int* gridItems, int* counts, float* misc, float* scaleTables)
{
    const FrameStereoOut o{keys, keysRight, keysUn, desc, descRight, uRight, depth, gridStart, gridItems, counts, misc, scaleTables};
    try {
        return frame_stereo_run((ORB_SLAM2::ORBextractor*)hL, (ORB_SLAM2::ORBextractor*)hR, imL, imR, rows, cols, stride, K4, bf, thDepth,
                                mbPreset, capacity, o);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "refgpu_frame_stereo: %s\n", e.what());
        return -3;
    }
}

}  // extern "C"
