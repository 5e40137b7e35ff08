/* ORBextractorGPU.h — header-only C++ shim with the reference's ORBextractor signature, forwarding to the
 * C-ABI of liborbx.so (include/orbx.h). Drop-in for include/ORBextractor.h:56-100 of
 * lreithmayr/ORB_SLAM2_2021: same constructor arguments, same operator(), same getters, and
 * ComputeStereoMatches as a free function taking what Frame::ComputeStereoMatches reads
 * (src/Frame.cc:522-700).
 *
 * Built two ways:
 *   - with OpenCV headers (define ORBX_WITH_OPENCV before including): operator()(cv::InputArray image,
 *     cv::InputArray mask, std::vector<cv::KeyPoint>&, cv::OutputArray descriptors) exactly as the reference;
 *   - without OpenCV (this image has none): the same calls on POD views — orbx_keypoint has the
 *     28-byte layout of cv::KeyPoint, images are (ptr, rows, cols, stride).
 */
#ifndef ORBEXTRACTOR_GPU_H
#define ORBEXTRACTOR_GPU_H
#include <stdexcept>
#include <string>
#include <vector>

#include "orbx.h"
#ifdef ORBX_WITH_OPENCV
#include <opencv2/core.hpp>
#endif

namespace ORB_SLAM2 {

class ORBextractorGPU {
public:
    /* include/ORBextractor.h:65-66 */
    ORBextractorGPU(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
        : nlevels_(nlevels), scaleFactor_(scaleFactor)
    {
        if (orbx_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device, &h_) != ORBX_OK)
            throw std::runtime_error(std::string("orbx_create: ") + orbx_last_error());
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        orbx_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(), nullptr);
    }
    ~ORBextractorGPU() { orbx_destroy(h_); }
    ORBextractorGPU(const ORBextractorGPU&) = delete;
    ORBextractorGPU& operator=(const ORBextractorGPU&) = delete;

    /* POD form of operator() (include/ORBextractor.h:75-77; the mask is ignored by the reference).
     * An empty image returns without touching the outputs, like src/ORBextractor.cc:1044-1045. */
    void operator()(const uint8_t* image, int rows, int cols, size_t stride, std::vector<orbx_keypoint>& keypoints,
                    std::vector<uint8_t>& descriptors)
    {
        if (!image || rows == 0 || cols == 0) return;
        const int cap = orbx_max_keypoints(h_, rows, cols);
        if (cap < 0) throw std::runtime_error(std::string("orbx_max_keypoints: ") + orbx_last_error());
        keypoints.resize(cap);
        descriptors.resize((size_t)cap * 32);
        int n = 0;
        if (orbx_extract(h_, image, rows, cols, stride, keypoints.data(), descriptors.data(), cap, &n) != ORBX_OK)
            throw std::runtime_error(std::string("orbx_extract: ") + orbx_last_error());
        keypoints.resize(n);
        descriptors.resize((size_t)n * 32);
    }

#ifdef ORBX_WITH_OPENCV
    /* exactly the reference's signature */
    void operator()(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors)
    {
        if (_image.empty()) return;
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
        const int cap = orbx_max_keypoints(h_, image.rows, image.cols);
        _keypoints.resize(cap);
        cv::Mat desc(cap, 32, CV_8U);
        int n = 0;
        if (orbx_extract(h_, image.data, image.rows, image.cols, image.step, reinterpret_cast<orbx_keypoint*>(_keypoints.data()),
                         desc.data, cap, &n) != ORBX_OK)
            throw std::runtime_error(std::string("orbx_extract: ") + orbx_last_error());
        _keypoints.resize(n);
        if (n == 0) _descriptors.release(); else desc.rowRange(0, n).copyTo(_descriptors);
    }
    /* mvImagePyramid (include/ORBextractor.h:100), fetched lazily from the device */
    cv::Mat ImagePyramidLevel(int level) const
    {
        int r = 0, c = 0;
        orbx_pyramid_level_size(h_, level, &r, &c);
        cv::Mat m(r, c, CV_8U);
        orbx_pyramid_level(h_, level, m.data, m.step);
        return m;
    }
#endif

    /* include/ORBextractor.h:79-98 */
    int GetLevels() const { return nlevels_; }
    float GetScaleFactor() const { return scaleFactor_; }
    std::vector<float> GetScaleFactors() const { return mvScaleFactor; }
    std::vector<float> GetInverseScaleFactors() const { return mvInvScaleFactor; }
    std::vector<float> GetScaleSigmaSquares() const { return mvLevelSigma2; }
    std::vector<float> GetInverseScaleSigmaSquares() const { return mvInvLevelSigma2; }

    orbx_extractor* handle() const { return h_; }

private:
    orbx_extractor* h_ = nullptr;
    int nlevels_;
    float scaleFactor_;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
};

/* Frame::ComputeStereoMatches (src/Frame.cc:522-700) on what the Frame holds: mvKeys/mvKeysRight,
 * mDescriptors/mDescriptorsRight (N x 32 bytes), mbf, mb -> mvuRight, mvDepth (sized N, -1 = no match).
 * keysAreTheExtractorsOwn: the arrays are the unmodified results of the extractors' last operator() (true inside
 * Frame::Frame); pass false after editing keypoints or descriptors. */
inline void ComputeStereoMatchesGPU(ORBextractorGPU& left, ORBextractorGPU& right, const orbx_keypoint* keysL, const uint8_t* descL,
                                    int nL, const orbx_keypoint* keysR, const uint8_t* descR, int nR, float mbf, float mb,
                                    std::vector<float>& mvuRight, std::vector<float>& mvDepth, bool keysAreTheExtractorsOwn = true)
{
    mvuRight.assign(nL, -1.0f);
    mvDepth.assign(nL, -1.0f);
    /* Inside Frame::Frame the keypoints and descriptors are exactly what the two operator() calls just returned and what
     * the extractors still hold on the device: match there without uploading anything. Anything else (other counts, an
     * extract in between) makes the resident call refuse, and the caller's arrays go up as usual. */
    if (keysAreTheExtractorsOwn &&
        orbx_stereo_match_resident(left.handle(), right.handle(), nL, nR, mbf, mb, mvuRight.data(), mvDepth.data()) == ORBX_OK)
        return;
    if (orbx_stereo_match(left.handle(), right.handle(), keysL, descL, nL, keysR, descR, nR, mbf, mb, mvuRight.data(),
                          mvDepth.data()) != ORBX_OK)
        throw std::runtime_error(std::string("orbx_stereo_match: ") + orbx_last_error());
}

/* What Frame::Frame(stereo) does at src/Frame.cc:113-125 — ExtractORB left, ExtractORB right, ComputeStereoMatches — as ONE
 * call (orbx_stereo_frame): both images go up together, every kernel runs on both, nothing comes back to the host in between.
 * For a maintainer who is willing to replace those lines of the constructor instead of keeping its two threads:
 *     ORB_SLAM2::StereoFrameGPU(*mpORBextractorLeft, *mpORBextractorRight, imLeft.data, imRight.data, imLeft.rows, imLeft.cols,
 *                               imLeft.step, imRight.step, mbf, mb, keysL, descL, keysR, descR, mvuRight, mvDepth);
 * POD outputs: keypoints as orbx_keypoint (the 28-byte cv::KeyPoint layout), descriptors n x 32 bytes. */
inline void StereoFrameGPU(ORBextractorGPU& left, ORBextractorGPU& right, const uint8_t* imLeft, const uint8_t* imRight, int rows, int cols,
                           size_t strideLeft, size_t strideRight, float mbf, float mb, std::vector<orbx_keypoint>& keysL,
                           std::vector<uint8_t>& descL, std::vector<orbx_keypoint>& keysR, std::vector<uint8_t>& descR,
                           std::vector<float>& mvuRight, std::vector<float>& mvDepth)
{
    if (!imLeft || !imRight || rows == 0 || cols == 0) return;
    const int cap = orbx_max_keypoints(left.handle(), rows, cols);
    if (cap < 0) throw std::runtime_error(std::string("orbx_max_keypoints: ") + orbx_last_error());
    keysL.resize(cap); keysR.resize(cap);
    descL.resize((size_t)cap * 32); descR.resize((size_t)cap * 32);
    mvuRight.resize(cap); mvDepth.resize(cap);
    int nL = 0, nR = 0;
    if (orbx_stereo_frame(left.handle(), right.handle(), imLeft, imRight, rows, cols, strideLeft, strideRight, mbf, mb, keysL.data(),
                          descL.data(), &nL, keysR.data(), descR.data(), &nR, cap, mvuRight.data(), mvDepth.data()) != ORBX_OK)
        throw std::runtime_error(std::string("orbx_stereo_frame: ") + orbx_last_error());
    keysL.resize(nL); keysR.resize(nR);
    descL.resize((size_t)nL * 32); descR.resize((size_t)nR * 32);
    mvuRight.resize(nL); mvDepth.resize(nL);
}

}  // namespace ORB_SLAM2
#endif
