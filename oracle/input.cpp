// oracle/input.cpp — CPU ORACLE (TEST INFRASTRUCTURE) for the input staging right before the hot path
// (SURVEY.md §8f N3):
//   stereo rectification  cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) with CV_32F maps from
//                         initUndistortRectifyMap      /root/reference/Examples/Stereo/arducam_images.cpp:127-128, 257-272
//   gray conversion       cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY)
//                                                      /root/reference/src/Tracking.cc:183-208
// Both are OpenCV primitives (un-vendored dependency); restated from OpenCV's 8-bit fixed-point paths and pinned
// bit for bit against cv2 4.13 in tests/test_input_oracle.py:
//   C-8  gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15                        (all 2^24 colours checked)
//   C-9  remap: sx = cvRound(mapx*32), ix = sx >> 5, fx = sx & 31 (same for y); weights (32-fx)(32-fy)*32 etc. as
//        shorts (32768 saturates to 32767, the table fix-up puts the missing 1 on the opposite corner — no visible
//        effect); out = (sum w*p + 2^14) >> 15; neighbours outside the image are the constant border value 0.
#include <cmath>
#include <cstdint>

#include "orb_oracle.h"
#include "../orb_slam2_2021_b200/csrc/orbx_math.h"

extern "C" void orc_cvt_gray(const uint8_t* src, int rows, int cols, int stride, int channels, int rgb_order, uint8_t* dst, int dstride)
{
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols; ++x) {
            const uint8_t* p = src + (size_t)y * stride + (size_t)x * channels;
            if (channels == 1) { dst[(size_t)y * dstride + x] = p[0]; continue; }
            const int r = rgb_order ? p[0] : p[2], g = p[1], b = rgb_order ? p[2] : p[0];
            dst[(size_t)y * dstride + x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15);
        }
}

extern "C" void orc_remap_linear(const uint8_t* src, int srows, int scols, int sstride, int channels, const float* mapx, const float* mapy,
                                 int mstride_floats, int drows, int dcols, uint8_t* dst, int dstride)
{
    for (int y = 0; y < drows; ++y)
        for (int x = 0; x < dcols; ++x) {
            int sx = orbx_cvround(mapx[(size_t)y * mstride_floats + x] * 32.0f), sy = orbx_cvround(mapy[(size_t)y * mstride_floats + x] * 32.0f);
            const int fx = sx & 31, fy = sy & 31;
            int ix = sx >> 5, iy = sy >> 5;
            ix = ix < -32768 ? -32768 : ix > 32767 ? 32767 : ix;  // saturate_cast<short>
            iy = iy < -32768 ? -32768 : iy > 32767 ? 32767 : iy;
            int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
            if (w00 > 32767) { w00 = 32767; w11 += 1; }
            for (int c = 0; c < channels; ++c) {
                auto px = [&](int yy, int xx) -> int {
                    return (xx >= 0 && xx < scols && yy >= 0 && yy < srows) ? src[(size_t)yy * sstride + (size_t)xx * channels + c] : 0;
                };
                const int v = w00 * px(iy, ix) + w01 * px(iy, ix + 1) + w10 * px(iy + 1, ix) + w11 * px(iy + 1, ix + 1);
                dst[(size_t)y * dstride + (size_t)x * channels + c] = (uint8_t)((v + (1 << 14)) >> 15);
            }
        }
}
