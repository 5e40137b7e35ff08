/* orbx_input.h — C-ABI for the input staging right before the ORB front-end:
 *
 *   stereo rectification  cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR), CV_32F maps from initUndistortRectifyMap
 *                         /root/reference/Examples/Stereo/arducam_images.cpp:127-128 (maps: :257-272), then the crop
 *                         imRect(cv::Rect(0, top, width, bottom - top)) :130-133
 *   gray conversion       cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY)
 *                         /root/reference/src/Tracking.cc:183-208 (stereo), :218-236 (RGB-D), :258-275 (mono)
 *
 * One fused kernel produces what ORBextractor::operator() is given: gray(crop(remap(raw))) — the remap runs on
 * the raw channels and the gray conversion on its result, in the reference's order, so every byte equals
 * OpenCV's. Raw camera frames become the ABI input of the batch engine: the rectified image never exists in
 * host memory and level 0 of the pyramid is written once.
 *
 * Same conventions as orbx.h. Part of liborbx.so; no CPU fallback.
 */
#ifndef ORBX_INPUT_H
#define ORBX_INPUT_H
#include "orbx_frame.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One image (host memory in and out): dst[crop_rows x out_cols] = gray(remap(src)) rows crop_top.. .
 * src: rows x cols x channels (1, 3 or 4) uint8, stride in bytes. rgb_order: 1 = RGB(A) (mbRGB), 0 = BGR(A).
 * map1/map2: CV_32FC1 x/y maps of map_rows x map_cols (stride in bytes), or both NULL for no rectification
 * (then out_cols = cols and the crop applies to src). out_cols = map_cols with maps. BORDER_CONSTANT 0. */
int orbx_rectify_gray(orbx_matcher* m, const uint8_t* src, int rows, int cols, size_t stride, int channels, int rgb_order,
                      const float* map1, const float* map2, int map_rows, int map_cols, size_t map_stride_bytes, int crop_top,
                      int crop_rows, uint8_t* dst, size_t dst_stride);

/* Batch engine: rectification maps of one side (0 = left, 1 = right), uploaded once per camera. The engine's image
 * size must be (crop rows) x map_cols with crop_top + rows <= map_rows. map1 == NULL clears the side's maps. */
int orbx_batch_set_rectification(orbx_batch* b, int side, const float* map1, const float* map2, int map_rows, int map_cols,
                                 size_t map_stride_bytes, int crop_top);

/* Batch engine: raw frames in, level 0 out. left/right: [n_frames][src_rows][src_cols][channels] uint8, tightly packed
 * host memory (right NULL for a mono engine). Replaces orbx_batch_upload for cameras that deliver unrectified or
 * colour frames; follow with orbx_batch_run. Without maps src_rows x src_cols must contain the engine's image
 * (rows crop_top.. are used, crop_top from orbx_batch_set_rectification or 0). */
int orbx_batch_upload_raw(orbx_batch* b, const uint8_t* left, const uint8_t* right, int n_frames, int src_rows, int src_cols,
                          int channels, int rgb_order);

#ifdef __cplusplus
}
#endif
#endif
