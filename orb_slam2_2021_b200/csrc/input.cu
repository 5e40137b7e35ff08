// input.cu — input staging right before the ORB front-end, declared in include/orbx_input.h:
// stereo rectification (cv::remap INTER_LINEAR with CV_32F maps, /root/reference/Examples/Stereo/arducam_images.cpp:127-133)
// and gray conversion (cvtColor, /root/reference/src/Tracking.cc:183-208), fused into ONE kernel that writes level 0.
//
// OpenCV's 8-bit arithmetic (pinned against cv2 in tests/test_input_oracle.py):
//   remap: sx = cvRound(mapx * 32), ix = sx >> 5, fx = sx & 31 (same for y); weights (32-fx)(32-fy)*32, fx(32-fy)*32,
//          (32-fx)fy*32, fx*fy*32 as shorts (32768 saturates to 32767 and the missing 1 goes to the opposite corner);
//          out = (sum w*p + 2^14) >> 15; neighbours outside the image are the constant border 0;
//   gray:  (R*9798 + G*19235 + B*3735 + 2^14) >> 15.
// The kernel is bandwidth work: per output pixel 8 B of maps, ~C bytes of raw image (neighbours overlap in L1/L2)
// and 1 B written. A thread produces 4 consecutive output pixels and one aligned 32-bit store.
#include <cstring>
#include <string>

#include "../../include/orbx_input.h"
#include "engine.h"
#include "matcher.h"

template <int C>
__device__ __forceinline__ int gray_of(const int* v, int rgbOrder)
{
    if (C == 1) return v[0];
    const int r = rgbOrder ? v[0] : v[2], g = v[1], b = rgbOrder ? v[2] : v[0];
    return (r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15;
}

#define STAGE_NI 8  // images per thread: the maps, weights and source offsets of a pixel are shared by all images of a side

template <int C>
__global__ void __launch_bounds__(128) orbx_stage_kernel(OrbxStageArgs a, int nImgs)
{
    const int x0 = (blockIdx.x * 128 + threadIdx.x) * 4, y = blockIdx.y, img0 = blockIdx.z * STAGE_NI;
    if (x0 >= a.cols) return;
    // per pixel: source offset of the top-left neighbour, the four weights, and which neighbours lie inside the image
    int ofs[4], w00[4], w01[4], w10[4], w11[4];
    unsigned in[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, a.cols - 1);  // lanes past the last column repeat it; their bytes land in the row padding
        if (a.map1) {
            const size_t mi = (size_t)(y + a.cropTop) * a.mapStride + x;
            const int sx = __float2int_rn(__fmul_rn(__ldg(a.map1 + mi), 32.0f)), sy = __float2int_rn(__fmul_rn(__ldg(a.map2 + mi), 32.0f));
            const int fx = sx & 31, fy = sy & 31;
            const int ix = min(max(sx >> 5, -32768), 32767), iy = min(max(sy >> 5, -32768), 32767);
            w00[k] = (32 - fx) * (32 - fy) * 32; w01[k] = fx * (32 - fy) * 32; w10[k] = (32 - fx) * fy * 32; w11[k] = fx * fy * 32;
            if (w00[k] > 32767) { w00[k] = 32767; w11[k] += 1; }
            const bool x0in = ix >= 0 && ix < a.srcCols, x1in = ix + 1 >= 0 && ix + 1 < a.srcCols;
            const bool y0in = iy >= 0 && iy < a.srcRows, y1in = iy + 1 >= 0 && iy + 1 < a.srcRows;
            in[k] = (x0in && y0in ? 1u : 0u) | (x1in && y0in ? 2u : 0u) | (x0in && y1in ? 4u : 0u) | (x1in && y1in ? 8u : 0u);
            ofs[k] = iy * a.srcStride + ix * C;
        } else {
            w00[k] = 32768; w01[k] = w10[k] = w11[k] = 0;  // identity: (32768 p + 2^14) >> 15 == p
            in[k] = 1u;
            ofs[k] = (y + a.cropTop) * a.srcStride + x * C;
        }
    }
    const int iEnd = min(img0 + STAGE_NI, nImgs);
    for (int img = img0; img < iEnd; ++img) {
        const uint8_t* src = a.src + (size_t)img * a.srcImgStride;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint8_t* p0 = src + ofs[k];
            const uint8_t* p1 = p0 + a.srcStride;
            int v[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const int p00 = (in[k] & 1u) ? p0[c] : 0, p01 = (in[k] & 2u) ? p0[C + c] : 0;
                const int p10 = (in[k] & 4u) ? p1[c] : 0, p11 = (in[k] & 8u) ? p1[C + c] : 0;
                v[c] = (w00[k] * p00 + w01[k] * p01 + w10[k] * p10 + w11[k] * p11 + (1 << 14)) >> 15;
            }
            packed |= (uint32_t)gray_of<C>(v, a.rgbOrder) << (8 * k);
        }
        // the pitch is a multiple of 16 and x0 of 4: the whole word lies inside the row (padding may be written)
        *reinterpret_cast<uint32_t*>(a.dst + (size_t)(a.imgStep * img + a.imgOfs) * a.dstImgStride + (size_t)y * a.dstPitch + x0) = packed;
    }
}

void orbx_launch_stage(const OrbxStageArgs& a, int channels, int nImgs, cudaStream_t s)
{
    dim3 grid((a.cols + 511) / 512, a.rows, (nImgs + STAGE_NI - 1) / STAGE_NI);
    if (channels == 1) orbx_stage_kernel<1><<<grid, 128, 0, s>>>(a, nImgs);
    else if (channels == 3) orbx_stage_kernel<3><<<grid, 128, 0, s>>>(a, nImgs);
    else orbx_stage_kernel<4><<<grid, 128, 0, s>>>(a, nImgs);
}

extern "C" int orbx_rectify_gray(orbx_matcher* m, const uint8_t* src, int rows, int cols, size_t stride, int channels, int rgbOrder,
                                 const float* map1, const float* map2, int mapRows, int mapCols, size_t mapStrideBytes, int cropTop,
                                 int cropRows, uint8_t* dst, size_t dstStride)
{
    if (!m || !src || !dst || rows < 1 || cols < 1 || (channels != 1 && channels != 3 && channels != 4) || stride < (size_t)cols * channels)
        return orbx_fail(ORBX_ERR_ARG, "bad source image");
    if ((map1 == nullptr) != (map2 == nullptr)) return orbx_fail(ORBX_ERR_ARG, "need both maps or none");
    const int outCols = map1 ? mapCols : cols, availRows = map1 ? mapRows : rows;
    if (map1 && (mapRows < 1 || mapCols < 1 || mapStrideBytes < (size_t)mapCols * 4 || (mapStrideBytes & 3))) return orbx_fail(ORBX_ERR_ARG, "bad maps");
    if (cropTop < 0 || cropRows < 1 || cropTop + cropRows > availRows || dstStride < (size_t)outCols) return orbx_fail(ORBX_ERR_ARG, "bad crop or destination");
    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));
    OrbxPacker pk;
    const size_t tight = (size_t)cols * channels;
    const size_t oSrc = pk.take(tight * rows), oM1 = pk.take(map1 ? (size_t)mapRows * mapCols * 4 : 0), oM2 = pk.take(map1 ? (size_t)mapRows * mapCols * 4 : 0);
    const size_t inBytes = pk.off;
    const int pitch = (outCols + 15) & ~15;
    const size_t oDst = pk.take((size_t)pitch * cropRows);
    TRY(orbx_matcher_grow(m, pk.off, inBytes, (size_t)pitch * cropRows));
    for (int r = 0; r < rows; ++r) std::memcpy(m->hIn + oSrc + r * tight, src + r * stride, tight);
    if (map1)
        for (int r = 0; r < mapRows; ++r) {
            std::memcpy(m->hIn + oM1 + (size_t)r * mapCols * 4, (const uint8_t*)map1 + r * mapStrideBytes, (size_t)mapCols * 4);
            std::memcpy(m->hIn + oM2 + (size_t)r * mapCols * 4, (const uint8_t*)map2 + r * mapStrideBytes, (size_t)mapCols * 4);
        }
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, m->stream), "H2D raw image", err));
    OrbxStageArgs a{};
    a.src = m->dBuf + oSrc; a.srcImgStride = 0; a.srcRows = rows; a.srcCols = cols; a.srcStride = (int)tight;
    a.map1 = map1 ? (const float*)(m->dBuf + oM1) : nullptr; a.map2 = map1 ? (const float*)(m->dBuf + oM2) : nullptr; a.mapStride = mapCols;
    a.cropTop = cropTop; a.rows = cropRows; a.cols = outCols;
    a.dst = m->dBuf + oDst; a.dstImgStride = 0; a.dstPitch = pitch; a.imgStep = 1; a.imgOfs = 0; a.rgbOrder = rgbOrder;
    orbx_launch_stage(a, channels, 1, m->stream);
    TRY(orbx_check(cudaGetLastError(), "stage launch", err));
    TRY(orbx_check(cudaMemcpy2DAsync(dst, dstStride, m->dBuf + oDst, pitch, outCols, cropRows, cudaMemcpyDeviceToHost, m->stream), "D2H gray", err));
    return orbx_check(cudaStreamSynchronize(m->stream), "sync", err);
}
