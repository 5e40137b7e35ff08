// blur.cu — 7x7 Gaussian (sigma 2) of every pyramid level, bit-for-bit cv::GaussianBlur on CV_8U.
// Replaces `GaussianBlur(workingMat, workingMat, Size(7,7), 2, 2, BORDER_REFLECT_101)` on the clone
// of each level (/root/reference/src/ORBextractor.cc:1083-1084).
//
// OpenCV's 8-bit path is fixed point: kernel {18,34,48,56,48,34,18}/256 per axis, 16-bit
// horizontal sums, one rounding at the end: out = (sum_v k * (sum_h k * I) + 2^15) >> 16, with
// reflect-101 indexing at the level's own edges (the reference blurs a clone of the ROI).
// One CTA per 64x32 output tile (all levels and images in one launch): the (70x38) source patch is
// staged in shared memory, the horizontal pass writes u16 rows, the vertical pass emits 4 px per
// thread as one aligned 32-bit store. HBM traffic: read 1.3x + write 1x the level.
#include "orbx_dev.cuh"

struct OrbxTile {
    short x0, y0;
    uint8_t level, pad[3];
};

#define BL_TW 64
#define BL_TH 32
#define BL_SW (BL_TW + 6)
#define BL_SH (BL_TH + 6)
#define BL_SP 72  // source patch row stride

__device__ __forceinline__ int reflect101(int p, int n)
{
    // n >= 62 here, one reflection suffices for |overshoot| <= 3
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return p;
}

__global__ void __launch_bounds__(256) orbx_blur_kernel(OrbxParams p, const OrbxTile* tiles)
{
    __shared__ __align__(16) uint8_t src[BL_SH * BL_SP];
    __shared__ __align__(16) uint16_t hor[BL_SH * BL_TW];
    const OrbxTile t = tiles[blockIdx.x];
    const int img = blockIdx.y;
    const OrbxLevel& L = p.lv[t.level];
    int pitch;
    const uint8_t* im = orbx_level_ptr(p, img, t.level, &pitch);
    for (int i = threadIdx.x; i < BL_SW * BL_SH; i += 256) {
        const int sy = i / BL_SW, sx = i - sy * BL_SW;
        const int gx = reflect101(t.x0 + sx - 3, L.w), gy = reflect101(t.y0 + sy - 3, L.h);
        // columns/rows beyond the level (tile overhang) are clamped into range; their outputs are not stored
        src[sy * BL_SP + sx] = im[(size_t)min(max(gy, 0), L.h - 1) * pitch + min(max(gx, 0), L.w - 1)];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < BL_SH * BL_TW; i += 256) {
        const int y = i / BL_TW, x = i - y * BL_TW;
        const uint8_t* s = src + y * BL_SP + x;
        hor[i] = (uint16_t)(18 * (s[0] + s[6]) + 34 * (s[1] + s[5]) + 48 * (s[2] + s[4]) + 56 * s[3]);
    }
    __syncthreads();
    uint8_t* dst = p.blur + (size_t)img * p.blurStride + L.boff;
    for (int i = threadIdx.x; i < BL_TH * (BL_TW / 4); i += 256) {
        const int y = i / (BL_TW / 4), x4 = (i - y * (BL_TW / 4)) * 4;
        const int gy = t.y0 + y, gx = t.x0 + x4;
        if (gy >= L.h || gx >= L.w) continue;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint16_t* h = hor + y * BL_TW + x4 + k;
            const uint32_t s = 18u * (h[0] + h[6 * BL_TW]) + 34u * (h[BL_TW] + h[5 * BL_TW]) +
                               48u * (h[2 * BL_TW] + h[4 * BL_TW]) + 56u * h[3 * BL_TW];
            packed |= ((s + 32768u) >> 16) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dst + (size_t)gy * L.bpitch + gx) = packed;  // bpitch % 16 == 0 absorbs the tail
    }
}

void orbx_launch_blur(const OrbxParams& p, const void* tileTable, int nTiles, cudaStream_t s)
{
    dim3 grid(nTiles, p.imgs);
    orbx_blur_kernel<<<grid, 256, 0, s>>>(p, (const OrbxTile*)tileTable);
}
