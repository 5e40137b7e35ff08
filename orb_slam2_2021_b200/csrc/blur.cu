// blur.cu — 7x7 Gaussian (sigma 2) of every pyramid level, bit-for-bit cv::GaussianBlur on CV_8U.
// Replaces `GaussianBlur(workingMat, workingMat, Size(7,7), 2, 2, BORDER_REFLECT_101)` on the clone
// of each level (/root/reference/src/ORBextractor.cc:1083-1084).
//
// OpenCV's 8-bit path is fixed point: kernel k = {18,34,48,56,48,34,18}/256 per axis, no intermediate
// rounding, one rounding at the end: out = (sum_ij k_i k_j I(x+i-3, y+j-3) + 2^15) >> 16 with
// reflect-101 indexing at the level's own edges (the reference blurs a clone of the ROI). Integer
// sums are exact, so the two 1-D passes may run in either order; here the VERTICAL pass runs first
// because its sums (<= 65280) fit 16 bits and two pixels share one IMAD on a packed u16x2 register.
//
// One CTA per 96 x 64 output tile, all levels and images in one launch:
//   * one thread issues a TMA box load (cp.async.bulk.tensor.3d, 128 x 70 x 1 u8, 8960 B) of the tile
//     plus its halo (x0-16 .. x0+111, y0-3 .. y0+66; the copy engine addresses the innermost dimension
//     in 16-byte units, so the box starts 16 px left of the tile, measured with tools/ubench/tma_test2.cu)
//     and the CTA waits on an mbarrier; the engine zero-fills what lies outside the level, and tiles on
//     a level edge then mirror the reflect-101 halo in shared memory (columns first, then whole rows);
//   * warp w owns output rows 8w..8w+7, lane L the four tile columns 4L..4L+3: it walks down 14 rows
//     with a 7-row window of packed (u16x2) pixel pairs in registers (1 LDS.32 + 2 PRMT per 4 px per
//     row), and the vertical sums are exchanged with the neighbouring lanes by warp shuffles, so no
//     second shared-memory pass and no second barrier are needed; every lane produces 4 output pixels
//     each for lanes 4..27 (one aligned 32-bit store, 96 B contiguous per warp and row).
// HBM traffic: reads 1.46x the level (halo; mostly L2 hits between neighbouring tiles), writes it once.
#include "orbx_dev.cuh"
#include "orbx_tma.cuh"

struct OrbxTile {
    short x0, y0;
    uint8_t level, pad[3];
};

#define BL_WARPS 8
#define BL_RS (ORBX_BLUR_TH / BL_WARPS)  // output rows per warp
#define BL_SH (ORBX_BLUR_TH + 6)         // tile rows incl. halo
#define BL_SW 128                        // tile columns incl. halo (TMA box width)
#define BL_XH 16                         // columns left of the tile in the box

__global__ void __launch_bounds__(BL_WARPS * 32) orbx_blur_kernel(OrbxParams p, const OrbxTile* tiles, const CUtensorMap* maps)
{
    __shared__ __align__(128) uint8_t src[BL_SH * BL_SW];
    __shared__ __align__(8) uint64_t bar;
    const OrbxTile t = tiles[blockIdx.x];
    const int img = blockIdx.y, tid = threadIdx.x;
    const OrbxLevel& L = p.lv[t.level];
    const int x0 = t.x0, y0 = t.y0, w = L.w, h = L.h;
    if (tid == 0) {
        orbx_mbar_init(&bar, 1);
        orbx_fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        orbx_mbar_expect_tx(&bar, BL_SH * BL_SW);
        orbx_tma_load_3d(src, maps + t.level, x0 - BL_XH, y0 - 3, img + p.imgBase, &bar);
    }
    orbx_mbar_wait(&bar, 0);

    // ---- reflect-101 halo on level edges (tile column c <-> level column x0-16+c, row r <-> y0-3+r)
    const bool edgeX = x0 == 0 || x0 + BL_SW - BL_XH > w, edgeY = y0 == 0 || y0 + BL_SH - 3 > h;
    if (edgeX) {
        for (int i = tid; i < BL_SH * 6; i += BL_WARPS * 32) {
            const int r = i / 6, k = i - r * 6;
            const int gx = k < 3 ? -(k + 1) : w + (k - 3);
            const int c = gx - x0 + BL_XH;
            if (c >= 0 && c < BL_SW) {
                const int gs = gx < 0 ? -gx : 2 * (w - 1) - gx;
                src[r * BL_SW + c] = src[r * BL_SW + gs - x0 + BL_XH];
            }
        }
    }
    if (edgeX || edgeY) __syncthreads();
    if (edgeY) {
        if (tid < 6 * 32) {
            const int k = tid >> 5, c4 = tid & 31;
            const int gy = k < 3 ? -(k + 1) : h + (k - 3);
            const int r = gy - y0 + 3;
            if (r >= 0 && r < BL_SH) {
                const int gs = gy < 0 ? -gy : 2 * (h - 1) - gy;
                reinterpret_cast<uint32_t*>(src)[r * (BL_SW / 4) + c4] = reinterpret_cast<const uint32_t*>(src)[(gs - y0 + 3) * (BL_SW / 4) + c4];
            }
        }
        __syncthreads();
    }

    // ---- vertical pass on packed pairs with a sliding 7-row window, horizontal pass through shuffles
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t* col = reinterpret_cast<const uint32_t*>(src) + (warp * BL_RS) * (BL_SW / 4) + lane;
    const int gx = x0 - BL_XH + 4 * lane;
    const bool store = lane >= BL_XH / 4 && lane < (BL_XH + ORBX_BLUR_TW) / 4 && gx < w;
    uint8_t* dst = p.blur + (size_t)img * p.blurStride + L.boff + (size_t)(y0 + warp * BL_RS) * L.bpitch + gx;
    uint32_t wa[7], wb[7];  // window: pixel pairs (0,1) and (2,3) of the lane's 4 columns
#pragma unroll
    for (int s = 0; s < BL_RS + 6; ++s) {
        const uint32_t v = col[s * (BL_SW / 4)];
#pragma unroll
        for (int k = 0; k < 6; ++k) { wa[k] = wa[k + 1]; wb[k] = wb[k + 1]; }
        wa[6] = __byte_perm(v, 0, 0x4140);
        wb[6] = __byte_perm(v, 0, 0x4342);
        if (s >= 6) {
            // V = sum_j k_j I(y+j-3): per lane <= 256*255, no carry between the two 16-bit lanes
            const uint32_t va = 18u * (wa[0] + wa[6]) + 34u * (wa[1] + wa[5]) + 48u * (wa[2] + wa[4]) + 56u * wa[3];
            const uint32_t vb = 18u * (wb[0] + wb[6]) + 34u * (wb[1] + wb[5]) + 48u * (wb[2] + wb[4]) + 56u * wb[3];
            const uint32_t la = __shfl_up_sync(0xffffffffu, va, 1), lb = __shfl_up_sync(0xffffffffu, vb, 1);
            const uint32_t ra = __shfl_down_sync(0xffffffffu, va, 1), rb = __shfl_down_sync(0xffffffffu, vb, 1);
            // columns -3 .. 6 relative to the lane's first pixel
            const uint32_t cm3 = la >> 16, cm2 = lb & 0xFFFFu, cm1 = lb >> 16;
            const uint32_t c0 = va & 0xFFFFu, c1 = va >> 16, c2 = vb & 0xFFFFu, c3 = vb >> 16;
            const uint32_t c4 = ra & 0xFFFFu, c5 = ra >> 16, c6 = rb & 0xFFFFu;
            const uint32_t s0 = 18u * (cm3 + c3) + 34u * (cm2 + c2) + 48u * (cm1 + c1) + 56u * c0 + 32768u;
            const uint32_t s1 = 18u * (cm2 + c4) + 34u * (cm1 + c3) + 48u * (c0 + c2) + 56u * c1 + 32768u;
            const uint32_t s2 = 18u * (cm1 + c5) + 34u * (c0 + c4) + 48u * (c1 + c3) + 56u * c2 + 32768u;
            const uint32_t s3 = 18u * (c0 + c6) + 34u * (c1 + c5) + 48u * (c2 + c4) + 56u * c3 + 32768u;
            // byte 2 of each sum (sums < 2^24)
            const uint32_t packed = __byte_perm(__byte_perm(s0, s1, 0x0062), __byte_perm(s2, s3, 0x0062), 0x5410);
            const int oy = s - 6;
            if (store && y0 + warp * BL_RS + oy < h) *reinterpret_cast<uint32_t*>(dst + (size_t)oy * L.bpitch) = packed;
        }
    }
}

void orbx_launch_blur(const OrbxParams& p, const void* tileTable, int nTiles, const void* levelMaps, cudaStream_t s)
{
    dim3 grid(nTiles, p.imgs);
    orbx_blur_kernel<<<grid, BL_WARPS * 32, 0, s>>>(p, (const OrbxTile*)tileTable, (const CUtensorMap*)levelMaps);
}
