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
// One CTA per 128 x 64 output tile, all levels and images in one launch:
//   * one thread issues a TMA box load (cp.async.bulk.tensor.3d, 160 x 70 x 1 u8, 11200 B) of the tile
//     plus its halo (x0-16 .. x0+143, y0-3 .. y0+66; the copy engine addresses the innermost dimension
//     in 16-byte units, so the box starts 16 px left of the tile, measured with tools/ubench/tma_test2.cu)
//     and the CTA waits on an mbarrier; the engine zero-fills what lies outside the level, and tiles on
//     a level edge then mirror the reflect-101 halo in shared memory (columns first, then whole rows);
//   * warp w owns output rows 16w..16w+15, lane L the four output columns 4L..4L+3. It walks down 22 rows;
//     per row the HORIZONTAL 7-tap sums of its 4 pixels come straight from the raw bytes of three
//     32-bit words with 10 byte dot products (IDP.4A: measured full rate, on the FMA pipe, so no unpacking
//     on the ALU pipe that bounds the other kernels); a 7-row register window of those sums feeds the
//     vertical pass (4 IMAD + 3 IADD per pixel) and one aligned 32-bit store per lane and row
//     (128 B contiguous per warp). No second shared-memory pass, no shuffles, every lane produces output.
// HBM traffic: reads 1.37x the level (halo; mostly L2 hits between neighbouring tiles), writes it once.
#include "orbx_dev.cuh"
#include "orbx_tma.cuh"

struct OrbxTile {
    short x0, y0;
    uint8_t level, pad[3];
};

#define BL_WARPS 4
#define BL_RS (ORBX_BLUR_TH / BL_WARPS)  // output rows per warp
#define BL_SH (ORBX_BLUR_TH + 6)         // tile rows incl. halo
#define BL_SW 160                        // tile columns incl. halo (TMA box width)
#define BL_XH 16                         // columns left of the tile in the box

__global__ void __launch_bounds__(BL_WARPS * 32) orbx_blur_kernel(OrbxParams p, const OrbxTile* tiles, const CUtensorMap* maps)
{
    __shared__ __align__(128) uint8_t src[BL_SH * BL_SW];
    __shared__ __align__(8) uint64_t bar;
    const OrbxTile t = tiles[blockIdx.x];
    const int img = blockIdx.y, tid = threadIdx.x;
    const OrbxLevel& L = p.lv[t.level];
    const int x0 = t.x0, y0 = t.y0, w = L.w, h_ = L.h;
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
    const bool edgeX = x0 == 0 || x0 + BL_SW - BL_XH > w, edgeY = y0 == 0 || y0 + BL_SH - 3 > h_;
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
        for (int i = tid; i < 6 * (BL_SW / 4); i += BL_WARPS * 32) {
            const int k = i / (BL_SW / 4), c4 = i - k * (BL_SW / 4);
            const int gy = k < 3 ? -(k + 1) : h_ + (k - 3);
            const int r = gy - y0 + 3;
            if (r >= 0 && r < BL_SH) {
                const int gs = gy < 0 ? -gy : 2 * (h_ - 1) - gy;
                reinterpret_cast<uint32_t*>(src)[r * (BL_SW / 4) + c4] = reinterpret_cast<const uint32_t*>(src)[(gs - y0 + 3) * (BL_SW / 4) + c4];
            }
        }
        __syncthreads();
    }

    // ---- horizontal pass by byte dot products, vertical pass over a sliding 7-row register window
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t* col = reinterpret_cast<const uint32_t*>(src) + (warp * BL_RS) * (BL_SW / 4) + (BL_XH / 4 - 1) + lane;
    const int gx = x0 + 4 * lane;
    const bool store = gx < w;
    uint8_t* dst = p.blur + (size_t)img * p.blurStride + L.boff + (size_t)(y0 + warp * BL_RS) * L.bpitch + gx;
#define PK(a, b, c, d) ((uint32_t)(a) | ((uint32_t)(b) << 8) | ((uint32_t)(c) << 16) | ((uint32_t)(d) << 24))
    uint32_t h[7][4];  // horizontal sums (<= 256*255) of the lane's 4 pixels, last 7 rows
#pragma unroll
    for (int s = 0; s < BL_RS + 6; ++s) {
        const uint32_t wm = col[s * (BL_SW / 4)], w0 = col[s * (BL_SW / 4) + 1], w1 = col[s * (BL_SW / 4) + 2];
#pragma unroll
        for (int k = 0; k < 6; ++k)
#pragma unroll
            for (int i = 0; i < 4; ++i) h[k][i] = h[k + 1][i];
        // pixel i of word w0 has its 7 taps at bytes i-3 .. i+3 relative to w0's byte 0
        h[6][0] = __dp4a(w0, PK(56, 48, 34, 18), __dp4a(wm, PK(0, 18, 34, 48), 0u));
        h[6][1] = __dp4a(w1, PK(18, 0, 0, 0), __dp4a(w0, PK(48, 56, 48, 34), __dp4a(wm, PK(0, 0, 18, 34), 0u)));
        h[6][2] = __dp4a(w1, PK(34, 18, 0, 0), __dp4a(w0, PK(34, 48, 56, 48), __dp4a(wm, PK(0, 0, 0, 18), 0u)));
        h[6][3] = __dp4a(w1, PK(48, 34, 18, 0), __dp4a(w0, PK(18, 34, 48, 56), 0u));
        if (s >= 6) {
            uint32_t v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                v[i] = 18u * (h[0][i] + h[6][i]) + 34u * (h[1][i] + h[5][i]) + 48u * (h[2][i] + h[4][i]) + 56u * h[3][i] + 32768u;
            // byte 2 of each sum (sums < 2^24)
            const uint32_t packed = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            const int oy = s - 6;
            if (store && y0 + warp * BL_RS + oy < h_) *reinterpret_cast<uint32_t*>(dst + (size_t)oy * L.bpitch) = packed;
        }
    }
#undef PK
}

void orbx_launch_blur(const OrbxParams& p, const void* tileTable, int nTiles, const void* levelMaps, cudaStream_t s)
{
    dim3 grid(nTiles, p.imgs);
    orbx_blur_kernel<<<grid, BL_WARPS * 32, 0, s>>>(p, (const OrbxTile*)tileTable, (const CUtensorMap*)levelMaps);
}
