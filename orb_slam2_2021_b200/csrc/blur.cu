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

struct __align__(8) OrbxTile {  // 8 bytes, loaded as one uniform 64-bit word
    short x0, y0;
    uint8_t level, pad[3];
};

#define BL_WARPS 4
#define BL_RS (ORBX_BLUR_TH / BL_WARPS)  // output rows per warp
#define BL_SH (ORBX_BLUR_TH + 6)         // tile rows incl. halo
#define BL_SW 160                        // tile columns incl. halo (TMA box width)
#define BL_XH 16                         // columns left of the tile in the box

// byte weights of the horizontal dot products, as constant-bank operands of IDP.4A (no moves into registers)
#define PK(a, b, c, d) ((uint32_t)(a) | ((uint32_t)(b) << 8) | ((uint32_t)(c) << 16) | ((uint32_t)(d) << 24))
__constant__ uint32_t c_blurW[10] = {PK(0, 18, 34, 48), PK(56, 48, 34, 18),                    // pixel 0: wm, w0
                                     PK(0, 0, 18, 34),  PK(48, 56, 48, 34), PK(18, 0, 0, 0),    // pixel 1: wm, w0, w1
                                     PK(0, 0, 0, 18),   PK(34, 48, 56, 48), PK(34, 18, 0, 0),   // pixel 2
                                     PK(18, 34, 48, 56), PK(48, 34, 18, 0)};                    // pixel 3: w0, w1
// vertical weights for IDP.2A (low / high byte pair of a word): rows 2t..2t+6 of the window -> {18,34 | 48,56} {48,34 | 18,0};
// rows 2t+1..2t+7 -> {0,18 | 34,48} {56,48 | 34,18}
__constant__ uint32_t c_blurV[4] = {PK(18, 34, 48, 56), PK(48, 34, 18, 0), PK(0, 18, 34, 48), PK(56, 48, 34, 18)};
#undef PK

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
    // rows this lane stores: none for columns right of the level, else up to the level's last row
    const int nst = gx < w ? min(BL_RS, h_ - (y0 + warp * BL_RS)) : 0;
    const int bpitch = L.bpitch;
    uint8_t* dst = p.blur + (size_t)img * p.blurStride + L.boff + (size_t)(y0 + warp * BL_RS) * bpitch + gx;
    // Two tile rows per step: their horizontal sums (+128, <= 65408) are packed per pixel as u16x2 (even row low, odd row high);
    // four such pairs P[0..3] = tile rows 2t .. 2t+7 are the window of output rows 2t (rows 2t..2t+6) and 2t+1 (rows
    // 2t+1..2t+7), and the vertical pass is four 2-term dot products (IDP.2A) per pixel instead of 3 additions + 4 multiply-adds.
    uint32_t P[4][4];
#pragma unroll
    for (int m = 0; m < (BL_RS + 6) / 2; ++m) {
        uint32_t hh[2][4];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int s = 2 * m + r;
            const uint32_t wm = col[s * (BL_SW / 4)], w0 = col[s * (BL_SW / 4) + 1], w1 = col[s * (BL_SW / 4) + 2];
            // pixel i of word w0 has its 7 taps at bytes i-3 .. i+3 relative to w0's byte 0 (weights: c_blurW). Every sum
            // starts at 128: the weights add up to 256, so the vertical pass comes out 2^15 high — the rounding term
            hh[r][0] = __dp4a(w0, c_blurW[1], __dp4a(wm, c_blurW[0], 128u));
            hh[r][1] = __dp4a(w1, c_blurW[4], __dp4a(w0, c_blurW[3], __dp4a(wm, c_blurW[2], 128u)));
            hh[r][2] = __dp4a(w1, c_blurW[7], __dp4a(w0, c_blurW[6], __dp4a(wm, c_blurW[5], 128u)));
            hh[r][3] = __dp4a(w1, c_blurW[9], __dp4a(w0, c_blurW[8], 128u));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            P[0][i] = P[1][i]; P[1][i] = P[2][i]; P[2][i] = P[3][i];
            P[3][i] = __byte_perm(hh[0][i], hh[1][i], 0x5410);
        }
        if (m >= 3) {
            uint32_t va[4], vb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                va[i] = __dp2a_hi(P[3][i], c_blurV[1], __dp2a_lo(P[2][i], c_blurV[1], __dp2a_hi(P[1][i], c_blurV[0], __dp2a_lo(P[0][i], c_blurV[0], 0u))));
                vb[i] = __dp2a_hi(P[3][i], c_blurV[3], __dp2a_lo(P[2][i], c_blurV[3], __dp2a_hi(P[1][i], c_blurV[2], __dp2a_lo(P[0][i], c_blurV[2], 0u))));
            }
            // byte 2 of each sum (sums < 2^24)
            const uint32_t pa = __byte_perm(__byte_perm(va[0], va[1], 0x0062), __byte_perm(va[2], va[3], 0x0062), 0x5410);
            const uint32_t pb = __byte_perm(__byte_perm(vb[0], vb[1], 0x0062), __byte_perm(vb[2], vb[3], 0x0062), 0x5410);
            if (2 * (m - 3) < nst) *reinterpret_cast<uint32_t*>(dst) = pa;
            if (2 * (m - 3) + 1 < nst) *reinterpret_cast<uint32_t*>(dst + bpitch) = pb;
            dst += 2 * bpitch;
        }
    }
}

void orbx_launch_blur(const OrbxParams& p, const void* tileTable, int nTiles, const void* levelMaps, cudaStream_t s)
{
    dim3 grid(nTiles, p.imgs);
    orbx_blur_kernel<<<grid, BL_WARPS * 32, 0, s>>>(p, (const OrbxTile*)tileTable, (const CUtensorMap*)levelMaps);
}
