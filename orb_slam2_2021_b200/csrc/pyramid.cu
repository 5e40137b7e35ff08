// pyramid.cu — scale pyramid: level l = cv::resize(level l-1, INTER_LINEAR) bit-for-bit.
// Replaces ORBextractor::ComputePyramid (/root/reference/src/ORBextractor.cc:1105-1135). The 19-px
// reflect border the reference adds around every level is never read on this path and is not stored.
//
// Arithmetic (OpenCV 8-bit bilinear, 11-bit coefficients): per destination column the source
// column sx and the pair (c0,c1) = (cvRound((1-f)*2048), cvRound(f*2048)) and per destination row
// sy,(b0,b1) are precomputed on the host in float64/float32 exactly as OpenCV does, then
//   H(r) = I[r][sx]*c0 + I[r][sx+1]*c1          (int, < 2^19)
//   out  = (((b0*(H(sy)>>4))>>16) + ((b1*(H(sy+1)>>4))>>16) + 2) >> 2
// Reads 1.44x and writes 1x the destination level; each thread produces 4 px x 16 rows (one aligned
// 32-bit store per row), source taps come through L1.
#include "orbx_dev.cuh"

struct ResizeTab {
    const int* xofs;   // [dw] source column
    const int* xcoef;  // [dw] c0 | c1<<16
    const int* yofs;   // [dh] source row (unclipped)
    const int* ycoef;  // [dh] b0 | b1<<16
};

#define RS_TX 32   // threads along x (4 columns each)
#define RS_TY 4    // row chunks per CTA
#define RS_ROWS 16  // destination rows per thread

// A thread owns 4 adjacent destination columns and walks down RS_ROWS destination rows: the column
// tables are read once, and a horizontally interpolated source row that serves two consecutive
// destination rows (scale 1.2: most of them) is computed once and kept in registers. The row pattern
// depends on y only, so the reuse test is warp-uniform.
__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_resize_kernel(OrbxParams p, int l, ResizeTab t)
{
    const OrbxLevel& D = p.lv[l];
    const int img = blockIdx.z;
    const int x0 = (blockIdx.x * RS_TX + threadIdx.x) * 4;
    if (x0 >= D.w) return;
    int sp;
    const uint8_t* src = orbx_level_ptr(p, img, l - 1, &sp);
    const int sw = p.lv[l - 1].w, sh = p.lv[l - 1].h;
    int sx[4], sx1[4], c0[4], c1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = min(x0 + i, D.w - 1);  // columns past the level land in the row padding; keep their reads in range
        sx[i] = t.xofs[x];
        sx1[i] = min(sx[i] + 1, sw - 1);
        const int xc = t.xcoef[x];
        c0[i] = xc & 0xFFFF;
        c1[i] = xc >> 16;
    }
    const int yBeg = (blockIdx.y * RS_TY + threadIdx.y) * RS_ROWS, yEnd = min(yBeg + RS_ROWS, D.h);
    uint8_t* dst = p.pyr + (size_t)img * p.pyrStride + D.off + (size_t)yBeg * D.pitch + x0;
    int hKeep[4] = {0, 0, 0, 0}, keepRow = -1;
#pragma unroll 1
    for (int y = yBeg; y < yEnd; ++y, dst += D.pitch) {
        const int yo = t.yofs[y], yc = t.ycoef[y];
        const int sy0 = min(max(yo, 0), sh - 1), sy1 = min(max(yo + 1, 0), sh - 1);
        const int b0 = yc & 0xFFFF, b1 = yc >> 16;
        int h0[4], h1[4];
        if (sy0 == keepRow) {
#pragma unroll
            for (int i = 0; i < 4; ++i) h0[i] = hKeep[i];
        } else {
            const uint8_t* r = src + (size_t)sy0 * sp;
#pragma unroll
            for (int i = 0; i < 4; ++i) h0[i] = r[sx[i]] * c0[i] + r[sx1[i]] * c1[i];
        }
        if (sy1 == sy0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) h1[i] = h0[i];
        } else {
            const uint8_t* r = src + (size_t)sy1 * sp;
#pragma unroll
            for (int i = 0; i < 4; ++i) h1[i] = r[sx[i]] * c0[i] + r[sx1[i]] * c1[i];
        }
        uint32_t packed = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = (((b0 * (h0[i] >> 4)) >> 16) + ((b1 * (h1[i] >> 4)) >> 16) + 2) >> 2;  // in [0,255]: b0 + b1 = 2048
            packed |= (uint32_t)v << (8 * i);
            hKeep[i] = h1[i];
        }
        keepRow = sy1;
        *reinterpret_cast<uint32_t*>(dst) = packed;  // pitch is a multiple of 16: in-row padding absorbs the tail
    }
}

// Level 0 ingest: tight rows (pitch = cols, any alignment) -> pitched rows (16-B multiple, TMA-legal).
// One thread per aligned 16-byte destination chunk; the source bytes come through L1 (rows are contiguous).
__global__ void __launch_bounds__(256) orbx_ingest_kernel(OrbxParams p, const uint8_t* src, int per, int side)
{
    const int chunksPerRow = p.lv[0].pitch >> 4;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= chunksPerRow * p.rows) return;
    const int y = i / chunksPerRow, cx = (i - y * chunksPerRow) << 4;
    const int f = blockIdx.y;
    const uint8_t* s = src + ((size_t)f * p.rows + y) * p.cols + cx;
    uint32_t w[4] = {0, 0, 0, 0};
    const int n = min(16, p.cols - cx);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        if (k < n) w[k >> 2] |= (uint32_t)s[k] << (8 * (k & 3));
    uint8_t* d = p.img0 + (size_t)(f * per + side) * p.img0Stride + (size_t)y * p.lv[0].pitch + cx;
    *reinterpret_cast<uint4*>(d) = make_uint4(w[0], w[1], w[2], w[3]);
}

void orbx_launch_ingest(const OrbxParams& p, const uint8_t* src, int nFrames, int per, int side, cudaStream_t s)
{
    const int chunks = (p.lv[0].pitch >> 4) * p.rows;
    dim3 grid((chunks + 255) / 256, nFrames);
    orbx_ingest_kernel<<<grid, 256, 0, s>>>(p, src, per, side);
}

// tabs[l] points at the level's table block: xofs[w] xcoef[w] yofs[h] ycoef[h]
void orbx_launch_pyramid(const OrbxParams& p, const int* const* tabs, cudaStream_t s)
{
    for (int l = 1; l < p.nlevels; ++l) {
        const OrbxLevel& D = p.lv[l];
        ResizeTab t;
        t.xofs = tabs[l];
        t.xcoef = t.xofs + D.w;
        t.yofs = t.xcoef + D.w;
        t.ycoef = t.yofs + D.h;
        dim3 block(RS_TX, RS_TY), grid((D.w + 4 * RS_TX - 1) / (4 * RS_TX), (D.h + RS_ROWS * RS_TY - 1) / (RS_ROWS * RS_TY), p.imgs);
        orbx_resize_kernel<<<grid, block, 0, s>>>(p, l, t);
    }
}
