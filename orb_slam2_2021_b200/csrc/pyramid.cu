// pyramid.cu — scale pyramid: level l = cv::resize(level l-1, INTER_LINEAR) bit-for-bit.
// Replaces ORBextractor::ComputePyramid (/root/reference/src/ORBextractor.cc:1105-1135). The 19-px
// reflect border the reference adds around every level is never read on this path and is not stored.
//
// Arithmetic (OpenCV 8-bit bilinear, 11-bit coefficients): per destination column the source
// column sx and the pair (c0,c1) = (cvRound((1-f)*2048), cvRound(f*2048)) and per destination row
// sy,(b0,b1) are precomputed on the host in float64/float32 exactly as OpenCV does, then
//   H(r) = I[r][sx]*c0 + I[r][sx+1]*c1          (int, < 2^19)
//   out  = (((b0*(H(sy)>>4))>>16) + ((b1*(H(sy+1)>>4))>>16) + 2) >> 2
// HBM-bound: reads 1.44x and writes 1x the destination level; each thread produces 4 px (one
// aligned 32-bit store), source taps come through L1.
#include "orbx_dev.cuh"

struct ResizeTab {
    const int* xofs;   // [dw] source column
    const int* xcoef;  // [dw] c0 | c1<<16
    const int* yofs;   // [dh] source row (unclipped)
    const int* ycoef;  // [dh] b0 | b1<<16
};

__global__ void __launch_bounds__(256) orbx_resize_kernel(OrbxParams p, int l, ResizeTab t)
{
    const OrbxLevel& D = p.lv[l];
    const int img = blockIdx.z;
    const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x0 >= D.w || y >= D.h) return;
    int sp;
    const uint8_t* src = orbx_level_ptr(p, img, l - 1, &sp);
    const int sw = p.lv[l - 1].w, sh = p.lv[l - 1].h;
    const int yo = t.yofs[y];
    const int sy0 = min(max(yo, 0), sh - 1), sy1 = min(max(yo + 1, 0), sh - 1);
    const int yc = t.ycoef[y];
    const int b0 = yc & 0xFFFF, b1 = yc >> 16;
    const uint8_t* r0 = src + (size_t)sy0 * sp;
    const uint8_t* r1 = src + (size_t)sy1 * sp;
    uint32_t packed = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = x0 + i;
        if (x < D.w) {
            const int sx = t.xofs[x];
            const int sx1 = min(sx + 1, sw - 1);
            const int xc = t.xcoef[x];
            const int c0 = xc & 0xFFFF, c1 = xc >> 16;
            const int h0 = r0[sx] * c0 + r0[sx1] * c1;
            const int h1 = r1[sx] * c0 + r1[sx1] * c1;
            int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            packed |= (uint32_t)v << (8 * i);
        }
    }
    uint8_t* dst = p.pyr + (size_t)img * p.pyrStride + D.off + (size_t)y * D.pitch;
    *reinterpret_cast<uint32_t*>(dst + x0) = packed;  // pitch is a multiple of 16: in-row padding absorbs the tail
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
        dim3 block(32, 8), grid((D.w + 127) / 128, (D.h + 7) / 8, p.imgs);
        orbx_resize_kernel<<<grid, block, 0, s>>>(p, l, t);
    }
}
