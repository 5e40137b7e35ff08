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
// 32-bit store per row).
#include "orbx_dev.cuh"

struct ResizeArgs {
    const int* xofs;    // [dw] source column
    const int* xcoef;   // [dw] c0 | c1<<16
    const int4* ytab;   // [dh] {source row (unclipped), b0 << 12, b1 << 12, 0}
    const uint8_t* src; // level l-1 of image 0
    uint8_t* dst;       // level l of image 0
    size_t srcStride, dstStride;  // bytes between images
    int sw, sh, sp;     // source size and row pitch
    int dw, dh, dp;     // destination size and row pitch
};

#define RS_TX 32   // threads along x (4 columns each)
#define RS_TY 4    // row chunks per CTA
#define RS_ROWS 16  // destination rows per thread

// A thread owns 4 adjacent destination columns and RS_ROWS destination rows. It walks DOWN THE SOURCE ROWS that
// those destination rows need (consecutive, because every destination row uses two adjacent source rows and the
// row step is below 2): each source row is interpolated horizontally exactly once and a destination row is emitted
// when its second source row arrives. Per source row the thread loads three aligned 32-bit words (they cover
// the <= 11 bytes its four column pairs can span for scale factors up to 2; every image buffer ends with 64 spare
// bytes, so the words may run past the end of a row), realigns them with two funnel shifts, picks the four byte
// pairs with two PRMT and gets the four horizontal taps from four IDP.2A (16-bit coefficient pair x byte pair, on
// the FMA pipe). Loads are non-coherent (__ldg) and issued two source rows ahead of their use, so their latency
// overlaps the arithmetic of the rows before. The walk is unrolled six times — the period of the three word
// buffers and the two H buffers — so that no register is ever moved, and all geometry comes as kernel arguments
// (constant-bank operands).
//
// Source rows are walked by their UNCLIPPED index r = yofs .. yofs_last + 1 and read at min(r, sh - 1) (scale factors
// above 1 never give a negative yofs): at the bottom edge, where OpenCV clips row sy + 1 onto the last row, the walk
// simply visits that row twice, so a destination row always combines the previous and the current row of the walk.
//
// WIDE: a level ratio above 2 (scaleFactor 2 with an odd source size, e.g. 701 -> 350 columns) lets the four columns span
// 9 source bytes and puts columns 2 and 3 at least 4 bytes past the first one; their pairs are then picked from bytes
// sx0+4 .. sx0+11 (the third word always holds byte sx0+8) instead of sx0 .. sx0+7.
template <bool WIDE>
__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_resize_kernel(const ResizeArgs A)
{
    const int img = blockIdx.z;
    const int x0 = (blockIdx.x * RS_TX + threadIdx.x) * 4;
    const int yBeg = (blockIdx.y * RS_TY + threadIdx.y) * RS_ROWS, yEnd = min(yBeg + RS_ROWS, A.dh);
    if (yBeg >= yEnd) return;
    // row table entries of this warp's destination rows, staged in shared memory: the emit step reads them with a
    // dependent compare right behind, which as global loads stalled every destination row
    __shared__ int4 ytabS[RS_TY][RS_ROWS];
    if (threadIdx.x < RS_ROWS) ytabS[threadIdx.y][threadIdx.x] = A.ytab[min(yBeg + (int)threadIdx.x, A.dh - 1)];
    __syncwarp();
    if (x0 >= A.dw) return;
    // column geometry: aligned word index of the first source column, byte shift, PRMT selectors, packed coefficients
    uint32_t coef[4];
    int d[4], sx0 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = min(x0 + i, A.dw - 1);  // columns past the level land in the row padding
        const int sx = A.xofs[x];
        if (i == 0) sx0 = sx;
        d[i] = sx - sx0;                      // 0 .. 6 for level ratios up to 2, 0 .. 7 up to 7/3 (WIDE)
        if (WIDE && i >= 2) d[i] = max(d[i] - 4, 0);  // relative to byte sx0+4 (only clamped columns fall below it)
        // the right neighbour is clamped to the last column (there c1 == 0 anyway, OpenCV sets f = 0 for sx >= sw-1)
        const int d1 = sx + 1 <= A.sw - 1 ? d[i] + 1 : d[i];
        d[i] |= d1 << 4;
        coef[i] = (uint32_t)A.xcoef[x];       // c0 | c1 << 16
    }
    const int sh8 = (sx0 & 3) * 8;
    const uint32_t selA = (uint32_t)(d[0] & 15) | ((uint32_t)(d[0] >> 4) << 4) | ((uint32_t)(d[1] & 15) << 8) | ((uint32_t)(d[1] >> 4) << 12);
    const uint32_t selB = (uint32_t)(d[2] & 15) | ((uint32_t)(d[2] >> 4) << 4) | ((uint32_t)(d[3] & 15) << 8) | ((uint32_t)(d[3] >> 4) << 12);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A.src + (size_t)img * A.srcStride) + (sx0 >> 2);
    uint8_t* dst = A.dst + (size_t)img * A.dstStride + (size_t)yBeg * A.dp + x0;
    const int4* yt = ytabS[threadIdx.y];
    int4 ye = yt[0];  // next destination row: {yofs, b0 << 12, b1 << 12}
    const int rBeg = ye.x, rEnd = yt[yEnd - 1 - yBeg].x + 1, spw = A.sp >> 2, shm1 = A.sh - 1;
    int y = yBeg, r = rBeg;
    uint32_t a0, a1, a2, b0, b1, b2, c0, c1, c2;  // source rows r, r+1, r+2 in flight
    uint32_t h0[4] = {0, 0, 0, 0}, h1[4];
    {
        const uint32_t* pa = src + (size_t)min(rBeg, shm1) * spw;
        const uint32_t* pb = src + (size_t)min(rBeg + 1, shm1) * spw;
        a0 = __ldg(pa); a1 = __ldg(pa + 1); a2 = __ldg(pa + 2);
        b0 = __ldg(pb); b1 = __ldg(pb + 1); b2 = __ldg(pb + 2);
    }
    // one source row: prefetch row r+2 into (N0,N1,N2), interpolate row r from (W0,W1,W2) into HC, emit the destination
    // row whose second source row this is from HP (row r-1) and HC
#define RS_STEP(W0, W1, W2, N0, N1, N2, HP, HC)                                                                          \
    {                                                                                                                    \
        const uint32_t* pn = src + (size_t)min(r + 2, shm1) * spw;                                                       \
        N0 = __ldg(pn); N1 = __ldg(pn + 1); N2 = __ldg(pn + 2);                                                          \
        const uint32_t lo = __funnelshift_r(W0, W1, sh8), hi = __funnelshift_r(W1, W2, sh8); /* bytes sx0 .. sx0+7 */    \
        const uint32_t pA = __byte_perm(lo, hi, selA);                                                                   \
        const uint32_t pB = WIDE ? __byte_perm(hi, W2 >> sh8, selB) : __byte_perm(lo, hi, selB);                         \
        /* H & ~15 = 16 (H >> 4): the vertical pass only ever uses H >> 4 */                                             \
        HC[0] = __dp2a_lo(coef[0], pA, 0u) & ~15u;                                                                       \
        HC[1] = __dp2a_hi(coef[1], pA, 0u) & ~15u;                                                                       \
        HC[2] = __dp2a_lo(coef[2], pB, 0u) & ~15u;                                                                       \
        HC[3] = __dp2a_hi(coef[3], pB, 0u) & ~15u;                                                                       \
        if (ye.x + 1 == r) { /* (one destination row per source row for scale factors > 1) */                            \
            uint32_t v[4];                                                                                               \
            _Pragma("unroll") for (int i = 0; i < 4; ++i) /* (b * (h >> 4)) >> 16 == mulhi(b << 12, h & ~15) */          \
                v[i] = (__umulhi((uint32_t)ye.y, HP[i]) + __umulhi((uint32_t)ye.z, HC[i]) + 2u) >> 2; /* in [0,255] */   \
            *reinterpret_cast<uint32_t*>(dst) = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410); \
            dst += A.dp; /* pitch is a multiple of 16: in-row padding absorbs the tail */                                 \
            ++yt;                                                                                                        \
            ye = ++y < yEnd ? yt[0] : make_int4(0x7ffffff0, 0, 0, 0);                                                    \
        }                                                                                                                \
        if (++r > rEnd) return;                                                                                          \
    }
#pragma unroll 1
    for (;;) {
        RS_STEP(a0, a1, a2, c0, c1, c2, h0, h1)
        RS_STEP(b0, b1, b2, a0, a1, a2, h1, h0)
        RS_STEP(c0, c1, c2, b0, b1, b2, h0, h1)
        RS_STEP(a0, a1, a2, c0, c1, c2, h1, h0)
        RS_STEP(b0, b1, b2, a0, a1, a2, h0, h1)
        RS_STEP(c0, c1, c2, b0, b1, b2, h1, h0)
    }
#undef RS_STEP
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

// tabs[l] points at the level's table block: xofs[w] xcoef[w], then (16-byte aligned) ytab[h] as int4
void orbx_launch_pyramid(const OrbxParams& p, const int* const* tabs, cudaStream_t s)
{
    for (int l = 1; l < p.nlevels; ++l) {
        const OrbxLevel &S = p.lv[l - 1], &D = p.lv[l];
        ResizeArgs A;
        A.xofs = tabs[l];
        A.xcoef = A.xofs + D.w;
        A.ytab = reinterpret_cast<const int4*>(A.xofs + ((2 * D.w + 3) & ~3));
        A.src = l == 1 ? p.img0 : p.pyr + S.off;
        A.srcStride = l == 1 ? p.img0Stride : p.pyrStride;
        A.dst = p.pyr + D.off;
        A.dstStride = p.pyrStride;
        A.sw = S.w; A.sh = S.h; A.sp = S.pitch;
        A.dw = D.w; A.dh = D.h; A.dp = D.pitch;
        dim3 block(RS_TX, RS_TY), grid((D.w + 4 * RS_TX - 1) / (4 * RS_TX), (D.h + RS_ROWS * RS_TY - 1) / (RS_ROWS * RS_TY), p.imgs);
        if (S.w > 2 * D.w) orbx_resize_kernel<true><<<grid, block, 0, s>>>(A);   // ratio in (2, 7/3): scaleFactor <= 2 by contract
        else orbx_resize_kernel<false><<<grid, block, 0, s>>>(A);
    }
}
