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

struct ResizeTab {
    const int* xofs;   // [dw] source column
    const int* xcoef;  // [dw] c0 | c1<<16
    const int* yofs;   // [dh] source row (unclipped)
    const int* ycoef;  // [dh] b0 | b1<<16
};

#define RS_TX 32   // threads along x (4 columns each)
#define RS_TY 4    // row chunks per CTA
#define RS_ROWS 16  // destination rows per thread

// A thread owns 4 adjacent destination columns and RS_ROWS destination rows. It walks DOWN THE SOURCE ROWS that
// those destination rows need (consecutive, because every destination row uses two adjacent source rows and the
// row step is below 2): each source row is interpolated horizontally exactly once and a destination row is emitted
// when its second source row arrives. Per source row the thread loads three aligned 32-bit words (they cover
// the <= 11 bytes its four column pairs can span for scale factors up to 2), realigns them with two funnel
// shifts, picks the four byte pairs with two PRMT and gets the four horizontal taps from four IDP.2A
// (16-bit coefficient pair x byte pair, on the FMA pipe). Loads are non-coherent (__ldg) and issued two source
// rows ahead of their use, so their latency overlaps the arithmetic of the rows before.
__device__ __forceinline__ void rs_load(const uint8_t* row, int aw, uint32_t& w0, uint32_t& w1, uint32_t& w2)
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(row) + aw;
    w0 = __ldg(p); w1 = __ldg(p + 1); w2 = __ldg(p + 2);
}

__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_resize_kernel(OrbxParams p, int l, ResizeTab t)
{
    const OrbxLevel& D = p.lv[l];
    const int img = blockIdx.z;
    const int x0 = (blockIdx.x * RS_TX + threadIdx.x) * 4;
    const int yBeg = (blockIdx.y * RS_TY + threadIdx.y) * RS_ROWS, yEnd = min(yBeg + RS_ROWS, D.h);
    if (yBeg >= yEnd) return;
    // row tables of this warp's RS_ROWS destination rows, staged in shared memory: the emit loop below reads them with a
    // dependent compare right behind, which as global loads stalled every destination row
    __shared__ int ytab[RS_TY][2 * RS_ROWS];
    for (int e = threadIdx.x; e < 2 * RS_ROWS; e += RS_TX) {
        const int k = e & (RS_ROWS - 1), yy = min(yBeg + k, D.h - 1);
        ytab[threadIdx.y][e] = e < RS_ROWS ? t.yofs[yy] : t.ycoef[yy];
    }
    __syncwarp();
    const int* yofsS = ytab[threadIdx.y];
    const int* ycoefS = yofsS + RS_ROWS;
    if (x0 >= D.w) return;
    int sp;
    const uint8_t* src = orbx_level_ptr(p, img, l - 1, &sp);
    const int sw = p.lv[l - 1].w, sh = p.lv[l - 1].h;
    // column geometry: aligned word index of the first source column, byte shift, PRMT selectors, packed coefficients
    uint32_t coef[4];
    int d[4], sx0 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = min(x0 + i, D.w - 1);  // columns past the level land in the row padding; keep their reads in range
        const int sx = t.xofs[x];
        if (i == 0) sx0 = sx;
        d[i] = sx - sx0;                      // 0 .. 6 for scale factors up to 2 (checked by the launcher)
        // the right neighbour is clamped to the last column (there c1 == 0 anyway, OpenCV sets f = 0 for sx >= sw-1):
        // selecting the same byte twice keeps the read inside the row
        const int d1 = sx + 1 <= sw - 1 ? d[i] + 1 : d[i];
        d[i] |= d1 << 4;
        coef[i] = (uint32_t)t.xcoef[x];       // c0 | c1 << 16
    }
    // words are read from the aligned address below sx0, but never past the last whole word of the source row pitch
    const int aw = min(sx0 >> 2, (sp >> 2) - 3);
    const int sh8 = (sx0 - 4 * aw) * 8;       // 0..24 normally; larger only in the clamped last words of a row
    const uint32_t selA = (uint32_t)(d[0] & 15) | ((uint32_t)(d[0] >> 4) << 4) | ((uint32_t)(d[1] & 15) << 8) | ((uint32_t)(d[1] >> 4) << 12);
    const uint32_t selB = (uint32_t)(d[2] & 15) | ((uint32_t)(d[2] >> 4) << 4) | ((uint32_t)(d[3] & 15) << 8) | ((uint32_t)(d[3] >> 4) << 12);
    uint8_t* dst = p.pyr + (size_t)img * p.pyrStride + D.off + (size_t)yBeg * D.pitch + x0;
    const int rBeg = min(max(yofsS[0], 0), sh - 1), rEnd = min(max(yofsS[yEnd - 1 - yBeg] + 1, 0), sh - 1);
    int y = yBeg, yo = yofsS[0], yc = ycoefS[0];
    int sy0 = min(max(yo, 0), sh - 1), sy1 = min(max(yo + 1, 0), sh - 1);
    uint32_t a0, a1, a2, b0, b1, b2;  // source rows r and r+1 in flight
    rs_load(src + (size_t)rBeg * sp, aw, a0, a1, a2);
    rs_load(src + (size_t)min(rBeg + 1, rEnd) * sp, aw, b0, b1, b2);
    uint32_t hPrev[4] = {0, 0, 0, 0};
#pragma unroll 1
    for (int r = rBeg; r <= rEnd; ++r) {
        uint32_t c0w, c1w, c2w;  // row r+2
        rs_load(src + (size_t)min(r + 2, rEnd) * sp, aw, c0w, c1w, c2w);
        // bytes sx0 .. sx0+7 of row r
        uint32_t lo, hi;
        if (sh8 < 32) { lo = __funnelshift_r(a0, a1, sh8); hi = __funnelshift_r(a1, a2, sh8); }
        else if (sh8 < 64) { lo = __funnelshift_r(a1, a2, sh8 - 32); hi = a2 >> (sh8 - 32); }
        else { lo = a2 >> (sh8 - 64); hi = 0; }  // (only in the clamped last words of a row)
        const uint32_t pA = __byte_perm(lo, hi, selA), pB = __byte_perm(lo, hi, selB);
        uint32_t hCur[4];
        hCur[0] = __dp2a_lo(coef[0], pA, 0u);
        hCur[1] = __dp2a_hi(coef[1], pA, 0u);
        hCur[2] = __dp2a_lo(coef[2], pB, 0u);
        hCur[3] = __dp2a_hi(coef[3], pB, 0u);
        while (y < yEnd && sy1 == r) {  // (at most one destination row per source row for scale factors > 1, two at the clamped edges)
            const uint32_t bb0 = (uint32_t)(yc & 0xFFFF) << 12, bb1 = (uint32_t)(yc >> 16) << 12;
            uint32_t v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t h0 = sy0 == sy1 ? hCur[i] : hPrev[i];
                // (b * (h >> 4)) >> 16  ==  mulhi(b << 12, h & ~15): h - (h & 15) = 16 (h >> 4), and b << 12 < 2^24
                v[i] = (__umulhi(bb0, h0 & ~15u) + __umulhi(bb1, hCur[i] & ~15u) + 2u) >> 2;  // in [0,255]: b0 + b1 = 2048
            }
            *reinterpret_cast<uint32_t*>(dst) = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
            dst += D.pitch;  // pitch is a multiple of 16: in-row padding absorbs the tail
            if (++y < yEnd) {
                yo = yofsS[y - yBeg]; yc = ycoefS[y - yBeg];
                sy0 = min(max(yo, 0), sh - 1); sy1 = min(max(yo + 1, 0), sh - 1);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) hPrev[i] = hCur[i];
        a0 = b0; a1 = b1; a2 = b2;
        b0 = c0w; b1 = c1w; b2 = c2w;
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
