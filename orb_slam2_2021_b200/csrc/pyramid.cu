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
#include "orbx_tma.cuh"

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
// CG: the source level was written earlier in the SAME kernel by other SMs (orbx_pyramid_chain_kernel): loads go to L2
// (ld.global.cg), the point of coherence, instead of through the non-coherent path.
// CG == 2: the source window sits in shared memory (orbx_resize_tma_kernel).
template <int CG>
__device__ __forceinline__ uint32_t rs_load(const uint32_t* p)
{
    if (CG == 2) {
        uint32_t v;
        asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(orbx_smem_addr(p)));
        return v;
    }
    return CG == 1 ? __ldcg(p) : __ldg(p);
}

// One destination tile of 128 columns x (4 x ROWS) rows, (bx, by), of image img by a (32, 4) thread block; ytabS is the
// block's staging of the row table.
template <bool WIDE, int CG, int ROWS>
__device__ __forceinline__ void resize_tile(const ResizeArgs& A, int img, int bx, int by, int4 (*ytabS)[ROWS])
{
    const int x0 = (bx * RS_TX + threadIdx.x) * 4;
    const int yBeg = (by * RS_TY + threadIdx.y) * ROWS, yEnd = min(yBeg + ROWS, A.dh);
    if (yBeg >= yEnd) return;
    // row table entries of this warp's destination rows, staged in shared memory: the emit step reads them with a
    // dependent compare right behind, which as global loads stalled every destination row
    if (threadIdx.x < ROWS) ytabS[threadIdx.y][threadIdx.x] = A.ytab[min(yBeg + (int)threadIdx.x, A.dh - 1)];
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
        a0 = rs_load<CG>(pa); a1 = rs_load<CG>(pa + 1); a2 = rs_load<CG>(pa + 2);
        b0 = rs_load<CG>(pb); b1 = rs_load<CG>(pb + 1); b2 = rs_load<CG>(pb + 2);
    }
    // one source row: prefetch row r+2 into (N0,N1,N2), interpolate row r from (W0,W1,W2) into HC, emit the destination
    // row whose second source row this is from HP (row r-1) and HC
#define RS_STEP(W0, W1, W2, N0, N1, N2, HP, HC)                                                                          \
    {                                                                                                                    \
        const uint32_t* pn = src + (size_t)min(r + 2, shm1) * spw;                                                       \
        N0 = rs_load<CG>(pn); N1 = rs_load<CG>(pn + 1); N2 = rs_load<CG>(pn + 2);                                        \
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

template <bool WIDE>
__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_resize_kernel(const ResizeArgs A)
{
    __shared__ int4 ytabS[RS_TY][RS_ROWS];
    resize_tile<WIDE, 0, RS_ROWS>(A, blockIdx.z, blockIdx.x, blockIdx.y, ytabS);
}

// ---- TMA-staged resize (the default of the per-level path; tuning key pyr_tma = 2 selects the word-load kernel above): the
// source window of a 128 x 64 destination tile arrives as ONE box load in shared memory (cp.async.bulk.tensor, mbarrier) and
// the same per-thread code reads its words from there (ld.shared): 0.445 against 0.468 ms per 512 KITTI images.
template <bool WIDE>
__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_resize_tma_kernel(const ResizeArgs A, const CUtensorMap* srcMap, int boxW, int boxH, int imgBase)
{
    extern __shared__ __align__(128) uint8_t rs_tile[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ int4 ytabS[RS_TY][ORBX_RS_TMA_ROWS];
    const int tid = threadIdx.y * RS_TX + threadIdx.x;
    const int x0d = blockIdx.x * (4 * RS_TX), y0d = blockIdx.y * (RS_TY * ORBX_RS_TMA_ROWS);
    const int xs = A.xofs[x0d] & ~15, r0 = A.ytab[y0d].x;  // window origin (the copy engine wants 16-byte column units)
    if (tid == 0) {
        orbx_mbar_init(&bar, 1);
        orbx_fence_mbar_init();
        orbx_mbar_expect_tx(&bar, boxW * boxH);
        orbx_tma_load_3d(rs_tile, srcMap, xs, r0, blockIdx.z + imgBase, &bar);
    }
    __syncthreads();
    orbx_mbar_wait(&bar, 0);
    // the tile code addresses the source as src + row * pitch + column: hand it the window under the coordinates of the level
    ResizeArgs B = A;
    B.src = rs_tile - ((size_t)r0 * boxW + xs);
    B.srcStride = 0;
    B.sp = boxW;
    resize_tile<WIDE, 2, ORBX_RS_TMA_ROWS>(B, blockIdx.z, blockIdx.x, blockIdx.y, ytabS);
}

// ---- the whole pyramid in ONE launch ---------------------------------------------------------------------------------
// For a handful of images (the single-frame path) seven dependent launches of a few dozen CTAs each are mostly launch
// and drain latency. Here every 128 x 16 tile of every level is a work item; CTAs draw items from an atomic ticket counter
// in an order in which an item only ever depends on items with SMALLER tickets (level-major inside a group of images), so
// whoever holds a ticket is running and a waiting CTA always waits for running ones: no deadlock, whatever the grid size.
// Readiness is tracked per (image, level, 16-row band): a band counter reaches the level's number of column tiles when the
// band is complete; a tile waits (one thread polls with acquire loads) for the source bands its rows come from. Writers
// fence before they count up, readers load the source through L2.
#define CH_ROWS 4                      // destination rows per thread in the chained kernel: short tiles, short dependency steps
#define CH_BAND (RS_TY * CH_ROWS)      // rows per tile = rows per readiness band
struct ChainLevel {
    ResizeArgs A;
    int tilesX, tilesY, tileBase, bandBase, wide;  // tileBase / bandBase: first tile / band of this level inside one image
};
struct ChainArgs {
    ChainLevel lv[ORBX_MAX_LEVELS];  // [1 .. nlevels-1] used
    int nlevels, imgs, tilesPerImg, bandsPerImg, group;
};

__global__ void __launch_bounds__(RS_TX * RS_TY) orbx_pyramid_chain_kernel(const __grid_constant__ ChainArgs C, int* state)
{
    __shared__ int4 ytabS[RS_TY][CH_ROWS];
    __shared__ int sh_ticket;
    const int tid = threadIdx.y * RS_TX + threadIdx.x;
    int* ticket = state;                  // state[0]: ticket counter, state[1]: watchdog flag, bands of image i from state[2 + i * bandsPerImg]
    int* ready = state + 2;
    const int total = C.imgs * C.tilesPerImg;
    for (;;) {
        if (tid == 0) sh_ticket = atomicAdd(ticket, 1);
        __syncthreads();
        const int t = sh_ticket;
        if (t >= total) return;
        // ticket -> (image, level, tile): groups of `group` images, level-major inside a group, image-minor inside a level
        const int perGroup = C.group * C.tilesPerImg;
        const int g = t / perGroup, r = t - g * perGroup;
        const int gn = min(C.group, C.imgs - g * C.group);  // images in this group
        int l = 1;
        while (l + 1 < C.nlevels && r >= gn * C.lv[l + 1].tileBase) ++l;
        const ChainLevel& L = C.lv[l];
        const int q = r - gn * L.tileBase, tile = q / gn, img = g * C.group + (q - tile * gn);
        const int by = tile / L.tilesX, bx = tile - by * L.tilesX;
        if (l > 1) {  // source rows of this tile's destination rows -> bands of level l-1 that must be complete
            if (tid == 0) {
                const ChainLevel& S = C.lv[l - 1];
                const int yB = by * CH_BAND, yE = min(yB + CH_BAND, L.A.dh) - 1;
                const int r0 = L.A.ytab[yB].x, r1 = min(L.A.ytab[yE].x + 1, L.A.sh - 1);
                const int b0 = r0 / CH_BAND, b1 = r1 / CH_BAND;
                const int* flags = ready + (size_t)img * C.bandsPerImg + S.bandBase;
                for (int b = b0; b <= b1; ++b) {
                    unsigned spins = 0;
                    for (;;) {
                        int v;
                        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flags + b) : "memory");
                        if (v >= S.tilesX) break;
                        if (++spins > (1u << 24)) { atomicExch(state + 1, 1); break; }  // watchdog: never hang the device
                        __nanosleep(64);
                    }
                }
            }
            __syncthreads();
        }
        if (L.wide) resize_tile<true, 1, CH_ROWS>(L.A, img, bx, by, ytabS);
        else resize_tile<false, 1, CH_ROWS>(L.A, img, bx, by, ytabS);
        __threadfence();   // this thread's pixels before the band counter
        __syncthreads();   // (also: everybody is done with ytabS and sh_ticket)
        if (tid == 0) atomicAdd(ready + (size_t)img * C.bandsPerImg + L.bandBase + by, 1);
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

static ResizeArgs resize_args(const OrbxParams& p, const int* const* tabs, int l)
{
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
    return A;
}

int orbx_pyramid_chain_state_ints(const OrbxParams& p)  // per image: its band counters; + 2 per launch (see the kernel)
{
    int bands = 0;
    for (int l = 1; l < p.nlevels; ++l) bands += (p.lv[l].h + CH_BAND - 1) / CH_BAND;
    return bands + 2;
}

// The whole pyramid of p.imgs images in one launch; `state` = this launch's private counters (orbx_pyramid_chain_state_ints
// per image), cleared here.
void orbx_launch_pyramid_chain(const OrbxParams& p, const int* const* tabs, int* state, cudaStream_t s)
{
    if (p.nlevels < 2) return;
    ChainArgs C{};
    int tiles = 0, bands = 0;
    for (int l = 1; l < p.nlevels; ++l) {
        ChainLevel& L = C.lv[l];
        L.A = resize_args(p, tabs, l);
        L.tilesX = (p.lv[l].w + 4 * RS_TX - 1) / (4 * RS_TX);
        L.tilesY = (p.lv[l].h + CH_BAND - 1) / CH_BAND;
        L.tileBase = tiles; L.bandBase = bands;
        L.wide = p.lv[l - 1].w > 2 * p.lv[l].w;
        tiles += L.tilesX * L.tilesY;
        bands += L.tilesY;
    }
    C.nlevels = p.nlevels; C.imgs = p.imgs; C.tilesPerImg = tiles; C.bandsPerImg = bands; C.group = 64;
    cudaMemsetAsync(state, 0, sizeof(int) * ((size_t)p.imgs * bands + 2), s);
    static int residentCtas[64] = {0};  // per device ordinal
    int dev = 0;
    cudaGetDevice(&dev);
    int& cap = residentCtas[dev & 63];
    if (!cap) {
        int perSm = 1, sms = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, orbx_pyramid_chain_kernel, RS_TX * RS_TY, 0);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cap = (perSm > 0 ? perSm : 1) * sms;
    }
    const long long total = (long long)p.imgs * tiles;
    orbx_pyramid_chain_kernel<<<(int)(total < cap ? total : cap), dim3(RS_TX, RS_TY), 0, s>>>(C, state);
}

void orbx_launch_pyramid_tma(const OrbxParams& p, const int* const* tabs, const void* srcMaps, const int* boxW, const int* boxH, cudaStream_t s)
{
    for (int l = 1; l < p.nlevels; ++l) {
        const OrbxLevel &S = p.lv[l - 1], &D = p.lv[l];
        const ResizeArgs A = resize_args(p, tabs, l);
        dim3 block(RS_TX, RS_TY), grid((D.w + 4 * RS_TX - 1) / (4 * RS_TX), (D.h + ORBX_RS_TMA_ROWS * RS_TY - 1) / (ORBX_RS_TMA_ROWS * RS_TY), p.imgs);
        dim3 gridW((D.w + 4 * RS_TX - 1) / (4 * RS_TX), (D.h + RS_ROWS * RS_TY - 1) / (RS_ROWS * RS_TY), p.imgs);
        const bool wide = S.w > 2 * D.w;
        if (boxW[l - 1] <= 0) {  // window too large for one box: the word-load kernel
            if (wide) orbx_resize_kernel<true><<<gridW, block, 0, s>>>(A);
            else orbx_resize_kernel<false><<<gridW, block, 0, s>>>(A);
            continue;
        }
        // (+2 rows: the row walk prefetches two rows ahead and drops what it does not need)
        const size_t smem = (size_t)boxW[l - 1] * (boxH[l - 1] + 2) + 16;
        const CUtensorMap* map = reinterpret_cast<const CUtensorMap*>(srcMaps) + (l - 1);
        if (wide) {
            if (smem > 32 * 1024) cudaFuncSetAttribute(orbx_resize_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            orbx_resize_tma_kernel<true><<<grid, block, smem, s>>>(A, map, boxW[l - 1], boxH[l - 1], p.imgBase);
        } else {
            if (smem > 32 * 1024) cudaFuncSetAttribute(orbx_resize_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            orbx_resize_tma_kernel<false><<<grid, block, smem, s>>>(A, map, boxW[l - 1], boxH[l - 1], p.imgBase);
        }
    }
}

// tabs[l] points at the level's table block: xofs[w] xcoef[w], then (16-byte aligned) ytab[h] as int4
void orbx_launch_pyramid(const OrbxParams& p, const int* const* tabs, cudaStream_t s)
{
    for (int l = 1; l < p.nlevels; ++l) {
        const OrbxLevel &S = p.lv[l - 1], &D = p.lv[l];
        const ResizeArgs A = resize_args(p, tabs, l);
        dim3 block(RS_TX, RS_TY), grid((D.w + 4 * RS_TX - 1) / (4 * RS_TX), (D.h + RS_ROWS * RS_TY - 1) / (RS_ROWS * RS_TY), p.imgs);
        if (S.w > 2 * D.w) orbx_resize_kernel<true><<<grid, block, 0, s>>>(A);   // ratio in (2, 7/3): scaleFactor <= 2 by contract
        else orbx_resize_kernel<false><<<grid, block, 0, s>>>(A);
    }
}
