// fast.cu — per-cell FAST-9/16 with non-max suppression and the per-cell threshold fallback.
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (/root/reference/src/ORBextractor.cc:792-832) and the two cv::FAST calls inside it (:812, :817).
//
// One CTA per 30-px cell (all levels and all images in one launch). Every interior pixel of the
// cell's ROI gets its corner strength
//     m = max over the 16 arcs of 9 contiguous ring pixels of min(+-(ring - centre))
// which is threshold independent: the pixel is a FAST corner at threshold t iff m > t and OpenCV's
// score is m-1. NMS keeps a pixel iff m is strictly greater than the strengths of its 8 neighbours
// (pixels outside the ROI interior count 0). Because a neighbour that is not a corner at t has
// m <= t < m_centre, the kept set at threshold t is {NMS maxima of the minTh map with m > t}: one
// pass yields both thresholds, and the cell falls back to minTh iff no maximum exceeds iniTh —
// exactly "FAST(ini) came back empty" (:814). Keypoints are written row-major into the cell's slot
// array, so cells concatenated in table order give the reference's candidate order.
//
// Blackwell mapping (measured with tools/ubench/pipes.cu: VIMNMX3.U16x2 issues at the same rate as a
// 32-bit IMNMX, so packing two pixels per register halves the ALU-pipe work that bounds this kernel):
//   * the ROI is staged in shared memory as 16-bit pixels, twice: A[y][x] = I(x,y) and
//     B[y][x] = I(x+1,y). A thread owns two horizontally adjacent pixels; for every ring offset the
//     pair (ring(x), ring(x+1)) is ONE aligned 32-bit LDS (from A for odd dx, from B for even dx)
//     that is already a u16x2 vector — no byte extraction;
//   * d'_k = ring_k + 256 - centre stays in [1,511] per lane, so the whole min/max network runs on
//     unsigned u16x2 with the 3-input DPX instructions (__vimin3_u16x2 / __vimax3_u16x2):
//     16+16 ops for the 9-windows and 8 for the reduction, per polarity, per pixel PAIR;
//   * lanes are laid out 8 pairs x 4 rows with a row stride of 24 (or 40) words, which makes every
//     ring load bank-conflict free;
//   * the ROI arrives by ONE TMA box load (cp.async.bulk.tensor.3d, 64 or 96 columns starting at x0 & ~15 —
//     the copy engine addresses the innermost dimension in 16-byte units — by rhMax rows) signalled on an
//     mbarrier; a short shared-to-shared pass widens it into A/B (the tile keeps the x0&3 offset and the
//     A/B roles swap with its parity); the strength map is kept as u16 so that NMS runs four pixels per
//     thread on the same packed max instructions.
#include "orbx_dev.cuh"
#include "orbx_tma.cuh"

struct OrbxCell {
    short x0, y0;        // ROI origin in level coordinates (iniX, iniY)
    uint8_t rw, rh;      // ROI size (maxX-iniX, maxY-iniY)
    uint8_t level, pad;
};

#define FAST_THREADS 128
#define FAST_WARPS (FAST_THREADS / 32)
#define FAST_MAX_ITERS 8   // 128-quad steps of one cell interior (<= 60x60 px = 900 quads)

__device__ __forceinline__ int imax3(int a, int b, int c) { return max(a, max(b, c)); }

// Corner strengths of the two pixels whose centre pair is *(u32*)b (see the layout note above).
// a points at A[row][c] with c odd, b at B[row][c-1]; SA = row stride in u16 elements.
template <int SA>
__device__ __forceinline__ void fast_strength2(const uint16_t* a, const uint16_t* b, int& m0, int& m1)
{
#define RING_A(dx, dy) (*reinterpret_cast<const uint32_t*>(a + (dy) * SA + (dx)))
#define RING_B(dx, dy) (*reinterpret_cast<const uint32_t*>(b + (dy) * SA + (dx)))
    const uint32_t nc = 0x01000100u - RING_B(0, 0);  // 256 - centre per lane, in [1,256]: no borrow between lanes
    uint32_t d[16];
    d[0] = RING_B(0, 3) + nc;    d[1] = RING_A(1, 3) + nc;    d[2] = RING_B(2, 2) + nc;    d[3] = RING_A(3, 1) + nc;
    d[4] = RING_A(3, 0) + nc;    d[5] = RING_A(3, -1) + nc;   d[6] = RING_B(2, -2) + nc;   d[7] = RING_A(1, -3) + nc;
    d[8] = RING_B(0, -3) + nc;   d[9] = RING_A(-1, -3) + nc;  d[10] = RING_B(-2, -2) + nc; d[11] = RING_A(-3, -1) + nc;
    d[12] = RING_A(-3, 0) + nc;  d[13] = RING_A(-3, 1) + nc;  d[14] = RING_B(-2, 2) + nc;  d[15] = RING_A(-1, 3) + nc;
#undef RING_A
#undef RING_B
    uint32_t lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = __vimin3_u16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        hi3[k] = __vimax3_u16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    uint32_t lo9[16], hi9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo9[k] = __vimin3_u16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
        hi9[k] = __vimax3_u16x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    }
    uint32_t br = __vimax3_u16x2(lo9[0], lo9[1], lo9[2]), dk = __vimin3_u16x2(hi9[0], hi9[1], hi9[2]);
#pragma unroll
    for (int k = 3; k < 15; k += 2) {
        br = __vimax3_u16x2(br, lo9[k], lo9[k + 1]);
        dk = __vimin3_u16x2(dk, hi9[k], hi9[k + 1]);
    }
    br = __vimax3_u16x2(br, lo9[15], lo9[15]);
    dk = __vimin3_u16x2(dk, hi9[15], hi9[15]);
    // bright = br - 256, dark = dk - 256 (per lane); strength = max(bright, -dark)
    m0 = max((int)(br & 0xFFFFu) - 256, 256 - (int)(dk & 0xFFFFu));
    m1 = max((int)(br >> 16) - 256, 256 - (int)(dk >> 16));
}

// One interior pixel quad (4 consecutive pixels of one row) per thread and step: bit k of the result
// = pixel k is a strict 3x3 maximum of the strength map; `ini` gets the subset above iniTh.
// s points at smap[row y][col 4q] (u16, 8-byte aligned); MS = row stride in elements.
template <int MS>
__device__ __forceinline__ unsigned nms_quad(const uint16_t* s, int tini, unsigned& ini)
{
    const uint2 r0a = *reinterpret_cast<const uint2*>(s), r0b = *reinterpret_cast<const uint2*>(s + 4);
    const uint2 r1a = *reinterpret_cast<const uint2*>(s + MS), r1b = *reinterpret_cast<const uint2*>(s + MS + 4);
    const uint2 r2a = *reinterpret_cast<const uint2*>(s + 2 * MS), r2b = *reinterpret_cast<const uint2*>(s + 2 * MS + 4);
    ini = 0;
    if ((r1a.y | r1b.x) == 0) return 0;  // no corner among the four centres
    const uint32_t cm0 = __vimax3_u16x2(r0a.x, r1a.x, r2a.x), cm1 = __vimax3_u16x2(r0a.y, r1a.y, r2a.y);
    const uint32_t cm2 = __vimax3_u16x2(r0b.x, r1b.x, r2b.x), cm3 = __vimax3_u16x2(r0b.y, r1b.y, r2b.y);
    const uint32_t L1 = __funnelshift_r(cm0, cm1, 16), R1 = __funnelshift_r(cm1, cm2, 16), R2 = __funnelshift_r(cm2, cm3, 16);
    const uint32_t ud1 = __vimax3_u16x2(r0a.y, r2a.y, r2a.y), ud2 = __vimax3_u16x2(r0b.x, r2b.x, r2b.x);
    const uint32_t nb1 = __vimax3_u16x2(L1, R1, ud1), nb2 = __vimax3_u16x2(R1, R2, ud2);
    const int m0 = r1a.y & 0xFFFFu, m1 = r1a.y >> 16, m2 = r1b.x & 0xFFFFu, m3 = r1b.x >> 16;
    const int n0 = nb1 & 0xFFFFu, n1 = nb1 >> 16, n2 = nb2 & 0xFFFFu, n3 = nb2 >> 16;
    const unsigned mx = (m0 > n0 ? 1u : 0u) | (m1 > n1 ? 2u : 0u) | (m2 > n2 ? 4u : 0u) | (m3 > n3 ? 8u : 0u);
    ini = mx & ((m0 > tini ? 1u : 0u) | (m1 > tini ? 2u : 0u) | (m2 > tini ? 4u : 0u) | (m3 > tini ? 8u : 0u));
    return mx;
}

template <int SA>
__global__ void __launch_bounds__(FAST_THREADS) orbx_fast_kernel(OrbxParams p, const OrbxCell* cells, const CUtensorMap* maps, int rhMax)
{
    constexpr int BW = SA == 48 ? 64 : 96;      // TMA box width in bytes
    constexpr int WSLOT = SA == 48 ? 16 : 32;   // word slots per row in the widening pass
    extern __shared__ __align__(128) uint8_t fast_smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ int listCnt;

    uint8_t* T = fast_smem;                                                     // raw box [rhMax][BW]
    uint16_t* A = reinterpret_cast<uint16_t*>(fast_smem + ((rhMax * BW + 127) & ~127));  // A[y][j] = I(xa + j, y0 + y), xa = x0 & ~3
    uint16_t* B = A + rhMax * SA;                                                // B[y][j] = I(xa + j + 1, y0 + y)
    uint16_t* smap = B + rhMax * SA;  // strength map: interior pixel (ix,iy) at [iy+1][ix+2], zero elsewhere

    const int cellId = blockIdx.x, img = blockIdx.y;
    const OrbxCell cell = cells[cellId];
    const OrbxLevel& L = p.lv[cell.level];
    const int rw = cell.rw, rh = cell.rh;
    const int iw = rw - 6, ih = rh - 6;  // interior = detection domain of cv::FAST
    int* cntOut = p.cellCnt + (size_t)img * p.cellsTotal + cellId;
    if (iw <= 0 || ih <= 0) {
        if (threadIdx.x == 0) *cntOut = 0;
        return;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        listCnt = 0;
        orbx_mbar_init(&bar, 1);
        orbx_fence_mbar_init();
        orbx_mbar_expect_tx(&bar, rhMax * BW);
        orbx_tma_load_3d(T, maps + cell.level, cell.x0 & ~15, cell.y0, img + p.imgBase, &bar);
    }
    {
        uint4* z = reinterpret_cast<uint4*>(smap);  // 16-byte aligned: 4*rhMax*SA bytes past A, SA % 8 == 0
        const int nz = (ih + 2) * (SA / 8);
        for (int i = threadIdx.x; i < nz; i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();  // barrier initialised (and armed) before anybody polls it
    orbx_mbar_wait(&bar, 0);

    // ---- widen the raw bytes into the two u16 pair layouts
    const int o = cell.x0 & 3;
    {
        const int nw = (o + rw + 4) >> 2;  // covers tile columns 0 .. o+rw (one past the ROI, read by odd-width pairs)
        const uint32_t* T32 = reinterpret_cast<const uint32_t*>(T) + ((cell.x0 & 15) >> 2);
        const int total = rh * WSLOT;
#pragma unroll 1
        for (int i = threadIdx.x; i < total; i += FAST_THREADS) {
            const int y = i / WSLOT, j = i % WSLOT;
            if (j < nw) {
                const uint32_t w = T32[y * (BW / 4) + j];
                const uint32_t wn = T32[y * (BW / 4) + j + 1];  // inside the box: (x0&15)/4 + nw + 1 <= BW/4
                const uint32_t t = __funnelshift_r(w, wn, 8);
                *reinterpret_cast<uint2*>(A + y * SA + 4 * j) = make_uint2(__byte_perm(w, 0, 0x4140), __byte_perm(w, 0, 0x4342));
                *reinterpret_cast<uint2*>(B + y * SA + 4 * j) = make_uint2(__byte_perm(t, 0, 0x4140), __byte_perm(t, 0, 0x4342));
            }
        }
    }
    __syncthreads();

    // ---- strength map, phase A: cheap necessary test on every pixel pair. Any 9 contiguous ring positions contain two
    // ADJACENT compass points (0,4,8,12), so a corner at minTh needs (p0|p8)&(p4|p12) for "brighter" or for "darker".
    // Lane = (pair lx, row ly) in an 8x4 patch (conflict-free LDS). Survivors (~29 % of the pairs on the synthetic
    // scenes, far fewer on low-texture images) go to a WARP-PRIVATE task list in the (now free) raw tile: no atomics,
    // and phases A and B of a warp need no CTA barrier between them.
    const int tmin = p.minTh, tini = p.iniTh;
    // centre tile column of pair xp is c = o + 3 + 2*xp. The array whose pair at (c + dx) is word aligned serves
    // dx: o even -> c odd: odd dx from A, even dx from B (at index c-1+dx); o odd -> c even: roles swap.
    const uint16_t* podd = (o & 1) ? B + (o + 2) : A + (o + 3);
    const uint16_t* pevn = (o & 1) ? A + (o + 3) : B + (o + 2);
    const int npairs = (iw + 1) >> 1;
    uint16_t* list = reinterpret_cast<uint16_t*>(T) + warp * (4 * ((ih + 15) >> 4) * npairs);
    int ntask = 0;  // warp-uniform
    {
        const int lx = lane & 7, ly = lane >> 3;
        const uint32_t cb = 0x7FFF7FFFu - 0x01000100u - 0x00010001u * (uint32_t)tmin;  // d' + cb has bit 15 set iff d' > 256 + t
        const uint32_t cd = 0x80008000u + 0x00FF00FFu - 0x00010001u * (uint32_t)tmin;  // cd - d' has bit 15 set iff d' < 256 - t
        const unsigned lt = (1u << lane) - 1;
#pragma unroll 1
        for (int y0 = warp * 4; y0 < ih; y0 += 4 * FAST_WARPS) {
            const int y = y0 + ly;
#pragma unroll 1
            for (int x0 = 0; x0 < npairs; x0 += 8) {
                const int xp = x0 + lx;
                bool keep = false;
                if (y < ih && xp < npairs) {
                    const uint16_t* a = podd + (y + 3) * SA + 2 * xp;
                    const uint16_t* b = pevn + (y + 3) * SA + 2 * xp;
                    const uint32_t nc = 0x01000100u - *reinterpret_cast<const uint32_t*>(b);
                    const uint32_t d0 = *reinterpret_cast<const uint32_t*>(b + 3 * SA) + nc, d8 = *reinterpret_cast<const uint32_t*>(b - 3 * SA) + nc;
                    const uint32_t d4 = *reinterpret_cast<const uint32_t*>(a + 3) + nc, d12 = *reinterpret_cast<const uint32_t*>(a - 3) + nc;
                    const uint32_t br = ((d0 + cb) | (d8 + cb)) & ((d4 + cb) | (d12 + cb));
                    const uint32_t dk = ((cd - d0) | (cd - d8)) & ((cd - d4) | (cd - d12));
                    keep = ((br | dk) & 0x80008000u) != 0;
                }
                const unsigned bal = __ballot_sync(0xffffffffu, keep);
                if (keep) list[ntask + __popc(bal & lt)] = (uint16_t)((y << 8) | xp);
                ntask += __popc(bal);
            }
        }
    }
    __syncwarp();
    // ---- phase B: the min/max network on the warp's surviving pairs (results go to the map by position)
#pragma unroll 1
    for (int i = lane; i < ntask; i += 32) {
        const int task = list[i], y = task >> 8, xp = task & 255;
        int m0, m1;
        fast_strength2<SA>(podd + (y + 3) * SA + 2 * xp, pevn + (y + 3) * SA + 2 * xp, m0, m1);
        m0 = m0 > tmin ? m0 : 0;
        m1 = (m1 > tmin && 2 * xp + 1 < iw) ? m1 : 0;
        *reinterpret_cast<uint32_t*>(smap + (y + 1) * SA + 2 + 2 * xp) = (uint32_t)m0 | ((uint32_t)m1 << 16);
    }
    __syncthreads();

    // ---- NMS, driven by the same task lists (only listed pairs can hold a corner): strict maximum over the 8
    // neighbours in the strength map. Maxima go to a CTA list kept[] = key | m << 16, key = y*64 + x (interior coords).
    uint32_t* kept = reinterpret_cast<uint32_t*>(A);  // A/B are dead after phase B
    bool anyIni = false;
#pragma unroll 1
    for (int i = lane; i < ntask; i += 32) {
        const int task = list[i], y = task >> 8, xp = task & 255;
        const uint16_t* c = smap + (y + 1) * SA + 2 + 2 * xp;  // centre pair, word aligned
        const uint32_t mm = *reinterpret_cast<const uint32_t*>(c);
        if (mm == 0) continue;
        const uint32_t u0 = *reinterpret_cast<const uint32_t*>(c - SA - 2), u1 = *reinterpret_cast<const uint32_t*>(c - SA), u2 = *reinterpret_cast<const uint32_t*>(c - SA + 2);
        const uint32_t l0 = *reinterpret_cast<const uint32_t*>(c - 2), l2 = *reinterpret_cast<const uint32_t*>(c + 2);
        const uint32_t d0 = *reinterpret_cast<const uint32_t*>(c + SA - 2), d1 = *reinterpret_cast<const uint32_t*>(c + SA), d2 = *reinterpret_cast<const uint32_t*>(c + SA + 2);
        // column maxima over the three rows, as pairs (x-2,x-1) (x,x+1) (x+2,x+3)
        const uint32_t cm0 = __vimax3_u16x2(u0, l0, d0), cm1 = __vimax3_u16x2(u1, mm, d1), cm2 = __vimax3_u16x2(u2, l2, d2);
        const uint32_t ud = __vimax3_u16x2(u1, d1, d1);
        // neighbours of pixel 0 (x): columns x-1, x+1 (all rows) and rows above/below in column x; pixel 1 likewise
        const int m0 = mm & 0xFFFFu, m1 = mm >> 16;
        const int n0 = max(max((int)(cm0 >> 16), (int)(cm1 >> 16)), (int)(ud & 0xFFFFu));
        const int n1 = max(max((int)(cm1 & 0xFFFFu), (int)(cm2 & 0xFFFFu)), (int)(ud >> 16));
        if (m0 > n0) {
            kept[atomicAdd(&listCnt, 1)] = (uint32_t)(y * 64 + 2 * xp) | ((uint32_t)m0 << 16);
            anyIni |= m0 > tini;
        }
        if (m1 > n1) {
            kept[atomicAdd(&listCnt, 1)] = (uint32_t)(y * 64 + 2 * xp + 1) | ((uint32_t)m1 << 16);
            anyIni |= m1 > tini;
        }
    }
    const bool useIni = __syncthreads_or(anyIni);  // :814 fallback to minTh iff FAST(ini) is empty; also publishes kept[]

    // ---- ordered emission: rank of a selected maximum = number of selected maxima with a smaller row-major key
    const int nkept = listCnt;
    const int thr = useIni ? tini : 0;
    const uint32_t slot0 = L.slotBase + (uint32_t)(cellId - L.cellBase) * L.slotCap;
    uint32_t* out = p.cellKp + (size_t)img * p.slotTotal + slot0;
#pragma unroll 1
    for (int i = threadIdx.x; i < nkept || i == 0; i += FAST_THREADS) {
        const uint32_t e = i < nkept ? kept[i] : 0u;
        const int key = e & 0xFFFFu, m = e >> 16;
        int rank = 0, total = 0;
        for (int j = 0; j < nkept; ++j) {
            const uint32_t f = kept[j];
            const bool sel = (int)(f >> 16) > thr;
            total += sel;
            rank += sel && (int)(f & 0xFFFFu) < key;
        }
        if (i < nkept && m > thr) out[rank] = (uint32_t)((key & 63) + 3) | ((uint32_t)((key >> 6) + 3) << 8) | ((uint32_t)(m - 1) << 16);
        if (i == 0) *cntOut = total;
    }
}

void orbx_fast_geometry(const OrbxParams& p, int* boxW, int* rhMaxOut)
{
    // largest ROI over the levels (ROI = cell + 6, src/ORBextractor.cc:797-808)
    int rwMax = 0, rhMax = 0;
    for (int l = 0; l < p.nlevels; ++l) {
        rwMax = p.lv[l].wCell + 6 > rwMax ? p.lv[l].wCell + 6 : rwMax;
        rhMax = p.lv[l].hCell + 6 > rhMax ? p.lv[l].hCell + 6 : rhMax;
    }
    // tile columns: 3 (alignment) + ROI + 1, rounded up to whole words; strength-map columns: interior + 8
    *boxW = rwMax + 8 <= 48 ? 64 : 96;
    *rhMaxOut = rhMax;
}

void orbx_launch_fast(const OrbxParams& p, const void* cellTable, const void* levelMaps, cudaStream_t s)
{
    int boxW, rhMax;
    orbx_fast_geometry(p, &boxW, &rhMax);
    dim3 grid(p.cellsTotal, p.imgs);
    const int sa = boxW == 64 ? 48 : 80;
    const size_t smem = (size_t)((rhMax * boxW + 127) & ~127) + (size_t)2 * rhMax * sa * 2 + (size_t)(rhMax - 4) * sa * 2;
    if (sa == 48)
        orbx_fast_kernel<48><<<grid, FAST_THREADS, smem, s>>>(p, (const OrbxCell*)cellTable, (const CUtensorMap*)levelMaps, rhMax);
    else
        orbx_fast_kernel<80><<<grid, FAST_THREADS, smem, s>>>(p, (const OrbxCell*)cellTable, (const CUtensorMap*)levelMaps, rhMax);
}
