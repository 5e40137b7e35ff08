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
//     ring load bank-conflict free.
#include "orbx_dev.cuh"

struct OrbxCell {
    short x0, y0;        // ROI origin in level coordinates (iniX, iniY)
    uint8_t rw, rh;      // ROI size (maxX-iniX, maxY-iniY)
    uint8_t level, pad;
};

#define FAST_THREADS 128
#define FAST_WARPS (FAST_THREADS / 32)
#define FAST_MAX_ITERS 32  // 128-pixel chunks of one cell interior (<= 60x60 px)

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

template <int SA>
__global__ void __launch_bounds__(FAST_THREADS) orbx_fast_kernel(OrbxParams p, const OrbxCell* cells, int rhMax)
{
    extern __shared__ __align__(16) uint8_t fast_smem[];
    __shared__ int warpCnt[FAST_MAX_ITERS][FAST_WARPS];

    uint16_t* A = reinterpret_cast<uint16_t*>(fast_smem);
    uint16_t* B = A + rhMax * SA;
    uint8_t* smap = reinterpret_cast<uint8_t*>(B + rhMax * SA);  // strength map with a 1-px zero frame

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
    int pitch;
    const uint8_t* lvl = orbx_level_ptr(p, img, cell.level, &pitch);
    const uint8_t* src = lvl + (size_t)cell.y0 * pitch + cell.x0;
    for (int y = warp; y < rh; y += FAST_WARPS) {
        const uint8_t* row = src + (size_t)y * pitch;
        for (int x = lane; x < rw; x += 32) {
            const uint16_t v = row[x];
            A[y * SA + x] = v;
            if (x) B[y * SA + x - 1] = v;
        }
    }
    const int mw = iw + 2;
    for (int i = threadIdx.x; i < mw * (ih + 2); i += FAST_THREADS) smap[i] = 0;
    __syncthreads();

    // ---- strength map: lane = (pair lx, row ly) in an 8x4 patch; warp w owns rows 4w.. (+16 per step)
    const int tmin = p.minTh, tini = p.iniTh;
    {
        const int lx = lane & 7, ly = lane >> 3;
        const int npairs = (iw + 1) >> 1;
        for (int y = warp * 4 + ly; y < ih; y += 4 * FAST_WARPS) {
            const uint16_t* arow = A + (y + 3) * SA + 3;
            const uint16_t* brow = B + (y + 3) * SA + 2;
            uint8_t* srow = smap + (y + 1) * mw + 1;
            for (int xp = lx; xp < npairs; xp += 8) {
                int m0, m1;
                fast_strength2<SA>(arow + 2 * xp, brow + 2 * xp, m0, m1);
                if (m0 > tmin) srow[2 * xp] = (uint8_t)m0;
                if (m1 > tmin && 2 * xp + 1 < iw) srow[2 * xp + 1] = (uint8_t)m1;
            }
        }
    }
    __syncthreads();

    // ---- NMS once; bit `it` of isMax / isIni = pixel (it*128 + tid) is a maximum (> minTh / > iniTh)
    const int npx = iw * ih;
    const float invw = __frcp_rn((float)iw);
    unsigned isMax = 0, isIni = 0;
    for (int it = 0, base = 0; base < npx; ++it, base += FAST_THREADS) {
        const int i = base + threadIdx.x;
        if (i < npx) {
            const int y = (int)(((float)i + 0.5f) * invw), x = i - y * iw;
            const uint8_t* s = smap + (y + 1) * mw + (x + 1);
            const int m = s[0];
            if (m) {
                const int nb = max(max(imax3(s[-mw - 1], s[-mw], s[-mw + 1]), imax3(s[mw - 1], s[mw], s[mw + 1])),
                                   max(s[-1], s[1]));
                if (m > nb) {
                    isMax |= 1u << it;
                    if (m > tini) isIni |= 1u << it;
                }
            }
        }
    }
    const unsigned sel = __syncthreads_or(isIni != 0) ? isIni : isMax;  // :814 fallback to minTh iff FAST(ini) is empty

    // ---- ordered (row-major) emission
    const int iters = (npx + FAST_THREADS - 1) / FAST_THREADS;
    for (int it = 0; it < iters; ++it) {
        const unsigned bal = __ballot_sync(0xffffffffu, (sel >> it) & 1u);
        if (lane == 0) warpCnt[it][warp] = __popc(bal);
    }
    __syncthreads();
    const uint32_t slot0 = L.slotBase + (uint32_t)(cellId - L.cellBase) * L.slotCap;
    uint32_t* out = p.cellKp + (size_t)img * p.slotTotal + slot0;
    int running = 0;
    for (int it = 0; it < iters; ++it) {
        const int4 c = *reinterpret_cast<const int4*>(warpCnt[it]);
        const unsigned keep = (sel >> it) & 1u;
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int i = it * FAST_THREADS + threadIdx.x;
            const int y = (int)(((float)i + 0.5f) * invw), x = i - y * iw;
            const int m = smap[(y + 1) * mw + (x + 1)];
            int ofs = running + __popc(bal & ((1u << lane) - 1));
            ofs += (warp > 0 ? c.x : 0) + (warp > 1 ? c.y : 0) + (warp > 2 ? c.z : 0);
            out[ofs] = (uint32_t)(x + 3) | ((uint32_t)(y + 3) << 8) | ((uint32_t)(m - 1) << 16);
        }
        running += c.x + c.y + c.z + c.w;
    }
    if (threadIdx.x == 0) *cntOut = running;
}

void orbx_launch_fast(const OrbxParams& p, const void* cellTable, cudaStream_t s)
{
    // largest ROI over the levels (ROI = cell + 6, src/ORBextractor.cc:797-808)
    int rwMax = 0, rhMax = 0;
    for (int l = 0; l < p.nlevels; ++l) {
        rwMax = p.lv[l].wCell + 6 > rwMax ? p.lv[l].wCell + 6 : rwMax;
        rhMax = p.lv[l].hCell + 6 > rhMax ? p.lv[l].hCell + 6 : rhMax;
    }
    rhMax += 1;  // pairs of an odd-width interior read one element past the ROI; keep that inside the buffer
    dim3 grid(p.cellsTotal, p.imgs);
    const size_t smapBytes = (size_t)(rwMax - 4) * (rhMax - 4);
    if (rwMax + 2 <= 48) {
        const size_t smem = (size_t)2 * rhMax * 48 * 2 + smapBytes;
        orbx_fast_kernel<48><<<grid, FAST_THREADS, smem, s>>>(p, (const OrbxCell*)cellTable, rhMax);
    } else {
        const size_t smem = (size_t)2 * rhMax * 80 * 2 + smapBytes;  // ROI <= 66 px (cells of levels narrower than 90 px)
        orbx_fast_kernel<80><<<grid, FAST_THREADS, smem, s>>>(p, (const OrbxCell*)cellTable, rhMax);
    }
}
