// fast.cu — per-cell FAST-9/16 with non-max suppression and the per-cell threshold fallback.
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (/root/reference/src/ORBextractor.cc:792-832) and the two cv::FAST calls inside it (:812, :817).
//
// One CTA per 30-px cell (all levels and all images in one launch). The cell's ROI (cell + 3-px ring
// margin, <= 66x66) is staged in shared memory; every interior pixel gets its corner strength
//     m = max over the 16 arcs of 9 contiguous ring pixels of min(+-(ring - centre))
// which is threshold independent: the pixel is a FAST corner at threshold t iff m > t and OpenCV's
// score is m-1. NMS keeps a pixel iff m is strictly greater than the strengths of its 8 neighbours
// (pixels outside the ROI interior count 0). Because a neighbour that is not a corner at t has
// m <= t < m_centre, the kept set at threshold t is {NMS maxima of the minTh map with m > t}: one
// pass yields both thresholds, and the cell falls back to minTh iff no maximum exceeds iniTh —
// exactly "FAST(ini) came back empty" (:814). Keypoints are written row-major into the cell's slot
// array, so cells concatenated in table order give the reference's candidate order.
#include "orbx_dev.cuh"

struct OrbxCell {
    short x0, y0;        // ROI origin in level coordinates (iniX, iniY)
    uint8_t rw, rh;      // ROI size (maxX-iniX, maxY-iniY)
    uint8_t level, pad;
};

#define FAST_THREADS 128
#define FAST_TILE_STRIDE 72
#define FAST_MAX_ROI 66

__device__ __forceinline__ int imin3(int a, int b, int c) { return min(a, min(b, c)); }
__device__ __forceinline__ int imax3(int a, int b, int c) { return max(a, max(b, c)); }

// corner strength of the pixel at c (shared-memory tile, row stride FAST_TILE_STRIDE); 0 if <= t
__device__ __forceinline__ int fast_strength(const uint8_t* c, int t)
{
    const int S = FAST_TILE_STRIDE;
    const int v = c[0];
    // every 9-arc contains ring position k or k+8: reject early on the two axis pairs
    const int d0 = c[3 * S] - v, d8 = c[-3 * S] - v;
    if (abs(d0) <= t && abs(d8) <= t) return 0;
    const int d4 = c[3] - v, d12 = c[-3] - v;
    if (abs(d4) <= t && abs(d12) <= t) return 0;
    int d[16];
    d[0] = d0; d[8] = d8; d[4] = d4; d[12] = d12;
    d[1] = c[3 * S + 1] - v;   d[2] = c[2 * S + 2] - v;   d[3] = c[S + 3] - v;
    d[5] = c[-S + 3] - v;      d[6] = c[-2 * S + 2] - v;  d[7] = c[-3 * S + 1] - v;
    d[9] = c[-3 * S - 1] - v;  d[10] = c[-2 * S - 2] - v; d[11] = c[-S - 3] - v;
    d[13] = c[S - 3] - v;      d[14] = c[2 * S - 2] - v;  d[15] = c[3 * S - 1] - v;
    // sliding minimum / maximum over windows of 9 (3-way steps)
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = imin3(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        hi3[k] = imax3(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    int bright = -256, dark = 256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        bright = max(bright, imin3(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]));
        dark = min(dark, imax3(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]));
    }
    const int m = max(bright, -dark);
    return m > t ? m : 0;
}

__global__ void __launch_bounds__(FAST_THREADS) orbx_fast_kernel(OrbxParams p, const OrbxCell* cells)
{
    __shared__ __align__(16) uint8_t tile[FAST_MAX_ROI * FAST_TILE_STRIDE];
    __shared__ uint8_t smap[(FAST_MAX_ROI - 4) * (FAST_MAX_ROI - 4)];  // strength map with a 1-px zero frame
    __shared__ int warpCnt[2][FAST_THREADS / 32];

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
    int pitch;
    const uint8_t* lvl = orbx_level_ptr(p, img, cell.level, &pitch);
    const uint8_t* src = lvl + (size_t)cell.y0 * pitch + cell.x0;
    for (int i = threadIdx.x; i < rw * rh; i += FAST_THREADS) {
        const int y = i / rw, x = i - y * rw;
        tile[y * FAST_TILE_STRIDE + x] = src[(size_t)y * pitch + x];
    }
    const int mw = iw + 2;
    for (int i = threadIdx.x; i < mw * (ih + 2); i += FAST_THREADS) smap[i] = 0;
    __syncthreads();

    const int npx = iw * ih;
    const int tmin = p.minTh, tini = p.iniTh;
    for (int i = threadIdx.x; i < npx; i += FAST_THREADS) {
        const int y = i / iw, x = i - y * iw;
        const int m = fast_strength(tile + (y + 3) * FAST_TILE_STRIDE + (x + 3), tmin);
        if (m) smap[(y + 1) * mw + (x + 1)] = (uint8_t)m;
    }
    __syncthreads();

    // pass 1: does any NMS maximum exceed iniTh?
    int any = 0;
    for (int i = threadIdx.x; i < npx; i += FAST_THREADS) {
        const int y = i / iw, x = i - y * iw;
        const uint8_t* s = smap + (y + 1) * mw + (x + 1);
        const int m = s[0];
        if (m > tini) {
            const int nb = max(max(imax3(s[-mw - 1], s[-mw], s[-mw + 1]), imax3(s[mw - 1], s[mw], s[mw + 1])),
                               max(s[-1], s[1]));
            any |= (m > nb);
        }
    }
    const int thr = __syncthreads_or(any) ? tini : tmin;

    // pass 2: ordered (row-major) emission of the maxima above the cell's threshold
    const uint32_t slot0 = L.slotBase + (uint32_t)(cellId - L.cellBase) * L.slotCap;
    uint32_t* out = p.cellKp + (size_t)img * p.slotTotal + slot0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int running = 0, buf = 0;  // every thread tracks the running count; warpCnt is double buffered
    for (int base = 0; base < npx; base += FAST_THREADS, buf ^= 1) {
        const int i = base + threadIdx.x;
        int keep = 0, m = 0, x = 0, y = 0;
        if (i < npx) {
            y = i / iw;
            x = i - y * iw;
            const uint8_t* s = smap + (y + 1) * mw + (x + 1);
            m = s[0];
            if (m > thr) {
                const int nb = max(max(imax3(s[-mw - 1], s[-mw], s[-mw + 1]), imax3(s[mw - 1], s[mw], s[mw + 1])),
                                   max(s[-1], s[1]));
                keep = m > nb;
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warpCnt[buf][warp] = __popc(bal);
        __syncthreads();
        int ofs = running, tot = 0;
#pragma unroll
        for (int w = 0; w < FAST_THREADS / 32; ++w) {
            const int c = warpCnt[buf][w];
            if (w < warp) ofs += c;
            tot += c;
        }
        if (keep) {
            ofs += __popc(bal & ((1u << lane) - 1));
            out[ofs] = (uint32_t)(x + 3) | ((uint32_t)(y + 3) << 8) | ((uint32_t)(m - 1) << 16);
        }
        running += tot;
    }
    if (threadIdx.x == 0) *cntOut = running;
}

void orbx_launch_fast(const OrbxParams& p, const void* cellTable, cudaStream_t s)
{
    dim3 grid(p.cellsTotal, p.imgs);
    orbx_fast_kernel<<<grid, FAST_THREADS, 0, s>>>(p, (const OrbxCell*)cellTable);
}
