// frame.cu — the rows right after the ORB front-end, declared in include/orbx_frame.h:
//   N4  Frame::ComputeStereoFromRGBD                      /root/reference/src/Frame.cc:702-723
//   N1  Frame::AssignFeaturesToGrid / PosInGrid           /root/reference/src/Frame.cc:279-294, 435-445
//       Frame::GetFeaturesInArea                          /root/reference/src/Frame.cc:376-433
//       ORBmatcher::SearchByProjection(cur, last, th, bMono)   /root/reference/src/ORBmatcher.cc:1348-1491
//
// SearchByProjection is sequential in the reference: a current-frame keypoint that already holds a map point
// with Observations() > 0 is skipped by every later last-frame point (:1418-1420). It is split here into
//   (1) grid kernel      — counting sort of the current keypoints into the 64x48 grid, ascending index per cell;
//   (2) search kernel    — warp per last-frame point: project (OpenCV's 32-bit gemm arithmetic), enumerate the grid
//                          cells of the window, Hamming distance per candidate, warp-min of
//                          (distance, enumeration order) = the reference's "first strict minimum";
//   (3) resolve kernel   — one warp per frame pair walks the points in order. The unconstrained best of (2) is final
//                          unless its keypoint was blocked by an earlier point (blocking only removes candidates);
//                          then the warp repeats the search with the blocked set. Rotation histogram, three maxima
//                          and the removal pass (:1455-1487) follow in the same kernel.
// All frame pairs of a call run in one launch each (blockIdx.y = pair); inputs travel in one packed H2D copy.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/orbx_frame.h"
#include "engine.h"
#include "matcher.h"
#include "handles.h"


namespace {

constexpr int GC = ORBX_GRID_COLS, GR = ORBX_GRID_ROWS, NCELL = GC * GR;
constexpr int TH_HIGH = 100, HISTO = 30;  // src/ORBmatcher.cc:37-39
constexpr unsigned long long NOKEY = ~0ull;

struct FrameDev {  // one (current, last) pair; pointers into the packed device buffer
    int nCur, nLast, levelMode /*0: +-1 window, 1: forward, 2: backward (:1402-1407)*/, checkOri;
    const orbx_keypoint* curKeys; const uint8_t* curDesc; const float* curUR;
    const orbx_keypoint* lastKeys; const uint8_t* hasPoint; const float* worldPos; const uint8_t* pointDesc; const uint8_t* observed;
    float Tcw[12], fx, fy, cx, cy, mbf, minX, maxX, minY, maxY, invW, invH, th;
    float scale[ORBX_MAX_LEVELS];
    int* cellStart; int* cellItems;        // CSR grid of the current frame
    unsigned long long* best;              // per last point: search key or NOKEY
    float4* proj;                          // per last point: u, v, radius, invzc (radius < 0: not searched)
    int* assigned; int* events; int* nmatches;
};

__device__ __forceinline__ int pos_in_grid(const FrameDev& P, float x, float y)
{
    // PosInGrid :435-445 (`round` = half away from zero)
    const int px = (int)roundf((x - P.minX) * P.invW), py = (int)roundf((y - P.minY) * P.invH);
    if (px < 0 || px >= GC || py < 0 || py >= GR) return -1;
    return px * GR + py;
}

// ---- (1) grid: one CTA per frame ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) orbx_grid_kernel(const FrameDev* frames)
{
    __shared__ int cnt[NCELL + 1];
    __shared__ int wsum[9];
    const FrameDev& P = frames[blockIdx.y];
    const int tid = threadIdx.x, n = P.nCur;
    for (int c = tid; c <= NCELL; c += 256) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        const int c = pos_in_grid(P, P.curKeys[i].x, P.curKeys[i].y);
        if (c >= 0) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    // exclusive scan of cnt[0..NCELL): 12 cells per thread
    constexpr int PER = NCELL / 256;
    int local[PER], s = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) { local[k] = cnt[tid * PER + k]; s += local[k]; }
    int inc = s;
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (tid == 0) { int run = 0; for (int w = 0; w < 8; ++w) { const int t = wsum[w]; wsum[w] = run; run += t; } wsum[8] = run; }
    __syncthreads();
    int run = wsum[warp] + inc - s;
#pragma unroll
    for (int k = 0; k < PER; ++k) { cnt[tid * PER + k] = run; P.cellStart[tid * PER + k] = run; run += local[k]; }
    if (tid == 0) P.cellStart[NCELL] = wsum[8];
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        const int c = pos_in_grid(P, P.curKeys[i].x, P.curKeys[i].y);
        if (c >= 0) P.cellItems[atomicAdd(&cnt[c], 1)] = i;
    }
    __syncthreads();
    // ascending keypoint index inside a cell (= the reference's push_back order); cells hold a handful of entries
    for (int c = tid; c < NCELL; c += 256) {
        const int b = P.cellStart[c], e = cnt[c];
        for (int i = b + 1; i < e; ++i) {
            const int v = P.cellItems[i];
            int j = i - 1;
            while (j >= b && P.cellItems[j] > v) { P.cellItems[j + 1] = P.cellItems[j]; --j; }
            P.cellItems[j + 1] = v;
        }
    }
}

// Cell window of GetFeaturesInArea :385-399; false = empty result
__device__ __forceinline__ bool area_cells(const FrameDev& P, float x, float y, float r, int& x0, int& x1, int& y0, int& y1)
{
    x0 = max(0, (int)floorf((x - P.minX - r) * P.invW));
    if (x0 >= GC) return false;
    x1 = min(GC - 1, (int)ceilf((x - P.minX + r) * P.invW));
    if (x1 < 0) return false;
    y0 = max(0, (int)floorf((y - P.minY - r) * P.invH));
    if (y0 >= GR) return false;
    y1 = min(GR - 1, (int)ceilf((y - P.minY + r) * P.invH));
    if (y1 < 0) return false;
    return true;
}

__device__ __forceinline__ bool in_area(const orbx_keypoint& kp, float x, float y, float r, int minLevel, int maxLevel)
{
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);  // :401
    if (bCheckLevels) {
        if (kp.octave < minLevel) return false;
        if (maxLevel >= 0 && kp.octave > maxLevel) return false;
    }
    return fabsf(kp.x - x) < r && fabsf(kp.y - y) < r;  // :424-428
}

// Best candidate of last-frame point i for the whole warp (:1396-1441). key = dist << 52 | cell ordinal << 40 |
// position in cell << 20 | keypoint index: the minimum is the reference's first strict minimum in candidate order.
__device__ unsigned long long search_point(const FrameDev& P, int i, float u, float v, float radius, float invzc, const uint32_t* blocked,
                                           int lane)
{
    const int oct = P.lastKeys[i].octave;
    int minL, maxL;
    if (P.levelMode == 1) { minL = oct; maxL = -1; }
    else if (P.levelMode == 2) { minL = 0; maxL = oct; }
    else { minL = oct - 1; maxL = oct + 1; }
    int x0, x1, y0, y1;
    unsigned long long key = NOKEY;
    if (area_cells(P, u, v, radius, x0, x1, y0, y1)) {
        const uint4* dm = reinterpret_cast<const uint4*>(P.pointDesc + 32 * (size_t)i);
        const uint4 a0 = dm[0], a1 = dm[1];
        const int ny = y1 - y0 + 1, ncell = (x1 - x0 + 1) * ny;
        const float ur = u - P.mbf * invzc;  // :1424
        for (int c = lane; c < ncell; c += 32) {
            const int ix = x0 + c / ny, iy = y0 + c % ny;
            const int b = P.cellStart[ix * GR + iy], e = P.cellStart[ix * GR + iy + 1];
            for (int j = b; j < e; ++j) {
                const int i2 = P.cellItems[j];
                if (!in_area(P.curKeys[i2], u, v, radius, minL, maxL)) continue;
                if (blocked && (blocked[i2 >> 5] >> (i2 & 31) & 1u)) continue;  // :1418-1420
                if (P.curUR) {
                    const float r2 = P.curUR[i2];
                    if (r2 > 0 && fabsf(ur - r2) > radius) continue;  // :1422-1428
                }
                const uint4* dc = reinterpret_cast<const uint4*>(P.curDesc + 32 * (size_t)i2);
                const uint4 b0 = dc[0], b1 = dc[1];
                const int dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                                 __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                const unsigned long long k = ((unsigned long long)dist << 52) | ((unsigned long long)c << 40) |
                                             ((unsigned long long)(j - b) << 20) | (unsigned long long)i2;
                key = k < key ? k : key;
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, key, o);
        key = t < key ? t : key;
    }
    return key;
}

// ---- (2) projection + unconstrained search: warp per last-frame point ------------------------------------
__global__ void __launch_bounds__(256) orbx_project_search_kernel(const FrameDev* frames)
{
    const FrameDev& P = frames[blockIdx.y];
    const int lane = threadIdx.x & 31, i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= P.nLast) return;
    float4 pr = make_float4(0.f, 0.f, -1.f, 0.f);
    unsigned long long key = NOKEY;
    if (P.hasPoint[i]) {
        // x3Dc = Rcw * x3Dw + tcw: cv::gemm on CV_32F 3x3 * 3x1 = float products, float sums left to right, then
        // (float)((double)sum + (double)t) (pinned against cv2 in tests/test_frame_oracle.py)
        const float X = P.worldPos[3 * i], Y = P.worldPos[3 * i + 1], Z = P.worldPos[3 * i + 2];
        float xc[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float s = __fadd_rn(__fadd_rn(__fmul_rn(P.Tcw[4 * r], X), __fmul_rn(P.Tcw[4 * r + 1], Y)), __fmul_rn(P.Tcw[4 * r + 2], Z));
            xc[r] = __double2float_rn(__dadd_rn((double)s, (double)P.Tcw[4 * r + 3]));
        }
        const float invzc = __double2float_rn(__ddiv_rn(1.0, (double)xc[2]));  // `1.0 / x3Dc.at<float>(2)` :1386
        if (!(invzc < 0)) {
            const float u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, xc[0]), invzc), P.cx);
            const float v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, xc[1]), invzc), P.cy);
            if (!(u < P.minX || u > P.maxX) && !(v < P.minY || v > P.maxY)) {
                const float radius = __fmul_rn(P.th, P.scale[P.lastKeys[i].octave]);  // :1400
                pr = make_float4(u, v, radius, invzc);
                key = search_point(P, i, u, v, radius, invzc, nullptr, lane);
            }
        }
    }
    if (lane == 0) { P.best[i] = key; P.proj[i] = pr; }
}

// ---- (3) in-order resolution, rotation consistency: one warp per frame pair -------------------------------
__global__ void __launch_bounds__(32) orbx_resolve_kernel(const FrameDev* frames)
{
    extern __shared__ uint32_t blocked[];  // bit i2: CurrentFrame.mvpMapPoints[i2] holds a point with Observations() > 0
    __shared__ int bins[HISTO];
    const FrameDev& P = frames[blockIdx.y];
    const int lane = threadIdx.x;
    for (int w = lane; w < (P.nCur + 31) / 32; w += 32) blocked[w] = 0;
    for (int i = lane; i < P.nCur; i += 32) P.assigned[i] = -1;
    if (lane < HISTO) bins[lane] = 0;
    __syncwarp();
    int nmatches = 0, nev = 0;
    const unsigned lt = (1u << lane) - 1;
    // One accepted match: CurrentFrame.mvpMapPoints[i2] = pMP (:1445), rotation bin (:1448-1458), and the blocking rule.
    auto commit = [&](int i, int i2, int slot) {
        P.assigned[i2] = i;
        if (P.checkOri) {
            float rot = P.lastKeys[i].angle - P.curKeys[i2].angle;
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * (1.0f / HISTO));
            if (bin == HISTO) bin = 0;
            P.events[slot] = (bin << 20) | i2;
            atomicAdd(&bins[bin], 1);
        }
        if (P.observed[i]) atomicOr(&blocked[i2 >> 5], 1u << (i2 & 31));
    };
    unsigned long long next = lane < P.nLast ? P.best[lane] : NOKEY;
    for (int i0 = 0; i0 < P.nLast; i0 += 32) {
        const unsigned long long mine = next;
        next = i0 + 32 + lane < P.nLast ? P.best[i0 + 32 + lane] : NOKEY;
        // Only points whose unconstrained best is within TH_HIGH can ever match (blocking removes candidates, the
        // best distance can only grow). 32 points are settled per step: those ahead of the first "problem" lane —
        // best keypoint already blocked, or shared with an earlier lane of the step — commit in parallel (distinct
        // keypoints, no order dependence); the problem lane is then handled alone, exactly like the reference's loop.
        unsigned pending = __ballot_sync(0xffffffffu, mine != NOKEY && (int)(mine >> 52) <= TH_HIGH);
        const int i2 = (int)(mine & 0xFFFFFu);
        while (pending) {
            const bool isPend = pending >> lane & 1u;
            const bool blk = isPend && (blocked[i2 >> 5] >> (i2 & 31) & 1u);
            const unsigned grp = __match_any_sync(0xffffffffu, isPend ? i2 : (int)(0x40000000u | lane));
            const bool dup = isPend && (grp & lt) != 0;
            const unsigned prob = __ballot_sync(0xffffffffu, blk || dup);
            const int first = prob ? __ffs(prob) - 1 : 32;
            const unsigned cm = first == 32 ? pending : pending & ((1u << first) - 1);
            if (cm >> lane & 1u) commit(i0 + lane, i2, nev + __popc(cm & lt));
            nmatches += __popc(cm);
            nev += __popc(cm);
            pending &= ~cm;
            __syncwarp();
            if (first < 32) {
                const int i = i0 + first;
                unsigned long long key = __shfl_sync(0xffffffffu, mine, first);
                int k2 = (int)(key & 0xFFFFFu);
                if (blocked[k2 >> 5] >> (k2 & 31) & 1u) {
                    const float4 pr = P.proj[i];
                    key = search_point(P, i, pr.x, pr.y, pr.z, pr.w, blocked, lane);
                    k2 = (int)(key & 0xFFFFFu);
                }
                if (key != NOKEY && (int)(key >> 52) <= TH_HIGH) {  // :1443
                    if (lane == 0) commit(i, k2, nev);
                    ++nmatches;
                    ++nev;
                }
                pending &= ~(1u << first);
                __syncwarp();
            }
        }
    }
    if (P.checkOri) {
        __syncwarp();
        int ind1 = -1, ind2 = -1, ind3 = -1;  // ComputeThreeMaxima (every lane computes the same)
        int max1 = 0, max2 = 0, max3 = 0;
        for (int b = 0; b < HISTO; ++b) {
            const int s = bins[b];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
            else if (s > max3) { max3 = s; ind3 = b; }
        }
        if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
        int removed = 0;
        for (int e = lane; e < nev; e += 32) {
            const int ev = P.events[e], bin = ev >> 20;
            if (bin != ind1 && bin != ind2 && bin != ind3) { P.assigned[ev & 0xFFFFF] = -1; ++removed; }
        }
        removed = __reduce_add_sync(0xffffffffu, removed);
        nmatches -= removed;
    }
    if (lane == 0) *P.nmatches = nmatches;
}

// ---- local-map SearchByProjection (src/ORBmatcher.cc:45-127): best AND second-best candidate per map point --------
struct MapDev {  // the map-point side of one (frame, local map) pair; the frame side is a FrameDev (grid, keys, descriptors)
    int nPts;
    const uint8_t* valid; const float* proj; const int* level; const float* viewCos; const uint8_t* pdesc; const uint8_t* observed;
    const uint8_t* occupied;  // per keypoint, may be NULL
    float th, nnratio;
    unsigned long long* best2;  // per point: second-best key (best is FrameDev::best)
};

__device__ __forceinline__ float map_radius(const FrameDev& P, const MapDev& X, int i)
{
    float r = ((double)X.viewCos[i] > 0.998) ? 2.5f : 4.0f;  // RadiusByViewingCos :134-140
    if (X.th != 1.0f) r = __fmul_rn(r, X.th);                // bFactor :49, :66-67
    return __fmul_rn(r, P.scale[X.level[i]]);
}

// Two smallest keys (distance, candidate order) of map point i over the candidates the reference's loop would visit.
__device__ void search_top2(const FrameDev& P, const MapDev& X, int i, const uint32_t* blocked, int lane, unsigned long long& k1,
                            unsigned long long& k2)
{
    const int lvl = X.level[i];
    const float x = X.proj[3 * i], y = X.proj[3 * i + 1], xr = X.proj[3 * i + 2], radius = map_radius(P, X, i);
    k1 = k2 = NOKEY;
    int x0, x1, y0, y1;
    if (area_cells(P, x, y, radius, x0, x1, y0, y1)) {
        const uint4* dm = reinterpret_cast<const uint4*>(X.pdesc + 32 * (size_t)i);
        const uint4 a0 = dm[0], a1 = dm[1];
        const int ny = y1 - y0 + 1, ncell = (x1 - x0 + 1) * ny;
        for (int c = lane; c < ncell; c += 32) {
            const int ix = x0 + c / ny, iy = y0 + c % ny;
            const int b = P.cellStart[ix * GR + iy], e = P.cellStart[ix * GR + iy + 1];
            for (int j = b; j < e; ++j) {
                const int i2 = P.cellItems[j];
                if (!in_area(P.curKeys[i2], x, y, radius, lvl - 1, lvl)) continue;
                if (blocked ? (blocked[i2 >> 5] >> (i2 & 31) & 1u) : (X.occupied && X.occupied[i2])) continue;  // :92-94
                if (P.curUR) {
                    const float r2 = P.curUR[i2];
                    if (r2 > 0 && fabsf(xr - r2) > radius) continue;  // :96-101
                }
                const uint4* dc = reinterpret_cast<const uint4*>(P.curDesc + 32 * (size_t)i2);
                const uint4 b0 = dc[0], b1 = dc[1];
                const int dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                                 __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                const unsigned long long k = ((unsigned long long)dist << 52) | ((unsigned long long)c << 40) |
                                             ((unsigned long long)(j - b) << 20) | (unsigned long long)i2;
                if (k < k1) { k2 = k1; k1 = k; } else if (k < k2) k2 = k;
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const unsigned long long lo = k1 < o1 ? k1 : o1, hi = k1 < o1 ? o1 : k1, s2 = k2 < o2 ? k2 : o2;
        k1 = lo;
        k2 = hi < s2 ? hi : s2;
    }
}

__global__ void __launch_bounds__(256) orbx_map_search_kernel(const FrameDev* frames, const MapDev* maps)
{
    const FrameDev& P = frames[blockIdx.y];
    const MapDev& X = maps[blockIdx.y];
    const int lane = threadIdx.x & 31, i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= X.nPts) return;
    unsigned long long k1 = NOKEY, k2 = NOKEY;
    if (X.valid[i]) search_top2(P, X, i, nullptr, lane, k1, k2);
    if (lane == 0) { P.best[i] = k1; X.best2[i] = k2; }
}

// :113-123 — accept the best candidate unless the second-best has the same level and is too close
__device__ __forceinline__ bool map_accept(const FrameDev& P, const MapDev& X, unsigned long long k1, unsigned long long k2)
{
    if (k1 == NOKEY) return false;
    const int bestDist = (int)(k1 >> 52);
    if (bestDist > TH_HIGH) return false;
    const int bestLevel = P.curKeys[(int)(k1 & 0xFFFFFu)].octave;
    const int bestDist2 = k2 == NOKEY ? 256 : (int)(k2 >> 52), bestLevel2 = k2 == NOKEY ? -1 : P.curKeys[(int)(k2 & 0xFFFFFu)].octave;
    return !(bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(X.nnratio, (float)bestDist2));
}

__global__ void __launch_bounds__(32) orbx_map_resolve_kernel(const FrameDev* frames, const MapDev* maps)
{
    extern __shared__ uint32_t blocked[];
    const FrameDev& P = frames[blockIdx.y];
    const MapDev& X = maps[blockIdx.y];
    const int lane = threadIdx.x;
    for (int w = lane; w < (P.nCur + 31) / 32; w += 32) {
        uint32_t bits = 0;
        if (X.occupied)
            for (int b = 0; b < 32 && 32 * w + b < P.nCur; ++b) bits |= X.occupied[32 * w + b] ? 1u << b : 0u;
        blocked[w] = bits;
    }
    for (int i = lane; i < P.nCur; i += 32) P.assigned[i] = -1;
    __syncwarp();
    auto isBlocked = [&](unsigned long long k) { const int q = (int)(k & 0xFFFFFu); return k != NOKEY && (blocked[q >> 5] >> (q & 31) & 1u); };
    auto commit = [&](int i, int i2) {
        P.assigned[i2] = i;  // F.mvpMapPoints[bestIdx] = pMP :120
        if (X.observed[i]) atomicOr(&blocked[i2 >> 5], 1u << (i2 & 31));
    };
    int nmatches = 0;
    for (int i0 = 0; i0 < X.nPts; i0 += 32) {
        const int i = i0 + lane;
        const unsigned long long k1 = i < X.nPts ? P.best[i] : NOKEY, k2 = i < X.nPts ? X.best2[i] : NOKEY;
        // only points whose unconstrained best is within TH_HIGH can ever match; a point's result is final when neither its
        // best nor its second-best keypoint is blocked or taken by an earlier point of this step
        unsigned pending = __ballot_sync(0xffffffffu, k1 != NOKEY && (int)(k1 >> 52) <= TH_HIGH);
        const int b1 = (int)(k1 & 0xFFFFFu), b2 = k2 == NOKEY ? -1 : (int)(k2 & 0xFFFFFu);
        while (pending) {
            const bool isPend = pending >> lane & 1u;
            bool prob = isPend && (isBlocked(k1) || isBlocked(k2));
            for (unsigned rest = pending; rest; rest &= rest - 1) {
                const int j = __ffs(rest) - 1;
                const int bj = __shfl_sync(0xffffffffu, b1, j);
                prob |= isPend && j < lane && (bj == b1 || bj == b2);
            }
            const unsigned pm = __ballot_sync(0xffffffffu, prob);
            const int first = pm ? __ffs(pm) - 1 : 32;
            const unsigned cm = first == 32 ? pending : pending & ((1u << first) - 1);
            const bool acc = (cm >> lane & 1u) && map_accept(P, X, k1, k2);
            if (acc) commit(i, b1);
            nmatches += __popc(__ballot_sync(0xffffffffu, acc));
            pending &= ~cm;
            __syncwarp();
            if (first < 32) {
                const int ip = i0 + first;
                unsigned long long q1 = __shfl_sync(0xffffffffu, k1, first), q2 = __shfl_sync(0xffffffffu, k2, first);
                if (isBlocked(q1) || isBlocked(q2)) search_top2(P, X, ip, blocked, lane, q1, q2);
                if (map_accept(P, X, q1, q2)) {
                    if (lane == 0) commit(ip, (int)(q1 & 0xFFFFFu));
                    ++nmatches;
                }
                pending &= ~(1u << first);
                __syncwarp();
            }
        }
    }
    if (lane == 0) *P.nmatches = nmatches;
}

// ---- GetFeaturesInArea as a query kernel (parity surface; SearchByProjection uses search_point) -----------
__global__ void __launch_bounds__(32) orbx_area_kernel(const FrameDev* frames, float x, float y, float r, int minLevel, int maxLevel,
                                                       int* out, int capacity, int* nOut)
{
    const FrameDev& P = frames[0];
    const int lane = threadIdx.x;
    int n = 0, x0, x1, y0, y1;
    if (area_cells(P, x, y, r, x0, x1, y0, y1)) {
        for (int ix = x0; ix <= x1; ++ix)
            for (int iy = y0; iy <= y1; ++iy) {
                const int b = P.cellStart[ix * GR + iy], e = P.cellStart[ix * GR + iy + 1];
                for (int j0 = b; j0 < e; j0 += 32) {
                    const int j = j0 + lane;
                    const int idx = j < e ? P.cellItems[j] : 0;
                    const bool ok = j < e && in_area(P.curKeys[idx], x, y, r, minLevel, maxLevel);
                    const unsigned m = __ballot_sync(0xffffffffu, ok);
                    const int pos = n + __popc(m & ((1u << lane) - 1));
                    if (ok && pos < capacity) out[pos] = idx;
                    n += __popc(m);
                }
            }
    }
    if (lane == 0) *nOut = n;
}

// ---- N4: ComputeStereoFromRGBD ----------------------------------------------------------------------------
__global__ void orbx_rgbd_kernel(const orbx_keypoint* keys, const float* keysUnX, int n, const float* depth, int rows, int cols,
                                 int strideFloats, float mbf, float* uRight, float* depthOut)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const orbx_keypoint kp = keys[i];
    const int v = (int)kp.y, u = (int)kp.x;  // imDepth.at<float>(v, u): float -> int truncation
    float ur = -1.f, dp = -1.f;
    if (v >= 0 && v < rows && u >= 0 && u < cols) {
        const float d = depth[(size_t)v * strideFloats + u];
        if (d > 0) {
            dp = d;
            ur = __fsub_rn(keysUnX ? keysUnX[i] : kp.x, __fdiv_rn(mbf, d));
        }
    }
    uRight[i] = ur;
    depthOut[i] = dp;
}

}  // namespace

// ======================================= host side / C-ABI ===================================================
namespace {

// z of tlc = Rlw * (-Rcw^T * tcw) + tlw (src/ORBmatcher.cc:1358-1367) in OpenCV's arithmetic: `-Rcw.t() * tcw` is the
// general gemm path (double accumulation, one rounding), `Rlw * twc + tlw` the small-matrix path (float products and
// sums, then (float)((double)s + (double)t)). Host code is compiled with -ffp-contract=off.
float tlc_z(const float* Tcw, const float* Tlw)
{
    float twc[3];
    for (int i = 0; i < 3; ++i) {
        double acc = 0;
        for (int k = 0; k < 3; ++k) acc += (double)Tcw[4 * k + i] * (double)Tcw[4 * k + 3];
        twc[i] = (float)(acc * -1.0);
    }
    const float p0 = Tlw[8] * twc[0], p1 = Tlw[9] * twc[1], p2 = Tlw[10] * twc[2];
    float s = p0 + p1;
    s = s + p2;
    return (float)((double)s + (double)Tlw[11]);
}

int check_frame(const orbx_frame_view* f, bool needDesc)
{
    if (!f || f->n < 0 || f->n >= (1 << 20) || (f->n && !f->keys_un) || (needDesc && f->n && !f->desc)) return orbx_fail(ORBX_ERR_ARG, "bad frame view");
    if (!f->scale_factors || f->nlevels < 1 || f->nlevels > ORBX_MAX_LEVELS) return orbx_fail(ORBX_ERR_ARG, "bad scale factor table");
    if (!(f->max_x > f->min_x) || !(f->max_y > f->min_y)) return orbx_fail(ORBX_ERR_ARG, "empty image bounds");
    return ORBX_OK;
}

// Fills the geometry part of a FrameDev from the CURRENT frame view.
void fill_frame(FrameDev& D, const orbx_frame_view& c)
{
    std::memset(&D, 0, sizeof(D));
    D.nCur = c.n;
    std::memcpy(D.Tcw, c.Tcw, sizeof(D.Tcw));
    D.fx = c.fx; D.fy = c.fy; D.cx = c.cx; D.cy = c.cy; D.mbf = c.mbf;
    D.minX = c.min_x; D.maxX = c.max_x; D.minY = c.min_y; D.maxY = c.max_y;
    D.invW = (float)GC / (c.max_x - c.min_x);  // src/Frame.cc:136-137
    D.invH = (float)GR / (c.max_y - c.min_y);
    for (int l = 0; l < c.nlevels; ++l) D.scale[l] = c.scale_factors[l];
}

// One frame with its grid on the device (grid and area entry points). Returns the FrameDev as laid out.
int stage_single_frame(orbx_matcher* m, const orbx_frame_view* f, FrameDev& D, size_t extraScratch, size_t* scratchOff)
{
    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));
    OrbxPacker pk;
    const size_t oDev = pk.take(sizeof(FrameDev)), oKeys = pk.take((size_t)f->n * 28);
    const size_t inBytes = pk.off;
    const size_t oStart = pk.take((NCELL + 1) * 4), oItems = pk.take((size_t)f->n * 4 + 4);
    *scratchOff = pk.take(extraScratch);
    TRY(orbx_matcher_grow(m, pk.off, inBytes, (NCELL + 1) * 4 + (size_t)f->n * 4 + 64));
    fill_frame(D, *f);
    D.curKeys = (const orbx_keypoint*)(m->dBuf + oKeys);
    D.cellStart = (int*)(m->dBuf + oStart);
    D.cellItems = (int*)(m->dBuf + oItems);
    std::memcpy(m->hIn + oDev, &D, sizeof(D));
    if (f->n) std::memcpy(m->hIn + oKeys, f->keys_un, (size_t)f->n * 28);
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, m->stream), "H2D frame", err));
    orbx_grid_kernel<<<dim3(1, 1), 256, 0, m->stream>>>((const FrameDev*)m->dBuf);
    return orbx_check(cudaGetLastError(), "grid launch", err);
}

}  // namespace

extern "C" int orbx_matcher_create(int device, orbx_matcher** out)
{
    if (!out) return orbx_fail(ORBX_ERR_ARG, "out is NULL");
    *out = nullptr;
    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    orbx_matcher* m = new orbx_matcher();
    m->device = device;
    const int rc = orbx_check(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking), "cudaStreamCreate", err);
    if (rc) { delete m; return rc; }
    *out = m;
    return ORBX_OK;
}

extern "C" void orbx_matcher_destroy(orbx_matcher* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->dBuf) cudaFree(m->dBuf);
    if (m->hIn) cudaFreeHost(m->hIn);
    if (m->hOut) cudaFreeHost(m->hOut);
    if (m->orderEvent) cudaEventDestroy(m->orderEvent);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

extern "C" int orbx_stereo_from_rgbd(orbx_matcher* m, const orbx_keypoint* keys, const float* keysUnX, int n, const float* depth, int rows,
                                     int cols, size_t strideBytes, float mbf, float* uRight, float* depthOut)
{
    if (!m || n < 0 || (n && (!keys || !uRight || !depthOut)) || !depth || rows < 1 || cols < 1 || strideBytes < (size_t)cols * 4 ||
        (strideBytes & 3))
        return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    if (n == 0) return ORBX_OK;
    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));
    OrbxPacker pk;
    const size_t oKeys = pk.take((size_t)n * 28), oUx = pk.take(keysUnX ? (size_t)n * 4 : 0), oDepth = pk.take((size_t)rows * cols * 4);
    const size_t inBytes = pk.off;
    const size_t oUr = pk.take((size_t)n * 4), oDp = pk.take((size_t)n * 4);
    TRY(orbx_matcher_grow(m, pk.off, inBytes, (size_t)n * 8 + 32));
    std::memcpy(m->hIn + oKeys, keys, (size_t)n * 28);
    if (keysUnX) std::memcpy(m->hIn + oUx, keysUnX, (size_t)n * 4);
    for (int r = 0; r < rows; ++r) std::memcpy(m->hIn + oDepth + (size_t)r * cols * 4, (const uint8_t*)depth + r * strideBytes, (size_t)cols * 4);
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, m->stream), "H2D rgbd", err));
    orbx_rgbd_kernel<<<(n + 255) / 256, 256, 0, m->stream>>>((const orbx_keypoint*)(m->dBuf + oKeys), keysUnX ? (const float*)(m->dBuf + oUx) : nullptr,
                                                             n, (const float*)(m->dBuf + oDepth), rows, cols, cols, mbf, (float*)(m->dBuf + oUr),
                                                             (float*)(m->dBuf + oDp));
    TRY(orbx_check(cudaGetLastError(), "rgbd launch", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, m->dBuf + oUr, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream), "D2H uRight", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut + (size_t)n * 4, m->dBuf + oDp, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream), "D2H depth", err));
    TRY(orbx_check(cudaStreamSynchronize(m->stream), "sync", err));
    std::memcpy(uRight, m->hOut, (size_t)n * 4);
    std::memcpy(depthOut, m->hOut + (size_t)n * 4, (size_t)n * 4);
    return ORBX_OK;
}

extern "C" int orbx_assign_features_to_grid(orbx_matcher* m, const orbx_frame_view* f, int* cellStart, int* cellItems, int* nItems)
{
    if (!m || !cellStart || !nItems || (f && f->n && !cellItems)) return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    TRY(check_frame(f, false));
    std::string& err = orbx_err_slot();
    FrameDev D;
    size_t so;
    TRY(stage_single_frame(m, f, D, 0, &so));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, D.cellStart, (NCELL + 1) * 4, cudaMemcpyDeviceToHost, m->stream), "D2H cellStart", err));
    if (f->n) TRY(orbx_check(cudaMemcpyAsync(m->hOut + (NCELL + 1) * 4, D.cellItems, (size_t)f->n * 4, cudaMemcpyDeviceToHost, m->stream), "D2H cellItems", err));
    TRY(orbx_check(cudaStreamSynchronize(m->stream), "sync", err));
    std::memcpy(cellStart, m->hOut, (NCELL + 1) * 4);
    *nItems = cellStart[NCELL];
    if (*nItems) std::memcpy(cellItems, m->hOut + (NCELL + 1) * 4, (size_t)*nItems * 4);
    return ORBX_OK;
}

extern "C" int orbx_get_features_in_area(orbx_matcher* m, const orbx_frame_view* f, float x, float y, float r, int minLevel, int maxLevel,
                                         int* indices, int capacity, int* nIndices)
{
    if (!m || !nIndices || capacity < 0 || (capacity && !indices)) return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    TRY(check_frame(f, false));
    std::string& err = orbx_err_slot();
    FrameDev D;
    size_t so;
    const int cap = capacity < f->n ? capacity : f->n;
    TRY(stage_single_frame(m, f, D, (size_t)cap * 4 + 16, &so));
    int* dOut = (int*)(m->dBuf + so);
    int* dN = dOut + cap + (cap & 1 ? 1 : 2);
    orbx_area_kernel<<<1, 32, 0, m->stream>>>((const FrameDev*)m->dBuf, x, y, r, minLevel, maxLevel, dOut, cap, dN);
    TRY(orbx_check(cudaGetLastError(), "area launch", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, dN, 4, cudaMemcpyDeviceToHost, m->stream), "D2H count", err));
    if (cap) TRY(orbx_check(cudaMemcpyAsync(m->hOut + 16, dOut, (size_t)cap * 4, cudaMemcpyDeviceToHost, m->stream), "D2H indices", err));
    TRY(orbx_check(cudaStreamSynchronize(m->stream), "sync", err));
    const int n = *(const int*)m->hOut;
    *nIndices = n;
    if (n > capacity) return orbx_fail(ORBX_ERR_CAPACITY, "more features in the area than the index buffer holds");
    if (n) std::memcpy(indices, m->hOut + 16, (size_t)n * 4);
    return ORBX_OK;
}

namespace {
// Keypoints / descriptors / uRight of a frame where a batch engine left them (orbx_frame_source): device pointers, no copy.
struct ResidentFrame {
    const orbx_keypoint* keys = nullptr; const uint8_t* desc = nullptr; const float* uRight = nullptr;
};
int resolve_source(orbx_matcher* m, const orbx_frame_source& src, const orbx_frame_view& v, ResidentFrame& out)
{
    orbx_batch* b = src.batch;
    const int per = b->stereo ? 2 : 1;
    if (src.frame < 0 || src.frame >= b->maxFrames || src.side < 0 || src.side >= per) return orbx_fail(ORBX_ERR_ARG, "frame source outside the batch");
    Engine& E = b->eng;
    if (E.device != m->device) return orbx_fail(ORBX_ERR_STATE, "batch and matcher live on different devices");
    if (v.n > E.P.maxKp) return orbx_fail(ORBX_ERR_ARG, "frame view has more keypoints than the batch engine can hold");
    if (v.nlevels < E.P.nlevels) return orbx_fail(ORBX_ERR_ARG, "scale factor table shorter than the extractor's levels");
    const size_t img = (size_t)src.frame * per + src.side, Pk = (size_t)E.P.maxKp;
    out.keys = reinterpret_cast<const orbx_keypoint*>((const uint8_t*)E.P.outKp + img * Pk * 28);
    out.desc = E.P.outDesc + img * Pk * 32;
    out.uRight = (b->stereo && src.side == 0) ? E.SP.uRight + (size_t)src.frame * Pk : nullptr;
    return ORBX_OK;
}
// the matcher's stream runs behind everything the batch engines have enqueued
int order_after_batch(orbx_matcher* m, orbx_batch* b, std::string& err)
{
    Engine& E = b->eng;
    if (!E.lastBounds.empty()) TRY(E.waitPipelined(err));
    if (!m->orderEvent) TRY(orbx_check(cudaEventCreateWithFlags(&m->orderEvent, cudaEventDisableTiming), "cudaEventCreate", err));
    TRY(orbx_check(cudaEventRecord(m->orderEvent, E.stream), "event", err));
    return orbx_check(cudaStreamWaitEvent(m->stream, m->orderEvent, 0), "wait", err);
}

int search_by_projection_impl(orbx_matcher* m, const orbx_frame_view* cur, const orbx_frame_source* curSrc, const orbx_frame_view* last,
                              const orbx_frame_source* lastSrc, const orbx_mappoint_view* lastPoints, int nPairs, float th, int bMono,
                              int checkOrientation, int* const* matchOfCur, int* nmatches)
{
    if (!m || !cur || !last || !lastPoints || nPairs < 1 || !matchOfCur || !nmatches) return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    std::string& err = orbx_err_slot();
    int maxLast = 0, maxCur = 0;
    std::vector<ResidentFrame> rc(nPairs), rl(nPairs);
    for (int p = 0; p < nPairs; ++p) {
        const bool curRes = curSrc && curSrc[p].batch, lastRes = lastSrc && lastSrc[p].batch;
        if (curRes) {  // keys_un / desc / u_right of the view are not read
            orbx_frame_view v = cur[p];
            v.keys_un = reinterpret_cast<const orbx_keypoint*>(&v); v.desc = reinterpret_cast<const uint8_t*>(&v);
            TRY(check_frame(&v, true));
            TRY(resolve_source(m, curSrc[p], cur[p], rc[p]));
        } else {
            TRY(check_frame(cur + p, true));
        }
        if (lastRes) {
            orbx_frame_view v = last[p];
            v.keys_un = reinterpret_cast<const orbx_keypoint*>(&v);
            TRY(check_frame(&v, false));
            TRY(resolve_source(m, lastSrc[p], last[p], rl[p]));
        } else {
            TRY(check_frame(last + p, false));
        }
        const orbx_mappoint_view& mp = lastPoints[p];
        if (last[p].n && (!mp.has_point || !mp.world_pos || !mp.point_desc || !mp.observed)) return orbx_fail(ORBX_ERR_ARG, "bad map point view");
        if (cur[p].n && !matchOfCur[p]) return orbx_fail(ORBX_ERR_ARG, "match_of_cur is NULL");
        if (lastRes) {  // the extractor's own octaves are below its level count
            if (lastSrc[p].batch->eng.P.nlevels > cur[p].nlevels) return orbx_fail(ORBX_ERR_ARG, "last-frame octaves outside the current frame's scale table");
        } else {
            for (int i = 0; i < last[p].n; ++i)
                if (last[p].keys_un[i].octave < 0 || last[p].keys_un[i].octave >= cur[p].nlevels)
                    return orbx_fail(ORBX_ERR_ARG, "last-frame keypoint octave outside the current frame's scale table");
        }
        maxLast = last[p].n > maxLast ? last[p].n : maxLast;
        maxCur = cur[p].n > maxCur ? cur[p].n : maxCur;
    }
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));

    // ---- layout: [FrameDev x nPairs | frame arrays of every pair that are not resident] (one H2D from pinned staging)
    //              | map point arrays, one block per array over all pairs | scratch | [assigned of every pair | nmatches] (one D2H)
    // The map point blocks are unpadded concatenations in pair order. A caller that keeps its map point arrays that way
    // (pair p+1's array starts where pair p's ends) has them copied straight from its memory with one cudaMemcpyAsync per
    // array; otherwise they pass through the pinned staging like everything else.
    struct Off { size_t curKeys, curDesc, curUR, lastKeys, has, pos, pdesc, obs, cellStart, cellItems, best, proj, events, assigned; };
    std::vector<Off> off(nPairs);
    OrbxPacker pk;
    const size_t oDev = pk.take(sizeof(FrameDev) * (size_t)nPairs);
    size_t totalLast = 0;
    bool contiguous = true;
    for (int p = 0; p < nPairs; ++p) {
        const size_t nc = cur[p].n, nl = last[p].n;
        const bool curRes = rc[p].keys != nullptr, lastRes = rl[p].keys != nullptr;
        off[p].curKeys = pk.take(curRes ? 0 : nc * 28); off[p].curDesc = pk.take(curRes ? 0 : nc * 32);
        off[p].curUR = pk.take(!curRes && cur[p].u_right ? nc * 4 : 0);
        off[p].lastKeys = pk.take(lastRes ? 0 : nl * 28);
        if (nl) {
            const orbx_mappoint_view &a = lastPoints[0], &q = lastPoints[p];
            contiguous = contiguous && q.has_point == a.has_point + totalLast && q.observed == a.observed + totalLast &&
                         q.world_pos == a.world_pos + 3 * totalLast && q.point_desc == a.point_desc + 32 * totalLast;
        }
        totalLast += nl;
    }
    if (!lastPoints[0].has_point || last[0].n == 0) contiguous = false;  // (the blocks are anchored at pair 0's arrays)
    const size_t stagedBytes = pk.off;
    const size_t oHas = pk.take(totalLast), oObs = pk.take(totalLast), oPos = pk.take(totalLast * 12), oDesc = pk.take(totalLast * 32);
    {
        size_t run = 0;
        for (int p = 0; p < nPairs; ++p) {
            off[p].has = oHas + run; off[p].obs = oObs + run; off[p].pos = oPos + 12 * run; off[p].pdesc = oDesc + 32 * run;
            run += last[p].n;
        }
    }
    const size_t inBytes = contiguous ? stagedBytes : pk.off;
    for (int p = 0; p < nPairs; ++p) {
        const size_t nc = cur[p].n, nl = last[p].n;
        off[p].cellStart = pk.take((NCELL + 1) * 4); off[p].cellItems = pk.take(nc * 4 + 4); off[p].best = pk.take(nl * 8 + 8);
        off[p].proj = pk.take(nl * 16 + 16); off[p].events = pk.take(nl * 4 + 4);
    }
    const size_t outBase = pk.off;
    for (int p = 0; p < nPairs; ++p) off[p].assigned = pk.take((size_t)cur[p].n * 4);
    const size_t oNm = pk.take((size_t)nPairs * 4);
    const size_t outBytes = pk.off - outBase;
    TRY(orbx_matcher_grow(m, pk.off, inBytes, outBytes));

    FrameDev* hd = reinterpret_cast<FrameDev*>(m->hIn + oDev);
    for (int p = 0; p < nPairs; ++p) {
        const orbx_frame_view &c = cur[p], &l = last[p];
        const orbx_mappoint_view& mp = lastPoints[p];
        const bool curRes = rc[p].keys != nullptr, lastRes = rl[p].keys != nullptr;
        FrameDev D;
        fill_frame(D, c);
        D.nLast = l.n;
        D.checkOri = checkOrientation ? 1 : 0;
        D.th = th;
        const float tz = tlc_z(c.Tcw, l.Tcw);
        const bool fwd = tz > c.mb && !bMono, bwd = -tz > c.mb && !bMono;  // :1369-1370
        D.levelMode = fwd ? 1 : bwd ? 2 : 0;
        uint8_t* d = m->dBuf;
        if (curRes) {
            D.curKeys = rc[p].keys; D.curDesc = rc[p].desc; D.curUR = rc[p].uRight;
        } else {
            D.curKeys = (const orbx_keypoint*)(d + off[p].curKeys); D.curDesc = d + off[p].curDesc;
            D.curUR = c.u_right ? (const float*)(d + off[p].curUR) : nullptr;
        }
        D.lastKeys = lastRes ? rl[p].keys : (const orbx_keypoint*)(d + off[p].lastKeys);
        D.hasPoint = d + off[p].has; D.worldPos = (const float*)(d + off[p].pos);
        D.pointDesc = d + off[p].pdesc; D.observed = d + off[p].obs;
        D.cellStart = (int*)(d + off[p].cellStart); D.cellItems = (int*)(d + off[p].cellItems);
        D.best = (unsigned long long*)(d + off[p].best); D.proj = (float4*)(d + off[p].proj); D.events = (int*)(d + off[p].events);
        D.assigned = (int*)(d + off[p].assigned); D.nmatches = (int*)(d + oNm) + p;
        hd[p] = D;
        uint8_t* h = m->hIn;
        const size_t nc = c.n, nl = l.n;
        if (nc && !curRes) {
            std::memcpy(h + off[p].curKeys, c.keys_un, nc * 28);
            std::memcpy(h + off[p].curDesc, c.desc, nc * 32);
            if (c.u_right) std::memcpy(h + off[p].curUR, c.u_right, nc * 4);
        }
        if (nl && !lastRes) std::memcpy(h + off[p].lastKeys, l.keys_un, nl * 28);
        if (nl && !contiguous) {
            std::memcpy(h + off[p].has, mp.has_point, nl);
            std::memcpy(h + off[p].pos, mp.world_pos, nl * 12);
            std::memcpy(h + off[p].pdesc, mp.point_desc, nl * 32);
            std::memcpy(h + off[p].obs, mp.observed, nl);
        }
    }
    cudaStream_t s = m->stream;
    // device-resident sources: behind whatever their batch engines still have in flight
    {
        std::vector<orbx_batch*> seen;
        for (int p = 0; p < nPairs; ++p)
            for (orbx_batch* b : {curSrc ? curSrc[p].batch : nullptr, lastSrc ? lastSrc[p].batch : nullptr}) {
                if (!b || std::find(seen.begin(), seen.end(), b) != seen.end()) continue;
                seen.push_back(b);
                TRY(order_after_batch(m, b, err));
            }
    }
    if (contiguous && totalLast) {
        const orbx_mappoint_view& a = lastPoints[0];
        TRY(orbx_check(cudaMemcpyAsync(m->dBuf + oHas, a.has_point, totalLast, cudaMemcpyHostToDevice, s), "H2D map points", err));
        TRY(orbx_check(cudaMemcpyAsync(m->dBuf + oObs, a.observed, totalLast, cudaMemcpyHostToDevice, s), "H2D map points", err));
        TRY(orbx_check(cudaMemcpyAsync(m->dBuf + oPos, a.world_pos, totalLast * 12, cudaMemcpyHostToDevice, s), "H2D map points", err));
        TRY(orbx_check(cudaMemcpyAsync(m->dBuf + oDesc, a.point_desc, totalLast * 32, cudaMemcpyHostToDevice, s), "H2D map points", err));
    }
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, s), "H2D frames", err));
    const FrameDev* dd = (const FrameDev*)(m->dBuf + oDev);
    orbx_grid_kernel<<<dim3(1, nPairs), 256, 0, s>>>(dd);
    if (maxLast) orbx_project_search_kernel<<<dim3((maxLast + 7) / 8, nPairs), 256, 0, s>>>(dd);
    const size_t smem = (size_t)((maxCur + 31) / 32 + 1) * 4;
    if (smem > 32 * 1024) cudaFuncSetAttribute(orbx_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    orbx_resolve_kernel<<<dim3(1, nPairs), 32, smem, s>>>(dd);
    TRY(orbx_check(cudaGetLastError(), "search launch", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, m->dBuf + outBase, outBytes, cudaMemcpyDeviceToHost, s), "D2H matches", err));
    TRY(orbx_check(cudaStreamSynchronize(s), "sync", err));
    for (int p = 0; p < nPairs; ++p) {
        if (cur[p].n) std::memcpy(matchOfCur[p], m->hOut + (off[p].assigned - outBase), (size_t)cur[p].n * 4);
        nmatches[p] = *(const int*)(m->hOut + (oNm - outBase) + (size_t)p * 4);
    }
    return ORBX_OK;
}
}  // namespace

extern "C" int orbx_search_by_projection(orbx_matcher* m, const orbx_frame_view* cur, const orbx_frame_view* last,
                                         const orbx_mappoint_view* lastPoints, int nPairs, float th, int bMono, int checkOrientation,
                                         int* const* matchOfCur, int* nmatches)
{
    return search_by_projection_impl(m, cur, nullptr, last, nullptr, lastPoints, nPairs, th, bMono, checkOrientation, matchOfCur, nmatches);
}

extern "C" int orbx_search_by_projection_resident(orbx_matcher* m, const orbx_frame_view* cur, const orbx_frame_source* curSrc,
                                                  const orbx_frame_view* last, const orbx_frame_source* lastSrc,
                                                  const orbx_mappoint_view* lastPoints, int nPairs, float th, int bMono,
                                                  int checkOrientation, int* const* matchOfCur, int* nmatches)
{
    return search_by_projection_impl(m, cur, curSrc, last, lastSrc, lastPoints, nPairs, th, bMono, checkOrientation, matchOfCur, nmatches);
}

static int search_local_points_impl(orbx_matcher* m, const orbx_frame_view* frames, const orbx_frame_source* frameSrc,
                                    const uint8_t* const* occupied, const orbx_trackpoint_view* points, int nPairs, float th,
                                    float nnratio, int* const* matchOfKey, int* nmatches)
{
    if (!m || !frames || !points || nPairs < 1 || !matchOfKey || !nmatches) return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    std::string& err = orbx_err_slot();
    int maxPts = 0, maxKeys = 0;
    std::vector<ResidentFrame> rf(nPairs);
    for (int p = 0; p < nPairs; ++p) {
        if (frameSrc && frameSrc[p].batch) {  // keys_un / desc / u_right of the view are not read
            orbx_frame_view v = frames[p];
            v.keys_un = reinterpret_cast<const orbx_keypoint*>(&v); v.desc = reinterpret_cast<const uint8_t*>(&v);
            TRY(check_frame(&v, true));
            TRY(resolve_source(m, frameSrc[p], frames[p], rf[p]));
        } else {
            TRY(check_frame(frames + p, true));
        }
        const orbx_trackpoint_view& t = points[p];
        if (t.n < 0 || (t.n && (!t.valid || !t.proj || !t.level || !t.view_cos || !t.point_desc || !t.observed))) return orbx_fail(ORBX_ERR_ARG, "bad map point view");
        if (frames[p].n && !matchOfKey[p]) return orbx_fail(ORBX_ERR_ARG, "match_of_key is NULL");
        for (int i = 0; i < t.n; ++i)
            if (t.valid[i] && (t.level[i] < 0 || t.level[i] >= frames[p].nlevels)) return orbx_fail(ORBX_ERR_ARG, "predicted level outside the frame's scale table");
        maxPts = t.n > maxPts ? t.n : maxPts;
        maxKeys = frames[p].n > maxKeys ? frames[p].n : maxKeys;
    }
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));
    struct Off { size_t keys, desc, ur, occ, valid, proj, level, vc, pdesc, obs, cellStart, cellItems, best, best2, assigned; };
    std::vector<Off> off(nPairs);
    OrbxPacker pk;
    const size_t oF = pk.take(sizeof(FrameDev) * (size_t)nPairs), oM = pk.take(sizeof(MapDev) * (size_t)nPairs);
    for (int p = 0; p < nPairs; ++p) {
        const size_t nk = frames[p].n, np = points[p].n;
        const bool res = rf[p].keys != nullptr;
        off[p].keys = pk.take(res ? 0 : nk * 28); off[p].desc = pk.take(res ? 0 : nk * 32); off[p].ur = pk.take(!res && frames[p].u_right ? nk * 4 : 0);
        off[p].occ = pk.take(occupied && occupied[p] ? nk : 0);
        off[p].valid = pk.take(np); off[p].proj = pk.take(np * 12); off[p].level = pk.take(np * 4); off[p].vc = pk.take(np * 4);
        off[p].pdesc = pk.take(np * 32); off[p].obs = pk.take(np);
    }
    const size_t inBytes = pk.off;
    for (int p = 0; p < nPairs; ++p) {
        const size_t nk = frames[p].n, np = points[p].n;
        off[p].cellStart = pk.take((NCELL + 1) * 4); off[p].cellItems = pk.take(nk * 4 + 4);
        off[p].best = pk.take(np * 8 + 8); off[p].best2 = pk.take(np * 8 + 8);
    }
    const size_t outBase = pk.off;
    for (int p = 0; p < nPairs; ++p) off[p].assigned = pk.take((size_t)frames[p].n * 4);
    const size_t oNm = pk.take((size_t)nPairs * 4);
    const size_t outBytes = pk.off - outBase;
    TRY(orbx_matcher_grow(m, pk.off, inBytes, outBytes));
    FrameDev* hf = reinterpret_cast<FrameDev*>(m->hIn + oF);
    MapDev* hm = reinterpret_cast<MapDev*>(m->hIn + oM);
    for (int p = 0; p < nPairs; ++p) {
        const orbx_frame_view& c = frames[p];
        const orbx_trackpoint_view& t = points[p];
        const bool occ = occupied && occupied[p];
        uint8_t *d = m->dBuf, *h = m->hIn;
        FrameDev D;
        fill_frame(D, c);
        const bool res = rf[p].keys != nullptr;
        if (res) {
            D.curKeys = rf[p].keys; D.curDesc = rf[p].desc; D.curUR = rf[p].uRight;
        } else {
            D.curKeys = (const orbx_keypoint*)(d + off[p].keys); D.curDesc = d + off[p].desc;
            D.curUR = c.u_right ? (const float*)(d + off[p].ur) : nullptr;
        }
        D.cellStart = (int*)(d + off[p].cellStart); D.cellItems = (int*)(d + off[p].cellItems);
        D.best = (unsigned long long*)(d + off[p].best); D.assigned = (int*)(d + off[p].assigned); D.nmatches = (int*)(d + oNm) + p;
        hf[p] = D;
        MapDev X{};
        X.nPts = t.n; X.th = th; X.nnratio = nnratio;
        X.valid = d + off[p].valid; X.proj = (const float*)(d + off[p].proj); X.level = (const int*)(d + off[p].level);
        X.viewCos = (const float*)(d + off[p].vc); X.pdesc = d + off[p].pdesc; X.observed = d + off[p].obs;
        X.occupied = occ ? d + off[p].occ : nullptr;
        X.best2 = (unsigned long long*)(d + off[p].best2);
        hm[p] = X;
        const size_t nk = c.n, np = t.n;
        if (nk && !res) {
            std::memcpy(h + off[p].keys, c.keys_un, nk * 28);
            std::memcpy(h + off[p].desc, c.desc, nk * 32);
            if (c.u_right) std::memcpy(h + off[p].ur, c.u_right, nk * 4);
        }
        if (nk && occ) std::memcpy(h + off[p].occ, occupied[p], nk);
        if (np) {
            std::memcpy(h + off[p].valid, t.valid, np); std::memcpy(h + off[p].proj, t.proj, np * 12);
            std::memcpy(h + off[p].level, t.level, np * 4); std::memcpy(h + off[p].vc, t.view_cos, np * 4);
            std::memcpy(h + off[p].pdesc, t.point_desc, np * 32); std::memcpy(h + off[p].obs, t.observed, np);
        }
    }
    cudaStream_t s = m->stream;
    {
        std::vector<orbx_batch*> seen;
        for (int p = 0; frameSrc && p < nPairs; ++p) {
            orbx_batch* b = frameSrc[p].batch;
            if (!b || std::find(seen.begin(), seen.end(), b) != seen.end()) continue;
            seen.push_back(b);
            TRY(order_after_batch(m, b, err));
        }
    }
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, s), "H2D local map", err));
    const FrameDev* df = (const FrameDev*)(m->dBuf + oF);
    const MapDev* dm = (const MapDev*)(m->dBuf + oM);
    orbx_grid_kernel<<<dim3(1, nPairs), 256, 0, s>>>(df);
    if (maxPts) orbx_map_search_kernel<<<dim3((maxPts + 7) / 8, nPairs), 256, 0, s>>>(df, dm);
    const size_t smem = (size_t)((maxKeys + 31) / 32 + 1) * 4;
    if (smem > 32 * 1024) cudaFuncSetAttribute(orbx_map_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    orbx_map_resolve_kernel<<<dim3(1, nPairs), 32, smem, s>>>(df, dm);
    TRY(orbx_check(cudaGetLastError(), "local map search launch", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, m->dBuf + outBase, outBytes, cudaMemcpyDeviceToHost, s), "D2H matches", err));
    TRY(orbx_check(cudaStreamSynchronize(s), "sync", err));
    for (int p = 0; p < nPairs; ++p) {
        if (frames[p].n) std::memcpy(matchOfKey[p], m->hOut + (off[p].assigned - outBase), (size_t)frames[p].n * 4);
        nmatches[p] = *(const int*)(m->hOut + (oNm - outBase) + (size_t)p * 4);
    }
    return ORBX_OK;
}

extern "C" int orbx_search_local_points(orbx_matcher* m, const orbx_frame_view* frames, const uint8_t* const* occupied,
                                        const orbx_trackpoint_view* points, int nPairs, float th, float nnratio, int* const* matchOfKey,
                                        int* nmatches)
{
    return search_local_points_impl(m, frames, nullptr, occupied, points, nPairs, th, nnratio, matchOfKey, nmatches);
}

extern "C" int orbx_search_local_points_resident(orbx_matcher* m, const orbx_frame_view* frames, const orbx_frame_source* frameSrc,
                                                 const uint8_t* const* occupied, const orbx_trackpoint_view* points, int nPairs, float th,
                                                 float nnratio, int* const* matchOfKey, int* nmatches)
{
    return search_local_points_impl(m, frames, frameSrc, occupied, points, nPairs, th, nnratio, matchOfKey, nmatches);
}
