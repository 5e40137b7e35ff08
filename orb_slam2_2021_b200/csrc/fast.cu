// fast.cu — per-cell FAST-9/16 with non-max suppression and the per-cell threshold fallback.
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (/root/reference/src/ORBextractor.cc:792-832) and the two cv::FAST calls inside it (:812, :817).
//
// One warp per 30-px cell (all levels and all images in one launch). Every interior pixel of the
// cell's ROI gets its corner strength
//     m = max over the 16 arcs of 9 contiguous ring pixels of min(+-(ring - centre))
// which is threshold independent: the pixel is a FAST corner at threshold t iff m > t and OpenCV's
// score is m-1. NMS keeps a pixel iff m is strictly greater than the strengths of its 8 neighbours
// (pixels outside the ROI interior count 0). Because a neighbour that is not a corner at t has
// m <= t < m_centre, the NMS at threshold t only needs the strengths above t. The cell is first run at
// iniTh (cheap pre-test, network, map and NMS all on the pixels that can be corners at iniTh); only a
// cell whose FAST(iniTh) comes back empty (:814) repeats at minTh. Keypoints are written row-major into
// the cell's slot array, so cells concatenated in table order give the reference's candidate order.
//
// Blackwell mapping (measured with tools/ubench/pipes.cu: VIMNMX3.U16x2 issues at the same rate as a 32-bit IMNMX, so
// packing two pixels per register halves the min/max work). The kernel is bound by the ALU pipe (min/max, LOP3, PRMT,
// SHF, ISETP, VABSDIFF4) with the issue slots close behind, while the FMA pipe (IMAD) and the load/store unit have room:
// every 1:1 move of an operation from the ALU pipe to IMAD paid, every 1:2 move lost (numbers at fast_strength2_raw).
//   * the ROI arrives by ONE TMA box load (cp.async.bulk.tensor.3d, 64 or 96 columns starting at x0 & ~15 — the copy
//     engine addresses the innermost dimension in 16-byte units — by rhMax rows) signalled on an mbarrier;
//   * pass 1 runs a byte-wise NECESSARY test for a corner at the attempt's threshold on the interior words of the raw
//     tile, 4 pixels at once (VABSDIFF4 + SWAR compare against the compass points); the surviving pixel
//     PAIRS are appended to the cell's task list step by step with ballots, a word's two pairs side by side, so the list
//     is sorted by (row, column); list positions are running shared-memory addresses advanced by multiply-adds;
//   * pass 2 runs the exact test on the listed pairs only. A thread owns two horizontally adjacent pixels (even tile
//     column first) and reads the 17 pairs (ring(x), ring(x+1)) straight from the raw tile (two byte loads and one
//     multiply-add each: a widened u16 copy of the ROI cost more than it saved). Min and max commute with
//     subtracting the centre, so the whole min/max network runs on the raw unsigned u16x2 ring values with the 3-input
//     DPX instructions (__vimin3_u16x2 / __vimax3_u16x2): 36 ops per polarity and pixel PAIR (pair and quad minima shared
//     by the two arcs that start at k and k+1);
//   * strengths stay packed to the end; m - thr goes to a byte map for m > thr; NMS is driven by the list of non-zero pairs
//     and uses packed column maxima; the list of non-zero pairs and the list of maxima are order-preserving compactions of
//     the sorted task list, so the maxima come out in the reference's row-major order and are emitted as they stand (the
//     rank computation of the earlier version — one compare per pair of maxima — is gone);
//   * one warp per cell, one cell per CTA at a time, long-lived warps: see the kernel.
#include "engine.h"
#include "orbx_dev.cuh"
#include "orbx_tma.cuh"

struct __align__(8) OrbxCell {  // 8 bytes, loaded as one 64-bit word
    short x0, y0;        // ROI origin in level coordinates (iniX, iniY)
    uint8_t rw, rh;      // ROI size (maxX-iniX, maxY-iniY)
    uint8_t level, pad;
};


// lo | hi << 16 as ONE multiply-add: the kernel is bound by the ALU pipe (min/max, PRMT, LOP3, SHF), the FMA pipe that
// executes IMAD is nearly idle — mad.lo keeps ptxas from turning the packing back into PRMT / LEA
__device__ __forceinline__ uint32_t orbx_pack_u16x2(uint32_t lo, uint32_t hi)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, 65536, %2;" : "=r"(r) : "r"(hi), "r"(lo));
    return r;
}
// one byte from shared memory at a compile-time offset (explicit, so that two neighbouring bytes stay two loads: merged into
// one 16-bit load they need a PRMT to come apart, and the ALU pipe is the one without room)
template <int OFF>
__device__ __forceinline__ uint32_t orbx_lds_u8(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}
__device__ __forceinline__ uint32_t orbx_mad(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ void orbx_sts_u16(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((uint16_t)v) : "memory");
}
__device__ __forceinline__ void orbx_sts_u8(uint32_t addr, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void orbx_sts_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// Corner strengths ABOVE thr (0 = not a corner at thr) of two horizontally adjacent pixels from the 17 u16x2 vectors cc
// (centre pair) and d[0..15] (ring pairs).
__device__ __forceinline__ uint32_t fast_network2(uint32_t cc, const uint32_t (&d)[16], uint32_t thr)
{
    // the network runs on the raw ring values (min and max commute with subtracting the centre): no per-value arithmetic
    // Arcs k and k+1 (k even) share the 8 values d[k+1..k+8]: max(min arc k, min arc k+1) = min(min8, max(d[k], d[k+9])).
    // Pair minima P (odd starts), quad minima Q = 4 contiguous values, then one 3-input op per even k: 36 operations per
    // polarity instead of 40 with 3/9-windows. The dark side is the dual (min <-> max).
    uint32_t pl[8], ph[8];  // j = 2i+1: min / max of d[j], d[j+1]
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        pl[i] = __vimin3_u16x2(d[2 * i + 1], d[(2 * i + 2) & 15], d[(2 * i + 2) & 15]);
        ph[i] = __vimax3_u16x2(d[2 * i + 1], d[(2 * i + 2) & 15], d[(2 * i + 2) & 15]);
    }
    uint32_t ql[8], qh[8];  // d[j .. j+3], j = 2i+1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        ql[i] = __vimin3_u16x2(pl[i], pl[(i + 1) & 7], pl[(i + 1) & 7]);
        qh[i] = __vimax3_u16x2(ph[i], ph[(i + 1) & 7], ph[(i + 1) & 7]);
    }
    uint32_t tl[8], th[8];  // k = 2i: best of arcs k and k+1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint32_t el = __vimax3_u16x2(d[2 * i], d[(2 * i + 9) & 15], d[(2 * i + 9) & 15]);
        const uint32_t eh = __vimin3_u16x2(d[2 * i], d[(2 * i + 9) & 15], d[(2 * i + 9) & 15]);
        tl[i] = __vimin3_u16x2(ql[i], ql[(i + 2) & 7], el);  // d[k+1..k+4], d[k+5..k+8]
        th[i] = __vimax3_u16x2(qh[i], qh[(i + 2) & 7], eh);
    }
    uint32_t br = __vimax3_u16x2(__vimax3_u16x2(tl[0], tl[1], tl[2]), __vimax3_u16x2(tl[3], tl[4], tl[5]), __vimax3_u16x2(tl[6], tl[7], tl[7]));
    uint32_t dk = __vimin3_u16x2(__vimin3_u16x2(th[0], th[1], th[2]), __vimin3_u16x2(th[3], th[4], th[5]), __vimin3_u16x2(th[6], th[7], th[7]));
    // strength above the threshold, max(brightest arc minimum - centre - thr, centre - darkest arc maximum - thr, 0), per half
    // and still packed. centre + thr and darkest + thr stay below 2^16 per half and each difference is clamped at 0 by the
    // max in front of it, so the 32-bit adds / subtracts cannot carry across the halves; a pixel cannot be a bright and a
    // dark corner at once (9 + 9 > 16 ring pixels), so the sum of the two terms is their maximum. The adds are multiply-adds
    // (FMA pipe); only the two max operations use the ALU pipe.
    const uint32_t hi = orbx_mad(thr, 0x10001u, cc), dkt = orbx_mad(thr, 0x10001u, dk);
    const uint32_t up = orbx_mad(hi, 0xFFFFFFFFu, __vimax3_u16x2(br, hi, hi));
    const uint32_t dn = orbx_mad(dkt, 0xFFFFFFFFu, __vimax3_u16x2(cc, dkt, dkt));
    return up + dn;
}

// The 17 vectors come straight from the raw tile (no widened copy of the ROI: building one cost more instructions and 7 KB
// of shared memory per cell than it saved here). t points at pixel 0 of the pair; every pair is two byte loads and one
// multiply-add (lo + 65536 * hi, FMA pipe). Measured on B200, 512 KITTI images: pairs at even dx as one 16-bit load + PRMT
// and the rest as two byte loads + PRMT (the form before): 1.322 ms; byte loads + IMAD for the odd ones only: 1.295;
// for all: 1.229 with the lean list bookkeeping below (1.234 with PRMT on the even ones) — the ALU pipe (min/max, PRMT,
// LOP3, SHF, ISETP) is what binds this kernel, the load/store unit and the FMA pipe have room. Replacing one PRMT by TWO
// multiplies (IMAD.HI + IMAD on a 16-bit load) loses (1.331): issue slots count too.
// BW = row stride of the tile.
template <int BW>
__device__ __forceinline__ uint32_t fast_strength2_raw(const uint8_t* t, uint32_t thr)
{
#define RING(dx, dy) orbx_pack_u16x2(orbx_lds_u8<(dy) * BW + (dx)>(ta), orbx_lds_u8<(dy) * BW + (dx) + 1>(ta))
    const uint32_t ta = orbx_smem_addr(t);
    const uint32_t cc = RING(0, 0);
    uint32_t d[16];
    d[0] = RING(0, 3);    d[1] = RING(1, 3);    d[2] = RING(2, 2);    d[3] = RING(3, 1);
    d[4] = RING(3, 0);    d[5] = RING(3, -1);   d[6] = RING(2, -2);   d[7] = RING(1, -3);
    d[8] = RING(0, -3);   d[9] = RING(-1, -3);  d[10] = RING(-2, -2); d[11] = RING(-3, -1);
    d[12] = RING(-3, 0);  d[13] = RING(-3, 1);  d[14] = RING(-2, 2);  d[15] = RING(-1, 3);
#undef RING
    return fast_network2(cc, d, thr);
}

// Pass-1 test of the 4 pixels of one tile word: bit 7 of byte i of the result is set if pixel i can be a corner at the
// threshold behind K (see the kernel): (|d0|>t or |d8|>t) and (|d4|>t or |d12|>t) on the compass points
template <int BW>
__device__ __forceinline__ uint32_t fast_pretest4(const uint32_t* tw, uint32_t K, uint32_t one)
{
    const uint32_t w = tw[0], wn = tw[1], wp = tw[-1];
    const uint32_t up = tw[-3 * (BW / 4)], dn = tw[3 * (BW / 4)];
    const uint32_t lf = __funnelshift_r(wp, w, 8), rt = __funnelshift_r(w, wn, 24);  // columns -3 / +3
    const uint32_t a0 = __vabsdiffu4(dn, w), a8 = __vabsdiffu4(up, w), a4 = __vabsdiffu4(rt, w), a12 = __vabsdiffu4(lf, w);
    // bit 7 of byte i of (a + K) | a, K = 127 - t in every byte, is set iff a_i > t — or a_i == t with a carry out of byte
    // i - 1 (a_{i-1} >= 129 + t): a byte that overflows has its own bit 7 restored by "| a" (it is >= 129), and the carry
    // it sends up can only turn a_i == t into a pass. The test stays NECESSARY (pass 2 is exact), so the four masking ANDs
    // of the carry-free form ((a & 0x7F7F7F7F) + K) | a are not needed: 8 ALU-pipe operations less per trip
    // ('one' is the kernel argument 1: a + K as the multiply-add a * one + K stays on the FMA pipe whatever ptxas thinks of
    // the pipe balance — as plain adds they came out as ALU-pipe VIADDs in some builds and as IMAD.IADDs in others)
    const uint32_t f0 = orbx_mad(a0, one, K) | a0, f8 = orbx_mad(a8, one, K) | a8;
    const uint32_t f4 = orbx_mad(a4, one, K) | a4, f12 = orbx_mad(a12, one, K) | a12;
    return (f0 | f8) & (f4 | f12);
}

// c_recip[n] = ceil(65536 / n): floor(t / n) = (t * c_recip[n]) >> 16 exactly while t * n < 65536 (here t <= 128, n <= 31)
__constant__ uint32_t c_recip[32] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4096, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115};

// One WARP per cell, CELLS cells per CTA (default 1). A cell used to belong to a 64-thread CTA; the fixed work per
// (warp, cell) — prologue, thread mapping, list flush, emission — is paid once per cell, there is no CTA barrier (only
// __syncwarp), and the list counters are warp-uniform. Warps never talk to each other; with CELLS = 1 the compiler can
// prove which values are warp-uniform (everything derived from blockIdx and the cell record) and keeps them on the uniform
// datapath; the SM then holds 28 one-warp CTAs of 7296 B (KITTI cells), the same 28 warps as seven 4-warp CTAs.
template <int SA, int CELLS>
__global__ void __launch_bounds__(CELLS * 32) orbx_fast_kernel(OrbxParams p, const OrbxCell* cells, const CUtensorMap* maps, int rhMax, int listCap,
                                                                    int sliceBytes, uint32_t one)
{
    constexpr int BW = SA == 48 ? 80 : 112;
    extern __shared__ __align__(128) uint8_t fast_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* T = fast_smem + (size_t)warp * sliceBytes;  // raw box [rhMax][BW]
    // strength map: ONE BYTE per pixel (m - thr <= 254). As u16 it cost 2 KB more per warp and capped the SM at 24 warps;
    // with 7.3 KB per warp 28 fit
    uint8_t* smap = T + ((rhMax * BW + 127) & ~127);
    uint16_t* tasks = reinterpret_cast<uint16_t*>(smap + (rhMax - 4) * SA);  // pass-1 survivors (pixel pairs)
    // pairs with a non-zero strength: written over the task list while pass 2 consumes it (entry k of nz is written in the
    // trip that read tasks[32*(trip)..], and k never runs ahead of what has been read)
    uint16_t* nz = tasks;
    // Shared-window addresses of the slice's parts, converted once: the lists, the map stores, the barrier and the TMA
    // destination are addressed through them. (Through an opaque move: otherwise the compiler re-derives the window base —
    // S2R SR_CgaCtaId, 0x400, LEA — at every use.) The warp's mbarrier lives in the last 8 bytes of its slice: no static
    // shared memory, so that with one warp per CTA the 7296-byte slice of a KITTI cell leaves room for exactly 28 CTAs per SM.
    uint32_t sT;
    asm volatile("mov.u32 %0, %1;" : "=r"(sT) : "r"(orbx_smem_addr(T)));
    const uint32_t sMap = sT + (uint32_t)(smap - T), sTasks = sMap + (uint32_t)((rhMax - 4) * SA);
    const uint32_t bar = sT + (uint32_t)sliceBytes - 8;
    if (lane == 0) {
        orbx_mbar_init_a(bar, 1);
        orbx_fence_mbar_init();
    }
    __syncwarp();  // barrier initialised before anybody polls it
    // long-lived warps: the grid is a small multiple of what the SMs hold and every warp walks over several (image, cell)
    // pairs — little CTA turnover (a third of the warp slots stood empty with one short-lived CTA per two cells), and
    // neighbouring warps work on neighbouring cells of the same image (their tiles overlap in L2)
    const int total = p.cellsTotal * p.imgs;
    unsigned phase = 0;
    // (image, cell) of the warp's work item, advanced without a division per cell
    const int stride = gridDim.x * CELLS, strideImg = stride / p.cellsTotal, strideCell = stride - strideImg * p.cellsTotal;
    int img = (blockIdx.x * CELLS + warp) / p.cellsTotal, cellId = (blockIdx.x * CELLS + warp) - img * p.cellsTotal;
#pragma unroll 1
    for (int work = blockIdx.x * CELLS + warp; work < total;
         work += stride, img += strideImg, cellId += strideCell, img += cellId >= p.cellsTotal, cellId -= cellId >= p.cellsTotal ? p.cellsTotal : 0) {
    const OrbxCell cell = cells[cellId];
    const OrbxLevel& L = p.lv[cell.level];
    const int iw = cell.rw - 6, ih = cell.rh - 6;  // interior = detection domain of cv::FAST
    int* cntOut = p.cellCnt + (size_t)img * p.cellsTotal + cellId;
    if (iw <= 0 || ih <= 0) {
        if (lane == 0) *cntOut = 0;
        continue;
    }
    if (lane == 0) {
        orbx_fence_mbar_init();  // the previous cell's kept[] (generic-proxy writes into the tile) before the TMA write
        orbx_mbar_expect_tx_a(bar, rhMax * BW);
        orbx_tma_load_3d_a(sT, maps + cell.level, cell.x0 & ~15, cell.y0, img + p.imgBase, bar);
    }
    {
        const int nzero = (ih + 2) * (SA / 16);
#pragma unroll 1
        for (int i = lane; i < nzero; i += 32) asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(sMap + 16u * i), "r"(0) : "memory");
    }
    __syncwarp();
    orbx_mbar_wait_a(bar, phase);
    phase ^= 1;

    // ---- threshold attempts (:812-819): first iniTh; only a cell whose FAST(iniTh) comes back empty repeats at minTh. A
    // neighbour that is not a corner at t has m <= t < m_centre, so the NMS at threshold t only needs the strengths m > t:
    // everything below (pre-test, network, map, NMS) runs on the pixels that can be corners at the attempt's threshold.
    // The strength map of a failed first attempt needs no clearing: the minTh attempt rewrites a superset of its entries.
    // Strength map, row stride SA: interior pixel (ix,iy) at [iy+1][A-column + SD], zero elsewhere. "A-column" = level
    // column minus (x0 & ~3): pixel pairs are aligned to even A-columns (= even tile bytes).
    const int tmin = p.minTh, tini = p.iniTh;
    const int o = cell.x0 & 3;
    const int ca0 = o + 3;                   // A-column of interior pixel 0
    const int SD = 2 + ((o + 1) & 1) - ca0;  // strength-map column = A-column + SD
    const uint32_t* tw0 = reinterpret_cast<const uint32_t*>(T) + ((cell.x0 & 15) >> 2);
    // kept[] = the NMS maxima, key | m << 16, written over the head of the tile (address sT; the raw tile is dead by then)
    const unsigned lt = (1u << lane) - 1;
    // pre-test mapping: fixed interior word column per lane, 32/nwi interior rows per step
    const int jmin = ca0 >> 2, nwi = ((ca0 + iw - 1) >> 2) - jmin + 1;
    const uint32_t rcpi = c_recip[nwi];
    const int pStep = (int)((32u * rcpi) >> 16);
    const int pyq = (int)(((uint32_t)lane * rcpi) >> 16);
    const int pj = jmin + lane - pyq * nwi;
    const unsigned vmask = pyq >= pStep ? 0u : ((4 * pj + 1 >= ca0 && 4 * pj < ca0 + iw) ? 1u : 0u) | ((4 * pj + 3 >= ca0 && 4 * pj + 2 < ca0 + iw) ? 2u : 0u);
    const uint32_t* ptw0 = tw0 + (pyq + 3) * (BW / 4) + pj;
    int thr = tini, nkept = 0;
#pragma unroll 1
    for (int attempt = 0;; ++attempt) {
        // ---- pass 1: a cheap NECESSARY test for "corner at thr" on the raw words: any 9 contiguous ring positions contain
        // two ADJACENT compass points (0,4,8,12), so a corner needs (|d0|>t or |d8|>t) and (|d4|>t or |d12|>t). Four
        // VABSDIFF4 + a SWAR byte compare test 4 pixels at once (fast_pretest4). The surviving pixel PAIRS (aligned to even
        // tile columns) are appended step by step with ballots: no atomics, and the list comes out in row-major order.
        // bit 7 of each byte of f(x) = (x + K) | x is set if x > min(thr, 127) (fast_pretest4): exact below 127 up to a rare
        // carry, still necessary above
        int ntAll = 0;
        const uint32_t tp0 = sTasks;
        uint32_t tp = tp0;  // shared-memory address of the list's end
        {
            // survivors are appended step by step with ballots, so the task list is in row-major order: the 32 tasks of a
            // pass-2 trip come from one to three tile rows and their ring loads spread over the banks
            const uint32_t K = 0x01010101u * (uint32_t)(0x7F - min(thr, 0x7F));
            const uint32_t* tw = ptw0;
            const uint32_t maskLo = (vmask & 1u) ? 0x00008080u : 0u, maskHi = (vmask & 2u) ? 0x80800000u : 0u;
            uint32_t code = (uint32_t)(pyq * 256 + 2 * pj);
            const uint32_t cs = (uint32_t)pStep << 8;
            // one step = the two pair flags of the lane's word. A word's two pairs go to adjacent list slots (lanes ascend
            // with (row, word column)), so the list is sorted by (row, column): the non-zero list and the NMS maxima inherit
            // the order and need no ranking at the end. The list position is a running shared-memory byte address; scaling
            // and adding are multiply-adds (FMA pipe)
            auto append = [&](uint32_t m, uint32_t cd) {
                const bool fl = (m & maskLo) != 0, fh = (m & maskHi) != 0;
                const unsigned bl = __ballot_sync(0xffffffffu, fl), bh = __ballot_sync(0xffffffffu, fh);
                uint32_t a = orbx_mad(__popc(bl & lt) + __popc(bh & lt), 2u, tp);
                if (fl) orbx_sts_u16(a, cd);
                if (fl) a += 2;
                if (fh) orbx_sts_u16(a, cd + 1);
                tp = orbx_mad(__popc(bl) + __popc(bh), 2u, tp);
            };
            // two steps per trip. Trips whose rows all lie inside the interior run without row predicates (idle lanes —
            // 32 is not a multiple of every row length — read words of their own slice and are masked by maskLo / maskHi);
            // the last, partial trip is guarded
            int y0s = 0;
#pragma unroll 1
            for (; y0s + 2 * pStep <= ih; y0s += 2 * pStep) {
                const uint32_t m0 = fast_pretest4<BW>(tw, K, one), m1 = fast_pretest4<BW>(tw + pStep * (BW / 4), K, one);
                tw += 2 * pStep * (BW / 4);
                append(m0, code);
                append(m1, code + cs);
                code += 2 * cs;
            }
            if (y0s < ih) {  // the partial trip: one or two guarded steps (warp-uniform choice)
                const bool p0 = vmask && y0s + pyq < ih;
                uint32_t m0 = 0;
                if (p0) m0 = fast_pretest4<BW>(tw, K, one);
                append(m0, code);
                if (y0s + pStep < ih) {
                    const bool p1 = vmask && y0s + pStep + pyq < ih;
                    uint32_t m1 = 0;
                    if (p1) m1 = fast_pretest4<BW>(tw + pStep * (BW / 4), K, one);
                    append(m1, code + cs);
                }
            }
        }
        ntAll = (int)((tp - tp0) >> 1);
        __syncwarp();  // task list complete
        // ---- pass 2: the min/max network on the surviving pairs
        int nzAll = 0;
        const uint32_t smapA = sMap + SD;  // shared-window address of strength-map column 0 of row -1
        uint32_t np = tp0;  // end of the non-zero list (it overwrites the task list from its start)
#pragma unroll 1
        for (int g0 = 0; g0 < ntAll; g0 += 32) {
            // no branch around the body: lanes past the end of the list repeat its last task and drop the result (the
            // branch made the flag a byte in a register, set and tested on both sides of the reconvergence point)
            const bool valid = g0 + lane < ntAll;
            const int task = tasks[min(g0 + lane, ntAll - 1)];
            const int y = task >> 8, c = 2 * (task & 255);
            const uint32_t mp = fast_strength2_raw<BW>(reinterpret_cast<const uint8_t*>(tw0) + (y + 3) * BW + c, (uint32_t)thr);
            // mp = m - thr per half, 0 = no corner. A pair at the edge of the interior has one pixel outside it (interior
            // column -1 or iw); its strength is cleared from the map after the loop instead of being masked here
            const bool hit = valid && mp != 0;
            if (hit) orbx_sts_u16(smapA + (y + 1) * SA + c, __byte_perm(mp, 0, 0x4420));  // map is pre-zeroed
            const unsigned bh = __ballot_sync(0xffffffffu, hit);
            if (hit) orbx_sts_u16(orbx_mad(__popc(bh & lt), 2u, np), task);
            np = orbx_mad(__popc(bh), 2u, np);
        }
        nzAll = (int)((np - tp0) >> 1);
        if (nzAll > 0 && ((ca0 | iw) & 1)) {
            // pixel pairs are aligned to even tile columns: with an odd first interior column the pair in front of it holds
            // interior column -1, with an odd end the last pair holds column iw. cv::FAST does not look at them: their map
            // entries are zeroed (as neighbours they count 0, as centres they cannot win)
            __syncwarp();
            const bool oddStart = ca0 & 1, oddEnd = (ca0 + iw) & 1;
#pragma unroll 1
            for (int r = lane; r < ih; r += 32) {
                const uint32_t a = smapA + (r + 1) * SA + ca0;
                if (oddStart) orbx_sts_u8(a - 1, 0);
                if (oddEnd) orbx_sts_u8(a + iw, 0);
            }
        }
        __syncwarp();  // strength map and list complete; the raw tile is dead from here (kept[] takes its place)

        // ---- NMS over the non-zero pairs: strict maximum over the 8 neighbours in the strength map
        nkept = 0;
        const uint32_t kp0 = sT;
        uint32_t kp = kp0;
#pragma unroll 1
        for (int g0 = 0; g0 < nzAll; g0 += 32) {
            const bool valid = g0 + lane < nzAll;  // (lanes past the end repeat the last entry, as in pass 2)
            const int task = nz[min(g0 + lane, nzAll - 1)], y = task >> 8, cc = 2 * (task & 255);
            const uint8_t* c = smap + (y + 1) * SA + cc + SD;  // centre pair (even byte address)
#define MAP2(q) __byte_perm((uint32_t)*reinterpret_cast<const uint16_t*>(q), 0, 0x4140)  /* two map bytes -> u16x2 */
            const uint32_t mm = MAP2(c);
            const uint32_t u0 = MAP2(c - SA - 2), u1 = MAP2(c - SA), u2 = MAP2(c - SA + 2);
            const uint32_t l0 = MAP2(c - 2), l2 = MAP2(c + 2);
            const uint32_t d0 = MAP2(c + SA - 2), d1 = MAP2(c + SA), d2 = MAP2(c + SA + 2);
#undef MAP2
            const uint32_t cm0 = __vimax3_u16x2(u0, l0, d0), cm1 = __vimax3_u16x2(u1, mm, d1), cm2 = __vimax3_u16x2(u2, l2, d2);
            const uint32_t ud = __vimax3_u16x2(u1, d1, d1);
            // neighbourhood maxima of both pixels, packed: n0 = max(cm0.hi, cm1.hi, ud.lo), n1 = max(cm2.lo, cm1.lo, ud.hi)
            const uint32_t nn = __vimax3_u16x2(__byte_perm(cm0, cm2, 0x5432), __byte_perm(cm1, 0, 0x1032), ud);
            // m > n per half  <=>  max(m, n + 1) == m  (n + 1 <= 255: no carry between the halves)
            const uint32_t n1 = nn + 0x10001u;
            const uint32_t ge = __vimax3_u16x2(mm, n1, n1) ^ mm;
            const bool k0 = valid && (ge & 0x0000FFFFu) == 0, k1 = valid && (ge & 0xFFFF0000u) == 0;
            // key | (m + thr) << 16 for both pixels; mm << 16 keeps the low half only
            const uint32_t kt = (uint32_t)(y * 64 + cc - ca0) + ((uint32_t)thr << 16);
            const uint32_t e0 = orbx_mad(mm, 65536u, kt);
            const uint32_t e1 = (mm & 0xFFFF0000u) + kt + 1;
            // (the first trip overwrites the head of the dead tile only after every lane has read its map words: the map
            // and the lists live behind the tile, so kept[] never aliases what this loop still reads)
            const unsigned b0 = __ballot_sync(0xffffffffu, k0), b1 = __ballot_sync(0xffffffffu, k1);
            uint32_t a = orbx_mad(__popc(b0 & lt) + __popc(b1 & lt), 4u, kp);  // a pair's two pixels side by side: kept[] stays row-major
            if (k0) orbx_sts_u32(a, e0);
            if (k0) a += 4;
            if (k1) orbx_sts_u32(a, e1);
            kp = orbx_mad(__popc(b0) + __popc(b1), 4u, kp);
        }
        nkept = (int)((kp - kp0) >> 2);
        __syncwarp();  // publishes kept[]
        if (nkept > 0 || attempt == 1 || tmin >= thr) break;
        thr = tmin;  // :814 FAST(iniTh) came back empty: once more with minTh
    }

    // ---- emission: kept[] is already in row-major order (the reference's candidate order inside a cell) — the task list is
    // built sorted, the non-zero list and the maxima are order-preserving compactions of it
    const uint32_t slot0 = L.slotBase + (uint32_t)(cellId - L.cellBase) * L.slotCap;
    uint32_t* out = p.cellKp + (size_t)img * p.slotTotal + slot0;
    if (lane == 0) *cntOut = nkept;
#pragma unroll 1
    for (int i = lane; i < nkept; i += 32) {
        uint32_t e;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(sT + 4u * i) : "memory");
        const int key = e & 0xFFFFu, m = e >> 16;
        out[i] = (uint32_t)((key & 63) + 3) | ((uint32_t)((key >> 6) + 3) << 8) | ((uint32_t)(m - 1) << 16);
    }
    __syncwarp();  // every lane is done with kept[] (= the tile) before the next cell's TMA lands in it
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
    *boxW = rwMax + 8 <= 48 ? 80 : 112;  // (needs 15 + ROI + 1 columns; the rest spreads the rows over the banks)
    *rhMaxOut = rhMax;
}

void orbx_launch_fast(const OrbxParams& p, const void* cellTable, const void* levelMaps, cudaStream_t s)
{
    int boxW, rhMax;
    orbx_fast_geometry(p, &boxW, &rhMax);
    const int sa = boxW == 80 ? 48 : 80;
    // list capacity: every pixel pair that can hold an interior pixel (interior = ROI - 6; one more pair for the alignment)
    const int iwMax = (sa == 48 ? 40 : 72) - 6, ihMax = rhMax - 6;
    const int listCap = (((iwMax + 4) / 2) * ihMax + 7) & ~7;
    // shared memory of one cell: tile + strength map + one list (the non-zero pairs overwrite the tasks)
    const int slice = (int)(((size_t)((rhMax * boxW + 127) & ~127) + (size_t)(rhMax - 4) * sa + (size_t)listCap * 2 + 8 /* mbarrier */ + 127) & ~(size_t)127);
    // cells (= warps) per CTA: tuning key fast_cells = 1 / 2 / 4 / 8. One warp per CTA is the default: everything that depends
    // on the cell alone is then provably warp-uniform for the compiler (uniform registers and the uniform datapath instead of
    // ALU-pipe instructions: 55 registers instead of 70, no divergence handling around the ballots) — 1.154 ms per 512 KITTI
    // images against 1.190 for 2 or 4 cells per CTA; 8 loses a third (tools/fast_cells_sweep.sh)
    int cellsPerCta = orbx_tune(ORBX_TUNE_FAST_CELLS);
    if (cellsPerCta != 1 && cellsPerCta != 2 && cellsPerCta != 4 && cellsPerCta != 8) cellsPerCta = 1;
    // CTAs launched per resident CTA slot (tuning key fast_grid_mult). 1 = fully persistent with static striding, which loses
    // to its own tail (cells differ in cost: 1.54 ms per 512 KITTI images); 16 lets the hardware scheduler balance
    // ~9-cell work packets (1.37 ms); one cell per warp pays the CTA turnover (1.43 ms): tools/fast_grid_sweep.sh
    const int gridMult = orbx_tune(ORBX_TUNE_FAST_GRID_MULT) > 0 ? orbx_tune(ORBX_TUNE_FAST_GRID_MULT) : 16;
    static int smCounts[64] = {0};  // per device ordinal
    int dev = 0;
    cudaGetDevice(&dev);
    int& smCount = smCounts[dev & 63];
    if (!smCount) cudaDeviceGetAttribute(&smCount, cudaDevAttrMultiProcessorCount, dev);
    size_t smem = (size_t)slice * cellsPerCta;
    // tuning key fast_ctas: cap the resident CTAs per SM (by asking for more shared memory than a CTA uses), so that CTAs of
    // kernels running beside this one on other streams find room on every SM
    if (const int capCtas = orbx_tune(ORBX_TUNE_FAST_CTAS)) {
        const size_t want = (size_t)(233472 / (capCtas + 1)) + 1024;  // just too large for capCtas + 1 CTAs
        if (want > smem && want <= 227 * 1024) smem = want;
    }
    const int wanted = (p.cellsTotal * p.imgs + cellsPerCta - 1) / cellsPerCta;
#define FAST_LAUNCH(SA_, C_)                                                                                                     \
    {                                                                                                                            \
        if (smem > 32 * 1024) /* static + dynamic may pass 48 KB; per-device attribute, set on every such launch */            \
            cudaFuncSetAttribute(orbx_fast_kernel<SA_, C_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);             \
        int perSm = 1;                                                                                                           \
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, orbx_fast_kernel<SA_, C_>, C_ * 32, smem);                         \
        const long long cap = (long long)(perSm > 0 ? perSm : 1) * smCount * gridMult;                                           \
        orbx_fast_kernel<SA_, C_><<<wanted < cap ? wanted : (int)cap, C_ * 32, smem, s>>>(p, (const OrbxCell*)cellTable,         \
                                                                                          (const CUtensorMap*)levelMaps, rhMax, listCap, slice, 1u); \
    }
    if (sa == 48) {
        if (cellsPerCta == 1) FAST_LAUNCH(48, 1) else if (cellsPerCta == 2) FAST_LAUNCH(48, 2) else if (cellsPerCta == 4) FAST_LAUNCH(48, 4) else FAST_LAUNCH(48, 8)
    } else {
        if (cellsPerCta == 1) FAST_LAUNCH(80, 1) else if (cellsPerCta == 2) FAST_LAUNCH(80, 2) else if (cellsPerCta == 4) FAST_LAUNCH(80, 4) else FAST_LAUNCH(80, 8)
    }
#undef FAST_LAUNCH
}
