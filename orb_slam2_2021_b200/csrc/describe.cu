// describe.cu — orientation, rBRIEF descriptor and final keypoint record: one warp per keypoint.
// Replaces IC_Angle / computeOrientation (/root/reference/src/ORBextractor.cc:75-102, 475-482),
// computeOrbDescriptor / computeDescriptors (:104-151, 1032-1039), the post-distribution fix-up
// (:836-850) and the output assembly of operator() (:1070-1101).
//
//  * both pixel neighbourhoods of a keypoint arrive by TMA box loads into the warp's shared-memory slice
//    (un-blurred 48x31 box for the disc, blurred 64x37 box for the samples), signalled on a per-warp mbarrier;
//  * IC_Angle: lane v (-15..15) owns one ROW of the radius-15 disc on the UN-blurred level: its span
//    |u| <= umax[|v|] is realigned with funnel shifts and summed with byte dot products (IDP.4A) —
//    sum(I) and sum(u*I) for 4 pixels per instruction; integer warp reduction gives m10, m01 exactly;
//    angle = fastAtan2((float)m01, (float)m10) with OpenCV's polynomial.
//  * rBRIEF: pair p = 32*j + lane; both sample points are rotated with the un-fused float
//    expressions of :116-117 and rounded half-to-even, read from the BLURRED level; the ballot of
//    "first < second" over the warp IS 32-bit word j of the descriptor (bit order of :119-148).
//  * keypoints are emitted level-major, inside a level in quadtree list order, pt scaled by
//    mvScaleFactor[level] for level > 0 (:1093-1099), size = (int)(31*scale), class_id = -1.
#include "orbx_dev.cuh"
#include "orbx_pattern.h"
#include "orbx_tma.cuh"

// the 256 point pairs as float4 (x0, y0, x1, y1): one coalesced 16-byte load per lane, no I2F in the loop
__device__ __align__(16) const float g_patternF[1024] = {ORBX_PATTERN_VALUES};
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

#define DESC_WARPS 8
#define DESC_P0_BYTES (ORBX_DESC_BOX0_W * ORBX_DESC_BOX0_H)  // 1488
#define DESC_P1_BYTES (ORBX_DESC_BOX1_W * ORBX_DESC_BOX1_H)  // 2368
#define DESC_P0_PITCH 1536                                   // 128-byte multiples per warp
#define DESC_P1_PITCH 2432

// cvRound for |v| < 2^22: adding 1.5*2^23 in round-to-nearest-even leaves round-half-even(v) in the low mantissa
// bits — the same result as __float2int_rn (orbx_cvround), but one FADD + IADD instead of an F2I on the slow XU pipe.
__device__ __forceinline__ int cvround_small(float v) { return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000; }

__global__ void __launch_bounds__(DESC_WARPS * 32) orbx_describe_kernel(OrbxParams p, const CUtensorMap* mapsPlain, const CUtensorMap* mapsBlur)
{
    __shared__ __align__(128) uint8_t patch0[DESC_WARPS * DESC_P0_PITCH];  // un-blurred disc boxes, row pitch 48
    __shared__ __align__(128) uint8_t patch1[DESC_WARPS * DESC_P1_PITCH];  // blurred boxes, row pitch 64
    __shared__ __align__(8) uint64_t bars[DESC_WARPS];
    const int img = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = blockIdx.x * DESC_WARPS + warp;
    // slot -> (level, k): keepBase/keepCap are multiples of 8, so a CTA never straddles two levels. Lane l looks at level l
    // (keepBase is ascending): the highest set bit of the ballot is the level — one indexed constant load instead of a loop.
    const int kbase = lane < p.nlevels ? p.lv[lane].keepBase : 0x7fffffff;
    const int level = 31 - __clz((int)__ballot_sync(0xffffffffu, (int)blockIdx.x * DESC_WARPS >= kbase));
    const OrbxLevel& L = p.lv[level];
    const int k = slot - L.keepBase;
    // the three loads are independent: issue them together, before the exit test (slot < keepTotal always)
    // (the exit test reads all three — offsets, coordinates and scores are never negative — or ptxas sinks two of the loads below the
    // exit and the round trips serialise)
    const int kcnt = p.keepCnt[img * p.nlevels + level];
    const int kofs = p.keepOfs[img * p.nlevels + level];
    const uint2 kp = p.keep[(size_t)img * p.keepTotal + slot];
    if (((kcnt - 1 - k) | kofs | (int)kp.x | (int)kp.y) < 0) return;  // k >= kcnt
    const int outIdx = k + kofs;
    const int cx = (int)(kp.x & 0xFFFF) + (ORBX_EDGE - 3), cy = (int)(kp.x >> 16) + (ORBX_EDGE - 3);  // :843-844

    // ---- both patches by TMA (the warp is the unit: its own mbarrier, no CTA barrier anywhere)
    uint8_t* P0 = patch0 + warp * DESC_P0_PITCH;
    uint8_t* P1 = patch1 + warp * DESC_P1_PITCH;
    uint64_t* bar = bars + warp;
    const int bx0 = (cx - ORBX_HALF_PATCH) & ~15, bx1 = (cx - 18) & ~15;
    if (lane == 0) {
        orbx_mbar_init(bar, 1);
        orbx_fence_mbar_init();
        orbx_mbar_expect_tx(bar, DESC_P0_BYTES + DESC_P1_BYTES);
        orbx_tma_load_3d(P0, mapsPlain + level, bx0, cy - ORBX_HALF_PATCH, img + p.imgBase, bar);
        orbx_tma_load_3d(P1, mapsBlur + level, bx1, cy - 18, img + p.imgBase, bar);
    }
    __syncwarp();
    orbx_mbar_wait(bar, 0);

    // ---- IC_Angle on the un-blurred level: lane = disc row v = lane-15, its contiguous span |u| <= umax[|v|].
    // The span is realigned to byte 0 with funnel shifts; sum(I) and sum((u+umax)*I) are byte dot products (IDP.4A).
    int m10 = 0, m01 = 0;
    {
        const int v = lane - ORBX_HALF_PATCH;
        const int um = lane < 31 ? c_umax[v < 0 ? -v : v] : -1;  // lane 31: empty span
        const int len8 = 8 * (2 * um + 1);                      // span length in bits of bytes
        const int c0 = cx - ORBX_HALF_PATCH - bx0 + ORBX_HALF_PATCH - (um < 0 ? 0 : um);  // first span column in the box
        const uint32_t* row = reinterpret_cast<const uint32_t*>(P0 + min(lane, 30) * ORBX_DESC_BOX0_W) + (c0 >> 2);
        const int sh = (c0 & 3) * 8;
        uint32_t w[9];
#pragma unroll
        for (int j = 0; j < 9; ++j) w[j] = row[j];  // may run into the next row / the pad: masked below
        uint32_t accS = 0, accU = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t mask = __funnelshift_lc(0xFFFFFFFFu, 0u, max(len8 - 32 * j, 0));  // low min(len-4j,4) bytes
            const uint32_t x = __funnelshift_r(w[j], w[j + 1], sh) & mask;
            accS = __dp4a(x, 0x01010101u, accS);
            accU = __dp4a(x, 0x03020100u + 0x04040404u * j, accU);  // weight = byte index in the span = u + um
        }
        m10 = (int)accU - um * (int)accS;
        m01 = v * (int)accS;
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);  // REDUX.SUM: one instruction instead of a 5-step butterfly
    m01 = __reduce_add_sync(0xffffffffu, m01);
    const float angle = orbx_fast_atan2((float)m01, (float)m10);

    // ---- rBRIEF on the blurred level (samples from the shared-memory box: 32-bit addresses, no 64-bit pointer math)
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    orbx_sincosf(orbx_fmul(angle, factorPI), &b, &a);
    // cvRound(v) = bits(v + 1.5*2^23) - 0x4B400000 (see cvround_small); the two constants of a (row, col) pair fold into
    // the base address: row*64 + col = bits_r*64 + bits_c - 65*0x4B400000 (mod 2^32, like the 32-bit shared address)
    const uint32_t bc = orbx_smem_addr(P1) + 18 * ORBX_DESC_BOX1_W + (cx - bx1) - 65u * 0x4B400000u;
    const float4* pat = reinterpret_cast<const float4*>(g_patternF) + lane;
    uint32_t myword = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 pt = pat[32 * j];
        const uint32_t r0 = __float_as_uint(__fadd_rn(orbx_fadd(orbx_fmul(pt.x, b), orbx_fmul(pt.y, a)), 12582912.0f));
        const uint32_t c0 = __float_as_uint(__fadd_rn(orbx_fsub(orbx_fmul(pt.x, a), orbx_fmul(pt.y, b)), 12582912.0f));
        const uint32_t r1 = __float_as_uint(__fadd_rn(orbx_fadd(orbx_fmul(pt.z, b), orbx_fmul(pt.w, a)), 12582912.0f));
        const uint32_t c1 = __float_as_uint(__fadd_rn(orbx_fsub(orbx_fmul(pt.z, a), orbx_fmul(pt.w, b)), 12582912.0f));
        uint32_t t0, t1;
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(bc + r0 * ORBX_DESC_BOX1_W + c0));
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(bc + r1 * ORBX_DESC_BOX1_W + c1));
        const uint32_t wd = __ballot_sync(0xffffffffu, t0 < t1);
        if (lane == j) myword = wd;
    }
    uint32_t* desc = reinterpret_cast<uint32_t*>(p.outDesc + ((size_t)img * p.maxKp + outIdx) * 32);
    if (lane < 8) desc[lane] = myword;

    // ---- keypoint record (28 bytes = 7 words, lanes 0..6)
    float fx = (float)cx, fy = (float)cy;
    if (level != 0) {
        fx = orbx_fmul(fx, L.scale);
        fy = orbx_fmul(fy, L.scale);
    }
    // (a select tree on the lane bits: a switch compiles to a divergent jump table)
    const uint32_t w01 = (lane & 1) ? __float_as_uint(fy) : __float_as_uint(fx);
    const uint32_t w23 = (lane & 1) ? __float_as_uint(angle) : __float_as_uint((float)L.patchSize);
    const uint32_t w45 = (lane & 1) ? (uint32_t)level : __float_as_uint((float)kp.y);
    const uint32_t w03 = (lane & 2) ? w23 : w01, w47 = (lane & 2) ? 0xFFFFFFFFu : w45;  // word 6: class_id = -1
    const uint32_t word = (lane & 4) ? w47 : w03;
    uint32_t* okp = reinterpret_cast<uint32_t*>((uint8_t*)p.outKp + ((size_t)img * p.maxKp + outIdx) * 28);
    if (lane < 7) okp[lane] = word;
}

// outCnt[img] = sum of the level counts (also correct when level 0 kept nothing)
__global__ void orbx_count_kernel(OrbxParams p)
{
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= p.imgs) return;
    int total = 0;
    for (int l = 0; l < p.nlevels; ++l) {
        p.keepOfs[img * p.nlevels + l] = total;
        total += p.keepCnt[img * p.nlevels + l];
    }
    p.outCnt[img] = total;
}

void orbx_launch_describe(const OrbxParams& p, const void* mapsPlain, const void* mapsBlur, cudaStream_t s)
{
    orbx_count_kernel<<<(p.imgs + 127) / 128, 128, 0, s>>>(p);
    dim3 grid((p.keepTotal + DESC_WARPS - 1) / DESC_WARPS, p.imgs);
    orbx_describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(p, (const CUtensorMap*)mapsPlain, (const CUtensorMap*)mapsBlur);
}
