// describe.cu — orientation, rBRIEF descriptor and final keypoint record: one warp per keypoint.
// Replaces IC_Angle / computeOrientation (/root/reference/src/ORBextractor.cc:75-102, 475-482),
// computeOrbDescriptor / computeDescriptors (:104-151, 1032-1039), the post-distribution fix-up
// (:836-850) and the output assembly of operator() (:1070-1101).
//
//  * IC_Angle: lane u (-15..15) owns one column of the radius-15 disc on the UN-blurred level and
//    accumulates u*I and v*I over the rows where |u| <= umax[|v|]; integer warp reduction gives
//    m10, m01 exactly; angle = fastAtan2((float)m01, (float)m10) with OpenCV's polynomial.
//  * rBRIEF: pair p = 32*j + lane; both sample points are rotated with the un-fused float
//    expressions of :116-117 and rounded half-to-even, read from the BLURRED level; the ballot of
//    "first < second" over the warp IS 32-bit word j of the descriptor (bit order of :119-148).
//  * keypoints are emitted level-major, inside a level in quadtree list order, pt scaled by
//    mvScaleFactor[level] for level > 0 (:1093-1099), size = (int)(31*scale), class_id = -1.
#include "orbx_dev.cuh"
#include "orbx_pattern.h"

__device__ const int8_t g_pattern[1024] = {ORBX_PATTERN_VALUES};
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

#define DESC_WARPS 8

__global__ void __launch_bounds__(DESC_WARPS * 32) orbx_describe_kernel(OrbxParams p)
{
    const int img = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);
    if (slot >= p.keepTotal) return;
    // slot -> (level, k)
    int level = 0;
#pragma unroll 1
    for (int l = 1; l < p.nlevels; ++l)
        if (slot >= p.lv[l].keepBase) level = l;
    const OrbxLevel& L = p.lv[level];
    const int k = slot - L.keepBase;
    const int* kc = p.keepCnt + img * p.nlevels;
    if (k >= kc[level]) return;
    int outIdx = k;
    for (int l = 0; l < level; ++l) outIdx += kc[l];

    const uint2 kp = p.keep[(size_t)img * p.keepTotal + slot];
    const int cx = (int)(kp.x & 0xFFFF) + (ORBX_EDGE - 3), cy = (int)(kp.x >> 16) + (ORBX_EDGE - 3);  // :843-844

    // ---- IC_Angle on the un-blurred level
    int pitch;
    const uint8_t* im = orbx_level_ptr(p, img, level, &pitch);
    const int u = lane - ORBX_HALF_PATCH;
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int au = abs(u);
        const uint8_t* c = im + (size_t)cy * pitch + cx + u;
        // fully unrolled: 31 independent loads in flight (the rolled loop was bound by load latency, ncu long_scoreboard)
        int colSum = 0;
#pragma unroll
        for (int v = -ORBX_HALF_PATCH; v <= ORBX_HALF_PATCH; ++v) {
            constexpr int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
            const int I = au <= kUmax[v < 0 ? -v : v] ? (int)c[v * pitch] : 0;
            colSum += I;
            m01 += v * I;
        }
        m10 = u * colSum;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = orbx_fast_atan2((float)m01, (float)m10);

    // ---- rBRIEF on the blurred level
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    orbx_sincosf(orbx_fmul(angle, factorPI), &b, &a);
    const uint8_t* bc = p.blur + (size_t)img * p.blurStride + L.boff + (size_t)cy * L.bpitch + cx;
    const int step = L.bpitch;
    uint32_t myword = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const char4 pt = *reinterpret_cast<const char4*>(g_pattern + 4 * (32 * j + lane));
        const float x0 = (float)pt.x, y0 = (float)pt.y, x1 = (float)pt.z, y1 = (float)pt.w;
        const int r0 = orbx_cvround(orbx_fadd(orbx_fmul(x0, b), orbx_fmul(y0, a)));
        const int c0 = orbx_cvround(orbx_fsub(orbx_fmul(x0, a), orbx_fmul(y0, b)));
        const int r1 = orbx_cvround(orbx_fadd(orbx_fmul(x1, b), orbx_fmul(y1, a)));
        const int c1 = orbx_cvround(orbx_fsub(orbx_fmul(x1, a), orbx_fmul(y1, b)));
        const int t0 = bc[r0 * step + c0], t1 = bc[r1 * step + c1];
        const uint32_t w = __ballot_sync(0xffffffffu, t0 < t1);
        if (lane == j) myword = w;
    }
    uint32_t* desc = reinterpret_cast<uint32_t*>(p.outDesc + ((size_t)img * p.maxKp + outIdx) * 32);
    if (lane < 8) desc[lane] = myword;

    // ---- keypoint record (28 bytes = 7 words, lanes 0..6)
    float fx = (float)cx, fy = (float)cy;
    if (level != 0) {
        fx = orbx_fmul(fx, L.scale);
        fy = orbx_fmul(fy, L.scale);
    }
    uint32_t word;
    switch (lane) {
        case 0: word = __float_as_uint(fx); break;
        case 1: word = __float_as_uint(fy); break;
        case 2: word = __float_as_uint((float)L.patchSize); break;
        case 3: word = __float_as_uint(angle); break;
        case 4: word = __float_as_uint((float)kp.y); break;
        case 5: word = (uint32_t)level; break;
        default: word = 0xFFFFFFFFu; break;  // class_id = -1
    }
    uint32_t* okp = reinterpret_cast<uint32_t*>((uint8_t*)p.outKp + ((size_t)img * p.maxKp + outIdx) * 28);
    if (lane < 7) okp[lane] = word;
}

// outCnt[img] = sum of the level counts (also correct when level 0 kept nothing)
__global__ void orbx_count_kernel(OrbxParams p)
{
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= p.imgs) return;
    int total = 0;
    for (int l = 0; l < p.nlevels; ++l) total += p.keepCnt[img * p.nlevels + l];
    p.outCnt[img] = total;
}

void orbx_launch_describe(const OrbxParams& p, cudaStream_t s)
{
    orbx_count_kernel<<<(p.imgs + 127) / 128, 128, 0, s>>>(p);
    dim3 grid((p.keepTotal + DESC_WARPS - 1) / DESC_WARPS, p.imgs);
    orbx_describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(p);
}
