// stereo.cu — Frame::ComputeStereoMatches on the device (/root/reference/src/Frame.cc:522-700) with
// ORBmatcher::DescriptorDistance (/root/reference/src/ORBmatcher.cc:1672-1688).
//
// Kernel 1, one warp per LEFT keypoint:
//   * row-band candidate test without materialising vRowIndices (:532-549): right keypoint iR is a
//     candidate of row (int)vL iff floor(yR - 2*s_oct) <= row <= ceil(yR + 2*s_oct); plus the octave
//     (:588) and disparity-range (:593) tests;
//   * 256-bit Hamming as 8 x __popc on two uint4 loads; the reference keeps the FIRST strictly
//     smaller distance while scanning iR ascending from TH_HIGH=100, i.e. the lexicographic minimum
//     of (dist, iR): each lane keeps min(dist<<16 | iR), then a warp shuffle-min;
//   * if best < 75: 11x11 SAD against 11 horizontal offsets on the un-blurred pyramid level of the
//     left keypoint (:608-652): 121 pixels spread over the lanes, 11 accumulators per lane, warp
//     reductions; then parabola sub-pixel fit, disparity and depth (:657-683) as single IEEE ops.
// Kernel 2, one CTA per frame: median of the accepted matches' SAD by a two-pass 8-bit radix select
// (SAD <= 121*510 < 2^16) and invalidation of SAD >= 1.5*1.4*median (:686-699).
#include "orbx_dev.cuh"

#define ST_WARPS 8

struct KpRec {
    float x, y, size, angle, response;
    int octave, class_id;
};

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint8_t* b)
{
    const uint4 b0 = *reinterpret_cast<const uint4*>(b), b1 = *reinterpret_cast<const uint4*>(b + 16);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// Row-band table (:532-549) in data-parallel form, one CTA per frame: counting sort of the right
// keypoints by (int)y. The exact band test floor(y-r) <= row <= ceil(y+r) is evaluated here once per
// right keypoint and carried in the record, so the matcher only touches ~19 rows' worth of candidates.
__global__ void __launch_bounds__(256) orbx_stereo_bin_kernel(OrbxParams pl, OrbxParams pr, OrbxStereoParams sp)
{
    extern __shared__ int hist[];  // rows0 + 1
    __shared__ int wsum[9];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int imgR = f * sp.imgStep + sp.imgOfsR;
    const int nR = min(pr.outCnt[imgR], sp.maxKp);
    const KpRec* kpsR = reinterpret_cast<const KpRec*>(pr.outKp) + (size_t)imgR * pr.maxKp;
    const int rows = sp.rows0;
    for (int i = tid; i <= rows; i += 256) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < nR; i += 256) {
        const int yb = min(max((int)floorf(kpsR[i].y), 0), rows - 1);
        atomicAdd(&hist[yb], 1);
    }
    __syncthreads();
    // exclusive scan of hist[0..rows), total in hist[rows]
    const int chunk = (rows + 255) / 256;
    const int b = min(tid * chunk, rows), e = min(b + chunk, rows);
    int s = 0;
    for (int i = b; i < e; ++i) s += hist[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int w = 0; w < 8; ++w) { const int v = wsum[w]; wsum[w] = run; run += v; }
        wsum[8] = run;
    }
    __syncthreads();
    int run = wsum[warp] + inc - s;
    for (int i = b; i < e; ++i) { const int v = hist[i]; hist[i] = run; run += v; }
    if (tid == 0) hist[rows] = wsum[8];
    __syncthreads();
    int* bs = sp.binStart + (size_t)f * (rows + 1);
    for (int i = tid; i <= rows; i += 256) bs[i] = hist[i];
    __syncthreads();
    uint4* rs = sp.rsorted + (size_t)f * sp.maxKp * 3;  // record = {x, row band, octave, iR} + the 32-byte descriptor
    const uint4* descR = reinterpret_cast<const uint4*>(pr.outDesc + (size_t)imgR * pr.maxKp * 32);
    for (int i = tid; i < nR; i += 256) {
        const KpRec kr = kpsR[i];
        const int yb = min(max((int)floorf(kr.y), 0), rows - 1);
        const int oc = min(max(kr.octave, 0), sp.nlevels - 1);
        const float r = orbx_fmul(2.0f, pl.lv[oc].scale);  // Frame copies the LEFT extractor's scale table (:104-110, :541)
        const int maxr = min(max((int)ceilf(orbx_fadd(kr.y, r)), -32768), 32767);
        const int minr = min(max((int)floorf(orbx_fsub(kr.y, r)), -32768), 32767);
        const int pos = atomicAdd(&hist[yb], 1);
        rs[3 * pos] = make_uint4(__float_as_uint(kr.x), ((unsigned)minr & 0xFFFFu) | ((unsigned)maxr << 16), (unsigned)kr.octave, (unsigned)i);
        rs[3 * pos + 1] = descR[2 * i];
        rs[3 * pos + 2] = descR[2 * i + 1];
    }
}

__global__ void __launch_bounds__(ST_WARPS * 32) orbx_stereo_match_kernel(OrbxParams pl, OrbxParams pr, OrbxStereoParams sp)
{
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int iL = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
    const int imgL = f * sp.imgStep + sp.imgOfsL, imgR = f * sp.imgStep + sp.imgOfsR;
    const int nL = pl.outCnt[imgL];
    if (iL >= nL) return;
    const size_t o = (size_t)f * sp.maxKp + iL;
    // defaults (:524-525)
    float outU = -1.f, outD = -1.f;
    int outSad = -1;

    const KpRec* kpsL = reinterpret_cast<const KpRec*>(pl.outKp) + (size_t)imgL * pl.maxKp;
    const uint8_t* descL = pl.outDesc + (size_t)imgL * pl.maxKp * 32;

    const float uL = kpsL[iL].x, vL = kpsL[iL].y;
    const int levelL = kpsL[iL].octave;
    const int row = (int)vL;
    const float maxD = orbx_fdiv(sp.mbf, sp.mb);  // :552-554 (mb == 0 -> +inf, as in the reference)
    const float minU = orbx_fsub(uL, maxD), maxU = uL;

    unsigned best = (100u << 16);  // TH_HIGH, iR 0
    float bestX = 0.f;             // uR of this lane's best candidate
    if (!(maxU < 0) && row >= 0 && row < sp.rows0) {
        const uint4 a0 = *reinterpret_cast<const uint4*>(descL + (size_t)iL * 32);
        const uint4 a1 = *reinterpret_cast<const uint4*>(descL + (size_t)iL * 32 + 16);
        const int rlo = max(row - sp.binR, 0), rhi = min(row + sp.binR + 1, sp.rows0);
        const int* bs = sp.binStart + (size_t)f * (sp.rows0 + 1);
        const uint4* rs = sp.rsorted + (size_t)f * sp.maxKp * 3;
        const int jEnd = bs[rhi];
        for (int j = bs[rlo] + lane; j < jEnd; j += 32) {
            const uint4 rec = rs[3 * j];
            const int octR = (int)rec.z;
            if (octR < levelL - 1 || octR > levelL + 1) continue;  // :588
            const float uR = __uint_as_float(rec.x);
            if (!(uR >= minU && uR <= maxU)) continue;  // :593
            const int minr = (int)(short)(rec.y & 0xFFFFu), maxr = (int)(short)(rec.y >> 16);
            if (row < minr || row > maxr) continue;  // iR is listed in vRowIndices[row] (:539-549)
            // the descriptor travels with the record (same or next cache line): no dependent trip to the descriptor array
            const unsigned d = (unsigned)hamming256(a0, a1, reinterpret_cast<const uint8_t*>(rs + 3 * j + 1));
            const unsigned key = (d << 16) | rec.w;
            if (key < best) { best = key; bestX = uR; }
        }
    }
    const unsigned mine = best;
    best = __reduce_min_sync(0xffffffffu, best);  // REDUX.MIN
    const int bestDist = (int)(best >> 16);
    // uR of the winner comes from the lane that found it (iR is unique, so exactly one lane holds the minimum)
    const unsigned won = __ballot_sync(0xffffffffu, mine == best);
    const float uR0 = __shfl_sync(0xffffffffu, bestX, __ffs(won) - 1);

    if (bestDist < 75) {  // thOrbDist = (TH_HIGH + TH_LOW)/2, :527, :606
        const float s = pl.lv[levelL].invScale;
        const float su = roundf(orbx_fmul(uL, s)), sv = roundf(orbx_fmul(vL, s)), sr = roundf(orbx_fmul(uR0, s));
        const int cu = (int)su, cv = (int)sv, cr = (int)sr;
        const int wl = pl.lv[levelL].w, hl = pl.lv[levelL].h, wr = pr.lv[levelL].w, hr = pr.lv[levelL].h;
        const bool ok = !(sr < 0.f || orbx_fadd(sr, 11.f) >= (float)wr) &&  // :632-635
                        !(cv - 5 < 0 || cv + 5 >= hl || cu - 5 < 0 || cu + 5 >= wl || cv + 5 >= hr || cr - 10 < 0 ||
                          cr + 10 >= wr);  // windows outside the level: the reference's rowRange/colRange would throw
        if (ok) {
            int pitchL, pitchR;
            const uint8_t* IL = orbx_level_ptr(pl, imgL, levelL, &pitchL);
            const uint8_t* IR = orbx_level_ptr(pr, imgR, levelL, &pitchR);
            // SAD(k) = sum |(IL - cl) - (IR_k - crv_k)| = sum |(IL + crv_k) - (IR_k + cl)|: both sides are non-negative 9-bit
            // numbers, so two offsets k share one register as u16x2 and |x - y| = max - min on the packed min/max
            // instructions (1.5 ALU operations per term instead of 4). The 11 right-image bytes of a pixel come from four
            // aligned words realigned by funnel shifts instead of 11 byte loads. Totals stay below 2^16 (121 * 510).
            const uint32_t cl2 = (uint32_t)IL[(size_t)cv * pitchL + cu] * 0x10001u;
            uint32_t crv2[6], acc[6];
            {
                // the 11 right-image centres IR[cv][cr-5 .. cr+5] from four aligned words (as the window rows below), not 11
                // byte loads: this kernel's time follows its memory instructions. (The upper half of crv2[5] holds the byte
                // after the last centre; the sum it feeds, the upper half of acc[5], is never read.)
                const uint8_t* c = IR + (size_t)cv * pitchR + cr - 5;
                const uint32_t* cw = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(c) & ~(uintptr_t)3);
                const int csh = (int)(reinterpret_cast<uintptr_t>(c) & 3) * 8;
                const uint32_t w0 = cw[0], w1 = cw[1], w2 = cw[2], w3 = cw[3];  // (buffers end with 64 spare bytes)
                const uint32_t b0 = __funnelshift_r(w0, w1, csh), b1 = __funnelshift_r(w1, w2, csh), b2 = __funnelshift_r(w2, w3, csh);
                crv2[0] = __byte_perm(b0, 0, 0x4140); crv2[1] = __byte_perm(b0, 0, 0x4342);
                crv2[2] = __byte_perm(b1, 0, 0x4140); crv2[3] = __byte_perm(b1, 0, 0x4342);
                crv2[4] = __byte_perm(b2, 0, 0x4140); crv2[5] = __byte_perm(b2, 0, 0x4342);
#pragma unroll
                for (int j = 0; j < 6; ++j) acc[j] = 0;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int pidx = lane + 32 * q;
                if (pidx < 121) {
                    const int dy = pidx / 11 - 5, dx = pidx % 11 - 5;
                    const uint32_t a2 = (uint32_t)IL[(size_t)(cv + dy) * pitchL + cu + dx] * 0x10001u;
                    const uint8_t* rrow = IR + (size_t)(cv + dy) * pitchR + cr + dx - 5;
                    const uint32_t* rw = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(rrow) & ~(uintptr_t)3);
                    const int sh = (int)(reinterpret_cast<uintptr_t>(rrow) & 3) * 8;
                    const uint32_t w0 = rw[0], w1 = rw[1], w2 = rw[2], w3 = rw[3];  // (buffers end with 64 spare bytes)
                    const uint32_t b0 = __funnelshift_r(w0, w1, sh), b1 = __funnelshift_r(w1, w2, sh), b2 = __funnelshift_r(w2, w3, sh);
                    uint32_t y[6];
                    y[0] = __byte_perm(b0, 0, 0x4140); y[1] = __byte_perm(b0, 0, 0x4342);
                    y[2] = __byte_perm(b1, 0, 0x4140); y[3] = __byte_perm(b1, 0, 0x4342);
                    y[4] = __byte_perm(b2, 0, 0x4140); y[5] = __byte_perm(b2, 0, 0x4342);  // (upper half of y[5] is unused)
#pragma unroll
                    for (int j = 0; j < 6; ++j) {
                        const uint32_t xx = a2 + crv2[j], yy = y[j] + cl2;
                        acc[j] += __vimax3_u16x2(xx, yy, yy) - __vimin3_u16x2(xx, yy, yy);  // per half: max >= min, no borrow
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < 6; ++j) acc[j] = __reduce_add_sync(0xffffffffu, acc[j]);  // REDUX.SUM; the halves cannot carry (see above)
            int sad[11];
#pragma unroll
            for (int k = 0; k < 11; ++k) sad[k] = (int)((k & 1) ? acc[k >> 1] >> 16 : acc[k >> 1] & 0xFFFFu);
            int bestSad = 0x7fffffff, bestInc = 0;
#pragma unroll
            for (int k = 0; k < 11; ++k)
                if (sad[k] < bestSad) { bestSad = sad[k]; bestInc = k - 5; }
            if (bestInc != -5 && bestInc != 5) {  // :654
                float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                for (int k = 1; k < 10; ++k)
                    if (k - 5 == bestInc) { d1 = (float)sad[k - 1]; d2 = (float)sad[k]; d3 = (float)sad[k + 1]; }
                const float denom = orbx_fmul(2.0f, orbx_fsub(orbx_fadd(d1, d3), orbx_fmul(2.0f, d2)));
                const float deltaR = orbx_fdiv(orbx_fsub(d1, d3), denom);  // :662
                if (!(deltaR < -1.f || deltaR > 1.f)) {
                    float bestuR = orbx_fmul(pl.lv[levelL].scale, orbx_fadd(orbx_fadd(sr, (float)bestInc), deltaR));
                    float disparity = orbx_fsub(uL, bestuR);
                    if (disparity >= 0.f && disparity < maxD) {  // :672
                        if (disparity <= 0.f) {
                            disparity = 0.01f;
                            bestuR = (float)((double)uL - 0.01);
                        }
                        outD = orbx_fdiv(sp.mbf, disparity);
                        outU = bestuR;
                        outSad = bestSad;
                    }
                }
            }
        }
    }
    if (lane == 0) {
        sp.uRight[o] = outU;
        sp.depth[o] = outD;
        sp.sad[o] = outSad;
    }
}

// Median cut (:686-699): one CTA per frame.
__global__ void __launch_bounds__(256) orbx_stereo_median_kernel(OrbxParams pl, OrbxStereoParams sp)
{
    __shared__ int hist[256];
    __shared__ int sh_n, sh_bin, sh_rank, sh_valid;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int nL = pl.outCnt[f * sp.imgStep + sp.imgOfsL];
    const int* sad = sp.sad + (size_t)f * sp.maxKp;
    hist[tid] = 0;
    if (tid == 0) { sh_n = 0; sh_valid = 0; }
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < nL; i += 256) {
        const int v = sad[i];
        if (v >= 0) { atomicAdd(&hist[v >> 8], 1); ++cnt; }
    }
    if (cnt) atomicAdd(&sh_n, cnt);
    __syncthreads();
    const int n = sh_n;
    if (n == 0) {  // reference: undefined behaviour on an empty vector; defined as "nothing to cut"
        if (tid == 0) sp.nMatched[f] = 0;
        return;
    }
    if (tid == 0) {
        int k = n / 2, b = 0;
        while (k >= hist[b]) { k -= hist[b]; ++b; }
        sh_bin = b;
        sh_rank = k;
    }
    __syncthreads();
    const int hb = sh_bin;
    __syncthreads();
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nL; i += 256) {
        const int v = sad[i];
        if (v >= 0 && (v >> 8) == hb) atomicAdd(&hist[v & 255], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int k = sh_rank, b = 0;
        while (k >= hist[b]) { k -= hist[b]; ++b; }
        sh_bin = (hb << 8) | b;
    }
    __syncthreads();
    const float median = (float)sh_bin;
    const float th = orbx_fmul(orbx_fmul(1.5f, 1.4f), median);
    int valid = 0;
    for (int i = tid; i < nL; i += 256) {
        const int v = sad[i];
        if (v >= 0) {
            if ((float)v >= th) {
                sp.uRight[(size_t)f * sp.maxKp + i] = -1.f;
                sp.depth[(size_t)f * sp.maxKp + i] = -1.f;
            } else {
                ++valid;
            }
        }
    }
    if (valid) atomicAdd(&sh_valid, valid);
    __syncthreads();
    if (tid == 0) sp.nMatched[f] = sh_valid;
}

void orbx_launch_stereo(const OrbxParams& pl, const OrbxParams& pr, const OrbxStereoParams& sp, cudaStream_t s)
{
    const size_t binSmem = (size_t)(sp.rows0 + 1) * sizeof(int);
    if (binSmem > 32 * 1024)  // static + dynamic above 48 KB needs the opt-in; per-device attribute, set on every such launch
        cudaFuncSetAttribute(orbx_stereo_bin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)binSmem);
    orbx_stereo_bin_kernel<<<sp.frames, 256, binSmem, s>>>(pl, pr, sp);
    dim3 grid((sp.maxKp + ST_WARPS - 1) / ST_WARPS, sp.frames);
    orbx_stereo_match_kernel<<<grid, ST_WARPS * 32, 0, s>>>(pl, pr, sp);
    orbx_stereo_median_kernel<<<sp.frames, 256, 0, s>>>(pl, sp);
}

// ---- orbx_stereo_match_static: the same match with its scratch in a static device array (no allocation) ----------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). No
// counterpart in the reference. The scratch of orbx_stereo_match (per-
// keypoint uRight / depth / SAD, the row-sorted right records, the row-bin table) lives in ONE file-scope array that every
// GPU loading the library holds until exit; all calls share it, so calls that overlap on a GPU race (see include/orbx.h).
#include <mutex>

#include "../../include/orbx.h"
#include "handles.h"

#define ST_STATIC_MAX_KP ORBX_STEREO_STATIC_MAX_KEYPOINTS
#define ST_STATIC_MAX_ROWS ORBX_STEREO_STATIC_MAX_ROWS
// layout (byte offsets, each a multiple of 256): uRight, depth, sad [MAX_KP] 4 B each; nMatched; rsorted [MAX_KP][3] uint4; binStart [MAX_ROWS + 1]
#define ST_OFF_U 0u
#define ST_OFF_D (ST_OFF_U + 4u * ST_STATIC_MAX_KP)
#define ST_OFF_SAD (ST_OFF_D + 4u * ST_STATIC_MAX_KP)
#define ST_OFF_NM (ST_OFF_SAD + 4u * ST_STATIC_MAX_KP)
#define ST_OFF_SORTED (ST_OFF_NM + 256u)
#define ST_OFF_BIN (ST_OFF_SORTED + 48u * ST_STATIC_MAX_KP)
#define ST_STATIC_BYTES (ST_OFF_BIN + ((4u * (ST_STATIC_MAX_ROWS + 1) + 255u) & ~255u))
static_assert(ST_STATIC_BYTES <= 1u << 20, "the static stereo scratch is meant to stay modest (<= 1 MiB per GPU)");

static __device__ __align__(256) uint8_t g_stereo_static_scratch[ST_STATIC_BYTES];

// In host code the array's name is only a stand-in for the symbol, not a device address: the address is fetched once per
// device, with that device current, and that pointer is what the kernels get.
static uint8_t* g_stereo_static_ptr[64] = {nullptr};
static std::mutex g_stereo_static_mutex;

static int stereo_static_scratch(int device, uint8_t** out, std::string& err)
{
    if (device < 0 || device >= 64) { err = "device ordinal outside 0..63"; return ORBX_ERR_ARG; }
    std::lock_guard<std::mutex> lock(g_stereo_static_mutex);
    if (!g_stereo_static_ptr[device]) {
        int rc = orbx_check(cudaSetDevice(device), "cudaSetDevice", err);
        if (rc) return rc;
        void* p = nullptr;
        rc = orbx_check(cudaGetSymbolAddress(&p, g_stereo_static_scratch), "cudaGetSymbolAddress(static stereo scratch)", err);
        if (rc) return rc;
        g_stereo_static_ptr[device] = (uint8_t*)p;
    }
    *out = g_stereo_static_ptr[device];
    return ORBX_OK;
}

extern "C" int orbx_stereo_match_static(orbx_extractor* l, orbx_extractor* r, const orbx_keypoint* kpsL, const uint8_t* descL, int nL,
                                        const orbx_keypoint* kpsR, const uint8_t* descR, int nR, float mbf, float mb, float* uRight,
                                        float* depth)
{
    std::string& err = orbx_err_slot();
    auto fail = [&](int code, const char* msg) { err = msg; return code; };
    if (!l || !r || nL < 0 || nR < 0 || (nL && (!kpsL || !descL || !uRight || !depth)) || (nR && (!kpsR || !descR)))
        return fail(ORBX_ERR_ARG, "bad arguments");
    if (!l->eng || !r->eng || !l->hasRun || !r->hasRun) return fail(ORBX_ERR_STATE, "both extractors must have run");
    Engine &EL = *l->eng, &ER = *r->eng;
    if (EL.rows != ER.rows || EL.cols != ER.cols || EL.P.nlevels != ER.P.nlevels || l->device != r->device)
        return fail(ORBX_ERR_STATE, "left/right extractors differ in image size, levels or device");
    if (nL > EL.P.maxKp || nR > ER.P.maxKp) return fail(ORBX_ERR_ARG, "more keypoints than the extractor can produce");
    if (nL > ST_STATIC_MAX_KP || nR > ST_STATIC_MAX_KP || EL.rows > ST_STATIC_MAX_ROWS)
        return fail(ORBX_ERR_ARG, "orbx_stereo_match_static: beyond the static scratch (at most 8192 keypoints per side and 4096 image rows); "
                                  "use orbx_stereo_match");
    for (int i = 0; i < nL; ++i)
        if (kpsL[i].octave < 0 || kpsL[i].octave >= l->cfg.nlevels) return fail(ORBX_ERR_ARG, "left keypoint octave outside 0..nlevels-1");
    for (int i = 0; i < nR; ++i)
        if (kpsR[i].octave < 0 || kpsR[i].octave >= l->cfg.nlevels) return fail(ORBX_ERR_ARG, "right keypoint octave outside 0..nlevels-1");
    for (int i = 0; i < nL; ++i) uRight[i] = depth[i] = -1.f;  // src/Frame.cc:524-525
    if (nL == 0) return ORBX_OK;
    uint8_t* scratch = nullptr;
    int rc = stereo_static_scratch(l->device, &scratch, err);
    if (rc) return rc;
#define ST_TRY(x) do { rc = (x); if (rc) return rc; } while (0)
    ST_TRY(orbx_check(cudaSetDevice(l->device), "cudaSetDevice", err));
    cudaStream_t s = EL.stream;
    EL.residentValid = ER.residentValid = false;  // the extractors' own results on the device are overwritten below
    // the array keeps the previous call's contents: the per-keypoint SAD and the match counter start from a defined state
    ST_TRY(orbx_check(cudaMemsetAsync(scratch + ST_OFF_SAD, 0xFF, 4u * ST_STATIC_MAX_KP + 256u, s), "clear static scratch", err));
    // the caller's keypoints/descriptors are authoritative: into the extractors' own (already allocated) output buffers
    ST_TRY(orbx_check(cudaMemcpyAsync(EL.P.outKp, kpsL, (size_t)nL * 28, cudaMemcpyHostToDevice, s), "H2D kpsL", err));
    ST_TRY(orbx_check(cudaMemcpyAsync(EL.P.outDesc, descL, (size_t)nL * 32, cudaMemcpyHostToDevice, s), "H2D descL", err));
    ST_TRY(orbx_check(cudaMemcpyAsync(EL.P.outCnt, &nL, 4, cudaMemcpyHostToDevice, s), "H2D nL", err));
    if (nR) {
        ST_TRY(orbx_check(cudaMemcpyAsync(ER.P.outKp, kpsR, (size_t)nR * 28, cudaMemcpyHostToDevice, s), "H2D kpsR", err));
        ST_TRY(orbx_check(cudaMemcpyAsync(ER.P.outDesc, descR, (size_t)nR * 32, cudaMemcpyHostToDevice, s), "H2D descR", err));
    }
    ST_TRY(orbx_check(cudaMemcpyAsync(ER.P.outCnt, &nR, 4, cudaMemcpyHostToDevice, s), "H2D nR", err));
    OrbxParams pl = EL.P, pr = ER.P;
    pl.imgs = pr.imgs = 1;
    OrbxStereoParams sp{};
    sp.frames = 1; sp.nlevels = pl.nlevels; sp.rows0 = EL.rows; sp.mbf = mbf; sp.mb = mb;
    sp.maxKp = nL > nR ? nL : nR;  // one frame: only a bound on the counts and the grid of the match kernel
    sp.imgStep = 1; sp.imgOfsL = 0; sp.imgOfsR = 0;
    sp.uRight = (float*)(scratch + ST_OFF_U); sp.depth = (float*)(scratch + ST_OFF_D); sp.sad = (int*)(scratch + ST_OFF_SAD);
    sp.nMatched = (int*)(scratch + ST_OFF_NM); sp.rsorted = (uint4*)(scratch + ST_OFF_SORTED); sp.binStart = (int*)(scratch + ST_OFF_BIN);
    sp.binR = orbx_stereo_bin_radius(l->cfg.scale.data(), l->cfg.nlevels);
    orbx_launch_stereo(pl, pr, sp, s);
    ST_TRY(orbx_check(cudaGetLastError(), "stereo launch", err));
    ST_TRY(orbx_check(cudaMemcpyAsync(uRight, sp.uRight, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H uRight", err));
    ST_TRY(orbx_check(cudaMemcpyAsync(depth, sp.depth, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H depth", err));
#undef ST_TRY
    return EL.sync(err);
}
