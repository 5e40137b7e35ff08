// orbx_math.h — deterministic scalar math shared by the CUDA kernels (device) and the CPU
// oracle (host). Every function is written so that host (g++ -ffp-contract=off) and device
// (nvcc -fmad=false, plus explicit *_rn intrinsics) evaluate the SAME sequence of IEEE-754
// operations: results are bit-identical on both sides by construction.
//
// What each routine restates (reference = /root/reference, OpenCV = un-vendored dependency):
//   orbx_cvround      cvRound (round-half-to-even)          used at src/ORBextractor.cc:79,116-117,445,1110
//   orbx_fast_atan2   cv::fastAtan2 (degree polynomial)      used at src/ORBextractor.cc:101
//   orbx_sincosf      cos/sin of the keypoint angle          used at src/ORBextractor.cc:109-110
//   orbx_roundf       ::round (half away from zero)          used at src/Frame.cc:613-615
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define ORBX_HD __host__ __device__ __forceinline__
#else
#define ORBX_HD static inline
#endif

// ---- single IEEE operations that must never be fused -------------------------------------
#if defined(__CUDA_ARCH__)
ORBX_HD float orbx_fmul(float a, float b) { return __fmul_rn(a, b); }
ORBX_HD float orbx_fadd(float a, float b) { return __fadd_rn(a, b); }
ORBX_HD float orbx_fsub(float a, float b) { return __fsub_rn(a, b); }
ORBX_HD float orbx_fdiv(float a, float b) { return __fdiv_rn(a, b); }
ORBX_HD double orbx_dmul(double a, double b) { return __dmul_rn(a, b); }
ORBX_HD double orbx_dadd(double a, double b) { return __dadd_rn(a, b); }
ORBX_HD double orbx_dfma(double a, double b, double c) { return __fma_rn(a, b, c); }
ORBX_HD int orbx_cvround(float v) { return __float2int_rn(v); }
ORBX_HD float orbx_roundf(float v) { return roundf(v); }
ORBX_HD float orbx_floorf(float v) { return floorf(v); }
ORBX_HD float orbx_ceilf(float v) { return ceilf(v); }
#else
ORBX_HD float orbx_fmul(float a, float b) { volatile float r = a * b; return r; }
ORBX_HD float orbx_fadd(float a, float b) { volatile float r = a + b; return r; }
ORBX_HD float orbx_fsub(float a, float b) { volatile float r = a - b; return r; }
ORBX_HD float orbx_fdiv(float a, float b) { volatile float r = a / b; return r; }
ORBX_HD double orbx_dmul(double a, double b) { volatile double r = a * b; return r; }
ORBX_HD double orbx_dadd(double a, double b) { volatile double r = a + b; return r; }
ORBX_HD double orbx_dfma(double a, double b, double c) { return fma(a, b, c); }
// lrintf honours the current rounding mode (default: to nearest even) == SSE cvtss2si == cvRound.
ORBX_HD int orbx_cvround(float v) { return (int)lrintf(v); }
ORBX_HD float orbx_roundf(float v) { return roundf(v); }
ORBX_HD float orbx_floorf(float v) { return floorf(v); }
ORBX_HD float orbx_ceilf(float v) { return ceilf(v); }
#endif

// ---- cv::fastAtan2 -------------------------------------------------------------------------
// OpenCV core/mathfuncs_core: degree-valued odd polynomial in min(|x|,|y|)/max(|x|,|y|),
// float32 throughout, Horner left to right, no FMA. The four coefficients are
// {0.9997878412794807,-0.3258083974640975,0.1555786518463281,-0.04432655554792128}f * (float)(180/pi)
// folded in float32 (values below verified with numpy float32).
ORBX_HD float orbx_fast_atan2(float y, float x)
{
    const float p1 = 0x1.ca44dep+5f, p3 = -0x1.2aaddcp+4f, p5 = 0x1.1d3f7ep+3f, p7 = -0x1.4515b2p+1f;
    const float eps = 0x1p-52f;  // (float)DBL_EPSILON
    float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = orbx_fdiv(ay, orbx_fadd(ax, eps));
        c2 = orbx_fmul(c, c);
        a = orbx_fmul(orbx_fadd(orbx_fmul(orbx_fadd(orbx_fmul(orbx_fadd(orbx_fmul(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = orbx_fdiv(ax, orbx_fadd(ay, eps));
        c2 = orbx_fmul(c, c);
        a = orbx_fsub(90.f,
                      orbx_fmul(orbx_fadd(orbx_fmul(orbx_fadd(orbx_fmul(orbx_fadd(orbx_fmul(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = orbx_fsub(180.f, a);
    if (y < 0) a = orbx_fsub(360.f, a);
    return a;
}

// ---- sin/cos of a float angle (radians) -----------------------------------------------------
// The reference writes cos(angle), sin(angle) on a float (src/ORBextractor.cc:109-110); g++ -O3 merges the pair into ONE call of
// glibc's sincosf (the only libm symbol the reference's object code imports for this path). glibc's float routines are not
// correctly rounded -- on [0, 2*pi] 0.13 % of all float angles give a sin or cos one ulp away from the correctly rounded value
// -- and one ulp in a or b can move a rotated pattern coordinate across a .5 boundary (seen: 1 descriptor bit in ~1e7
// keypoints). Bit parity therefore needs glibc's own arithmetic, and this is a restatement of it: the sincosf of glibc >= 2.28
// (sysdeps/ieee754/flt-32/s_sincosf.c, s_sincosf.h, x86_64 sincosf_poly.h; the algorithm published in ARM's optimized-routines):
// quadrant n and reduced argument in double (x - n*pi/2, no Cody-Waite split: |x| < 120 only), a degree-7 sine and a degree-8
// cosine polynomial in double, one rounding to float, swap / negate by quadrant. Operation order and the fused multiply-adds
// are those of the FMA build that glibc's ifunc selects on every AVX2 host; tests/test_math.py compares this routine with the
// host's sincosf on EVERY float in [0, 2*pi*1.0001] (1.09e9 inputs, 0 differences) and the same sweep shows the unfused (SSE2)
// build of glibc returns the same floats on that whole domain, so one variant serves both kinds of host.
// Domain: host |x| < 120; device 0 <= x < 7.0685 (the reference's angle is fastAtan2 * pi/180, in [0, 2*pi]).
// (device: the constants sit in constant memory so that DMUL/DFMA take them as c[][] operands; as literals each costs two
// moves into a register pair, and the whole warp runs this once per keypoint)
#define ORBX_SINCOS_CONSTANTS                                                                                         \
    0x1.45f306dc9c883p+23, 0x1.921fb54442d18p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, \
        0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13
#if defined(__CUDACC__)
static __constant__ double orbx_sincos_dev[9] = {ORBX_SINCOS_CONSTANTS};
#endif
ORBX_HD void orbx_sincosf(float xf, float* s_out, float* c_out)
{
#if defined(__CUDA_ARCH__)
    const double* K = orbx_sincos_dev;
    const uint32_t bits = __float_as_uint(xf);
#else
    static const double K[9] = {ORBX_SINCOS_CONSTANTS};
    uint32_t bits;
    {
        union { float f; uint32_t u; } cv;
        cv.f = xf;
        bits = cv.u;
    }
#endif
    const double HPI_INV24 = K[0], HPI = K[1];            // 2/pi * 2^24, pi/2
    const double C1 = K[2], C2 = K[3], C3 = K[4], C4 = K[5];  // cos: 1 + C1 z + C2 z^2 + C3 z^3 + C4 z^4
    const double S1 = K[6], S2 = K[7], S3 = K[8];             // sin: x + S1 x^3 + S2 x^5 + S3 x^7
    double x = (double)xf;
    bool swap, sneg, cneg;  // quadrant: results exchanged (n odd), sine negated (n & 3 in {1, 2}), cosine negated (n & 2)
#if !defined(__CUDA_ARCH__)
    if (bits >= 0x40e231d6u) {  // negative or >= 7.0685835: the general fast path of glibc (|x| < 120). The extractor never
                                // gets here, and the device build leaves it out to stay free of branches.
        const uint32_t top = (bits >> 20) & 0x7ffu;  // abstop12
        if (top <= 0x397u) {                         // |x| < 2^-12
            *s_out = xf;
            *c_out = 1.0f;
            return;
        }
        int n = 0;
        if (top > 0x3f3u) {
            const double r = orbx_dmul(x, HPI_INV24);
            n = ((int32_t)r + 0x800000) >> 24;
            x = orbx_dfma(-(double)n, HPI, x);
        }
        swap = (n & 1) != 0;
        sneg = ((n + 1) & 2) != 0;
        cneg = (n & 2) != 0;
    } else
#endif
    {
        // 0 <= x < 7.0685835: every angle the extractor produces. glibc's reduce_fast takes the quadrant from
        // n = ((int32_t)(x * 2/pi * 2^24) + 2^23) >> 24, a double multiply and two conversions at the head of the dependency
        // chain. n is monotonic in x, so on this interval it is the number of thresholds not above x; the four floats are the
        // first inputs on which glibc's expression steps (found by walking every float; positive floats order like their bit
        // patterns). Below pi/4 glibc skips the reduction, which is n = 0 here as well, and its |x| < 2^-12 shortcut
        // (sin = x, cos = 1) returns what the polynomials round to anyway (checked on every float below 2^-11).
        const bool p1 = bits >= 0x3f490fdbu, p2 = bits >= 0x4016cbe4u, p3 = bits >= 0x407b53d2u, p4 = bits >= 0x40afede0u;
#if defined(__CUDA_ARCH__)
        // (double)n for n = 0..4 as the high word of the double: no I2F.F64, four selects
        const int hi = p4 ? 0x40100000 : p3 ? 0x40080000 : p2 ? 0x40000000 : p1 ? 0x3FF00000 : 0;
        const double nd = __hiloint2double(hi, 0);
#else
        const double nd = (double)((int)p1 + (int)p2 + (int)p3 + (int)p4);
#endif
        x = orbx_dfma(-nd, HPI, x);  // n == 0: x - 0 * pi/2 = x exactly
        swap = (p1 && !p2) || (p3 && !p4);  // n = 1, 3
        sneg = p1 && !p3;                   // n = 1, 2
        cneg = p2 && !p4;                   // n = 2, 3
    }
    const double x2 = orbx_dmul(x, x);
    const double x3 = orbx_dmul(x2, x), x4 = orbx_dmul(x2, x2);
    const double s1v = orbx_dfma(x2, S3, S2), c2v = orbx_dfma(x2, C4, C3);
    const double c1v = orbx_dfma(x2, C1, 1.0);
    const double x5 = orbx_dmul(x2, x3), x6 = orbx_dmul(x2, x4);
    const double sv = orbx_dfma(x3, S1, x), cv2 = orbx_dfma(x4, C2, c1v);
    // glibc multiplies x by sign[n & 3] before the odd polynomial and takes the negated cosine table for n & 2; both are exact
    // sign changes of the results (round-to-nearest is symmetric), applied here to the two floats instead.
    float fs = (float)orbx_dfma(s1v, x5, sv), fc = (float)orbx_dfma(c2v, x6, cv2);
    fs = sneg ? -fs : fs;
    fc = cneg ? -fc : fc;
    *s_out = swap ? fc : fs;
    *c_out = swap ? fs : fc;
}

// 256-bit Hamming distance, descriptors read as 8 x uint32 (src/ORBmatcher.cc:1672-1688).
ORBX_HD int orbx_popc32(uint32_t v)
{
#if defined(__CUDA_ARCH__)
    return __popc(v);
#else
    return __builtin_popcount(v);
#endif
}
