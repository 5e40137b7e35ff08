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
// The reference calls cos(float)/sin(float) (libm cosf/sinf, src/ORBextractor.cc:109-110). libm is
// not reproducible across host and device, so both sides use this routine: Cody-Waite reduction
// by pi/2 and the classic degree-13/14 minimax kernels, all in float64 with explicit fma, then one
// rounding to float32. The float64 result is accurate to <1 ulp(double), so the float32 value is
// the correctly rounded sin/cos except for ~1e-8 of inputs (tests/test_math.py measures it).
// (device: the 15 double constants sit in constant memory so that DMUL/DFMA take them as c[][] operands; as literals each
// costs two moves into a register pair, and the whole warp runs this once per keypoint)
#define ORBX_SINCOS_CONSTANTS                                                                                            \
    6.36619772367581382433e-01, 1.57079632679489655800e+00, 6.12323399573676603587e-17, -1.66666666666666324348e-01,     \
        8.33333333332248946124e-03, -1.98412698298579493134e-04, 2.75573137070700676789e-06, -2.50507602534068634195e-08, \
        1.58969099521155010221e-10, 4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05,  \
        -2.75573143513906633035e-07, 2.08757232129817482790e-09, -1.13596475577881948265e-11
#if defined(__CUDACC__)
static __constant__ double orbx_sincos_dev[15] = {ORBX_SINCOS_CONSTANTS};
#endif
ORBX_HD void orbx_sincosf(float xf, float* s_out, float* c_out)
{
#if defined(__CUDA_ARCH__)
    const double* K = orbx_sincos_dev;
#else
    static const double K[15] = {ORBX_SINCOS_CONSTANTS};
#endif
    const double TWO_OVER_PI = K[0], PIO2_HI = K[1], PIO2_LO = K[2];
    const double S1 = K[3], S2 = K[4], S3 = K[5], S4 = K[6], S5 = K[7], S6 = K[8];
    const double C1 = K[9], C2 = K[10], C3 = K[11], C4 = K[12], C5 = K[13], C6 = K[14];
    double x = (double)xf;
    double kd = rint(orbx_dmul(x, TWO_OVER_PI));
    int k = (int)kd;
    double r = orbx_dfma(-kd, PIO2_HI, x);
    r = orbx_dfma(-kd, PIO2_LO, r);
    double z = orbx_dmul(r, r);
    // sin kernel: r + r*z*(S1 + z*(S2 + ...))
    double ps = orbx_dfma(z, S6, S5);
    ps = orbx_dfma(z, ps, S4);
    ps = orbx_dfma(z, ps, S3);
    ps = orbx_dfma(z, ps, S2);
    ps = orbx_dfma(z, ps, S1);
    double sr = orbx_dfma(orbx_dmul(r, z), ps, r);
    // cos kernel: 1 - z/2 + z*z*(C1 + z*(C2 + ...))
    double pc = orbx_dfma(z, C6, C5);
    pc = orbx_dfma(z, pc, C4);
    pc = orbx_dfma(z, pc, C3);
    pc = orbx_dfma(z, pc, C2);
    pc = orbx_dfma(z, pc, C1);
    double cr = orbx_dfma(orbx_dmul(z, z), pc, orbx_dfma(z, -0.5, 1.0));
    double s, c;
    switch (k & 3) {
        case 0: s = sr; c = cr; break;
        case 1: s = cr; c = -sr; break;
        case 2: s = -sr; c = -cr; break;
        default: s = -cr; c = sr; break;
    }
    *s_out = (float)s;
    *c_out = (float)c;
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
