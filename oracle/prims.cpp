// oracle/prims.cpp — CPU restatement of the OpenCV primitives on the hot path. TEST INFRASTRUCTURE
// (see orb_oracle.h). OpenCV is not vendored by the reference; the arithmetic below restates the
// published 8-bit fixed-point algorithms (OpenCV 4.x imgproc/resize.cpp, smooth.dispatch.cpp
// fixed-point path, features2d/fast.cpp + fast_score.cpp) and is pinned bit-for-bit against
// cv2 4.13.0 in tests/test_oracle_kat.py.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "../orb_slam2_2021_b200/csrc/orbx_math.h"
#include "orb_oracle.h"

// BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba
extern "C" int orc_reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

// ---- cv::resize(INTER_LINEAR), CV_8UC1. Called at src/ORBextractor.cc:1118 --------------------
// Coefficients: 11-bit fixed point (2048). Horizontal pass keeps 19-bit ints; the vertical pass
// is ((b0*(H0>>4))>>16 + (b1*(H1>>4))>>16 + 2) >> 2.
namespace {
struct AxisTab {
    std::vector<int> ofs;
    std::vector<short> c0, c1;
};

AxisTab make_axis(int src, int dst, bool clamp_frac)
{
    AxisTab t;
    t.ofs.resize(dst);
    t.c0.resize(dst);
    t.c1.resize(dst);
    const double inv_scale = (double)dst / src;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dst; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (clamp_frac) {  // x axis: taps outside the row collapse onto the edge pixel
            if (s < 0) { s = 0; f = 0.f; }
            if (s >= src - 1) { s = src - 1; f = 0.f; }
        }
        t.ofs[d] = s;
        t.c0[d] = (short)orbx_cvround((1.f - f) * 2048.f);
        t.c1[d] = (short)orbx_cvround(f * 2048.f);
    }
    return t;
}
}  // namespace

extern "C" void orc_resize_linear(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh,
                                  int dstride)
{
    AxisTab xt = make_axis(sw, dw, true);
    AxisTab yt = make_axis(sh, dh, false);  // y axis: row indices are clipped, the fraction is kept
    std::vector<int> h0(dw), h1(dw);
    for (int dy = 0; dy < dh; ++dy) {
        int sy0 = std::min(std::max(yt.ofs[dy], 0), sh - 1);
        int sy1 = std::min(std::max(yt.ofs[dy] + 1, 0), sh - 1);
        const uint8_t* r0 = src + (size_t)sy0 * sstride;
        const uint8_t* r1 = src + (size_t)sy1 * sstride;
        for (int dx = 0; dx < dw; ++dx) {
            int sx = xt.ofs[dx];
            int sx1 = std::min(sx + 1, sw - 1);
            h0[dx] = r0[sx] * xt.c0[dx] + r0[sx1] * xt.c1[dx];
            h1[dx] = r1[sx] * xt.c0[dx] + r1[sx1] * xt.c1[dx];
        }
        int b0 = yt.c0[dy], b1 = yt.c1[dy];
        uint8_t* out = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (h0[dx] >> 4)) >> 16) + ((b1 * (h1[dx] >> 4)) >> 16) + 2) >> 2;
            out[dx] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---- cv::GaussianBlur(Size(7,7), 2, 2, BORDER_REFLECT_101), CV_8UC1 ------------------------------
// Called at src/ORBextractor.cc:1084 on a clone of the level (so the reflection is at the level's
// own edges). 8-bit fixed-point kernel {18,34,48,56,48,34,18}/256 per axis, 16-bit intermediate,
// one rounding at the end: (sum + 2^15) >> 16.
extern "C" void orc_gaussian7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint32_t> hbuf((size_t)w * h);
    std::vector<int> xi(w + 6);
    for (int x = -3; x < w + 3; ++x) xi[x + 3] = orc_reflect101(x, w);
    for (int y = 0; y < h; ++y) {
        const uint8_t* r = src + (size_t)y * sstride;
        uint32_t* o = hbuf.data() + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            uint32_t s = 0;
            for (int k = 0; k < 7; ++k) s += (uint32_t)K[k] * r[xi[x + k]];
            o[x] = s;
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint32_t* rows[7];
        for (int k = 0; k < 7; ++k) rows[k] = hbuf.data() + (size_t)orc_reflect101(y + k - 3, h) * w;
        uint8_t* o = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            uint32_t s = 0;
            for (int k = 0; k < 7; ++k) s += (uint32_t)K[k] * rows[k][x];
            o[x] = (uint8_t)((s + 32768u) >> 16);
        }
    }
}

// ---- cv::FAST(roi, kps, threshold, nonmaxSuppression=true) == TYPE_9_16 -------------------------
// Called per 30-px cell at src/ORBextractor.cc:812,817. A pixel is a corner iff 9 contiguous ring
// pixels are all > v+t or all < v-t. Its score is max over the 16 arcs of min |v - ring| (same
// sign) minus 1, kept iff strictly greater than the scores of its 8 neighbours (non-corners and
// pixels outside [3,n-3) score 0). Emission is row-major; KeyPoint(x, y, 7.f, -1, score).
namespace {
const int RING_DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
const int RING_DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

inline bool has_arc9(uint32_t m16)
{
    uint32_t m = m16 | (m16 << 16);
    m &= m >> 1;  // runs of >=2
    m &= m >> 2;  // >=4
    m &= m >> 4;  // >=8
    m &= m >> 1;  // >=9
    return (m & 0xFFFFu) != 0;
}

// score (0 if not a corner at threshold t)
inline int fast_score(const uint8_t* p, int stride, int t)
{
    int v = p[0];
    // early out, exact: any 9 contiguous ring positions contain position 0 or 8 and position 4 or 12, and every pixel of a
    // corner's arc differs from the centre by more than t (OpenCV's detector starts with tests of this kind as well)
    {
        const int d0 = (int)p[3 * stride] - v, d8 = (int)p[-3 * stride] - v;
        if (d0 <= t && d0 >= -t && d8 <= t && d8 >= -t) return 0;
        const int d4 = (int)p[3] - v, d12 = (int)p[-3] - v;
        if (d4 <= t && d4 >= -t && d12 <= t && d12 >= -t) return 0;
    }
    int d[16];
    uint32_t bright = 0, dark = 0;
    for (int k = 0; k < 16; ++k) {
        d[k] = (int)p[RING_DY[k] * stride + RING_DX[k]] - v;
        if (d[k] > t) bright |= 1u << k;
        if (d[k] < -t) dark |= 1u << k;
    }
    bool cb = has_arc9(bright), cd = has_arc9(dark);
    if (!cb && !cd) return 0;
    int best = 0;
    for (int s = 0; s < 16; ++s) {
        int mn = 1 << 30, mx = -(1 << 30);
        for (int j = 0; j < 9; ++j) {
            int dv = d[(s + j) & 15];
            mn = std::min(mn, dv);
            mx = std::max(mx, dv);
        }
        best = std::max(best, std::max(mn, -mx));  // bright arc: min d ; dark arc: min -d
    }
    return best - 1;
}
}  // namespace

extern "C" int orc_fast9(const uint8_t* roi, int w, int h, int stride, int threshold, orc_keypoint* out, int capacity)
{
    if (w < 7 || h < 7) return 0;
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) score[(size_t)y * w + x] = fast_score(roi + (size_t)y * stride + x, stride, threshold);
    int n = 0;
    for (int y = 3; y < h - 3; ++y) {
        for (int x = 3; x < w - 3; ++x) {
            int s = score[(size_t)y * w + x];
            if (s <= 0) continue;  // a corner always has arc minimum > t >= 0, hence score >= t >= 1... see note
            bool keep = true;
            for (int dy = -1; dy <= 1 && keep; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    if (!dx && !dy) continue;
                    if (score[(size_t)(y + dy) * w + (x + dx)] >= s) { keep = false; break; }
                }
            if (!keep) continue;
            if (n < capacity) {
                out[n].x = (float)x;
                out[n].y = (float)y;
                out[n].size = 7.f;
                out[n].angle = -1.f;
                out[n].response = (float)s;
                out[n].octave = 0;
                out[n].class_id = -1;
            }
            ++n;
        }
    }
    return n;
}

extern "C" float orc_fast_atan2(float y, float x) { return orbx_fast_atan2(y, x); }
extern "C" void orc_sincos(float rad, float* s, float* c) { orbx_sincosf(rad, s, c); }

// Checker for tests/test_math.py: how many floats with bit patterns lo, lo+step, ... <= hi give a sin or cos that differs from
// the host's glibc sincosf (the symbol the reference's object code calls for src/ORBextractor.cc:110).
extern "C" long orc_sincos_vs_libm(uint32_t lo, uint32_t hi, uint32_t step)
{
    long bad = 0;
    for (uint64_t u = lo; u <= hi; u += step) {
        const uint32_t b = (uint32_t)u;
        float x, s0, c0, s1, c1;
        memcpy(&x, &b, 4);
        sincosf(x, &s0, &c0);
        orbx_sincosf(x, &s1, &c1);
        bad += (memcmp(&s0, &s1, 4) != 0) | (memcmp(&c0, &c1, 4) != 0);
    }
    return bad;
}

// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1672-1688 (8 x 32-bit popcount)
extern "C" int orc_hamming256(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t wa, wb;
        std::memcpy(&wa, a + 4 * i, 4);
        std::memcpy(&wb, b + 4 * i, 4);
        dist += orbx_popc32(wa ^ wb);
    }
    return dist;
}
