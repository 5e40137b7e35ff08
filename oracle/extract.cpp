// oracle/extract.cpp — CPU restatement of ORBextractor (TEST INFRASTRUCTURE, see orb_oracle.h).
// Follows /root/reference/src/ORBextractor.cc; each function cites the lines it restates.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>
#include <vector>

#include "../orb_slam2_2021_b200/csrc/orbx_math.h"
#include "../orb_slam2_2021_b200/csrc/orbx_pattern.h"
#include "oracle_internal.h"

static const int8_t kPattern[1024] = {ORBX_PATTERN_VALUES};

enum { PATCH_SIZE = 31, HALF_PATCH = 15, EDGE = 19 };

// ---- constructor: src/ORBextractor.cc:413-473 ----------------------------------------------------
extern "C" orc_extractor* orc_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th)
{
    if (nlevels < 1 || nlevels > 32 || nfeatures < 0) return nullptr;
    orc_extractor* e = new orc_extractor();
    e->nfeatures = nfeatures;
    e->scale_factor = (double)scale_factor;  // the member is a double initialised from a float (ORBextractor.h:113)
    e->nlevels = nlevels;
    e->ini_th = ini_th;
    e->min_th = min_th;
    e->scale.assign(nlevels, 1.f);
    e->sigma2.assign(nlevels, 1.f);
    for (int i = 1; i < nlevels; ++i) {
        e->scale[i] = (float)((double)e->scale[i - 1] * e->scale_factor);  // :424 float*double -> float
        e->sigma2[i] = e->scale[i] * e->scale[i];                          // :425
    }
    e->inv_scale.resize(nlevels);
    e->inv_sigma2.resize(nlevels);
    for (int i = 0; i < nlevels; ++i) {
        e->inv_scale[i] = 1.0f / e->scale[i];
        e->inv_sigma2[i] = 1.0f / e->sigma2[i];
    }
    // per-level quotas :438-449
    e->quota.resize(nlevels);
    float factor = (float)(1.0 / e->scale_factor);
    float desired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; ++l) {
        e->quota[l] = orbx_cvround(desired);
        sum += e->quota[l];
        desired *= factor;
    }
    e->quota[nlevels - 1] = std::max(nfeatures - sum, 0);
    // umax :457-472 — end of each row of the radius-15 disc
    e->umax.assign(HALF_PATCH + 1, 0);
    int vmax = (int)std::floor(HALF_PATCH * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH * HALF_PATCH;
    for (int v = 0; v <= vmax; ++v) e->umax[v] = (int)std::lrint(std::sqrt(hp2 - v * v));
    for (int v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
    e->levels.resize(nlevels);
    e->blurred.resize(nlevels);
    e->candidates.resize(nlevels);
    return e;
}

extern "C" void orc_destroy(orc_extractor* e) { delete e; }

extern "C" void orc_tables(const orc_extractor* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                           int* quotas, int* umax)
{
    for (int i = 0; i < e->nlevels; ++i) {
        if (scale) scale[i] = e->scale[i];
        if (inv_scale) inv_scale[i] = e->inv_scale[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
        if (quotas) quotas[i] = e->quota[i];
    }
    if (umax) for (int i = 0; i <= HALF_PATCH; ++i) umax[i] = e->umax[i];
}

// ---- ComputePyramid: src/ORBextractor.cc:1105-1135 -------------------------------------------------
// The 19-px reflect border the reference adds is never read on this path, so levels are stored tight.
static void compute_pyramid(orc_extractor* e, const uint8_t* img, int rows, int cols, int stride)
{
    for (int l = 0; l < e->nlevels; ++l) {
        float s = e->inv_scale[l];
        int w = orbx_cvround((float)cols * s), h = orbx_cvround((float)rows * s);  // :1110
        OrcImage& L = e->levels[l];
        L.w = w;
        L.h = h;
        L.px.assign((size_t)std::max(w, 0) * std::max(h, 0), 0);
        if (w <= 0 || h <= 0) continue;
        if (l == 0) {
            for (int y = 0; y < h; ++y) std::memcpy(&L.px[(size_t)y * w], img + (size_t)y * stride, w);
        } else {
            const OrcImage& P = e->levels[l - 1];
            orc_resize_linear(P.px.data(), P.w, P.h, P.w, L.px.data(), w, h, w);  // :1118
        }
    }
}

// ---- cell-wise FAST: src/ORBextractor.cc:768-832 ---------------------------------------------------
static void detect_level(const orc_extractor* e, const OrcImage& im, std::vector<orc_keypoint>& cand)
{
    cand.clear();
    const int minBX = EDGE - 3, minBY = minBX;
    const int maxBX = im.w - EDGE + 3, maxBY = im.h - EDGE + 3;
    const float W = 30;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    if (nCols < 1 || nRows < 1) return;  // reference divides by zero here (ceil(width/0)); define: no cells
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    std::vector<orc_keypoint> cell(4096);
    for (int i = 0; i < nRows; ++i) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; ++j) {
            const float iniX = (float)(minBX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = (float)maxBX;
            const int x0 = (int)iniX, y0 = (int)iniY, rw = (int)maxX - x0, rh = (int)maxY - y0;
            const uint8_t* roi = im.px.data() + (size_t)y0 * im.w + x0;
            int n = orc_fast9(roi, rw, rh, im.w, e->ini_th, cell.data(), (int)cell.size());  // :812
            if (n == 0) n = orc_fast9(roi, rw, rh, im.w, e->min_th, cell.data(), (int)cell.size());  // :817
            for (int k = 0; k < n; ++k) {
                orc_keypoint kp = cell[k];
                kp.x += j * wCell;  // :825-826 (relative to minBorder)
                kp.y += i * hCell;
                cand.push_back(kp);
            }
        }
    }
}

// ---- IC_Angle: src/ORBextractor.cc:75-102 -----------------------------------------------------------
static float ic_angle(const OrcImage& im, float px, float py, const std::vector<int>& umax)
{
    int m01 = 0, m10 = 0;
    const uint8_t* c = im.px.data() + (size_t)orbx_cvround(py) * im.w + orbx_cvround(px);
    for (int u = -HALF_PATCH; u <= HALF_PATCH; ++u) m10 += u * c[u];
    for (int v = 1; v <= HALF_PATCH; ++v) {
        int vsum = 0, d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int plus = c[u + v * im.w], minus = c[u - v * im.w];
            vsum += plus - minus;
            m10 += u * (plus + minus);
        }
        m01 += v * vsum;
    }
    return orbx_fast_atan2((float)m01, (float)m10);
}

// ---- computeOrbDescriptor: src/ORBextractor.cc:104-151 ---------------------------------------------
static void orb_descriptor(const OrcImage& blur, const orc_keypoint& kp, uint8_t* desc)
{
    const float factorPI = (float)(3.14159265358979323846 / 180.f);  // :104 double division, one rounding
    float angle = orbx_fmul(kp.angle, factorPI);
    float a, b;
    orbx_sincosf(angle, &b, &a);  // a = cos, b = sin
    const uint8_t* c = blur.px.data() + (size_t)orbx_cvround(kp.y) * blur.w + orbx_cvround(kp.x);
    const int step = blur.w;
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int j = 0; j < 8; ++j) {
            const int8_t* p = kPattern + 4 * (8 * i + j);
            int t[2];
            for (int q = 0; q < 2; ++q) {
                float fx = (float)p[2 * q], fy = (float)p[2 * q + 1];
                int ry = orbx_cvround(orbx_fadd(orbx_fmul(fx, b), orbx_fmul(fy, a)));  // :116
                int rx = orbx_cvround(orbx_fsub(orbx_fmul(fx, a), orbx_fmul(fy, b)));  // :117
                t[q] = c[ry * step + rx];
            }
            val |= (t[0] < t[1]) << j;
        }
        desc[i] = (uint8_t)val;
    }
}

// ---- operator(): src/ORBextractor.cc:1041-1103 -------------------------------------------------------
extern "C" int orc_extract(orc_extractor* e, const uint8_t* img, int rows, int cols, int stride, orc_keypoint* kps,
                           uint8_t* desc, int capacity)
{
    if (!e) return -2;
    if (rows == 0 || cols == 0 || !img) return 0;  // :1044 empty image -> silent return
    if (rows < 0 || cols < 0 || stride < cols) return -2;
    compute_pyramid(e, img, rows, cols, stride);
    std::vector<std::vector<orc_keypoint>> all(e->nlevels);
    for (int l = 0; l < e->nlevels; ++l) {
        const OrcImage& im = e->levels[l];
        const int minBX = EDGE - 3, minBY = minBX, maxBX = im.w - EDGE + 3, maxBY = im.h - EDGE + 3;
        if (maxBX - minBX < 30 || maxBY - minBY < 30) return -3;
        detect_level(e, im, e->candidates[l]);
        int nIni = (int)std::round((float)(maxBX - minBX) / (maxBY - minBY));  // :546
        if (nIni < 1) return -3;
        orc_distribute_impl(e->candidates[l], minBX, maxBX, minBY, maxBY, e->quota[l], all[l]);
        const int scaledPatch = (int)(PATCH_SIZE * e->scale[l]);  // :836
        for (orc_keypoint& k : all[l]) {  // :840-850
            k.x += minBX;
            k.y += minBY;
            k.octave = l;
            k.size = (float)scaledPatch;
        }
    }
    for (int l = 0; l < e->nlevels; ++l)  // :853-855
        for (orc_keypoint& k : all[l]) k.angle = ic_angle(e->levels[l], k.x, k.y, e->umax);

    int total = 0;
    for (int l = 0; l < e->nlevels; ++l) total += (int)all[l].size();
    if (total > capacity) return -1;
    int off = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        const OrcImage& im = e->levels[l];
        OrcImage& B = e->blurred[l];
        B.w = im.w;
        B.h = im.h;
        B.px.resize(im.px.size());
        orc_gaussian7(im.px.data(), im.w, im.h, im.w, B.px.data(), im.w);  // :1083-1084 (harmless when the level is empty)
        for (orc_keypoint& k : all[l]) {
            orb_descriptor(B, k, desc + (size_t)off * 32);  // :1088
            if (l != 0) {  // :1093-1099
                k.x *= e->scale[l];
                k.y *= e->scale[l];
            }
            kps[off++] = k;
        }
    }
    return total;
}

extern "C" int orc_level_size(const orc_extractor* e, int level, int* rows, int* cols)
{
    if (!e || level < 0 || level >= e->nlevels) return -2;
    *rows = e->levels[level].h;
    *cols = e->levels[level].w;
    return 0;
}

static int copy_out(const OrcImage& im, uint8_t* dst, int dst_stride)
{
    for (int y = 0; y < im.h; ++y) std::memcpy(dst + (size_t)y * dst_stride, &im.px[(size_t)y * im.w], im.w);
    return 0;
}

extern "C" int orc_level_copy(const orc_extractor* e, int level, uint8_t* dst, int dst_stride)
{
    if (!e || level < 0 || level >= e->nlevels) return -2;
    return copy_out(e->levels[level], dst, dst_stride);
}

extern "C" int orc_level_blur_copy(const orc_extractor* e, int level, uint8_t* dst, int dst_stride)
{
    if (!e || level < 0 || level >= e->nlevels) return -2;
    return copy_out(e->blurred[level], dst, dst_stride);
}

extern "C" int orc_level_candidates(const orc_extractor* e, int level, orc_keypoint* out, int capacity)
{
    if (!e || level < 0 || level >= e->nlevels) return -2;
    const std::vector<orc_keypoint>& c = e->candidates[level];
    for (int i = 0; i < (int)c.size() && i < capacity; ++i) out[i] = c[i];
    return (int)c.size();
}
