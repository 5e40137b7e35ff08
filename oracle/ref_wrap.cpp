// oracle/ref_wrap.cpp — TEST INFRASTRUCTURE ONLY (see orb_oracle.h).
//
// C entry points around the REFERENCE'S OWN ORB_SLAM2::ORBextractor, compiled from /root/reference/src/ORBextractor.cc
// where it lies (oracle/Makefile, target _ref -> oracle/_ref/libref_orbextractor.so). The reference needs the OpenCV C++
// API, which this image lacks; oracle/cvshim/ stands in for the API and this file supplies the five OpenCV functions the
// reference calls, on top of the oracle's restatements in prims.cpp (pinned bit for bit against cv2 4.13 by
// tests/test_oracle_kat.py). Everything above those primitives — pyramid loop, cell loop with the threshold fallback,
// DistributeOctTree/DivideNode, IC_Angle, computeOrbDescriptor with its pattern, the output assembly — is the reference's
// own object code, so tests that compare against this library pin the oracle's restatement of those parts (and the CUDA
// path) against the reference itself.
//
// Determinism of DistributeOctTree: the reference sorts (size, ExtractorNode*) pairs (src/ORBextractor.cc:687), i.e. ties
// are broken by heap address. This library replaces operator new by a bump arena that is active while a call runs and
// never reuses memory inside a call, so addresses grow in allocation order and the tie-break becomes "later-created node
// first" — one legitimate execution of the reference (an allocator that hands out increasing addresses), and the canonical
// rule of the oracle and the CUDA path. The library is linked with -Bsymbolic so that only allocations made
// by this library's own code (the reference's std::list / std::vector instantiations) use the arena.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <cstring>
#include <mutex>
#include <new>
#include <sys/mman.h>
#include <vector>

#include <opencv2/core/core.hpp>

#define private public  // Frame::AssignFeaturesToGrid is private; access control changes neither layout nor symbol names
#include "Frame.h"      // with cvshim/ref_stubs.h force-included (see the Makefile): MapPoint, KeyFrame, ... read as stubs
#undef private
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "orb_oracle.h"

// ---------------------------------------------------------------- arena behind operator new (active inside a call)
namespace {
struct Arena {
    size_t used = 0;
    bool active = false;
    char* lo = nullptr;  // ONE address range per thread, so that addresses grow monotonically inside a call
    char* hi = nullptr;
};
thread_local Arena g_arena;
constexpr size_t kReserve = (size_t)4 << 30;  // address space only (MAP_NORESERVE); pages are touched on demand

void* arena_alloc(size_t n)
{
    Arena& a = g_arena;
    if (!a.lo) {
        void* m = mmap(nullptr, kReserve, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (m == MAP_FAILED) { std::fprintf(stderr, "ref_wrap: arena reservation failed\n"); std::abort(); }
        a.lo = (char*)m;
        a.hi = a.lo + kReserve;
        a.used = 0;
    }
    n = (n + 15) & ~(size_t)15;
    if (a.used + n > kReserve) { std::fprintf(stderr, "ref_wrap: arena exhausted\n"); std::abort(); }
    void* p = a.lo + a.used;
    a.used += n;
    return p;
}
inline bool in_arena(const void* p) { return g_arena.lo && (const char*)p >= g_arena.lo && (const char*)p < g_arena.hi; }

// A second, process-wide arena for ref_frame_stereo: the reference's stereo constructor runs its two extractions on
// std::threads of its own (src/Frame.cc:113-116), where the calling thread's arena is out of reach. While it is active,
// every thread of this library allocates from one range with an atomic bump, so addresses still grow in allocation order
// inside each thread — which is all the (size, node pointer) sort of one DistributeOctTree call needs.
std::atomic<bool> g_sharedActive{false};
std::atomic<size_t> g_sharedUsed{0};
char* g_sharedLo = nullptr;
std::mutex g_sharedMutex;  // one ref_frame_stereo at a time
void* shared_alloc(size_t n)
{
    n = (n + 15) & ~(size_t)15;
    const size_t at = g_sharedUsed.fetch_add(n);
    if (at + n > kReserve) { std::fprintf(stderr, "ref_wrap: shared arena exhausted\n"); std::abort(); }
    return g_sharedLo + at;
}
inline bool in_shared(const void* p) { return g_sharedLo && (const char*)p >= g_sharedLo && (const char*)p < g_sharedLo + kReserve; }
}  // namespace

void* operator new(size_t n)
{
    if (g_arena.active) return arena_alloc(n);
    if (g_sharedActive.load(std::memory_order_acquire)) return shared_alloc(n);
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept
{
    if (!p || in_arena(p) || in_shared(p)) return;  // arena memory is dropped wholesale at the end of the call
    std::free(p);
}
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

// ---------------------------------------------------------------- the OpenCV functions the reference calls
namespace cv {

// cv::resize, INTER_LINEAR, CV_8UC1 (called at src/ORBextractor.cc:1118; dst is a preallocated ROI of the right size)
void resize(const Mat& src, Mat& dst, Size dsize, double, double, int interpolation)
{
    if (interpolation != INTER_LINEAR) std::abort();
    dst.create(dsize.height, dsize.width, CV_8UC1);
    orc_resize_linear(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

// cv::copyMakeBorder with BORDER_REFLECT_101 (:1120, :1130). src may be a ROI of dst (the in-place case of :1120): the
// interior is then already in place and only the border is written, from interior pixels. BORDER_ISOLATED = do not look
// outside the ROI, which is what this does in every case (for :1130 the source is a whole image).
void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType)
{
    if ((borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101) std::abort();
    dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    const bool inplace = src.data == dst.data + (size_t)top * dst.step + left && (size_t)src.step == (size_t)dst.step;
    for (int y = 0; y < dst.rows; ++y) {
        const int sy = orc_reflect101(y - top, src.rows);
        const uchar* s = src.data + (size_t)sy * src.step;
        uchar* d = dst.data + (size_t)y * dst.step;
        const bool interiorRow = y >= top && y < top + src.rows;
        for (int x = 0; x < dst.cols; ++x) {
            const bool interior = interiorRow && x >= left && x < left + src.cols;
            if (interior && inplace) continue;
            d[x] = s[orc_reflect101(x - left, src.cols)];
        }
    }
}

// cv::FAST, TYPE_9_16 (:812, :817)
void FAST(const Mat& image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression)
{
    if (!nonmaxSuppression) std::abort();
    keypoints.clear();
    const int cap = image.rows * image.cols;
    static thread_local orc_keypoint* tmp = nullptr;  // malloc-backed scratch: the arena never reuses memory inside a call
    static thread_local int tmpCap = 0;
    if (cap > tmpCap) {
        tmp = (orc_keypoint*)std::realloc(tmp, (size_t)cap * sizeof(orc_keypoint));
        tmpCap = cap;
    }
    const int n = orc_fast9(image.data, image.cols, image.rows, (int)image.step, threshold, tmp, cap);
    keypoints.reserve(n);
    for (int i = 0; i < n; ++i)
        keypoints.push_back(KeyPoint(tmp[i].x, tmp[i].y, tmp[i].size, tmp[i].angle, tmp[i].response, tmp[i].octave, tmp[i].class_id));
}

// cv::GaussianBlur 7x7, sigma 2, BORDER_REFLECT_101, in place on a clone (:1084)
void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY, int borderType)
{
    if (ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 || borderType != BORDER_REFLECT_101) std::abort();
    Mat out(src.rows, src.cols, CV_8UC1);
    orc_gaussian7(src.data, src.cols, src.rows, (int)src.step, out.data, (int)out.step);
    dst.create(src.rows, src.cols, CV_8UC1);
    for (int y = 0; y < src.rows; ++y) std::memcpy(dst.ptr(y), out.ptr(y), (size_t)src.cols);
}

float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

void KeyPointsFilter::retainBest(std::vector<KeyPoint>&, int) { std::abort(); }  // only ComputeKeyPointsOld (never called)

}  // namespace cv

// ---------------------------------------------------------------- C entry points (ctypes)
namespace {
struct ArenaScope {  // operator new comes from the arena while the reference runs; everything is dropped afterwards
    ArenaScope() { g_arena.active = true; g_arena.used = 0; }
    ~ArenaScope() { g_arena.active = false; g_arena.used = 0; }
};
}  // namespace

#include "ref_frame_entry.inc"

extern "C" {

// The reference's stereo Frame constructor (src/Frame.cc:80-152), unchanged, on the reference's CPU extractors: see
// ref_frame_entry.inc. Returns N (>= 0), -1 when `capacity` is too small, -2 when the stereo vectors were not sized.
int ref_frame_stereo(void* hL, void* hR, const uint8_t* imL, const uint8_t* imR, int rows, int cols, int stride, const float* K4,
                     float bf, float thDepth, float mbPreset, int capacity, orc_keypoint* keys, orc_keypoint* keysRight,
                     orc_keypoint* keysUn, uint8_t* desc, uint8_t* descRight, float* uRight, float* depth, int* gridStart,
                     int* gridItems, int* counts, float* misc, float* scaleTables)
{
    std::lock_guard<std::mutex> lock(g_sharedMutex);
    if (!g_sharedLo) {
        void* m = mmap(nullptr, kReserve, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (m == MAP_FAILED) { std::fprintf(stderr, "ref_wrap: shared arena reservation failed\n"); std::abort(); }
        g_sharedLo = (char*)m;
    }
    g_sharedUsed.store(0);
    g_sharedActive.store(true, std::memory_order_release);
    const FrameStereoOut o{keys, keysRight, keysUn, desc, descRight, uRight, depth, gridStart, gridItems, counts, misc, scaleTables};
    const int rc = frame_stereo_run((ORB_SLAM2::ORBextractor*)hL, (ORB_SLAM2::ORBextractor*)hR, imL, imR, rows, cols, stride, K4, bf,
                                    thDepth, mbPreset, capacity, o);
    g_sharedActive.store(false, std::memory_order_release);
    g_sharedUsed.store(0);
    return rc;
}

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
{
    return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);  // malloc-backed: arena inactive
}

void ref_destroy(void* h) { delete (ORB_SLAM2::ORBextractor*)h; }

// scale tables through the reference's own getters; arrays of nlevels, any pointer may be NULL
int ref_tables(void* h, float* scale, float* invScale, float* sigma2, float* invSigma2)
{
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)h;
    const int n = e->GetLevels();
    const std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                             d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; ++i) {
        if (scale) scale[i] = a[i];
        if (invScale) invScale[i] = b[i];
        if (sigma2) sigma2[i] = c[i];
        if (invSigma2) invSigma2[i] = d[i];
    }
    return n;
}

// ORBextractor::operator() (src/ORBextractor.cc:1041-1103). Returns the number of keypoints, or -1 if capacity is too
// small (nothing is written then). kps has the cv::KeyPoint layout; desc is capacity x 32 bytes.
int ref_extract(void* h, const uint8_t* img, int rows, int cols, int stride, orc_keypoint* kps, uint8_t* desc, int capacity)
{
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)h;
    int n;
    {
        ArenaScope scope;
        {
            cv::Mat image(rows, cols, CV_8UC1, (void*)img, (size_t)stride), mask, descriptors;
            std::vector<cv::KeyPoint> keypoints;
            (*e)(image, mask, keypoints, descriptors);
            n = (int)keypoints.size();
            if (n <= capacity) {
                static_assert(sizeof(cv::KeyPoint) == sizeof(orc_keypoint), "cv::KeyPoint layout");
                if (n) std::memcpy(kps, keypoints.data(), (size_t)n * sizeof(orc_keypoint));
                for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
            } else {
                n = -1;
            }
        }  // the call's containers die before the arena is dropped
    }
    return n;
}

// mvImagePyramid[level] of the last call (the reference's public member), without its border
int ref_level_size(void* h, int level, int* rows, int* cols)
{
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)h;
    if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
    *rows = e->mvImagePyramid[level].rows;
    *cols = e->mvImagePyramid[level].cols;
    return 0;
}

int ref_level_copy(void* h, int level, uint8_t* dst, int dstStride)
{
    ORB_SLAM2::ORBextractor* e = (ORB_SLAM2::ORBextractor*)h;
    if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
    const cv::Mat& m = e->mvImagePyramid[level];
    for (int y = 0; y < m.rows; ++y) std::memcpy(dst + (size_t)y * dstStride, m.ptr(y), (size_t)m.cols);
    return 0;
}

// Frame::ComputeStereoMatches (src/Frame.cc:522-700) on the reference's own Frame object. The members the function reads
// are set the way the stereo constructor sets them (src/Frame.cc:98-125): the two extractors (hL, hR must have extracted
// the frame's images last: the function reads their pyramids), keypoints, descriptors, scale tables, N, mbf — and mb,
// which the constructor leaves uninitialised at that point (:552 reads it, :149 assigns it), is an explicit argument.
// Outputs: mvuRight / mvDepth (nL floats each).
int ref_stereo_match(void* hL, void* hR, const orc_keypoint* kpsL, const uint8_t* descL, int nL, const orc_keypoint* kpsR,
                     const uint8_t* descR, int nR, float mbf, float mb, float* uRight, float* depth)
{
    ORB_SLAM2::ORBextractor* eL = (ORB_SLAM2::ORBextractor*)hL;
    ORB_SLAM2::ORBextractor* eR = (ORB_SLAM2::ORBextractor*)hR;
    ORB_SLAM2::Frame F;
    F.mpORBextractorLeft = eL;
    F.mpORBextractorRight = eR;
    F.mvScaleFactors = eL->GetScaleFactors();
    F.mvInvScaleFactors = eL->GetInverseScaleFactors();
    F.mvKeys.resize(nL);
    F.mvKeysRight.resize(nR);
    if (nL) std::memcpy(F.mvKeys.data(), kpsL, (size_t)nL * sizeof(orc_keypoint));
    if (nR) std::memcpy(F.mvKeysRight.data(), kpsR, (size_t)nR * sizeof(orc_keypoint));
    F.mDescriptors = cv::Mat(nL, 32, CV_8U, (void*)descL, 32).clone();
    F.mDescriptorsRight = cv::Mat(nR, 32, CV_8U, (void*)descR, 32).clone();
    F.N = nL;
    F.mbf = mbf;
    F.mb = mb;
    F.ComputeStereoMatches();
    if ((int)F.mvuRight.size() != nL || (int)F.mvDepth.size() != nL) return -1;
    for (int i = 0; i < nL; ++i) {
        uRight[i] = F.mvuRight[i];
        depth[i] = F.mvDepth[i];
    }
    return 0;
}

// Frame::ComputeStereoFromRGBD (src/Frame.cc:702-723): mvKeys index the depth map, mvKeysUn give the x of uRight.
int ref_stereo_from_rgbd(const orc_keypoint* kps, const orc_keypoint* kpsUn, int n, const float* depthMap, int rows, int cols,
                         int strideBytes, float mbf, float* uRight, float* depth)
{
    ORB_SLAM2::Frame F;
    F.N = n;
    F.mbf = mbf;
    F.mvKeys.resize(n);
    F.mvKeysUn.resize(n);
    if (n) std::memcpy(F.mvKeys.data(), kps, (size_t)n * sizeof(orc_keypoint));
    if (n) std::memcpy(F.mvKeysUn.data(), kpsUn, (size_t)n * sizeof(orc_keypoint));
    cv::Mat im(rows, cols, CV_32F, (void*)depthMap, (size_t)strideBytes);
    F.ComputeStereoFromRGBD(im);
    for (int i = 0; i < n; ++i) {
        uRight[i] = F.mvuRight[i];
        depth[i] = F.mvDepth[i];
    }
    return 0;
}

// Frame::AssignFeaturesToGrid (src/Frame.cc:279-294) + PosInGrid (:435-445) on the reference's Frame; the image bounds
// and cell sizes are the statics the constructors set (:129-137). Result as CSR over cells ix * 48 + iy.
static void ref_frame_grid_setup(ORB_SLAM2::Frame& F, const orc_keypoint* keysUn, int n, float minX, float maxX, float minY, float maxY)
{
    ORB_SLAM2::Frame::mnMinX = minX;
    ORB_SLAM2::Frame::mnMaxX = maxX;
    ORB_SLAM2::Frame::mnMinY = minY;
    ORB_SLAM2::Frame::mnMaxY = maxY;
    ORB_SLAM2::Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / (maxX - minX);
    ORB_SLAM2::Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / (maxY - minY);
    F.N = n;
    F.mvKeysUn.resize(n);
    if (n) std::memcpy(F.mvKeysUn.data(), keysUn, (size_t)n * sizeof(orc_keypoint));
    F.AssignFeaturesToGrid();
}

int ref_assign_features_to_grid(const orc_keypoint* keysUn, int n, float minX, float maxX, float minY, float maxY,
                                int* cellStart, int* items)
{
    ORB_SLAM2::Frame F;
    ref_frame_grid_setup(F, keysUn, n, minX, maxX, minY, maxY);
    int k = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
        for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
            cellStart[ix * FRAME_GRID_ROWS + iy] = k;
            for (size_t j = 0; j < F.mGrid[ix][iy].size(); ++j) items[k++] = (int)F.mGrid[ix][iy][j];
        }
    cellStart[FRAME_GRID_COLS * FRAME_GRID_ROWS] = k;
    return k;
}

// Frame::GetFeaturesInArea (src/Frame.cc:376-433). Returns the number of indices (written to out, capacity n).
int ref_features_in_area(const orc_keypoint* keysUn, int n, float minX, float maxX, float minY, float maxY, float x, float y,
                         float r, int minLevel, int maxLevel, int* out)
{
    ORB_SLAM2::Frame F;
    ref_frame_grid_setup(F, keysUn, n, minX, maxX, minY, maxY);
    const std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size(); ++i) out[i] = (int)v[i];
    return (int)v.size();
}

// ---- Frame::ComputeBoW (src/Frame.cc:447-454) over the vendored DBoW2, which compiles against the OpenCV stand-in:
// ORBVocabulary::loadFromTextFile (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1351-1439) reads the vocabulary, and
// TemplatedVocabulary::transform(features, BowVector, FeatureVector, 4) (:1140-1272) with FORB::distance runs unchanged.
void* ref_voc_load(const char* textFile)
{
    ORB_SLAM2::ORBVocabulary* v = new ORB_SLAM2::ORBVocabulary();
    if (!v->loadFromTextFile(textFile)) {
        delete v;
        return nullptr;
    }
    return v;
}

void ref_voc_destroy(void* v) { delete (ORB_SLAM2::ORBVocabulary*)v; }

int ref_voc_words(void* v) { return (int)((ORB_SLAM2::ORBVocabulary*)v)->size(); }

// BowVector as (word id, value) in map order; FeatureVector as CSR (node id, start, feature indices) in map order.
// Capacities: n entries each (start: n + 1).
int ref_compute_bow(void* v, const uint8_t* desc, int n, int* wordIds, double* wordVals, int* nWords, int* fvNodes, int* fvStart,
                    int* fvFeats, int* nNodes)
{
    ORB_SLAM2::Frame F;
    F.mpORBvocabulary = (ORB_SLAM2::ORBVocabulary*)v;
    F.mDescriptors = cv::Mat(n, 32, CV_8U, (void*)desc, 32).clone();
    F.N = n;
    F.ComputeBoW();
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = F.mBowVec.begin(); it != F.mBowVec.end(); ++it, ++k) {
        wordIds[k] = (int)it->first;
        wordVals[k] = it->second;
    }
    *nWords = k;
    int m = 0, f = 0;
    for (DBoW2::FeatureVector::const_iterator it = F.mFeatVec.begin(); it != F.mFeatVec.end(); ++it, ++m) {
        fvNodes[m] = (int)it->first;
        fvStart[m] = f;
        for (size_t j = 0; j < it->second.size(); ++j) fvFeats[f++] = (int)it->second[j];
    }
    fvStart[m] = f;
    *nNodes = m;
    return 0;
}

// ---- ORBmatcher::SearchByProjection, both overloads, as the reference's own src/ORBmatcher.cc on its own Frame objects.
// The inputs use the oracle's structs (orb_oracle.h: what the functions read from a Frame and from the map points).
static void ref_fill_frame(ORB_SLAM2::Frame& F, const orc_frame* f, bool current)
{
    ORB_SLAM2::Frame::fx = f->fx;
    ORB_SLAM2::Frame::fy = f->fy;
    ORB_SLAM2::Frame::cx = f->cx;
    ORB_SLAM2::Frame::cy = f->cy;
    F.mbf = f->mbf;
    F.mb = f->mb;
    F.mvKeys.resize(f->n);
    if (f->n) std::memcpy(F.mvKeys.data(), f->keys_un, (size_t)f->n * sizeof(orc_keypoint));  // :1398 reads mvKeys[i].octave
    F.mTcw = cv::Mat(4, 4, CV_32F);
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) F.mTcw.at<float>(r, c) = r < 3 ? f->Tcw[4 * r + c] : (c == 3 ? 1.f : 0.f);
    if (current) {
        ref_frame_grid_setup(F, f->keys_un, f->n, f->min_x, f->max_x, f->min_y, f->max_y);  // N, mvKeysUn, statics, mGrid
        F.mDescriptors = cv::Mat(f->n, 32, CV_8U, (void*)f->desc, 32).clone();
        F.mvuRight.assign(f->n, -1.f);
        if (f->u_right) F.mvuRight.assign(f->u_right, f->u_right + f->n);
        F.mvScaleFactors.assign(f->scale_factors, f->scale_factors + 16);  // callers pass >= 16 entries (see ref.py)
        F.mvpMapPoints.assign(f->n, static_cast<ORB_SLAM2::MapPoint*>(NULL));
    } else {
        F.N = f->n;
        F.mvKeysUn = F.mvKeys;
    }
}

int ref_search_by_projection(const orc_frame* cur, const orc_frame* last, const orc_mappoints* mp, float th, int bMono,
                             int checkOrientation, int* matchOfCur)
{
    ORB_SLAM2::Frame C, L;
    ref_fill_frame(L, last, false);
    ref_fill_frame(C, cur, true);  // last: the statics (camera, bounds) are the current frame's
    std::vector<ORB_SLAM2::MapPoint> pts(last->n);
    L.mvpMapPoints.assign(last->n, static_cast<ORB_SLAM2::MapPoint*>(NULL));
    L.mvbOutlier.assign(last->n, false);
    for (int i = 0; i < last->n; ++i) {
        if (!mp->has_point[i]) continue;
        pts[i].mWorldPos = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; ++k) pts[i].mWorldPos.at<float>(k) = mp->world_pos[3 * i + k];
        pts[i].mDescriptor = cv::Mat(1, 32, CV_8U, (void*)(mp->point_desc + 32 * (size_t)i), 32).clone();
        pts[i].nObs = mp->observed[i] ? 1 : 0;
        L.mvpMapPoints[i] = &pts[i];
    }
    ORB_SLAM2::ORBmatcher matcher(0.9f, checkOrientation != 0);
    const int n = matcher.SearchByProjection(C, L, th, bMono != 0);
    for (int i = 0; i < cur->n; ++i) matchOfCur[i] = C.mvpMapPoints[i] ? (int)(C.mvpMapPoints[i] - pts.data()) : -1;
    return n;
}

int ref_search_local_points(const orc_frame* f, const uint8_t* occupied, const orc_trackpoints* tp, float th, float nnratio,
                            int* matchOfKey)
{
    ORB_SLAM2::Frame F;
    ref_fill_frame(F, f, true);
    ORB_SLAM2::MapPoint held;  // stands for "a point with Observations() > 0" in the occupied entries
    held.nObs = 1;
    if (occupied)
        for (int i = 0; i < f->n; ++i)
            if (occupied[i]) F.mvpMapPoints[i] = &held;
    std::vector<ORB_SLAM2::MapPoint> pts(tp->n);
    std::vector<ORB_SLAM2::MapPoint*> vp(tp->n);
    for (int i = 0; i < tp->n; ++i) {
        pts[i].mbTrackInView = tp->valid[i] != 0;
        pts[i].mTrackProjX = tp->proj[3 * i];
        pts[i].mTrackProjY = tp->proj[3 * i + 1];
        pts[i].mTrackProjXR = tp->proj[3 * i + 2];
        pts[i].mnTrackScaleLevel = tp->level[i];
        pts[i].mTrackViewCos = tp->view_cos[i];
        pts[i].mDescriptor = cv::Mat(1, 32, CV_8U, (void*)(tp->point_desc + 32 * (size_t)i), 32).clone();
        pts[i].nObs = tp->observed[i] ? 1 : 0;
        vp[i] = &pts[i];
    }
    ORB_SLAM2::ORBmatcher matcher(nnratio, true);
    const int n = matcher.SearchByProjection(F, vp, th);
    for (int i = 0; i < f->n; ++i) {
        ORB_SLAM2::MapPoint* q = F.mvpMapPoints[i];
        matchOfKey[i] = (q && q != &held) ? (int)(q - pts.data()) : -1;
    }
    return n;
}

}  // extern "C"
