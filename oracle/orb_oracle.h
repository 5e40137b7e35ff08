/* orb_oracle.h — CPU ORACLE for the ORB front-end hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This library is the checker, never the product: only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it. The shipped path
 * (orb_slam2_2021_b200/, liborbx.so) never links, imports or calls anything in oracle/.
 *
 * It is a from-scratch sequential restatement of
 *   ORBextractor::ORBextractor / operator() / ComputePyramid / ComputeKeyPointsOctTree /
 *   DistributeOctTree / ExtractorNode::DivideNode / IC_Angle / computeOrbDescriptor
 *                                               (/root/reference/src/ORBextractor.cc:75-151, 413-856, 1032-1135)
 *   Frame::ComputeStereoMatches                 (/root/reference/src/Frame.cc:522-700)
 *   ORBmatcher::DescriptorDistance              (/root/reference/src/ORBmatcher.cc:1672-1688)
 * and of the OpenCV primitives those call (OpenCV is an un-vendored dependency of the reference,
 * CMakeLists.txt:15-21 "OpenCV 4.2", README.md:12 "tested with 4.5.4"; absent from /root/reference):
 *   cv::resize INTER_LINEAR 8UC1, cv::GaussianBlur 7x7 sigma 2 8UC1, cv::FAST TYPE_9_16 + NMS,
 *   cv::fastAtan2, cvRound, BORDER_REFLECT_101.
 *
 * PARITY PIN: the reference has no tests or golden vectors for this path. The restatement is
 * pinned against the reference's OWN object code: oracle/_ref (src/ORBextractor.cc and src/Frame.cc
 * compiled where they lie against the OpenCV stand-in of oracle/cvshim/, see ref_wrap.cpp;
 * tests/test_reference_pin.py: keypoints, angles, descriptors, mvuRight and mvDepth identical). The OpenCV primitives underneath are pinned against
 * the real OpenCV kernels through cv2 4.13.0 (tests/test_oracle_kat.py: resize, blur, FAST,
 * fastAtan2 bit-exact) and the whole path against golden vectors produced by a cv2-driven pipeline
 * (tests/golden/make_golden.py).
 * One choice is pinned by decision because the reference itself is not deterministic there
 * (DESIGN.md §oracle): the quadtree tie-break (creation order instead of heap address).
 * cos/sin: the reference's object code calls glibc sincosf; orbx_sincosf restates glibc's algorithm and
 * equals it on every float in [0, 2*pi] (tests/test_math.py, exhaustive).
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same 28-byte layout as cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id). */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orc_keypoint;

typedef struct orc_extractor orc_extractor;

/* ORBextractor::ORBextractor, src/ORBextractor.cc:413-473 */
orc_extractor* orc_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast);
void orc_destroy(orc_extractor* e);
/* scale tables / quotas / umax; any pointer may be NULL. Arrays of nlevels (umax: 16). */
void orc_tables(const orc_extractor* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                int* quotas, int* umax);

/* ORBextractor::operator(), src/ORBextractor.cc:1041-1103. Returns the number of keypoints
 * (>=0) or a negative error: -1 capacity too small, -2 bad arguments, -3 degenerate geometry
 * (level too small or nIni == 0: the reference has undefined behaviour there). An empty image
 * (rows==0 or cols==0) returns 0 and touches nothing. */
int orc_extract(orc_extractor* e, const uint8_t* img, int rows, int cols, int stride,
                orc_keypoint* kps, uint8_t* desc, int capacity);

/* mvImagePyramid[level] of the last extract (without the unused 19-px border). */
int orc_level_size(const orc_extractor* e, int level, int* rows, int* cols);
int orc_level_copy(const orc_extractor* e, int level, uint8_t* dst, int dst_stride);
/* blurred working copy of a level (src/ORBextractor.cc:1083-1084), for tests */
int orc_level_blur_copy(const orc_extractor* e, int level, uint8_t* dst, int dst_stride);
/* vToDistributeKeys of a level of the last extract (x,y relative to minBorder), for tests */
int orc_level_candidates(const orc_extractor* e, int level, orc_keypoint* out, int capacity);

/* Frame::ComputeStereoMatches, src/Frame.cc:522-700. left/right are the extractors that produced
 * the keypoints (their un-blurred pyramids are read, :620-639). mb is explicit because the
 * reference reads it uninitialised (:552 vs :149). Returns the number of valid matches. */
int orc_stereo_match(const orc_extractor* left, const orc_extractor* right,
                     const orc_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                     const orc_keypoint* kps_r, const uint8_t* desc_r, int n_r,
                     float mbf, float mb, float* u_right, float* depth);

/* ---- primitives, exported for the known-answer tests against cv2 ---- */
void orc_resize_linear(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orc_gaussian7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
int orc_fast9(const uint8_t* roi, int w, int h, int stride, int threshold, orc_keypoint* out, int capacity);
float orc_fast_atan2(float y, float x);
void orc_sincos(float rad, float* s, float* c);
long orc_sincos_vs_libm(uint32_t lo, uint32_t hi, uint32_t step); /* mismatches against the host's glibc sincosf */
int orc_hamming256(const uint8_t* a, const uint8_t* b);
int orc_reflect101(int p, int n);
/* DistributeOctTree on arbitrary candidates (x,y,response used), src/ORBextractor.cc:542-766 */
int orc_distribute(const orc_keypoint* cand, int n, int min_x, int max_x, int min_y, int max_y, int n_target,
                   orc_keypoint* out, int capacity);

/* ---- rows right after the hot path, oracle/frame.cpp ---- */

/* What ORBmatcher::SearchByProjection / Frame::GetFeaturesInArea read from a Frame. */
typedef struct {
    int n;                       /* Frame::N */
    const orc_keypoint* keys_un; /* mvKeysUn (x, y, octave, angle are read) */
    const uint8_t* desc;         /* mDescriptors, n x 32 (current frame only) */
    const float* u_right;        /* mvuRight (current frame only; NULL = no stereo information) */
    float Tcw[12];               /* rows 0..2 of mTcw: [R | t] */
    float fx, fy, cx, cy, mbf, mb;
    float min_x, max_x, min_y, max_y; /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;  /* mvScaleFactors */
} orc_frame;

/* Per keypoint i of the LAST frame: its map point, as far as SearchByProjection reads it. */
typedef struct {
    const uint8_t* has_point;  /* mvpMapPoints[i] != NULL && !mvbOutlier[i] */
    const float* world_pos;    /* n x 3, MapPoint::GetWorldPos() */
    const uint8_t* point_desc; /* n x 32, MapPoint::GetDescriptor() */
    const uint8_t* observed;   /* MapPoint::Observations() > 0 */
} orc_mappoints;

/* Frame::ComputeStereoFromRGBD, src/Frame.cc:702-723. keys_un_x may be NULL (no distortion: mvKeysUn == mvKeys).
 * Keypoints outside the depth image are left at -1 (the reference would read out of bounds). Returns #valid. */
int orc_stereo_from_rgbd(const orc_keypoint* keys, const float* keys_un_x, int n, const float* depth, int rows, int cols,
                         int stride_floats, float mbf, float* u_right, float* depth_out);
/* Frame::AssignFeaturesToGrid, src/Frame.cc:279-294: CSR form of mGrid, cell (ix, iy) at index ix*48+iy. */
int orc_assign_features_to_grid(const orc_frame* f, int* cell_start, int* cell_items);
/* Frame::GetFeaturesInArea, src/Frame.cc:376-433. Returns the count or -1 if capacity is too small. */
int orc_features_in_area(const orc_frame* f, float x, float y, float r, int min_level, int max_level, int* out, int capacity);
/* x3Dc = Rcw * x3Dw + tcw with OpenCV's 32-bit gemm arithmetic (src/ORBmatcher.cc:1381-1382) */
void orc_project(const float* Tcw, const float* xw, float* xc);
float orc_tlc_z(const float* Tcw, const float* Tlw);
/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono), src/ORBmatcher.cc:1348-1491, with
 * CurrentFrame.mvpMapPoints all NULL on entry (src/Tracking.cc:897). match_of_cur[i2] = index of the last-frame
 * keypoint whose map point ends up in CurrentFrame.mvpMapPoints[i2], or -1. Returns nmatches. */
int orc_search_by_projection(const orc_frame* cur, const orc_frame* last, const orc_mappoints* mp, float th, int b_mono,
                             int check_orientation, int* match_of_cur);

/* Per map point of the local map, as far as the local-map SearchByProjection reads it. */
typedef struct {
    int n;
    const uint8_t* valid;      /* pMP->mbTrackInView && !pMP->isBad() */
    const float* proj;         /* n x 3: mTrackProjX, mTrackProjY, mTrackProjXR */
    const int* level;          /* mnTrackScaleLevel */
    const float* view_cos;     /* mTrackViewCos */
    const uint8_t* point_desc; /* n x 32, GetDescriptor() */
    const uint8_t* observed;   /* Observations() > 0 */
} orc_trackpoints;
/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th), src/ORBmatcher.cc:45-127; nnratio = mfNNratio.
 * occupied[i] = F.mvpMapPoints[i] holds a point with Observations() > 0 on entry (NULL = none). match_of_key[i] = index
 * of the map point newly stored in F.mvpMapPoints[i], or -1 (entry left as it was). Returns nmatches. */
int orc_search_local_points(const orc_frame* F, const uint8_t* occupied, const orc_trackpoints* tp, float th, float nnratio,
                            int* match_of_key);

/* ---- input staging right before the hot path, oracle/input.cpp ---- */
/* cvtColor(..., CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY), src/Tracking.cc:183-208; channels 1 copies. */
void orc_cvt_gray(const uint8_t* src, int rows, int cols, int stride, int channels, int rgb_order, uint8_t* dst, int dstride);
/* cv::remap(src, dst, mapx, mapy, INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT 0, 8-bit, 1/3/4 channels
 * (Examples/Stereo/arducam_images.cpp:127-128). */
void orc_remap_linear(const uint8_t* src, int srows, int scols, int sstride, int channels, const float* mapx, const float* mapy,
                      int mstride_floats, int drows, int dcols, uint8_t* dst, int dstride);

/* ---- bag of words, oracle/bow.cpp ---- */
typedef struct orc_vocabulary orc_vocabulary;
/* The vocabulary as TemplatedVocabulary::loadFromTextFile builds it (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1351-1436):
 * entry i describes node id i+1 (the root is node 0): parent id, leaf flag, 32-byte descriptor, weight. NULL on bad parents. */
orc_vocabulary* orc_voc_create(int L, int n_nodes, const int* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight);
void orc_voc_destroy(orc_vocabulary* v);
int orc_voc_words(const orc_vocabulary* v);
/* Frame::ComputeBoW = transform(features, BowVector, FeatureVector, levelsup), TF_IDF + L1_NORM (src/Frame.cc:447-454).
 * BowVector -> word_ids/word_vals ascending by word id (capacity n); FeatureVector -> CSR ascending by node id:
 * fv_nodes[m], fv_start[m+1], fv_feats (capacity n, n+1, n). */
int orc_compute_bow(const orc_vocabulary* voc, const uint8_t* desc, int n, int levelsup, int* word_ids, double* word_vals,
                    int* n_words, int* fv_nodes, int* fv_start, int* fv_feats, int* n_fv_nodes);

#ifdef __cplusplus
}
#endif
#endif
