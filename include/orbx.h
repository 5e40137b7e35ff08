/* orbx.h — C-ABI of the B200-native ORB front-end (liborbx.so).
 *
 * Drop-in boundary for ONE hot path of lreithmayr/ORB_SLAM2_2021: ORBextractor::operator() and
 * Frame::ComputeStereoMatches(). Plain pointers and sizes only; no C++/torch/OpenCV types. Every
 * entry point names the reference interface it replaces (paths relative to /root/reference).
 * All functions return ORBX_OK (0) or a negative orbx_status; none throws. A handle is
 * single-threaded; different handles may be used from different threads concurrently (each owns
 * its CUDA stream), which is how Frame::Frame runs left and right (src/Frame.cc:113-116).
 * There is no CPU fallback: every call fails with ORBX_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef ORBX_H
#define ORBX_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    ORBX_OK = 0,
    ORBX_ERR_ARG = -1,      /* null pointer, bad size, bad level ...                              */
    ORBX_ERR_CAPACITY = -2, /* caller's output buffers too small (need orbx_max_keypoints())       */
    ORBX_ERR_GEOMETRY = -3, /* a pyramid level is smaller than 62 px or taller than 2x its width:  */
                            /* the reference divides by zero there (src/ORBextractor.cc:546,786)   */
    ORBX_ERR_CUDA = -4,     /* CUDA runtime error; text in orbx_last_error()                       */
    ORBX_ERR_STATE = -5     /* e.g. stereo match before both extractors ran on equal-size images   */
} orbx_status;

/* Same 28-byte layout as cv::KeyPoint {Point2f pt; float size, angle, response; int octave, class_id}. */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbx_keypoint;

typedef struct orbx_extractor orbx_extractor; /* one ORBextractor instance                       */
typedef struct orbx_batch orbx_batch;         /* frame-batch engine for throughput                */

const char* orbx_last_error(void);            /* thread-local text of the last failure            */
int orbx_device_count(void);
/* Debug/experiment knobs (no reference counterpart; the library never reads the environment). Keys: run_streams,
 * run_chunks, e2e_streams, e2e_chunks, e2e_first_div, e2e_trace, fast_cells, fast_grid_mult, fast_ctas, numa_off, pyr_chain, pyr_tma; value 0 restores
 * the built-in default. Process-wide, not thread-safe against running calls. Unknown key: ORBX_ERR_ARG. */
int orbx_set_tuning(const char* key, int value);

/* ---- ORBextractor ------------------------------------------------------------------------------ */

/* Replaces ORBextractor::ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST,
 * int minThFAST) — include/ORBextractor.h:65-66, src/ORBextractor.cc:413-473 (constructed at
 * src/Tracking.cc:125-131). `device` is the CUDA ordinal.
 * Where this library is narrower than the reference (ORBX_ERR_ARG with a text): 1 <= nlevels <= 12, 1 < scaleFactor <= 2.0,
 * thresholds 1..254, and at most 65000 quadtree nodes per level (a per-level feature quota far beyond any ORB-SLAM2 setting;
 * quotas above ~2600 per level run from a global-memory node list instead of shared memory). */
int orbx_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast, int device,
                orbx_extractor** out);
void orbx_destroy(orbx_extractor* e);

/* Replaces GetLevels/GetScaleFactor(s)/GetInverseScaleFactors/GetScaleSigmaSquares/
 * GetInverseScaleSigmaSquares — include/ORBextractor.h:75-98 — plus mnFeaturesPerLevel (:117).
 * Each array has nlevels entries; any pointer may be NULL. */
int orbx_get_levels(const orbx_extractor* e);
int orbx_get_tables(const orbx_extractor* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int* features_per_level);
/* Upper bound of keypoints one extract can return: sum over levels of max(quota, 4*nIni) + 4. */
int orbx_max_keypoints(const orbx_extractor* e, int rows, int cols);

/* Replaces void ORBextractor::operator()(InputArray image, InputArray mask, vector<KeyPoint>&,
 * OutputArray descriptors) — include/ORBextractor.h:75-77, src/ORBextractor.cc:1041-1103; called from
 * Frame::ExtractORB, src/Frame.cc:296-302. `image` is host CV_8UC1, `stride` bytes per row; the mask
 * is ignored by the reference and has no parameter here. Writes *n keypoints (levels concatenated
 * 0..L-1, coordinates scaled to level 0) and n x 32 descriptor bytes. An empty image (rows or cols
 * == 0) returns ORBX_OK with *n = 0 and leaves kps/desc untouched, like :1044-1045. */
int orbx_extract(orbx_extractor* e, const uint8_t* image, int rows, int cols, size_t stride, orbx_keypoint* kps,
                 uint8_t* desc, int capacity, int* n);

/* Replaces the public member std::vector<cv::Mat> mvImagePyramid — include/ORBextractor.h:100 —
 * read by Frame::ComputeStereoMatches (src/Frame.cc:529,620,634,639). Levels live on the device until
 * the next extract on this handle; this copies one level (without the unused 19-px border) to the host. */
int orbx_pyramid_level_size(const orbx_extractor* e, int level, int* rows, int* cols);
int orbx_pyramid_level(const orbx_extractor* e, int level, uint8_t* dst, size_t dst_stride);

/* ---- Frame::ComputeStereoMatches ------------------------------------------------------------------ */

/* Replaces void Frame::ComputeStereoMatches() — include/Frame.h:92, src/Frame.cc:522-700 — with
 * ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1672-1688) inside. left/right are the extractors
 * that produced kps_l/kps_r (their device-resident pyramids are read). mbf = baseline*fx, mb = baseline
 * (explicit because the reference reads Frame::mb before assigning it, :552 vs :149). Writes mvuRight /
 * mvDepth (n_l floats each, -1 = no match). */
int orbx_stereo_match(orbx_extractor* left, orbx_extractor* right, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                      int n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r, float mbf, float mb,
                      float* u_right, float* depth);

/* The same match on what the two extractors left on the device: nothing is uploaded. For the reference's own flow —
 * Frame::ExtractORB on two threads (src/Frame.cc:113-116), then Frame::ComputeStereoMatches (:125) on the keypoints and
 * descriptors those calls returned, unmodified. n_l / n_r must equal the counts of the handles' last orbx_extract
 * (ORBX_ERR_STATE otherwise, also after an orbx_stereo_match that uploaded other keypoints: use orbx_stereo_match then). */
int orbx_stereo_match_resident(orbx_extractor* left, orbx_extractor* right, int n_l, int n_r, float mbf, float mb,
                               float* u_right, float* depth);

/* One stereo frame in ONE call: what Frame::Frame(stereo) does at src/Frame.cc:113-125 — ExtractORB on the left and the right
 * image (both at once, as the reference's two threads) and ComputeStereoMatches — for a caller that can hand over both
 * images together. Both images go up through pinned staging, every kernel runs on both images in one launch, the match
 * reads the device-resident keypoints and descriptors, all results come down in one go, and from the second call on the
 * whole sequence replays as one CUDA graph. `left` and `right` must have been created with the same parameters on the same
 * device; the left handle owns the frame's device memory. Outputs as orbx_extract (per side, `capacity` records each) and
 * orbx_stereo_match (n_l floats each). Afterwards orbx_pyramid_level on either handle returns this frame's levels;
 * orbx_stereo_match* need a new orbx_extract on both handles first (ORBX_ERR_STATE). */
int orbx_stereo_frame(orbx_extractor* left, orbx_extractor* right, const uint8_t* image_left, const uint8_t* image_right,
                      int rows, int cols, size_t stride_left, size_t stride_right, float mbf, float mb,
                      orbx_keypoint* kps_l, uint8_t* desc_l, int* n_l, orbx_keypoint* kps_r, uint8_t* desc_r, int* n_r,
                      int capacity, float* u_right, float* depth);

/* orbx_stereo_match without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results as orbx_stereo_match; its device scratch is a static array inside the library
 * (about 0.5 MiB that every GPU loading liborbx.so holds until exit) instead of memory allocated on first use, and the call
 * issues no cudaMalloc / cudaFree / allocating routine. Limits: n_l, n_r <= ORBX_STEREO_STATIC_MAX_KEYPOINTS and image rows
 * <= ORBX_STEREO_STATIC_MAX_ROWS, beyond them ORBX_ERR_ARG (use orbx_stereo_match). Every call on a GPU shares the one
 * array: calls that overlap on the same GPU (two threads, two handle pairs) race; serialise them. */
#define ORBX_STEREO_STATIC_MAX_KEYPOINTS 8192
#define ORBX_STEREO_STATIC_MAX_ROWS 4096
int orbx_stereo_match_static(orbx_extractor* left, orbx_extractor* right, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                             int n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r, float mbf, float mb,
                             float* u_right, float* depth);

/* ---- frame batches (throughput path; what Frame::Frame(stereo) does, N frames per call) ----------- */

/* One engine for `max_frames` frames of rows x cols. stereo != 0: each frame is a left/right pair and
 * the stereo match runs; stereo == 0: mono extraction only (src/Frame.cc:216-260). */
int orbx_batch_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast, int rows,
                      int cols, int max_frames, int stereo, int device, orbx_batch** out);
void orbx_batch_destroy(orbx_batch* b);
int orbx_batch_max_keypoints(const orbx_batch* b); /* per image: row pitch of the output arrays   */

/* Host -> device copy of n_frames tightly packed images (left[n][rows][cols], right likewise or NULL). */
int orbx_batch_upload(orbx_batch* b, const uint8_t* left, const uint8_t* right, int n_frames);
/* Same, but the sources are DEVICE pointers (device-to-device copy). */
int orbx_batch_upload_device(orbx_batch* b, const void* left_dev, const void* right_dev, int n_frames);
/* Run extract(L) [+ extract(R) + stereo match] on the resident frames. Asynchronous on the engine's
 * stream unless `sync` != 0. */
int orbx_batch_run(orbx_batch* b, int n_frames, float mbf, float mb, int sync);
/* Device -> host copy of the results. Row pitch of every array is orbx_batch_max_keypoints():
 *   n_left[n], kps_left[n][P], desc_left[n][P][32], (right likewise), u_right[n][P], depth[n][P].
 * Any pointer may be NULL (skipped). Synchronises the stream. */
int orbx_batch_download(orbx_batch* b, int n_frames, int* n_left, orbx_keypoint* kps_left, uint8_t* desc_left,
                        int* n_right, orbx_keypoint* kps_right, uint8_t* desc_right, float* u_right, float* depth);
/* Sum of keypoint counts of the last run (left, right) and of valid stereo matches; synchronises. */
int orbx_batch_stats(orbx_batch* b, int n_frames, long long* n_left, long long* n_right, long long* n_matched);
/* upload + run + download in one call (the end-to-end path). */
int orbx_batch_process(orbx_batch* b, const uint8_t* left, const uint8_t* right, int n_frames, float mbf, float mb,
                       int* n_left, orbx_keypoint* kps_left, uint8_t* desc_left, int* n_right,
                       orbx_keypoint* kps_right, uint8_t* desc_right, float* u_right, float* depth);
/* Streaming form of orbx_batch_process for a continuous feed of batches: returns as soon as the copies and kernels are
 * enqueued, so that the uploads of call k+1 overlap the kernels of call k (per chunk, call k+1 waits only for call k's use
 * of the same staging area and result rows). The host buffers of a call (inputs AND outputs, pinned for real overlap) must
 * not be touched until orbx_batch_wait() returns; use two output sets to alternate. Any other orbx_batch_* call on the
 * same batch drains the pipeline first. */
int orbx_batch_process_async(orbx_batch* b, const uint8_t* left, const uint8_t* right, int n_frames, float mbf, float mb,
                             int* n_left, orbx_keypoint* kps_left, uint8_t* desc_left, int* n_right,
                             orbx_keypoint* kps_right, uint8_t* desc_right, float* u_right, float* depth);
/* Blocks until every orbx_batch_process_async of this batch has finished and its outputs are in host memory. */
int orbx_batch_wait(orbx_batch* b);
/* The CUDA stream the engine launches on (cudaStream_t as void*), for event timing by the caller. */
void* orbx_batch_stream(orbx_batch* b);
/* Number of kernels the most recent orbx_batch_run launched (a batch of 32+ frames runs as 2 chunks on two streams;
 * before any run: the launches of one chunk). For bench.py's gpu_launches. */
int orbx_batch_launches_per_run(const orbx_batch* b);
/* Per-stage device timing with CUDA events on the engine's stream (bench.py's roofline leg). While on,
 * every orbx_batch_run records events at the stage boundaries (up to 512 runs are kept);
 * orbx_batch_stage_times() synchronises, returns the summed milliseconds per stage over those runs and
 * the number of runs, and resets. Stages: orbx_batch_stage_name(0..orbx_batch_stage_count()-1). */
int orbx_batch_set_profiling(orbx_batch* b, int on);
int orbx_batch_stage_count(void);
const char* orbx_batch_stage_name(int i);
int orbx_batch_stage_times(orbx_batch* b, double* ms, int* runs);
/* Bytes of HBM the engine holds. */
long long orbx_batch_device_bytes(const orbx_batch* b);
/* Page-locked host memory for the end-to-end path (cudaHostAlloc / cudaFreeHost), placed on the NUMA node of the
 * current CUDA device when the host allows it (orbx_set_tuning("numa_off", 1): plain allocation). orbx_host_numa_node reports the node the
 * last allocation aimed at (-1 unknown, -2 none yet) and whether the placement policy was in force. */
void* orbx_host_alloc(size_t bytes);
void orbx_host_free(void* p);
int orbx_host_numa_node(int* bound);
/* Debug/test taps on frame f, side s (0 left, 1 right): pyramid level, blurred level, FAST candidates
 * of a level in reference order (x,y relative to minBorder; response). */
int orbx_batch_level_size(const orbx_batch* b, int level, int* rows, int* cols);
int orbx_batch_level(orbx_batch* b, int frame, int side, int level, int blurred, uint8_t* dst, size_t dst_stride);
int orbx_batch_candidates(orbx_batch* b, int frame, int side, int level, orbx_keypoint* out, int capacity, int* n);

#ifdef __cplusplus
}
#endif
#endif
