/* orbx_frame.h — C-ABI for the rows right after the ORB front-end hot path:
 *
 *   N4  Frame::ComputeStereoFromRGBD                          /root/reference/src/Frame.cc:702-723
 *   N1  Frame::AssignFeaturesToGrid / PosInGrid               /root/reference/src/Frame.cc:279-294, 435-445
 *       Frame::GetFeaturesInArea                              /root/reference/src/Frame.cc:376-433
 *       ORBmatcher::SearchByProjection(Frame&, const Frame&, float th, bool bMono)
 *                                                             /root/reference/src/ORBmatcher.cc:1348-1491
 *       ORBmatcher::SearchByProjection(Frame&, const std::vector<MapPoint*>&, float th)
 *                                                             /root/reference/src/ORBmatcher.cc:45-127
 *
 * Same conventions as orbx.h: plain pointers and sizes, int status codes (ORBX_OK / ORBX_ERR_*),
 * orbx_last_error() for the message, no CPU fallback (ORBX_ERR_CUDA without a device). Part of liborbx.so.
 * All arrays are HOST memory; the matcher owns its device scratch and one CUDA stream. A matcher is
 * single-threaded; different matchers may be used concurrently.
 */
#ifndef ORBX_FRAME_H
#define ORBX_FRAME_H
#include "orbx.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_GRID_COLS 64 /* FRAME_GRID_COLS, include/Frame.h:39 */
#define ORBX_GRID_ROWS 48 /* FRAME_GRID_ROWS, include/Frame.h:38 */

typedef struct orbx_matcher orbx_matcher;

/* What SearchByProjection / GetFeaturesInArea read from a Frame (include/Frame.h). */
typedef struct {
    int n;                        /* Frame::N */
    const orbx_keypoint* keys_un; /* mvKeysUn: x, y, octave, angle are read */
    const uint8_t* desc;          /* mDescriptors, n x 32 (current frame only; may be NULL for the last frame) */
    const float* u_right;         /* mvuRight (current frame only; NULL = no stereo information, as if all -1) */
    float Tcw[12];                /* rows 0..2 of mTcw: [R | t] */
    float fx, fy, cx, cy, mbf, mb;
    float min_x, max_x, min_y, max_y; /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;   /* mvScaleFactors; every keys_un[i].octave must index it */
    int nlevels;                  /* length of scale_factors */
} orbx_frame_view;

/* Per keypoint i of the LAST frame: its map point, as far as SearchByProjection reads it. */
typedef struct {
    const uint8_t* has_point;  /* LastFrame.mvpMapPoints[i] != NULL && !LastFrame.mvbOutlier[i] */
    const float* world_pos;    /* n x 3, MapPoint::GetWorldPos() */
    const uint8_t* point_desc; /* n x 32, MapPoint::GetDescriptor() */
    const uint8_t* observed;   /* MapPoint::Observations() > 0 */
} orbx_mappoint_view;

int orbx_matcher_create(int device, orbx_matcher** out);
void orbx_matcher_destroy(orbx_matcher* m);

/* Frame::ComputeStereoFromRGBD (src/Frame.cc:702-723): for keypoint i, d = depth[(int)y][(int)x]; if d > 0 then
 * depth_out[i] = d and u_right[i] = keys_un_x[i] - mbf / d, else both -1. keys_un_x may be NULL (no distortion:
 * mvKeysUn == mvKeys). depth is the CV_32F depth map AFTER the caller's mDepthMapFactor scaling
 * (src/Tracking.cc:238-239). Keypoints outside the map give -1 (the reference would read out of bounds). */
int orbx_stereo_from_rgbd(orbx_matcher* m, const orbx_keypoint* keys, const float* keys_un_x, int n, const float* depth, int rows,
                          int cols, size_t stride_bytes, float mbf, float* u_right, float* depth_out);

/* Frame::AssignFeaturesToGrid (src/Frame.cc:279-294) as a CSR grid: cell (ix, iy) holds
 * cell_items[cell_start[ix*48+iy] .. cell_start[ix*48+iy+1]) in ascending keypoint index (the reference's push_back
 * order). cell_start has 64*48+1 entries, cell_items capacity f->n; *n_items = keypoints that fell inside the grid. */
int orbx_assign_features_to_grid(orbx_matcher* m, const orbx_frame_view* f, int* cell_start, int* cell_items, int* n_items);

/* Frame::GetFeaturesInArea (src/Frame.cc:376-433): indices in the reference's order (ix, then iy, then cell order).
 * min_level / max_level as in the reference (-1 = unbounded). ORBX_ERR_CAPACITY if more than `capacity` indices. */
int orbx_get_features_in_area(orbx_matcher* m, const orbx_frame_view* f, float x, float y, float r, int min_level, int max_level,
                              int* indices, int capacity, int* n_indices);

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) with mbCheckOrientation = check_orientation
 * (src/ORBmatcher.cc:1348-1491), for n_pairs independent (current, last) frame pairs in one call.
 * CurrentFrame.mvpMapPoints is all NULL on entry, as at both call sites (src/Tracking.cc:897-911).
 * match_of_cur[p][i2] = index i of the last-frame keypoint whose map point ends up in
 * CurrentFrame.mvpMapPoints[i2], or -1 (capacity cur[p].n each); nmatches[p] = the reference's return value. */
int orbx_search_by_projection(orbx_matcher* m, const orbx_frame_view* cur, const orbx_frame_view* last,
                              const orbx_mappoint_view* last_points, int n_pairs, float th, int b_mono, int check_orientation,
                              int* const* match_of_cur, int* nmatches);

/* Device-resident source of a frame's keypoints, descriptors and uRight: frame `frame`, side `side` (0 = left / mono,
 * 1 = right) of a batch engine after orbx_batch_run / orbx_batch_process*. batch == NULL: read the view's host arrays. */
typedef struct {
    orbx_batch* batch;
    int frame, side;
} orbx_frame_source;

/* orbx_search_by_projection without the round trip through the host between the rows: where cur_src[p] / last_src[p] name a
 * batch engine, that frame's keys_un, desc and u_right (current frame) or keys_un (last frame) are read where the extractor
 * and the stereo match left them on the device — as the reference's flow does in memory (src/Tracking.cc:897-911 works on
 * the Frame that Frame::Frame just filled); the view's pointers to them are ignored and its n must not exceed what the
 * engine produced (mvKeysUn == mvKeys: rectified input, as in the stereo examples). cur_src / last_src may be NULL (all
 * host). Map point arrays that are one concatenation in pair order (pair p+1's array starts where pair p's ends, for all
 * four arrays) are copied straight from the caller's memory — page-locked for full speed — instead of being repacked.
 * Everything else as orbx_search_by_projection, results identical. */
int orbx_search_by_projection_resident(orbx_matcher* m, const orbx_frame_view* cur, const orbx_frame_source* cur_src,
                                       const orbx_frame_view* last, const orbx_frame_source* last_src,
                                       const orbx_mappoint_view* last_points, int n_pairs, float th, int b_mono,
                                       int check_orientation, int* const* match_of_cur, int* nmatches);

/* Per map point of the local map, as far as the local-map SearchByProjection reads it (MapPoint members set by
 * Frame::isInFrustum, src/Frame.cc:304-374). */
typedef struct {
    int n;
    const uint8_t* valid;      /* pMP->mbTrackInView && !pMP->isBad() */
    const float* proj;         /* n x 3: mTrackProjX, mTrackProjY, mTrackProjXR */
    const int* level;          /* mnTrackScaleLevel; must index the frame's scale table */
    const float* view_cos;     /* mTrackViewCos */
    const uint8_t* point_desc; /* n x 32, MapPoint::GetDescriptor() */
    const uint8_t* observed;   /* MapPoint::Observations() > 0 */
} orbx_trackpoint_view;

/* ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) with
 * mfNNratio = nnratio (src/ORBmatcher.cc:45-127; caller Tracking::SearchLocalPoints, src/Tracking.cc:1205-1215), for
 * n_pairs independent (frame, local map) pairs. occupied[p][i] != 0 iff F.mvpMapPoints[i] holds a point with
 * Observations() > 0 on entry (occupied or occupied[p] may be NULL = none). match_of_key[p][i] = index of the map point
 * the call stores in F.mvpMapPoints[i], or -1 (entry unchanged); nmatches[p] = the reference's return value. */
int orbx_search_local_points(orbx_matcher* m, const orbx_frame_view* frames, const uint8_t* const* occupied,
                             const orbx_trackpoint_view* points, int n_pairs, float th, float nnratio, int* const* match_of_key,
                             int* nmatches);

/* orbx_search_local_points with the frames' keypoints / descriptors / uRight read where a batch engine left them
 * (frame_src[p], see orbx_search_by_projection_resident; frame_src or frame_src[p].batch NULL = the view's host arrays).
 * The local map (points) is host state of the SLAM system and is passed as before. Results identical. */
int orbx_search_local_points_resident(orbx_matcher* m, const orbx_frame_view* frames, const orbx_frame_source* frame_src,
                                      const uint8_t* const* occupied, const orbx_trackpoint_view* points, int n_pairs, float th,
                                      float nnratio, int* const* match_of_key, int* nmatches);

#ifdef __cplusplus
}
#endif
#endif
