// oracle/frame.cpp — CPU ORACLE (TEST INFRASTRUCTURE) for the rows right after the hot path:
//   N4  Frame::ComputeStereoFromRGBD                 /root/reference/src/Frame.cc:702-723
//   N1  Frame::AssignFeaturesToGrid / PosInGrid      /root/reference/src/Frame.cc:279-294, 435-445
//       Frame::GetFeaturesInArea                     /root/reference/src/Frame.cc:376-433
//       ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono)   /root/reference/src/ORBmatcher.cc:1348-1491
//       ORBmatcher::ComputeThreeMaxima               /root/reference/src/ORBmatcher.cc (same file, helper)
// Sequential restatement; every function follows the cited lines statement by statement.
//
// OpenCV arithmetic the reference reaches through cv::Mat expressions (pinned against cv2 4.13 in
// tests/test_frame_oracle.py):
//   C-6  `Rcw * x3Dw + tcw` and `Rlw * twc + tlw`  = cv::gemm(A, b, 1, c, 1) on CV_32F 3x3 * 3x1: OpenCV's small-matrix
//        path — float products, float sums left to right, then (float)((double)sum * alpha + (double)c * beta).
//   C-7  `-Rcw.t() * tcw` = cv::gemm(Rcw, tcw, -1, noArray, 0, GEMM_1_T): the general path — double accumulation over k,
//        times alpha, one rounding to float.
// `fx * xc * invzc + cx` is evaluated as single float operations without contraction (the same decision as for
// the descriptor sampling).
#include <cmath>
#include <cstring>
#include <vector>

#include "orb_oracle.h"

namespace {

const int GRID_ROWS = 48, GRID_COLS = 64;  // include/Frame.h:38-39
const int TH_HIGH = 100, HISTO_LENGTH = 30;  // src/ORBmatcher.cc:37-39

struct Grid {
    float invW, invH;
    std::vector<int> cell[GRID_COLS][GRID_ROWS];
};

void build_grid(const orc_frame* f, Grid& g)
{
    // src/Frame.cc:136-137
    g.invW = (float)GRID_COLS / (f->max_x - f->min_x);
    g.invH = (float)GRID_ROWS / (f->max_y - f->min_y);
    for (int i = 0; i < f->n; ++i) {  // :286-293
        const orc_keypoint& kp = f->keys_un[i];
        // PosInGrid :435-445; `round` is std::round (half away from zero) on a float
        const int px = (int)std::round((kp.x - f->min_x) * g.invW);
        const int py = (int)std::round((kp.y - f->min_y) * g.invH);
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
        g.cell[px][py].push_back(i);
    }
}

// Frame::GetFeaturesInArea :376-433
void features_in_area(const orc_frame* f, const Grid& g, float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out)
{
    out.clear();
    const int nMinCellX = std::max(0, (int)std::floor((x - f->min_x - r) * g.invW));
    if (nMinCellX >= GRID_COLS) return;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - f->min_x + r) * g.invW));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - f->min_y - r) * g.invH));
    if (nMinCellY >= GRID_ROWS) return;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - f->min_y + r) * g.invH));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (int idx : g.cell[ix][iy]) {
                const orc_keypoint& kp = f->keys_un[idx];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float distx = kp.x - x, disty = kp.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) out.push_back(idx);
            }
}

// C-6: one row of gemm(A, b, 1, c, 1) for 3x3 * 3x1 (volatile keeps every product and sum a rounded float)
float gemm3_row(const float* a, const float* b, float c)
{
    volatile float p0 = a[0] * b[0], p1 = a[1] * b[1], p2 = a[2] * b[2];
    volatile float s = p0 + p1;
    s = s + p2;
    return (float)((double)s * 1.0 + (double)c * 1.0);
}

}  // namespace

extern "C" int orc_stereo_from_rgbd(const orc_keypoint* keys, const float* keys_un_x, int n, const float* depth, int rows, int cols,
                                    int stride_floats, float mbf, float* u_right, float* depth_out)
{
    // src/Frame.cc:702-723; imDepth.at<float>(v, u) converts the float coordinates to int by truncation
    int valid = 0;
    for (int i = 0; i < n; ++i) {
        u_right[i] = -1.f;
        depth_out[i] = -1.f;
        const int v = (int)keys[i].y, u = (int)keys[i].x;
        if (v < 0 || v >= rows || u < 0 || u >= cols) continue;  // the reference would read out of bounds
        const float d = depth[(size_t)v * stride_floats + u];
        if (d > 0) {
            depth_out[i] = d;
            u_right[i] = (keys_un_x ? keys_un_x[i] : keys[i].x) - mbf / d;
            ++valid;
        }
    }
    return valid;
}

extern "C" int orc_assign_features_to_grid(const orc_frame* f, int* cell_start /*[64*48+1]*/, int* cell_items /*[n]*/)
{
    Grid g;
    build_grid(f, g);
    int k = 0;
    for (int ix = 0; ix < GRID_COLS; ++ix)
        for (int iy = 0; iy < GRID_ROWS; ++iy) {
            cell_start[ix * GRID_ROWS + iy] = k;
            for (int idx : g.cell[ix][iy]) cell_items[k++] = idx;
        }
    cell_start[GRID_COLS * GRID_ROWS] = k;
    return k;
}

extern "C" int orc_features_in_area(const orc_frame* f, float x, float y, float r, int min_level, int max_level, int* out, int capacity)
{
    Grid g;
    build_grid(f, g);
    std::vector<int> v;
    features_in_area(f, g, x, y, r, min_level, max_level, v);
    if ((int)v.size() > capacity) return -1;
    for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
    return (int)v.size();
}

extern "C" void orc_project(const float* Tcw /*3x4 rows*/, const float* xw, float* xc)
{
    const float R[3][3] = {{Tcw[0], Tcw[1], Tcw[2]}, {Tcw[4], Tcw[5], Tcw[6]}, {Tcw[8], Tcw[9], Tcw[10]}};
    for (int r = 0; r < 3; ++r) xc[r] = gemm3_row(R[r], xw, Tcw[4 * r + 3]);
}

// z component of tlc = Rlw * (-Rcw^T * tcw) + tlw  (src/ORBmatcher.cc:1358-1367)
extern "C" float orc_tlc_z(const float* Tcw, const float* Tlw)
{
    float twc[3];
    for (int i = 0; i < 3; ++i) {  // C-7
        double acc = 0;
        for (int k = 0; k < 3; ++k) acc += (double)Tcw[4 * k + i] * (double)Tcw[4 * k + 3];
        twc[i] = (float)(acc * -1.0);
    }
    const float Rl2[3] = {Tlw[8], Tlw[9], Tlw[10]};
    return gemm3_row(Rl2, twc, Tlw[11]);
}

extern "C" int orc_search_by_projection(const orc_frame* cur, const orc_frame* last, const orc_mappoints* mp, float th, int b_mono,
                                        int check_orientation, int* match_of_cur)
{
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    Grid g;
    build_grid(cur, g);
    const float tlcz = orc_tlc_z(cur->Tcw, last->Tcw);
    const bool bForward = tlcz > cur->mb && !b_mono;
    const bool bBackward = -tlcz > cur->mb && !b_mono;
    std::vector<int> assigned(cur->n, -1);  // CurrentFrame.mvpMapPoints (index into last), all NULL on entry (src/Tracking.cc:897)
    std::vector<int> vIndices2;
    for (int i = 0; i < last->n; i++) {
        if (!mp->has_point[i]) continue;  // pMP && !mvbOutlier[i]
        float x3Dc[3];
        orc_project(cur->Tcw, mp->world_pos + 3 * i, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = (float)(1.0 / (double)x3Dc[2]);
        if (invzc < 0) continue;
        volatile float ux = cur->fx * xc;
        ux = ux * invzc;
        const float u = ux + cur->cx;
        volatile float vy = cur->fy * yc;
        vy = vy * invzc;
        const float v = vy + cur->cy;
        if (u < cur->min_x || u > cur->max_x) continue;
        if (v < cur->min_y || v > cur->max_y) continue;
        const int nLastOctave = last->keys_un[i].octave;
        const float radius = th * cur->scale_factors[nLastOctave];
        if (bForward) features_in_area(cur, g, u, v, radius, nLastOctave, -1, vIndices2);
        else if (bBackward) features_in_area(cur, g, u, v, radius, 0, nLastOctave, vIndices2);
        else features_in_area(cur, g, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mp->point_desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (assigned[i2] >= 0 && mp->observed[assigned[i2]]) continue;  // mvpMapPoints[i2]->Observations() > 0
            if (cur->u_right && cur->u_right[i2] > 0) {
                volatile float bz = cur->mbf * invzc;
                const float ur = u - bz;
                const float er = std::fabs(ur - cur->u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = orc_hamming256(dMP, cur->desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            assigned[bestIdx2] = i;
            nmatches++;
            if (check_orientation) {
                float rot = last->keys_un[i].angle - cur->keys_un[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        {  // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < HISTO_LENGTH; i++) {
                const int s = (int)rotHist[i].size();
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
                else if (s > max3) { max3 = s; ind3 = i; }
            }
            if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
            else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        }
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { assigned[idx] = -1; nmatches--; }
    }
    for (int i = 0; i < cur->n; ++i) match_of_cur[i] = assigned[i];
    return nmatches;
}

// ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th)
// /root/reference/src/ORBmatcher.cc:45-127 (caller Tracking::SearchLocalPoints, src/Tracking.cc:1205-1215).
// occupied[i] = F.mvpMapPoints[i] != NULL && F.mvpMapPoints[i]->Observations() > 0 on entry (may be NULL).
extern "C" int orc_search_local_points(const orc_frame* F, const uint8_t* occupied, const orc_trackpoints* tp, float th, float nnratio,
                                       int* match_of_key)
{
    int nmatches = 0;
    Grid g;
    build_grid(F, g);
    const bool bFactor = th != 1.0;
    std::vector<int> assigned(F->n, -1);
    std::vector<uint8_t> blocked(F->n, 0);
    for (int i = 0; i < F->n; ++i) blocked[i] = occupied ? occupied[i] : 0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < tp->n; iMP++) {
        if (!tp->valid[iMP]) continue;  // mbTrackInView && !isBad()
        const int nPredictedLevel = tp->level[iMP];
        float r = ((double)tp->view_cos[iMP] > 0.998) ? 2.5f : 4.0f;  // RadiusByViewingCos :134-140
        if (bFactor) r *= th;
        const float radius = r * F->scale_factors[nPredictedLevel];
        features_in_area(F, g, tp->proj[3 * iMP], tp->proj[3 * iMP + 1], radius, nPredictedLevel - 1, nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = tp->point_desc + 32 * (size_t)iMP;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (blocked[idx]) continue;  // F.mvpMapPoints[idx] && Observations() > 0
            if (F->u_right && F->u_right[idx] > 0) {
                const float er = std::fabs(tp->proj[3 * iMP + 2] - F->u_right[idx]);
                if (er > radius) continue;
            }
            const int dist = orc_hamming256(dMP, F->desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist;
                bestLevel2 = bestLevel; bestLevel = F->keys_un[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F->keys_un[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) continue;
            assigned[bestIdx] = iMP;
            blocked[bestIdx] = tp->observed[iMP];
            nmatches++;
        }
    }
    for (int i = 0; i < F->n; ++i) match_of_key[i] = assigned[i];
    return nmatches;
}
