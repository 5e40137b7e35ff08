// oracle/stereo.cpp — CPU restatement of Frame::ComputeStereoMatches
// (/root/reference/src/Frame.cc:522-700). TEST INFRASTRUCTURE (see orb_oracle.h).
#include <algorithm>
#include <climits>
#include <cmath>
#include <utility>
#include <vector>

#include "../orb_slam2_2021_b200/csrc/orbx_math.h"
#include "oracle_internal.h"

extern "C" int orc_stereo_match(const orc_extractor* left, const orc_extractor* right, const orc_keypoint* kps_l,
                                const uint8_t* desc_l, int n_l, const orc_keypoint* kps_r, const uint8_t* desc_r,
                                int n_r, float mbf, float mb, float* u_right, float* depth)
{
    for (int i = 0; i < n_l; ++i) u_right[i] = depth[i] = -1.0f;  // :524-525
    if (!left || !right || n_l <= 0) return 0;
    const int TH_HIGH = 100, TH_LOW = 50;           // src/ORBmatcher.cc:37-38
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;   // :527
    const int nRows = left->levels[0].h;            // :529
    const std::vector<float>& sf = left->scale;     // Frame copies the LEFT extractor's tables, :104-110
    const std::vector<float>& isf = left->inv_scale;

    std::vector<std::vector<int>> rowIdx(nRows);  // :532-549
    for (int iR = 0; iR < n_r; ++iR) {
        const float y = kps_r[iR].y;
        const float r = orbx_fmul(2.0f, sf[kps_r[iR].octave]);
        const int maxr = (int)orbx_ceilf(orbx_fadd(y, r));
        const int minr = (int)orbx_floorf(orbx_fsub(y, r));
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);  // reference indexes unchecked
    }

    const float minZ = mb, minD = 0;      // :552-554
    const float maxD = orbx_fdiv(mbf, minZ);
    std::vector<std::pair<int, int>> distIdx;

    for (int iL = 0; iL < n_l; ++iL) {  // :560
        const orc_keypoint& kl = kps_l[iL];
        const int levelL = kl.octave;
        const float vL = kl.y, uL = kl.x;
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        const std::vector<int>& cands = rowIdx[row];
        if (cands.empty()) continue;
        const float minU = orbx_fsub(uL, maxD), maxU = orbx_fsub(uL, minD);
        if (maxU < 0) continue;

        int bestDist = TH_HIGH, bestR = 0;
        for (int iR : cands) {  // :582-603
            const orc_keypoint& kr = kps_r[iR];
            if (kr.octave < levelL - 1 || kr.octave > levelL + 1) continue;
            if (kr.x >= minU && kr.x <= maxU) {
                int d = orc_hamming256(desc_l + (size_t)iL * 32, desc_r + (size_t)iR * 32);
                if (d < bestDist) { bestDist = d; bestR = iR; }
            }
        }
        if (bestDist >= thOrbDist) continue;  // :606

        // :608-655 SAD refinement on the un-blurred level of the LEFT keypoint
        const float uR0 = kps_r[bestR].x;
        const float s = isf[levelL];
        const float su = orbx_roundf(orbx_fmul(kl.x, s));
        const float sv = orbx_roundf(orbx_fmul(kl.y, s));
        const float sr = orbx_roundf(orbx_fmul(uR0, s));
        const int w = 5, L = 5;
        const OrcImage& IL = left->levels[levelL];
        const OrcImage& IR = right->levels[levelL];
        const int cu = (int)su, cv = (int)sv, cr = (int)sr;
        const float iniu = sr + L - w, endu = sr + L + w + 1;
        if (iniu < 0 || endu >= (float)IR.w) continue;  // :632-635
        // windows that leave the level would make the reference's rowRange/colRange throw
        if (cv - w < 0 || cv + w >= IL.h || cu - w < 0 || cu + w >= IL.w || cv + w >= IR.h || cr - L - w < 0 ||
            cr + L + w >= IR.w)
            continue;
        int bestSad = INT_MAX, bestInc = 0;
        float dists[2 * 5 + 1];
        const int cl = IL.px[(size_t)cv * IL.w + cu];
        for (int inc = -L; inc <= L; ++inc) {
            const int crv = IR.px[(size_t)cv * IR.w + cr + inc];
            int sad = 0;  // sum |(IL - IL_c) - (IR - IR_c)|: small integers, exact in float as well
            for (int dy = -w; dy <= w; ++dy)
                for (int dx = -w; dx <= w; ++dx) {
                    int a = (int)IL.px[(size_t)(cv + dy) * IL.w + cu + dx] - cl;
                    int b = (int)IR.px[(size_t)(cv + dy) * IR.w + cr + inc + dx] - crv;
                    sad += std::abs(a - b);
                }
            float dist = (float)sad;
            if (dist < (float)bestSad) { bestSad = (int)dist; bestInc = inc; }
            dists[L + inc] = dist;
        }
        if (bestInc == -L || bestInc == L) continue;  // :654

        const float d1 = dists[L + bestInc - 1], d2 = dists[L + bestInc], d3 = dists[L + bestInc + 1];  // :657-662
        const float denom = orbx_fmul(2.0f, orbx_fsub(orbx_fadd(d1, d3), orbx_fmul(2.0f, d2)));
        const float deltaR = orbx_fdiv(orbx_fsub(d1, d3), denom);
        if (deltaR < -1 || deltaR > 1) continue;  // NaN passes, then fails disparity >= 0 below

        float bestuR = orbx_fmul(sf[levelL], orbx_fadd(orbx_fadd(sr, (float)bestInc), deltaR));  // :668
        float disparity = orbx_fsub(uL, bestuR);
        if (disparity >= minD && disparity < maxD) {  // :672-682
            if (disparity <= 0) {
                disparity = 0.01f;
                bestuR = (float)((double)uL - 0.01);
            }
            depth[iL] = orbx_fdiv(mbf, disparity);
            u_right[iL] = bestuR;
            distIdx.push_back({bestSad, iL});
        }
    }

    if (distIdx.empty()) return 0;  // reference: UB (indexes an empty vector, :687); defined as no-op
    std::sort(distIdx.begin(), distIdx.end());  // :686-699
    const float median = (float)distIdx[distIdx.size() / 2].first;
    const float thDist = orbx_fmul(orbx_fmul(1.5f, 1.4f), median);
    int valid = (int)distIdx.size();
    for (int i = (int)distIdx.size() - 1; i >= 0; --i) {
        if ((float)distIdx[i].first < thDist) break;
        u_right[distIdx[i].second] = -1;
        depth[distIdx[i].second] = -1;
        --valid;
    }
    return valid;
}
