// oracle/oracle_internal.h — shared declarations of the CPU oracle (TEST INFRASTRUCTURE).
#pragma once
#include <cstdint>
#include <vector>

#include "orb_oracle.h"

struct OrcImage {
    int w = 0, h = 0;
    std::vector<uint8_t> px;  // tight rows, stride == w
};

struct orc_extractor {
    int nfeatures, nlevels, ini_th, min_th;
    double scale_factor;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota, umax;
    std::vector<OrcImage> levels;   // mvImagePyramid (un-blurred)
    std::vector<OrcImage> blurred;  // workingMat per level
    std::vector<std::vector<orc_keypoint>> candidates;  // vToDistributeKeys per level
};

// DistributeOctTree (src/ORBextractor.cc:542-766) with the canonical tie-break.
void orc_distribute_impl(const std::vector<orc_keypoint>& cand, int minX, int maxX, int minY, int maxY, int N,
                         std::vector<orc_keypoint>& out);
