// shim_main.cpp — drives include/ORBextractorGPU.h (the C++ drop-in with the reference's signature) the way
// Frame::Frame(stereo) does (src/Frame.cc:113-125): two extractors on two threads, then the stereo match.
// usage: shim_main left.raw right.raw rows cols nfeatures scale nlevels ini min mbf mb out.bin
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/ORBextractorGPU.h"

static std::vector<uint8_t> slurp(const char* path, size_t n)
{
    std::vector<uint8_t> v(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    fclose(f);
    return v;
}

int main(int argc, char** argv)
{
    if (argc != 13) return 2;
    const int rows = atoi(argv[3]), cols = atoi(argv[4]), nf = atoi(argv[5]), nl = atoi(argv[7]), ini = atoi(argv[8]), mn = atoi(argv[9]);
    const float sf = (float)atof(argv[6]), mbf = (float)atof(argv[10]), mb = (float)atof(argv[11]);
    const std::vector<uint8_t> imL = slurp(argv[1], (size_t)rows * cols), imR = slurp(argv[2], (size_t)rows * cols);
    try {
        ORB_SLAM2::ORBextractorGPU exL(nf, sf, nl, ini, mn), exR(nf, sf, nl, ini, mn);
        std::vector<orbx_keypoint> kL, kR;
        std::vector<uint8_t> dL, dR;
        std::thread tl([&] { exL(imL.data(), rows, cols, cols, kL, dL); });
        std::thread tr([&] { exR(imR.data(), rows, cols, cols, kR, dR); });
        tl.join();
        tr.join();
        std::vector<float> uR, depth;
        ORB_SLAM2::ComputeStereoMatchesGPU(exL, exR, kL.data(), dL.data(), (int)kL.size(), kR.data(), dR.data(), (int)kR.size(), mbf, mb, uR, depth);
        FILE* f = fopen(argv[12], "wb");
        const int n[2] = {(int)kL.size(), (int)kR.size()};
        fwrite(n, sizeof(int), 2, f);
        fwrite(kL.data(), sizeof(orbx_keypoint), kL.size(), f);
        fwrite(dL.data(), 1, dL.size(), f);
        fwrite(kR.data(), sizeof(orbx_keypoint), kR.size(), f);
        fwrite(dR.data(), 1, dR.size(), f);
        fwrite(uR.data(), sizeof(float), uR.size(), f);
        fwrite(depth.data(), sizeof(float), depth.size(), f);
        fclose(f);
        printf("levels %d scale %.3f nL %d nR %d\n", exL.GetLevels(), exL.GetScaleFactor(), n[0], n[1]);
        // the one-call form must leave exactly what the reference-shaped sequence above left
        std::vector<orbx_keypoint> k2L, k2R;
        std::vector<uint8_t> d2L, d2R;
        std::vector<float> u2, depth2;
        for (int rep = 0; rep < 3; ++rep)  // direct launches, graph capture, graph replay
            ORB_SLAM2::StereoFrameGPU(exL, exR, imL.data(), imR.data(), rows, cols, cols, cols, mbf, mb, k2L, d2L, k2R, d2R, u2, depth2);
        const bool same = k2L.size() == kL.size() && k2R.size() == kR.size() && d2L == dL && d2R == dR &&
                          !memcmp(k2L.data(), kL.data(), kL.size() * sizeof(orbx_keypoint)) &&
                          !memcmp(k2R.data(), kR.data(), kR.size() * sizeof(orbx_keypoint)) &&
                          !memcmp(u2.data(), uR.data(), uR.size() * 4) && !memcmp(depth2.data(), depth.data(), depth.size() * 4);
        if (!same) { fprintf(stderr, "StereoFrameGPU differs from operator() x2 + ComputeStereoMatchesGPU\n"); return 3; }
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
