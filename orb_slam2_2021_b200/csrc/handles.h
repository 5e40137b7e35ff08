// handles.h — the single-image handle behind orbx_extractor* (include/orbx.h); shared by abi.cu and stereo.cu.
#pragma once
#include <memory>
#include <string>

#include "engine.h"

struct orbx_extractor {
    ExtractorConfig cfg;
    int device = 0;
    std::unique_ptr<Engine> eng;  // capacity 1, rebuilt when the image size changes
    bool hasRun = false;
    // stereo scratch (used when this handle is the LEFT one)
    float* dU = nullptr; float* dD = nullptr; int* dSad = nullptr; int* dNm = nullptr;
    uint4* dSorted = nullptr; int* dBin = nullptr;
    int scratchRows = 0;
    int scratchKp = 0;
    uint8_t* hOut = nullptr; size_t hOutCap = 0;  // pinned staging of orbx_stereo_match_resident's results
    // orbx_stereo_frame: the LEFT handle owns an engine for the frame's two images (image 0 = left, 1 = right); both handles
    // then read their pyramid levels from it (view) until their next orbx_extract
    std::unique_ptr<Engine> pair;
    orbx_extractor* pairPeer = nullptr;  // the other handle whose view points into this handle's pair engine
    orbx_extractor* pairOwner = nullptr; // (on that other handle) the handle that owns the engine its view points into
    Engine* view = nullptr;              // where the most recent pyramid of this handle lives
    int viewImg = 0;
};

struct orbx_batch {
    ExtractorConfig cfg;
    Engine eng;
    int maxFrames = 0;
    bool stereo = false;
};

std::string& orbx_err_slot();  // thread-local text behind orbx_last_error() (abi.cu)
