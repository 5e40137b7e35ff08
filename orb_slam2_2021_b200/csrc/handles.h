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
    uint8_t* hOut = nullptr; size_t hOutCap = 0;  // pinned staging of orbx_extract's results
};

std::string& orbx_err_slot();  // thread-local text behind orbx_last_error() (abi.cu)
