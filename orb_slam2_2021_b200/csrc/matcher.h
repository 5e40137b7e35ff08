// matcher.h — the orbx_matcher context shared by frame.cu (include/orbx_frame.h) and input.cu
// (include/orbx_input.h): one CUDA stream, one growable device buffer, pinned host staging for the packed
// inputs and outputs of a call.
#pragma once
#include <string>

#include "engine.h"

#ifndef TRY
#define TRY(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)
#endif
std::string& orbx_err_slot();  // abi.cu: the thread's orbx_last_error() text

struct orbx_matcher {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint8_t* dBuf = nullptr; size_t dCap = 0;   // packed inputs | scratch | outputs
    uint8_t* hIn = nullptr; size_t hInCap = 0;  // pinned staging of the inputs
    uint8_t* hOut = nullptr; size_t hOutCap = 0;
    cudaEvent_t orderEvent = nullptr;  // orders the matcher's stream behind a batch engine's (device-resident sources)
};

inline int orbx_fail(int code, const char* msg) { orbx_err_slot() = msg; return code; }

inline int orbx_matcher_grow(orbx_matcher* m, size_t dBytes, size_t inBytes, size_t outBytes)
{
    std::string& err = orbx_err_slot();
    if (dBytes > m->dCap) {
        if (m->dBuf) cudaFree(m->dBuf);
        m->dBuf = nullptr; m->dCap = 0;
        TRY(orbx_check(cudaMalloc(&m->dBuf, dBytes + dBytes / 2), "matcher device buffer", err));
        m->dCap = dBytes + dBytes / 2;
    }
    if (inBytes > m->hInCap) {
        if (m->hIn) cudaFreeHost(m->hIn);
        m->hIn = nullptr; m->hInCap = 0;
        TRY(orbx_check(cudaHostAlloc(&m->hIn, inBytes + inBytes / 2, cudaHostAllocDefault), "matcher pinned input", err));
        m->hInCap = inBytes + inBytes / 2;
    }
    if (outBytes > m->hOutCap) {
        if (m->hOut) cudaFreeHost(m->hOut);
        m->hOut = nullptr; m->hOutCap = 0;
        TRY(orbx_check(cudaHostAlloc(&m->hOut, outBytes + outBytes / 2, cudaHostAllocDefault), "matcher pinned output", err));
        m->hOutCap = outBytes + outBytes / 2;
    }
    return ORBX_OK;
}

// 16-byte aligned sub-allocation inside one buffer
struct OrbxPacker {
    size_t off = 0;
    size_t take(size_t bytes) { const size_t o = off; off = (off + bytes + 15) & ~(size_t)15; return o; }
};
