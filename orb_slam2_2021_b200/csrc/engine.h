// engine.h — host side of the device pipeline: geometry tables, HBM buffers, launch sequence.
// One Engine serves a fixed (extractor parameters, image size, max images) configuration. The
// single-image handle (orbx_extractor) owns an Engine of capacity 1 that is rebuilt when the image
// size changes; the batch handle (orbx_batch) owns one of capacity max_frames*(1|2).
#pragma once
#include <string>
#include <vector>

#include "orbx_dev.cuh"

struct ExtractorConfig {
    int nfeatures, nlevels, iniTh, minTh;
    float scaleFactor;
    // ORBextractor::ORBextractor tables (src/ORBextractor.cc:413-449)
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota;
    void init(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh);
};

class Engine {
public:
    Engine() = default;
    ~Engine();
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;

    // returns 0 or an orbx_status; err receives the text
    int init(const ExtractorConfig& cfg, int rows, int cols, int maxImgs, bool stereo, int device, std::string& err);
    void release();

    // level geometry without allocating anything (for orbx_max_keypoints)
    // TMA descriptors of the un-blurred levels with the given box (device array of nlevels CUtensorMap)
    int makeLevelMaps(int boxW, int boxH, bool blurred, void** dMaps, std::string& err, const int* boxWs = nullptr, const int* boxHs = nullptr);
    static int plan(const ExtractorConfig& cfg, int rows, int cols, OrbxParams& p, std::string& err);

    int uploadImages(const uint8_t* left, const uint8_t* right, int nFrames, bool fromDevice, std::string& err);
    int uploadSingle(const uint8_t* img, size_t stride, std::string& err);
    int runExtract(int nImgs, std::string& err);                         // pyramid .. describe
    int runStereo(int nFrames, float mbf, float mb, std::string& err);   // both sides in this engine
    int runBatch(int nFrames, float mbf, float mb, int nChunks, std::string& err);  // extract (+ stereo), chunks on two streams
    // The same on images [imgBase, imgBase+nImgs) / frames [f0, f0+nFrames) only (chunked pipeline; no stage timing)
    OrbxParams slice(int imgBase, int nImgs) const;
    void launchExtract(const OrbxParams& p, cudaStream_t s);
    void launchPyramid(const OrbxParams& p, cudaStream_t s);
    void launchStereo(int f0, int nFrames, float mbf, float mb, cudaStream_t s);
    // upload + run + download with H2D, kernels and D2H of successive frame chunks overlapped on three streams
    struct HostOutputs {
        int* nLeft; void* kpsLeft; uint8_t* descLeft; int* nRight; void* kpsRight; uint8_t* descRight; float* uRight; float* depth;
    };
    int processPipelined(const uint8_t* left, const uint8_t* right, int nFrames, float mbf, float mb, const HostOutputs& out,
                         int chunkFrames, std::string& err, bool wait = true);
    int waitPipelined(std::string& err);  // drains an asynchronous processPipelined
    cudaStream_t upStream = nullptr, downStream = nullptr, stream2 = nullptr;
    cudaStream_t stream3 = nullptr, stream4 = nullptr;  // further compute streams of the end-to-end pipeline
    // raw-frame input path (include/orbx_input.h): raw frames and per-side rectification maps on the device
    uint8_t* dRaw = nullptr; size_t rawCap = 0;
    float* dMap1[2] = {nullptr, nullptr}; float* dMap2[2] = {nullptr, nullptr};
    int mapRows[2] = {0, 0}, mapCols[2] = {0, 0}, cropTop[2] = {0, 0};
    std::vector<cudaEvent_t> chunkEvents;
    // cross-call ordering of the asynchronous end-to-end pipeline: per chunk, "staging consumed" and "results copied out"
    std::vector<cudaEvent_t> ingestEvents, downEvents;
    std::vector<int> lastBounds;  // chunk boundaries of the previous pipelined call still in flight (empty = none)
    int lastRunChunks = 1;  // chunks of the most recent runBatch (kernel launches = chunks x launches per chunk)
    int sync(std::string& err);

    // Single-frame path (orbx_extract: capacity 1; orbx_stereo_frame: capacity 2, stereo): the images of ONE frame go up
    // through pinned staging with one flat copy each, the whole launch sequence (ingest, extract [, stereo match]) and the
    // copies of every result into pinned host memory replay as ONE CUDA graph (captured on the second call for a given
    // mbf/mb), and one synchronisation ends the call. Results: hOut + frameOff* (full pitch maxKp per image).
    uint8_t* hIn = nullptr;   // pinned, [maxImgs][rows*cols]
    uint8_t* hOut = nullptr;  // pinned: counts | keypoints | descriptors | uRight | depth
    size_t frameOffKp = 0, frameOffDesc = 0, frameOffU = 0, frameOffD = 0, hOutBytes = 0;
    cudaGraphExec_t frameGraph = nullptr;
    cudaEvent_t frameFork = nullptr, frameJoin = nullptr;
    float graphMbf = 0.f, graphMb = 0.f;
    int frameCalls = 0;
    bool graphBroken = false;
    bool residentValid = false;  // outKp/outDesc/outCnt on the device are what the last processFrame produced (hOut holds the counts)
    int processFrame(const uint8_t* const* images, const size_t* strides, float mbf, float mb, std::string& err);
    void enqueueFrame(float mbf, float mb, cudaStream_t s);

    OrbxParams P{};          // device pointers + geometry (imgs = capacity)
    OrbxStereoParams SP{};   // stereo outputs (batch engines with stereo == true)
    ExtractorConfig cfg;
    int device = 0, maxImgs = 0, rows = 0, cols = 0;
    bool stereo = false;
    cudaStream_t stream = nullptr;
    int nTiles = 0;
    void* dCells = nullptr;
    void* dTiles = nullptr;
    int* dResizeTab = nullptr;
    uint8_t* dStage = nullptr;  // tight-row landing buffer of host uploads ([side][frame][rows*cols])
    void* dResizeMaps = nullptr;  // un-blurred level l with the source window of a destination tile of level l + 1 as its box
    std::vector<int> rsBoxW, rsBoxH;
    void* dFastMaps = nullptr;  // same tensors, box = (64|96) x (largest cell ROI height) x 1
    void* dDescMaps0 = nullptr;  // un-blurred levels, box 48 x 31 (IC_Angle disc)
    void* dDescMaps1 = nullptr;  // BLURRED levels, box 64 x 37 (rBRIEF samples)
    void* dBlurMaps = nullptr;  // CUtensorMap[nlevels]: un-blurred level l as a (w, h, image) u8 tensor, box 128 x 70 x 1
    std::vector<const int*> resizeTabs;  // per level device pointers into dResizeTab
    std::vector<void*> allocs;
    size_t bytesAllocated = 0;
    int launchesPerExtract = 0;

    // optional per-stage timing with CUDA events on `stream` (bench.py's roofline leg)
    enum { kStages = 7, kMaxProfiledRuns = 512 };
    bool profiling = false;
    std::vector<cudaEvent_t> events;  // (kStages+1) per profiled run
    int profiledRuns = 0;
    int setProfiling(bool on, std::string& err);
    int stageTimes(double* ms, int* runs, std::string& err);  // accumulates + resets
    void mark(int stage);
    static const char* stageName(int i);
};

int orbx_check(cudaError_t e, const char* what, std::string& err);

// Debug/experiment knobs, set through the one entry point orbx_set_tuning() (include/orbx.h); 0 = built-in default.
// Product code never reads the environment.
enum OrbxTune {
    ORBX_TUNE_RUN_STREAMS,     // compute streams of a device-resident run (2..4)
    ORBX_TUNE_RUN_CHUNKS,      // chunks of a device-resident run
    ORBX_TUNE_E2E_STREAMS,     // compute streams of the end-to-end pipeline (1..4)
    ORBX_TUNE_E2E_CHUNKS,      // chunks of an end-to-end call
    ORBX_TUNE_E2E_FIRST_DIV,   // split of the first chunk of a blocking end-to-end call
    ORBX_TUNE_E2E_TRACE,       // per-chunk copy/compute times on stderr
    ORBX_TUNE_FAST_CELLS,      // cells (warps) per FAST CTA: 1 (default) / 2 / 4 / 8
    ORBX_TUNE_FAST_GRID_MULT,  // FAST CTAs launched per resident CTA slot
    ORBX_TUNE_NUMA_OFF,        // 1: orbx_host_alloc skips the NUMA placement
    ORBX_TUNE_PYR_CHAIN,       // pyramid: 0 = one launch for <= 8 images, else one per level; 1 = always one launch; 2 = always per level
    ORBX_TUNE_PYR_TMA,         // per-level resize: 0 / 1 = TMA-staged source tiles (default), 2 = the word-load kernel
    ORBX_TUNE_FAST_CTAS,       // FAST: at most this many CTAs per SM (0 = as many as fit), leaves room for kernels on other streams
    ORBX_TUNE_COUNT
};
int orbx_tune(OrbxTune key);
int orbx_stereo_bin_radius(const float* scale, int nlevels);
