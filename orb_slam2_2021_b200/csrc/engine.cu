// engine.cu — host side: extractor tables, level/cell/tile geometry, HBM buffers, launch sequence.
#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include <cuda.h>  // CUtensorMap types only; the encoder is fetched from the driver at run time (no -lcuda)

#include "../../include/orbx.h"

int orbx_check(cudaError_t e, const char* what, std::string& err)
{
    if (e == cudaSuccess) return ORBX_OK;
    err = std::string(what) + ": " + cudaGetErrorString(e);
    return ORBX_ERR_CUDA;
}

// ORBextractor::ORBextractor, /root/reference/src/ORBextractor.cc:413-449. scaleFactor is stored in a
// double member initialised from the float argument (include/ORBextractor.h:113), so products and
// quotients below are taken in double and rounded to float exactly where the reference does.
void ExtractorConfig::init(int nf, float sf, int nl, int ini, int mn)
{
    nfeatures = nf; scaleFactor = sf; nlevels = nl; iniTh = ini; minTh = mn;
    const double sfd = (double)sf;
    scale.assign(nl, 1.f); sigma2.assign(nl, 1.f); invScale.assign(nl, 1.f); invSigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        scale[i] = (float)((double)scale[i - 1] * sfd);
        sigma2[i] = scale[i] * scale[i];
    }
    for (int i = 0; i < nl; ++i) {
        invScale[i] = 1.0f / scale[i];
        invSigma2[i] = 1.0f / sigma2[i];
    }
    quota.assign(nl, 0);
    const float factor = (float)(1.0 / sfd);
    float desired = nf * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        quota[l] = orbx_cvround(desired);
        sum += quota[l];
        desired *= factor;
    }
    quota[nl - 1] = nf - sum > 0 ? nf - sum : 0;
}

// Rows a right keypoint can serve: floor(y-r)..ceil(y+r), r = 2*scale[octave] (src/Frame.cc:539-549); a left
// keypoint on row v therefore only needs right keypoints with |(int)y - v| <= ceil(2*max scale) + 1.
int orbx_stereo_bin_radius(const float* scale, int nlevels)
{
    float m = 1.f;
    for (int i = 0; i < nlevels; ++i) m = scale[i] > m ? scale[i] : m;
    return (int)std::ceil(2.0f * m) + 1;
}

// ---- tuning table (orbx_set_tuning) ---------------------------------------------------------------------
static int g_tune[ORBX_TUNE_COUNT] = {0};
static const char* const g_tuneNames[ORBX_TUNE_COUNT] = {"run_streams", "run_chunks", "e2e_streams", "e2e_chunks", "e2e_first_div",
                                                         "e2e_trace", "fast_cells", "fast_grid_mult", "numa_off", "pyr_chain", "pyr_tma", "fast_ctas"};
int orbx_tune(OrbxTune key) { return g_tune[key]; }
extern "C" int orbx_set_tuning(const char* key, int value)
{
    for (int i = 0; key && i < ORBX_TUNE_COUNT; ++i)
        if (!std::strcmp(key, g_tuneNames[i])) { g_tune[i] = value; return ORBX_OK; }
    return ORBX_ERR_ARG;
}

static inline unsigned align_up(unsigned v, unsigned a) { return (v + a - 1) / a * a; }

struct HostCell {
    short x0, y0;
    uint8_t rw, rh, level, pad;
};
struct HostTile {
    short x0, y0;
    uint8_t level, pad[3];
};

// Cell grid of one level, /root/reference/src/ORBextractor.cc:776-809 (float arithmetic kept).
static void level_cells(int w, int h, int level, OrbxLevel& L, std::vector<HostCell>* out)
{
    const int minBX = ORBX_EDGE - 3, minBY = minBX, maxBX = w - ORBX_EDGE + 3, maxBY = h - ORBX_EDGE + 3;
    const float W = 30;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    L.nCols = (int)(width / W);
    L.nRows = (int)(height / W);
    L.wCell = (int)std::ceil(width / L.nCols);
    L.hCell = (int)std::ceil(height / L.nRows);
    L.nCells = 0;
    for (int i = 0; i < L.nRows; ++i) {
        const float iniY = (float)(minBY + i * L.hCell);
        float maxY = iniY + L.hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < L.nCols; ++j) {
            const float iniX = (float)(minBX + j * L.wCell);
            float maxX = iniX + L.wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = (float)maxBX;
            if (out) {
                HostCell c;
                c.x0 = (short)iniX; c.y0 = (short)iniY;
                c.rw = (uint8_t)((int)maxX - (int)iniX); c.rh = (uint8_t)((int)maxY - (int)iniY);
                c.level = (uint8_t)level; c.pad = 0;
                out->push_back(c);
            }
            ++L.nCells;
        }
    }
}

int Engine::plan(const ExtractorConfig& cfg, int rows, int cols, OrbxParams& p, std::string& err)
{
    if (cfg.nlevels < 1 || cfg.nlevels > ORBX_MAX_LEVELS) { err = "nlevels out of range (1..12)"; return ORBX_ERR_ARG; }
    if (rows <= 0 || cols <= 0 || cols > 32000 || rows > 32000) { err = "bad image size"; return ORBX_ERR_ARG; }
    std::memset(&p, 0, sizeof(p));
    p.nlevels = cfg.nlevels; p.iniTh = cfg.iniTh; p.minTh = cfg.minTh; p.rows = rows; p.cols = cols;
    unsigned off = 0, boff = 0, slot = 0, cand = 0;
    int cellBase = 0, keepBase = 0, spillCap = 0;
    for (int l = 0; l < cfg.nlevels; ++l) {
        OrbxLevel& L = p.lv[l];
        L.w = orbx_cvround((float)cols * cfg.invScale[l]);  // :1110
        L.h = orbx_cvround((float)rows * cfg.invScale[l]);
        if (L.w - 32 < 30 || L.h - 32 < 30) {
            err = "pyramid level " + std::to_string(l) + " smaller than 62 px: the reference divides by zero here";
            return ORBX_ERR_GEOMETRY;
        }
        L.pitch = (int)align_up(L.w, 16);
        L.bpitch = (int)align_up(L.w, 16);
        L.off = off;
        if (l > 0) off += align_up((unsigned)L.pitch * L.h, 256);
        L.boff = boff;
        boff += align_up((unsigned)L.bpitch * L.h, 256);
        level_cells(L.w, L.h, l, L, nullptr);
        L.cellBase = cellBase; cellBase += L.nCells;
        L.slotCap = ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);
        L.slotBase = slot; slot += (unsigned)L.nCells * L.slotCap;
        L.candBase = cand; L.candCap = L.nCells * L.slotCap; cand += L.candCap;
        if (L.candCap >= (1 << 24)) { err = "level too large (candidate index needs 24 bits)"; return ORBX_ERR_GEOMETRY; }
        L.quota = cfg.quota[l];
        const int bw = L.w - 32, bh = L.h - 32;
        L.nIni = (int)std::round((float)bw / bh);  // :546
        if (L.nIni < 1) { err = "portrait level: nIni == 0, the reference divides by zero here"; return ORBX_ERR_GEOMETRY; }
        L.hX = (float)bw / L.nIni;  // :548
        int cap = (L.quota > 4 * L.nIni ? L.quota : 4 * L.nIni) + 4;
        L.keepCap = (cap + 7) & ~7;  // whole describe CTAs (8 warps = 8 slots) per level
        // the quadtree keeps a level's node list in the shared memory of one CTA (227 KB opt-in maximum per CTA on sm_100) and
        // moves it to global memory above that (~2600 nodes per level); list positions are 16 bits wide
        if (L.keepCap > 65000) {
            err = "per-level feature quota " + std::to_string(L.quota) + " (level " + std::to_string(l) + ") too large: the quadtree holds at most 65000 nodes per level";
            return ORBX_ERR_ARG;
        }
        if (orbx_octree_smem_total(L.keepCap) > 227u * 1024u) spillCap = std::max(spillCap, L.keepCap);
        L.keepBase = keepBase; keepBase += L.keepCap;
        L.scale = cfg.scale[l]; L.invScale = cfg.invScale[l];
        L.patchSize = (int)(31 * cfg.scale[l]);  // :836
    }
    p.img0Stride = (size_t)rows * p.lv[0].pitch;
    p.pyrStride = off; p.blurStride = boff;
    p.cellsTotal = cellBase; p.slotTotal = slot; p.candTotal = cand;
    p.keepTotal = keepBase; p.maxKp = keepBase;
    p.octSpill = nullptr;
    p.octSpillStride = spillCap ? align_up(orbx_octree_list_bytes(spillCap), 256) : 0;
    if (p.maxKp > 65535) { err = "more than 65535 keypoints per image not supported"; return ORBX_ERR_ARG; }
    return ORBX_OK;
}

// cv::resize(INTER_LINEAR) coefficient tables for one axis (OpenCV imgproc/resize.cpp): 11-bit fixed
// point, computed in the same float64/float32 steps on the host so the kernel only does integer math.
static void axis_table(int src, int dst, bool clampFrac, int* ofs, int* coef)
{
    const double scale = 1.0 / ((double)dst / src);
    for (int d = 0; d < dst; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (clampFrac) {
            if (s < 0) { s = 0; f = 0.f; }
            if (s >= src - 1) { s = src - 1; f = 0.f; }
        }
        const int c0 = (short)orbx_cvround((1.f - f) * 2048.f), c1 = (short)orbx_cvround(f * 2048.f);
        ofs[d] = s;
        coef[d] = (c0 & 0xFFFF) | (c1 << 16);
    }
}

Engine::~Engine() { release(); }

void Engine::release()
{
    if (allocs.empty() && !stream) return;
    cudaSetDevice(device);
    for (void* a : allocs) cudaFree(a);
    allocs.clear();
    if (stream) cudaStreamDestroy(stream);
    if (dRaw) { cudaFree(dRaw); dRaw = nullptr; rawCap = 0; }
    for (int s = 0; s < 2; ++s) {
        if (dMap1[s]) cudaFree(dMap1[s]);
        if (dMap2[s]) cudaFree(dMap2[s]);
        dMap1[s] = dMap2[s] = nullptr;
    }
    if (upStream) cudaStreamDestroy(upStream);
    if (downStream) cudaStreamDestroy(downStream);
    if (stream2) cudaStreamDestroy(stream2);
    if (stream3) cudaStreamDestroy(stream3);
    if (stream4) cudaStreamDestroy(stream4);
    for (cudaEvent_t e : chunkEvents) cudaEventDestroy(e);
    chunkEvents.clear();
    for (cudaEvent_t e : ingestEvents) cudaEventDestroy(e);
    for (cudaEvent_t e : downEvents) cudaEventDestroy(e);
    ingestEvents.clear(); downEvents.clear(); lastBounds.clear();
    if (frameGraph) { cudaGraphExecDestroy(frameGraph); frameGraph = nullptr; }
    if (hIn) { cudaFreeHost(hIn); hIn = nullptr; }
    if (hOut) { cudaFreeHost(hOut); hOut = nullptr; }
    if (frameFork) { cudaEventDestroy(frameFork); frameFork = nullptr; }
    if (frameJoin) { cudaEventDestroy(frameJoin); frameJoin = nullptr; }
    frameCalls = 0; graphBroken = false; residentValid = false;
    for (cudaEvent_t e : events) if (e) cudaEventDestroy(e);
    events.clear(); profiling = false; profiledRuns = 0;
    stream = stream2 = stream3 = stream4 = upStream = downStream = nullptr;
    bytesAllocated = 0;
}

template <typename T>
static int dalloc(Engine* e, T** ptr, size_t count, const char* what, std::string& err)
{
    void* d = nullptr;
    const size_t bytes = (count ? count : 1) * sizeof(T);
    int rc = orbx_check(cudaMalloc(&d, bytes), what, err);
    if (rc) return rc;
    e->allocs.push_back(d);
    e->bytesAllocated += bytes;
    *ptr = (T*)d;
    return ORBX_OK;
}

#define TRY(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int Engine::makeLevelMaps(int boxW, int boxH, bool blurred, void** dMaps, std::string& err, const int* boxWs, const int* boxHs)
{
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    TRY(orbx_check(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q), "cudaGetDriverEntryPoint", err));
    if (!fn || q != cudaDriverEntryPointSuccess) { err = "driver has no cuTensorMapEncodeTiled"; return ORBX_ERR_CUDA; }
    std::vector<CUtensorMap> maps(P.nlevels);
    for (int l = 0; l < P.nlevels; ++l) {
        const OrbxLevel& L = P.lv[l];
        if (boxWs) {  // per-level boxes (resize: entry l = box over level l as the SOURCE of level l + 1; 0 = no box)
            boxW = boxWs[l] > 0 ? boxWs[l] : 16;
            boxH = boxHs[l] > 0 ? boxHs[l] : 1;
        }
        void* base = blurred ? (void*)(P.blur + L.boff) : l == 0 ? (void*)P.img0 : (void*)(P.pyr + L.off);
        const cuuint64_t dims[3] = {(cuuint64_t)L.w, (cuuint64_t)L.h, (cuuint64_t)maxImgs};
        const cuuint64_t strides[2] = {(cuuint64_t)(blurred ? L.bpitch : L.pitch),
                                       (cuuint64_t)(blurred ? P.blurStride : l == 0 ? P.img0Stride : P.pyrStride)};
        const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1u};
        const cuuint32_t es[3] = {1u, 1u, 1u};
        const CUresult r = ((EncodeTiledFn)fn)(&maps[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, es,
                                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { err = "cuTensorMapEncodeTiled failed for level " + std::to_string(l) + " (CUresult " + std::to_string((int)r) + ")"; return ORBX_ERR_CUDA; }
    }
    CUtensorMap* d = nullptr;
    TRY(dalloc(this, &d, maps.size(), "tensor maps", err));
    TRY(orbx_check(cudaMemcpy(d, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice), "tensor maps H2D", err));
    *dMaps = d;
    return ORBX_OK;
}

int Engine::init(const ExtractorConfig& c, int r, int w, int nImgs, bool st, int dev, std::string& err)
{
    release();
    cfg = c; rows = r; cols = w; maxImgs = nImgs; stereo = st; device = dev;
    if (nImgs < 1 || nImgs > 65535) { err = "image capacity must be 1..65535"; return ORBX_ERR_ARG; }
    TRY(plan(cfg, rows, cols, P, err));
    P.imgs = nImgs;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { err = "no CUDA device: liborbx has no CPU path"; return ORBX_ERR_CUDA; }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking), "cudaStreamCreate", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&upStream, cudaStreamNonBlocking), "cudaStreamCreate", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&downStream, cudaStreamNonBlocking), "cudaStreamCreate", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&stream2, cudaStreamNonBlocking), "cudaStreamCreate", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&stream3, cudaStreamNonBlocking), "cudaStreamCreate", err));
    TRY(orbx_check(cudaStreamCreateWithFlags(&stream4, cudaStreamNonBlocking), "cudaStreamCreate", err));

    const size_t n = (size_t)nImgs;
    TRY(dalloc(this, &P.img0, n * P.img0Stride + 64, "img0", err));
    TRY(dalloc(this, &dStage, n * rows * cols + 64, "stage", err));
    TRY(dalloc(this, &P.pyr, n * P.pyrStride + 64, "pyr", err));
    TRY(dalloc(this, &P.blur, n * P.blurStride + 64, "blur", err));
    TRY(dalloc(this, &P.cellCnt, n * P.cellsTotal, "cellCnt", err));
    TRY(dalloc(this, &P.cellKp, n * P.slotTotal, "cellKp", err));
    TRY(dalloc(this, &P.candXY, n * P.candTotal, "candXY", err));
    TRY(dalloc(this, &P.candRs, n * P.candTotal, "candRs", err));
    TRY(dalloc(this, &P.label, n * P.candTotal, "label", err));
    TRY(dalloc(this, &P.keep, n * P.keepTotal, "keep", err));
    TRY(dalloc(this, &P.keepCnt, n * P.nlevels, "keepCnt", err));
    TRY(dalloc(this, &P.candCnt, n * P.nlevels, "candCnt", err));
    TRY(dalloc(this, &P.keepOfs, n * P.nlevels, "keepOfs", err));
    uint8_t* okp = nullptr;
    TRY(dalloc(this, &okp, n * P.maxKp * 28, "outKp", err));
    P.outKp = okp;
    TRY(dalloc(this, &P.outDesc, n * P.maxKp * 32, "outDesc", err));
    TRY(dalloc(this, &P.outCnt, n, "outCnt", err));
    P.chainInts = orbx_pyramid_chain_state_ints(P);
    TRY(dalloc(this, &P.chainState, n * P.chainInts, "chainState", err));
    if (P.octSpillStride) TRY(dalloc(this, &P.octSpill, n * P.nlevels * P.octSpillStride, "quadtree node lists", err));
    TRY(orbx_check(cudaMemsetAsync(P.outCnt, 0, n * sizeof(int), stream), "memset", err));
    TRY(orbx_check(cudaMemsetAsync(P.keepCnt, 0, n * P.nlevels * sizeof(int), stream), "memset", err));
    TRY(orbx_check(cudaMemsetAsync(P.chainState, 0, n * P.chainInts * sizeof(int), stream), "memset", err));
    TRY(orbx_check(cudaMemsetAsync(P.candCnt, 0, n * P.nlevels * sizeof(int), stream), "memset", err));

    // cell table and blur tile table
    std::vector<HostCell> cells;
    std::vector<HostTile> tiles;
    for (int l = 0; l < P.nlevels; ++l) {
        OrbxLevel tmp = P.lv[l];
        level_cells(P.lv[l].w, P.lv[l].h, l, tmp, &cells);
        for (int y = 0; y < P.lv[l].h; y += ORBX_BLUR_TH)
            for (int x = 0; x < P.lv[l].w; x += ORBX_BLUR_TW) {
                HostTile t; t.x0 = (short)x; t.y0 = (short)y; t.level = (uint8_t)l; t.pad[0] = t.pad[1] = t.pad[2] = 0;
                tiles.push_back(t);
            }
    }
    nTiles = (int)tiles.size();
    HostCell* dc = nullptr; HostTile* dt = nullptr;
    TRY(dalloc(this, &dc, cells.size(), "cells", err));
    TRY(dalloc(this, &dt, tiles.size(), "tiles", err));
    dCells = dc; dTiles = dt;
    TRY(orbx_check(cudaMemcpy(dc, cells.data(), cells.size() * sizeof(HostCell), cudaMemcpyHostToDevice), "cells H2D", err));
    TRY(orbx_check(cudaMemcpy(dt, tiles.data(), tiles.size() * sizeof(HostTile), cudaMemcpyHostToDevice), "tiles H2D", err));

    // resize tables: per level >= 1: xofs[w] xcoef[w] yofs[h] ycoef[h]
    std::vector<int> tab;
    std::vector<size_t> tabOff(P.nlevels, 0);
    for (int l = 1; l < P.nlevels; ++l) {
        const OrbxLevel &S = P.lv[l - 1], &D = P.lv[l];
        tab.resize((tab.size() + 3) & ~(size_t)3);  // int4 alignment of the row table below
        tabOff[l] = tab.size();
        const size_t yOff = (2 * (size_t)D.w + 3) & ~(size_t)3;
        tab.resize(tab.size() + yOff + 4 * (size_t)D.h);
        int* base = tab.data() + tabOff[l];
        axis_table(S.w, D.w, true, base, base + D.w);
        std::vector<int> yofs(D.h), ycoef(D.h);
        axis_table(S.h, D.h, false, yofs.data(), ycoef.data());
        for (int y = 0; y < D.h; ++y) {  // {source row, b0 << 12, b1 << 12, 0}: the kernel multiplies with mulhi
            int* e4 = base + yOff + 4 * (size_t)y;
            e4[0] = yofs[y];
            e4[1] = (ycoef[y] & 0xFFFF) << 12;
            e4[2] = ((unsigned)ycoef[y] >> 16) << 12;
            e4[3] = 0;
        }
    }
    // TMA-staged resize (the default for batches; tuning key pyr_tma = 2 selects the word-load kernel): per destination level the largest source window of a 128 x 64 tile
    rsBoxW.assign(P.nlevels, 0); rsBoxH.assign(P.nlevels, 0);
    for (int l = 1; l < P.nlevels; ++l) {
        const OrbxLevel& D = P.lv[l];
        const int* xo = tab.data() + tabOff[l];
        const int* yt = xo + ((2 * (size_t)D.w + 3) & ~(size_t)3);
        int bw = 0, bh = 0;
        for (int x0 = 0; x0 < D.w; x0 += 128) {
            const int xl = std::min(x0 + 127, D.w - 1);
            bw = std::max(bw, xo[xl] + 12 - (xo[x0] & ~15));
        }
        for (int y0 = 0; y0 < D.h; y0 += 4 * ORBX_RS_TMA_ROWS) {
            const int yl = std::min(y0 + 4 * ORBX_RS_TMA_ROWS - 1, D.h - 1);
            bh = std::max(bh, yt[4 * yl] + 2 - yt[4 * y0]);
        }
        bw = (bw + 15) & ~15;
        if (bw <= 256 && bh <= 256) { rsBoxW[l - 1] = bw; rsBoxH[l - 1] = bh; }  // stored at the SOURCE level's index
    }
    TRY(dalloc(this, &dResizeTab, tab.size(), "resizeTab", err));
    TRY(orbx_check(cudaMemcpy(dResizeTab, tab.data(), tab.size() * sizeof(int), cudaMemcpyHostToDevice), "tab H2D", err));
    resizeTabs.assign(P.nlevels, nullptr);
    for (int l = 1; l < P.nlevels; ++l) resizeTabs[l] = dResizeTab + tabOff[l];

    TRY(makeLevelMaps(0, 0, false, &dResizeMaps, err, rsBoxW.data(), rsBoxH.data()));
    TRY(makeLevelMaps(ORBX_BLUR_TW + 32, ORBX_BLUR_TH + 6, false, &dBlurMaps, err));
    TRY(makeLevelMaps(ORBX_DESC_BOX0_W, ORBX_DESC_BOX0_H, false, &dDescMaps0, err));
    TRY(makeLevelMaps(ORBX_DESC_BOX1_W, ORBX_DESC_BOX1_H, true, &dDescMaps1, err));
    {
        int bw, rh;
        orbx_fast_geometry(P, &bw, &rh);
        TRY(makeLevelMaps(bw, rh, false, &dFastMaps, err));
    }

    if (stereo) {
        const size_t fr = n / 2;
        SP.maxKp = P.maxKp; SP.nlevels = P.nlevels; SP.rows0 = rows;
        SP.imgStep = 2; SP.imgOfsL = 0; SP.imgOfsR = 1;
        TRY(dalloc(this, &SP.uRight, fr * P.maxKp, "uRight", err));
        TRY(dalloc(this, &SP.depth, fr * P.maxKp, "depth", err));
        TRY(dalloc(this, &SP.sad, fr * P.maxKp, "sad", err));
        TRY(dalloc(this, &SP.nMatched, fr, "nMatched", err));
        TRY(dalloc(this, &SP.rsorted, fr * P.maxKp * 3, "rsorted", err));  // record + descriptor: 3 x uint4
        TRY(dalloc(this, &SP.binStart, fr * (rows + 1), "binStart", err));
        SP.binR = orbx_stereo_bin_radius(cfg.scale.data(), cfg.nlevels);
    }
    launchesPerExtract = (P.nlevels - 1) + 1 + 1 + 1 + 2;
    return orbx_check(cudaStreamSynchronize(stream), "init sync", err);
}

int Engine::uploadImages(const uint8_t* left, const uint8_t* right, int nFrames, bool fromDevice, std::string& err)
{
    const int per = stereo ? 2 : 1;
    if (nFrames < 1 || nFrames * per > maxImgs || !left || (stereo && !right)) { err = "bad upload arguments"; return ORBX_ERR_ARG; }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    // image 2f+s <- side s of frame f. Host images arrive with ONE linear copy per side into a staging buffer
    // (a strided 3-D copy runs far below PCIe rate); a kernel then re-pitches the tight rows into level 0.
    const size_t tight = (size_t)rows * cols * nFrames;
    for (int side = 0; side < per; ++side) {
        const uint8_t* src = side ? right : left;
        if (!fromDevice) {
            uint8_t* st = dStage + (size_t)side * rows * cols * (maxImgs / per);
            TRY(orbx_check(cudaMemcpyAsync(st, src, tight, cudaMemcpyHostToDevice, stream), side ? "upload right" : "upload left", err));
            src = st;
        }
        orbx_launch_ingest(P, src, nFrames, per, side, stream);
    }
    TRY(orbx_check(cudaGetLastError(), "ingest launch", err));
    return ORBX_OK;
}

int Engine::uploadSingle(const uint8_t* image, size_t stride, std::string& err)
{
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    return orbx_check(cudaMemcpy2DAsync(P.img0, (size_t)P.lv[0].pitch, image, stride, cols, rows, cudaMemcpyHostToDevice, stream), "upload image", err);
}

int Engine::runExtract(int nImgs, std::string& err)
{
    if (nImgs < 1 || nImgs > maxImgs) { err = "bad image count"; return ORBX_ERR_ARG; }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    OrbxParams p = slice(0, nImgs);
    mark(0);
    launchPyramid(p, stream);
    mark(1);
    orbx_launch_fast(p, dCells, dFastMaps, stream);
    mark(2);
    orbx_launch_octree(p, dCells, stream);
    mark(3);
    orbx_launch_blur(p, dTiles, nTiles, dBlurMaps, stream);
    mark(4);
    orbx_launch_describe(p, dDescMaps0, dDescMaps1, stream);
    mark(5);
    if (!stereo) { mark(6); mark(7); if (profiling && profiledRuns < kMaxProfiledRuns) ++profiledRuns; }
    return orbx_check(cudaGetLastError(), "extract launch", err);
}

int Engine::runStereo(int nFrames, float mbf, float mb, std::string& err)
{
    if (!stereo || nFrames < 1 || nFrames * 2 > maxImgs) { err = "bad stereo arguments"; return ORBX_ERR_ARG; }
    launchStereo(0, nFrames, mbf, mb, stream);
    mark(6);
    mark(7);
    if (profiling && profiledRuns < kMaxProfiledRuns) ++profiledRuns;
    return orbx_check(cudaGetLastError(), "stereo launch", err);
}

OrbxParams Engine::slice(int b, int nImgs) const
{
    OrbxParams p = P;
    const size_t o = (size_t)b;
    p.imgs = nImgs; p.imgBase = b;
    p.img0 += o * P.img0Stride; p.pyr += o * P.pyrStride; p.blur += o * P.blurStride;
    p.cellCnt += o * P.cellsTotal; p.cellKp += o * P.slotTotal;
    p.candXY += o * P.candTotal; p.candRs += o * P.candTotal; p.label += o * P.candTotal;
    p.keep += o * P.keepTotal; p.keepCnt += o * P.nlevels; p.candCnt += o * P.nlevels; p.keepOfs += o * P.nlevels;
    p.outKp = (uint8_t*)P.outKp + o * P.maxKp * 28; p.outDesc += o * P.maxKp * 32; p.outCnt += o;
    p.chainState += o * P.chainInts;
    if (P.octSpill) p.octSpill += o * P.nlevels * P.octSpillStride;
    return p;
}

// Few images: the whole pyramid as one launch (tiles chained by readiness counters); many images: one launch per level,
// each wide enough by itself. Tuning key pyr_chain: 1 forces the one-launch form, 2 the per-level form.
void Engine::launchPyramid(const OrbxParams& p, cudaStream_t s)
{
    const int mode = orbx_tune(ORBX_TUNE_PYR_CHAIN);
    if (mode == 1 || (mode == 0 && p.imgs <= 8)) orbx_launch_pyramid_chain(p, resizeTabs.data(), p.chainState, s);
    else if (orbx_tune(ORBX_TUNE_PYR_TMA) != 2) orbx_launch_pyramid_tma(p, resizeTabs.data(), dResizeMaps, rsBoxW.data(), rsBoxH.data(), s);
    else orbx_launch_pyramid(p, resizeTabs.data(), s);
}

void Engine::launchExtract(const OrbxParams& p, cudaStream_t s)
{
    launchPyramid(p, s);
    orbx_launch_fast(p, dCells, dFastMaps, s);
    orbx_launch_octree(p, dCells, s);
    orbx_launch_blur(p, dTiles, nTiles, dBlurMaps, s);
    orbx_launch_describe(p, dDescMaps0, dDescMaps1, s);
}

void Engine::launchStereo(int f0, int nFrames, float mbf, float mb, cudaStream_t s)
{
    const OrbxParams p = slice(2 * f0, 2 * nFrames);
    OrbxStereoParams sp = SP;
    const size_t o = (size_t)f0;
    sp.frames = nFrames; sp.mbf = mbf; sp.mb = mb;
    sp.uRight += o * SP.maxKp; sp.depth += o * SP.maxKp; sp.sad += o * SP.maxKp; sp.nMatched += o;
    sp.rsorted += o * SP.maxKp * 3; sp.binStart += o * (rows + 1);
    orbx_launch_stereo(p, p, sp, s);
}

// Device-resident run of a whole batch as interleaved chunks on the compute streams (two; ORBX_RUN_STREAMS up to four): the low-occupancy stages of
// one chunk (quadtree, stereo search, top pyramid levels) fill the issue slots the wide stages of the other chunk leave
// idle. `stream` forks to the other streams and joins them again, so callers keep ordering and timing on `stream` alone.
int Engine::runBatch(int nFrames, float mbf, float mb, int nChunks, std::string& err)
{
    const int per = stereo ? 2 : 1;
    if (nFrames < 1 || nFrames * per > maxImgs) { err = "bad frame count"; return ORBX_ERR_ARG; }
    lastRunChunks = 1;
    if (profiling || nChunks < 2 || nFrames < 2 * nChunks) {  // per-stage events need the serial order
        TRY(runExtract(nFrames * per, err));
        return stereo ? runStereo(nFrames, mbf, mb, err) : ORBX_OK;
    }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    const int envStreams = orbx_tune(ORBX_TUNE_RUN_STREAMS);
    const int nStreams = std::min(std::max(envStreams > 0 ? envStreams : 2, 2), std::min(4, nChunks));
    cudaStream_t pool[4] = {stream, stream2, stream3, stream4};
    while ((int)chunkEvents.size() < 4) {
        cudaEvent_t e;
        TRY(orbx_check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate", err));
        chunkEvents.push_back(e);
    }
    TRY(orbx_check(cudaEventRecord(chunkEvents[0], stream), "event", err));
    for (int i = 1; i < nStreams; ++i) TRY(orbx_check(cudaStreamWaitEvent(pool[i], chunkEvents[0], 0), "wait", err));
    const int chunkFrames = (nFrames + nChunks - 1) / nChunks;
    for (int c = 0, f0 = 0; f0 < nFrames; ++c, f0 += chunkFrames) {
        const int n = std::min(chunkFrames, nFrames - f0);
        cudaStream_t cs = pool[c % nStreams];
        launchExtract(slice(per * f0, per * n), cs);
        if (stereo) launchStereo(f0, n, mbf, mb, cs);
    }
    lastRunChunks = (nFrames + chunkFrames - 1) / chunkFrames;
    TRY(orbx_check(cudaGetLastError(), "batch launch", err));
    for (int i = 1; i < nStreams; ++i) {
        TRY(orbx_check(cudaEventRecord(chunkEvents[i], pool[i]), "event", err));
        TRY(orbx_check(cudaStreamWaitEvent(stream, chunkEvents[i], 0), "wait", err));
    }
    return ORBX_OK;
}

// Chunked end-to-end pipeline: chunk c's images go up on `upStream` while chunk c-1 computes on `stream` and chunk
// c-2's results come down on `downStream` (H2D and D2H use separate copy engines; PCIe is full duplex).
// With wait == false the call returns once everything is enqueued: the next call's uploads then overlap this call's
// compute (chunk c of the next call waits only for chunk c of this one: its staging area consumed, its results copied out),
// and waitPipelined() drains. Host buffers must stay untouched until then.
int Engine::processPipelined(const uint8_t* left, const uint8_t* right, int nFrames, float mbf, float mb, const HostOutputs& out,
                             int chunkFrames, std::string& err, bool wait)
{
    const int per = stereo ? 2 : 1;
    if (nFrames < 1 || nFrames * per > maxImgs || !left || (stereo && !right)) { err = "bad process arguments"; return ORBX_ERR_ARG; }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    if (chunkFrames < 1) chunkFrames = nFrames;
    // chunk boundaries: equal chunks, except that the first one may be split so that compute starts after a short copy
    std::vector<int> bounds{0};
    {
        // (a streamed call has no ramp to shorten: the previous call is still computing while its first chunk uploads)
        const int firstDivEnv = orbx_tune(ORBX_TUNE_E2E_FIRST_DIV) > 0 ? orbx_tune(ORBX_TUNE_E2E_FIRST_DIV) : 2;
        const int firstDiv = wait ? firstDivEnv : 1;
        if (firstDiv > 1 && chunkFrames >= 2 * firstDiv && nFrames > chunkFrames) {
            for (int k = 1; k < firstDiv; ++k) bounds.push_back(chunkFrames * k / firstDiv);
        }
        for (int f = chunkFrames; f < nFrames; f += chunkFrames) bounds.push_back(f);
        bounds.push_back(nFrames);
    }
    const int nChunks = (int)bounds.size() - 1;
    // a call still in flight with a different chunk layout: no per-chunk correspondence, drain it first
    const bool chained = !lastBounds.empty() && lastBounds == bounds;
    if (!lastBounds.empty() && !chained) TRY(waitPipelined(err));
    while ((int)ingestEvents.size() < nChunks) {
        cudaEvent_t e1, e2;
        TRY(orbx_check(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming), "cudaEventCreate", err));
        TRY(orbx_check(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming), "cudaEventCreate", err));
        ingestEvents.push_back(e1);
        downEvents.push_back(e2);
    }
    while ((int)chunkEvents.size() < 2 * nChunks) {
        cudaEvent_t e;
        TRY(orbx_check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate", err));
        chunkEvents.push_back(e);
    }
    const size_t tight = (size_t)rows * cols, Pk = (size_t)P.maxKp;
    const size_t sideStride = tight * (maxImgs / per);
    // orbx_set_tuning("e2e_trace", 1): timed events around every chunk's copy-in, compute and copy-out, printed to stderr after the call
    const bool trace = orbx_tune(ORBX_TUNE_E2E_TRACE) != 0;
    std::vector<cudaEvent_t> tev;
    auto tmark = [&](cudaStream_t s) {
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, s);
        tev.push_back(e);
    };
    const int nStreams = [] { const int t = orbx_tune(ORBX_TUNE_E2E_STREAMS), v = t > 0 ? t : 2; return v > 4 ? 4 : v; }();
    cudaStream_t pool[4] = {stream, stream2, stream3, stream4};
    if (!chained) {
        // work still pending on `stream` (an unsynchronised orbx_batch_run / upload) reads the buffers this call overwrites:
        // the copy stream and the other compute streams start behind it
        TRY(orbx_check(cudaEventRecord(chunkEvents[0], stream), "event", err));
        TRY(orbx_check(cudaStreamWaitEvent(upStream, chunkEvents[0], 0), "wait", err));
        for (int i = 1; i < nStreams; ++i) TRY(orbx_check(cudaStreamWaitEvent(pool[i], chunkEvents[0], 0), "wait", err));
    }
    tmark(upStream);  // t0
    for (int c = 0; c < nChunks; ++c) {
        const int f0 = bounds[c], n = bounds[c + 1] - bounds[c];
        cudaEvent_t up = chunkEvents[2 * c], done = chunkEvents[2 * c + 1];
        tmark(upStream);
        if (chained) TRY(orbx_check(cudaStreamWaitEvent(upStream, ingestEvents[c], 0), "wait", err));  // staging area free again
        for (int side = 0; side < per; ++side) {
            const uint8_t* src = (side ? right : left) + f0 * tight;
            TRY(orbx_check(cudaMemcpyAsync(dStage + side * sideStride + f0 * tight, src, n * tight, cudaMemcpyHostToDevice, upStream), "H2D chunk", err));
        }
        TRY(orbx_check(cudaEventRecord(up, upStream), "event", err));
        tmark(upStream);
        // chunks alternate between two compute streams: the low-occupancy kernels of one chunk (quadtree, small
        // pyramid levels) overlap with the wide kernels of the next
        cudaStream_t cs = pool[c % nStreams];
        TRY(orbx_check(cudaStreamWaitEvent(cs, up, 0), "wait", err));
        tmark(cs);
        if (chained) TRY(orbx_check(cudaStreamWaitEvent(cs, downEvents[c], 0), "wait", err));  // previous results of these frames are out
        const OrbxParams p = slice(per * f0, per * n);
        for (int side = 0; side < per; ++side)
            orbx_launch_ingest(p, dStage + side * sideStride + f0 * tight, n, per, side, cs);
        TRY(orbx_check(cudaEventRecord(ingestEvents[c], cs), "event", err));
        launchExtract(p, cs);
        if (stereo) launchStereo(f0, n, mbf, mb, cs);
        TRY(orbx_check(cudaGetLastError(), "chunk launch", err));
        TRY(orbx_check(cudaEventRecord(done, cs), "event", err));
        tmark(cs);
        TRY(orbx_check(cudaStreamWaitEvent(downStream, done, 0), "wait", err));
        tmark(downStream);
        auto rowsD2H = [&](void* dst, const void* src, size_t rowBytes, size_t devStride) -> int {
            if (!dst) return ORBX_OK;
            return orbx_check(cudaMemcpy2DAsync(dst, rowBytes, src, devStride, rowBytes, n, cudaMemcpyDeviceToHost, downStream), "D2H chunk", err);
        };
        const uint8_t* kp = (const uint8_t*)P.outKp + (size_t)per * f0 * Pk * 28;
        const uint8_t* ds = P.outDesc + (size_t)per * f0 * Pk * 32;
        const int* cn = P.outCnt + per * f0;
        TRY(rowsD2H(out.nLeft ? out.nLeft + f0 : nullptr, cn, 4, per * 4));
        TRY(rowsD2H(out.kpsLeft ? (uint8_t*)out.kpsLeft + f0 * Pk * 28 : nullptr, kp, Pk * 28, per * Pk * 28));
        TRY(rowsD2H(out.descLeft ? out.descLeft + f0 * Pk * 32 : nullptr, ds, Pk * 32, per * Pk * 32));
        if (stereo) {
            TRY(rowsD2H(out.nRight ? out.nRight + f0 : nullptr, cn + 1, 4, per * 4));
            TRY(rowsD2H(out.kpsRight ? (uint8_t*)out.kpsRight + f0 * Pk * 28 : nullptr, kp + Pk * 28, Pk * 28, per * Pk * 28));
            TRY(rowsD2H(out.descRight ? out.descRight + f0 * Pk * 32 : nullptr, ds + Pk * 32, Pk * 32, per * Pk * 32));
            TRY(rowsD2H(out.uRight ? out.uRight + f0 * Pk : nullptr, SP.uRight + f0 * Pk, Pk * 4, Pk * 4));
            TRY(rowsD2H(out.depth ? out.depth + f0 * Pk : nullptr, SP.depth + f0 * Pk, Pk * 4, Pk * 4));
        }
        TRY(orbx_check(cudaEventRecord(downEvents[c], downStream), "event", err));
        tmark(downStream);
    }
    lastBounds = bounds;
    if (!wait && !trace) return ORBX_OK;
    lastBounds.clear();
    TRY(orbx_check(cudaStreamSynchronize(downStream), "down sync", err));
    for (int i = nStreams - 1; i >= 0; --i) TRY(orbx_check(cudaStreamSynchronize(pool[i]), "stream sync", err));
    if (trace) {  // 6 events per chunk after t0: up begin/end, compute begin/end, down begin/end
        auto at = [&](size_t i) { float t = 0; cudaEventElapsedTime(&t, tev[0], tev[i]); return t; };
        for (int c = 0; c < nChunks; ++c)
            fprintf(stderr, "[orbx e2e] chunk %d frames %d..%d  up %.3f-%.3f  compute %.3f-%.3f  down %.3f-%.3f ms\n", c, bounds[c], bounds[c + 1],
                    at(1 + 6 * c), at(2 + 6 * c), at(3 + 6 * c), at(4 + 6 * c), at(5 + 6 * c), at(6 + 6 * c));
        for (auto e : tev) cudaEventDestroy(e);
    }
    return ORBX_OK;
}

// ---- single-frame path -------------------------------------------------------------------------------------
void Engine::enqueueFrame(float mbf, float mb, cudaStream_t s)
{
    const int per = stereo ? 2 : 1;
    const size_t tight = (size_t)rows * cols, sideStride = tight * (maxImgs / per), Pk = (size_t)P.maxKp;
    for (int side = 0; side < per; ++side) orbx_launch_ingest(P, dStage + side * sideStride, maxImgs / per, per, side, s);
    // the blur needs only the pyramid: it runs beside FAST and the quadtree (a second branch of the graph) and joins before
    // the descriptors
    const OrbxParams p = slice(0, maxImgs);
    launchPyramid(p, s);
    cudaEventRecord(frameFork, s);
    cudaStreamWaitEvent(stream2, frameFork, 0);
    orbx_launch_blur(p, dTiles, nTiles, dBlurMaps, stream2);
    cudaEventRecord(frameJoin, stream2);
    orbx_launch_fast(p, dCells, dFastMaps, s);
    orbx_launch_octree(p, dCells, s);
    cudaStreamWaitEvent(s, frameJoin, 0);
    orbx_launch_describe(p, dDescMaps0, dDescMaps1, s);
    if (stereo) launchStereo(0, maxImgs / 2, mbf, mb, s);
    cudaMemcpyAsync(hOut, P.outCnt, (size_t)maxImgs * sizeof(int), cudaMemcpyDeviceToHost, s);
    cudaMemcpyAsync(hOut + 32, P.chainState + 1, sizeof(int), cudaMemcpyDeviceToHost, s);  // watchdog flag of the one-launch pyramid
    cudaMemcpyAsync(hOut + frameOffKp, P.outKp, (size_t)maxImgs * Pk * 28, cudaMemcpyDeviceToHost, s);
    cudaMemcpyAsync(hOut + frameOffDesc, P.outDesc, (size_t)maxImgs * Pk * 32, cudaMemcpyDeviceToHost, s);
    if (stereo) {
        cudaMemcpyAsync(hOut + frameOffU, SP.uRight, (size_t)(maxImgs / 2) * Pk * 4, cudaMemcpyDeviceToHost, s);
        cudaMemcpyAsync(hOut + frameOffD, SP.depth, (size_t)(maxImgs / 2) * Pk * 4, cudaMemcpyDeviceToHost, s);
    }
}

int Engine::processFrame(const uint8_t* const* images, const size_t* strides, float mbf, float mb, std::string& err)
{
    if (maxImgs > 2) { err = "the single-frame path serves engines of one frame"; return ORBX_ERR_STATE; }
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    const int per = stereo ? 2 : 1;
    const size_t tight = (size_t)rows * cols, Pk = (size_t)P.maxKp;
    if (!hIn) {
        frameOffKp = 64;
        frameOffDesc = frameOffKp + (size_t)maxImgs * Pk * 28;
        frameOffU = frameOffDesc + (size_t)maxImgs * Pk * 32;
        frameOffD = frameOffU + Pk * 4;
        hOutBytes = frameOffD + Pk * 4;
        TRY(orbx_check(cudaHostAlloc(&hIn, tight * maxImgs, cudaHostAllocDefault), "pinned input staging", err));
        TRY(orbx_check(cudaHostAlloc(&hOut, hOutBytes, cudaHostAllocDefault), "pinned output staging", err));
        TRY(orbx_check(cudaEventCreateWithFlags(&frameFork, cudaEventDisableTiming), "cudaEventCreate", err));
        TRY(orbx_check(cudaEventCreateWithFlags(&frameJoin, cudaEventDisableTiming), "cudaEventCreate", err));
    }
    // image i = side (i % per) of the frame: one flat copy each into the staging area the ingest kernel reads. A caller's
    // buffer that is already page-locked and tight goes up directly; anything else passes through the pinned buffer.
    for (int i = 0; i < maxImgs; ++i) {
        const uint8_t* src = images[i];
        bool direct = strides[i] == (size_t)cols;
        if (direct) {
            cudaPointerAttributes at{};
            direct = cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost;
            cudaGetLastError();
        }
        if (!direct) {
            uint8_t* d = hIn + (size_t)i * tight;
            if (strides[i] == (size_t)cols) std::memcpy(d, src, tight);
            else for (int y = 0; y < rows; ++y) std::memcpy(d + (size_t)y * cols, src + (size_t)y * strides[i], (size_t)cols);
            src = d;
        }
        const int side = i % per, f = i / per;
        TRY(orbx_check(cudaMemcpyAsync(dStage + side * tight * (maxImgs / per) + f * tight, src, tight, cudaMemcpyHostToDevice, stream), "H2D frame", err));
    }
    if (frameGraph && (std::memcmp(&graphMbf, &mbf, 4) || std::memcmp(&graphMb, &mb, 4))) {  // stereo parameters are baked into the graph
        cudaGraphExecDestroy(frameGraph);
        frameGraph = nullptr;
    }
    if (!frameGraph && !graphBroken && frameCalls > 0) {
        // (the first call ran the sequence directly: per-function attributes are set, lazily loaded modules are resident)
        cudaGraph_t g = nullptr;
        if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
            enqueueFrame(mbf, mb, stream);
            const cudaError_t ec = cudaStreamEndCapture(stream, &g);
            if (ec != cudaSuccess || !g || cudaGraphInstantiate(&frameGraph, g, 0) != cudaSuccess) { frameGraph = nullptr; graphBroken = true; }
            if (g) cudaGraphDestroy(g);
            graphMbf = mbf; graphMb = mb;
        } else {
            graphBroken = true;
        }
        cudaGetLastError();
    }
    ++frameCalls;
    if (frameGraph) TRY(orbx_check(cudaGraphLaunch(frameGraph, stream), "graph launch", err));
    else enqueueFrame(mbf, mb, stream);
    TRY(orbx_check(cudaGetLastError(), "frame launch", err));
    TRY(sync(err));
    if (*reinterpret_cast<const int*>(hOut + 32) != 0) {  // a tile of the one-launch pyramid gave up waiting for its source rows
        err = "pyramid chain watchdog: a tile waited too long for its source level (results of this frame are invalid)";
        return ORBX_ERR_CUDA;
    }
    residentValid = true;
    return ORBX_OK;
}

int Engine::waitPipelined(std::string& err)
{
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    lastBounds.clear();
    TRY(orbx_check(cudaStreamSynchronize(upStream), "up sync", err));
    TRY(orbx_check(cudaStreamSynchronize(downStream), "down sync", err));
    cudaStream_t pool[4] = {stream, stream2, stream3, stream4};
    for (int i = 3; i >= 0; --i) TRY(orbx_check(cudaStreamSynchronize(pool[i]), "stream sync", err));
    return ORBX_OK;
}

int Engine::sync(std::string& err)
{
    return orbx_check(cudaStreamSynchronize(stream), "stream sync", err);
}

// ---- per-stage timing -------------------------------------------------------------------------------
const char* Engine::stageName(int i)
{
    static const char* names[kStages] = {"pyramid", "fast", "octree", "blur", "describe", "stereo", "-"};
    return i >= 0 && i < kStages ? names[i] : "?";
}

int Engine::setProfiling(bool on, std::string& err)
{
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    if (on && events.empty()) {
        events.resize((size_t)kMaxProfiledRuns * (kStages + 1));
        for (auto& e : events) TRY(orbx_check(cudaEventCreate(&e), "cudaEventCreate", err));
    }
    profiling = on;
    profiledRuns = 0;
    return ORBX_OK;
}

void Engine::mark(int stage)
{
    if (!profiling || profiledRuns >= kMaxProfiledRuns) return;
    cudaEventRecord(events[(size_t)profiledRuns * (kStages + 1) + stage], stream);
}

int Engine::stageTimes(double* ms, int* runs, std::string& err)
{
    TRY(sync(err));
    for (int s = 0; s < kStages; ++s) ms[s] = 0;
    for (int r = 0; r < profiledRuns; ++r)
        for (int s = 0; s < kStages; ++s) {
            float t = 0;
            TRY(orbx_check(cudaEventElapsedTime(&t, events[(size_t)r * (kStages + 1) + s], events[(size_t)r * (kStages + 1) + s + 1]), "event time", err));
            ms[s] += t;
        }
    *runs = profiledRuns;
    profiledRuns = 0;
    return ORBX_OK;
}
