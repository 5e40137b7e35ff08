// abi.cu — the C-ABI of liborbx.so (declared in include/orbx.h). No CPU fallback anywhere: every
// compute entry point needs a CUDA device and reports ORBX_ERR_CUDA otherwise.
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>

#include "../../include/orbx.h"
#include "engine.h"
#include "handles.h"

static thread_local std::string g_err;
#define TRY(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)
static int fail(int code, const char* msg) { g_err = msg; return code; }
std::string& orbx_err_slot() { return g_err; }  // for frame.cu



extern "C" const char* orbx_last_error(void) { return g_err.c_str(); }

extern "C" int orbx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

static int check_cfg(int nfeatures, float sf, int nlevels, int ini, int mn)
{
    if (nfeatures < 1 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !(sf > 1.0f) || !(sf <= 2.0f) || ini < 1 || mn < 1 || ini > 254 || mn > 254)
        return fail(ORBX_ERR_ARG, "bad extractor parameters (need nfeatures>=1, 1<=nlevels<=12, 1<scaleFactor<=2, 1<=iniTh<=254, 1<=minTh<=254)");
    return ORBX_OK;
}

extern "C" int orbx_create(int nfeatures, float sf, int nlevels, int ini, int mn, int device, orbx_extractor** out)
{
    if (!out) return fail(ORBX_ERR_ARG, "out is NULL");
    *out = nullptr;
    TRY(check_cfg(nfeatures, sf, nlevels, ini, mn));
    orbx_extractor* e = new orbx_extractor();
    e->cfg.init(nfeatures, sf, nlevels, ini, mn);
    e->device = device;
    *out = e;
    return ORBX_OK;
}

extern "C" void orbx_destroy(orbx_extractor* e)
{
    if (!e) return;
    if (e->dU) { cudaSetDevice(e->device); cudaFree(e->dU); cudaFree(e->dD); cudaFree(e->dSad); cudaFree(e->dNm); cudaFree(e->dSorted); cudaFree(e->dBin); }
    if (e->hOut) { cudaSetDevice(e->device); cudaFreeHost(e->hOut); }
    if (e->pairPeer) {  // the peer's view dies with this handle's frame engine
        if (e->pairPeer->view == e->pair.get()) e->pairPeer->view = nullptr;
        if (e->pairPeer->pairOwner == e) e->pairPeer->pairOwner = nullptr;
    }
    if (e->pairOwner && e->pairOwner->pairPeer == e) e->pairOwner->pairPeer = nullptr;  // and nobody keeps a pointer to this handle
    delete e;
}

extern "C" int orbx_get_levels(const orbx_extractor* e) { return e ? e->cfg.nlevels : ORBX_ERR_ARG; }

extern "C" int orbx_get_tables(const orbx_extractor* e, float* scale, float* inv, float* sig2, float* invSig2, int* quotas)
{
    if (!e) return fail(ORBX_ERR_ARG, "handle is NULL");
    for (int i = 0; i < e->cfg.nlevels; ++i) {
        if (scale) scale[i] = e->cfg.scale[i];
        if (inv) inv[i] = e->cfg.invScale[i];
        if (sig2) sig2[i] = e->cfg.sigma2[i];
        if (invSig2) invSig2[i] = e->cfg.invSigma2[i];
        if (quotas) quotas[i] = e->cfg.quota[i];
    }
    return ORBX_OK;
}

extern "C" int orbx_max_keypoints(const orbx_extractor* e, int rows, int cols)
{
    if (!e) return fail(ORBX_ERR_ARG, "handle is NULL");
    OrbxParams p;
    TRY(Engine::plan(e->cfg, rows, cols, p, g_err));
    return p.maxKp;
}

extern "C" int orbx_extract(orbx_extractor* e, const uint8_t* image, int rows, int cols, size_t stride, orbx_keypoint* kps,
                            uint8_t* desc, int capacity, int* n)
{
    if (!e) return fail(ORBX_ERR_ARG, "handle is NULL");
    if (rows == 0 || cols == 0 || !image) {  // src/ORBextractor.cc:1044-1045: silent return; an out-count must not stay undefined
        if (n) *n = 0;
        return ORBX_OK;
    }
    if (rows < 0 || cols < 0 || stride < (size_t)cols || !kps || !desc || !n) return fail(ORBX_ERR_ARG, "bad arguments");
    if (!e->eng || e->eng->rows != rows || e->eng->cols != cols) {
        e->hasRun = false;
        if (e->view == e->eng.get()) e->view = nullptr;
        std::unique_ptr<Engine> ne(new Engine());
        TRY(ne->init(e->cfg, rows, cols, 1, false, e->device, g_err));
        e->eng = std::move(ne);
    }
    Engine& E = *e->eng;
    // pinned staging up, the launch sequence and the copies of the count and of every record into pinned memory as one CUDA
    // graph, one synchronisation (Engine::processFrame)
    const uint8_t* imgs[1] = {image};
    const size_t strides[1] = {stride};
    TRY(E.processFrame(imgs, strides, 0.f, 0.f, g_err));
    e->hasRun = true;
    e->view = &E; e->viewImg = 0;
    const int cnt = *reinterpret_cast<const int*>(E.hOut);
    if (cnt > capacity) return fail(ORBX_ERR_CAPACITY, "keypoint capacity too small (see orbx_max_keypoints)");
    if (cnt > 0) {
        std::memcpy(kps, E.hOut + E.frameOffKp, (size_t)cnt * 28);
        std::memcpy(desc, E.hOut + E.frameOffDesc, (size_t)cnt * 32);
    }
    *n = cnt;
    return ORBX_OK;
}

extern "C" int orbx_pyramid_level_size(const orbx_extractor* e, int level, int* rows, int* cols)
{
    if (!e || !rows || !cols) return fail(ORBX_ERR_ARG, "bad arguments");
    if (!e->view) return fail(ORBX_ERR_STATE, "no image extracted yet");
    if (level < 0 || level >= e->cfg.nlevels) return fail(ORBX_ERR_ARG, "bad level");
    *rows = e->view->P.lv[level].h;
    *cols = e->view->P.lv[level].w;
    return ORBX_OK;
}

static int copy_level(Engine& E, int img, int level, bool blurred, uint8_t* dst, size_t dstStride)
{
    const OrbxLevel& L = E.P.lv[level];
    const uint8_t* src;
    size_t pitch;
    if (blurred) { src = E.P.blur + (size_t)img * E.P.blurStride + L.boff; pitch = L.bpitch; }
    else if (level == 0) { src = E.P.img0 + (size_t)img * E.P.img0Stride; pitch = E.P.lv[0].pitch; }
    else { src = E.P.pyr + (size_t)img * E.P.pyrStride + L.off; pitch = L.pitch; }
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    TRY(orbx_check(cudaMemcpy2DAsync(dst, dstStride, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost, E.stream), "D2H level", g_err));
    return E.sync(g_err);
}

extern "C" int orbx_pyramid_level(const orbx_extractor* e, int level, uint8_t* dst, size_t dstStride)
{
    if (!e || !dst) return fail(ORBX_ERR_ARG, "bad arguments");
    if (!e->view) return fail(ORBX_ERR_STATE, "no image extracted yet");
    if (level < 0 || level >= e->cfg.nlevels || dstStride < (size_t)e->view->P.lv[level].w) return fail(ORBX_ERR_ARG, "bad level or stride");
    return copy_level(*e->view, e->viewImg, level, false, dst, dstStride);
}

// Device scratch of the two-handle stereo match (owned by the LEFT handle; grown when the image size does): per-keypoint
// uRight / depth / SAD, the match counter, the row-sorted right records and the row-bin table.
static int ensure_stereo_scratch(orbx_extractor* l, const Engine& EL)
{
    if (l->scratchKp >= EL.P.maxKp && l->scratchRows >= EL.rows) return ORBX_OK;
    if (l->dU) { cudaFree(l->dU); cudaFree(l->dD); cudaFree(l->dSad); cudaFree(l->dNm); cudaFree(l->dSorted); cudaFree(l->dBin); }
    l->dU = l->dD = nullptr; l->dSad = l->dNm = nullptr; l->dSorted = nullptr; l->dBin = nullptr;
    l->scratchKp = l->scratchRows = 0;
    const size_t k = EL.P.maxKp;
    TRY(orbx_check(cudaMalloc(&l->dU, k * 4), "stereo scratch", g_err));
    TRY(orbx_check(cudaMalloc(&l->dD, k * 4), "stereo scratch", g_err));
    TRY(orbx_check(cudaMalloc(&l->dSad, k * 4), "stereo scratch", g_err));
    TRY(orbx_check(cudaMalloc(&l->dNm, 4), "stereo scratch", g_err));
    TRY(orbx_check(cudaMalloc(&l->dSorted, 3 * k * sizeof(uint4)), "stereo scratch", g_err));
    TRY(orbx_check(cudaMalloc(&l->dBin, ((size_t)EL.rows + 1) * 4), "stereo scratch", g_err));
    l->scratchKp = EL.P.maxKp;
    l->scratchRows = EL.rows;
    return ORBX_OK;
}

extern "C" int orbx_stereo_match(orbx_extractor* l, orbx_extractor* r, const orbx_keypoint* kpsL, const uint8_t* descL, int nL,
                                 const orbx_keypoint* kpsR, const uint8_t* descR, int nR, float mbf, float mb, float* uRight,
                                 float* depth)
{
    if (!l || !r || nL < 0 || nR < 0 || (nL && (!kpsL || !descL || !uRight || !depth)) || (nR && (!kpsR || !descR)))
        return fail(ORBX_ERR_ARG, "bad arguments");
    if (!l->eng || !r->eng || !l->hasRun || !r->hasRun) return fail(ORBX_ERR_STATE, "both extractors must have run");
    Engine &EL = *l->eng, &ER = *r->eng;
    if (EL.rows != ER.rows || EL.cols != ER.cols || EL.P.nlevels != ER.P.nlevels || l->device != r->device)
        return fail(ORBX_ERR_STATE, "left/right extractors differ in image size, levels or device");
    if (nL > EL.P.maxKp || nR > ER.P.maxKp) return fail(ORBX_ERR_ARG, "more keypoints than the extractor can produce");
    // the octave indexes the level tables and the pyramid on the device (src/Frame.cc:541, :617-620)
    for (int i = 0; i < nL; ++i)
        if (kpsL[i].octave < 0 || kpsL[i].octave >= l->cfg.nlevels) return fail(ORBX_ERR_ARG, "left keypoint octave outside 0..nlevels-1");
    for (int i = 0; i < nR; ++i)
        if (kpsR[i].octave < 0 || kpsR[i].octave >= l->cfg.nlevels) return fail(ORBX_ERR_ARG, "right keypoint octave outside 0..nlevels-1");
    for (int i = 0; i < nL; ++i) uRight[i] = depth[i] = -1.f;  // src/Frame.cc:524-525
    if (nL == 0) return ORBX_OK;
    TRY(orbx_check(cudaSetDevice(l->device), "cudaSetDevice", g_err));
    TRY(ensure_stereo_scratch(l, EL));
    cudaStream_t s = EL.stream;
    EL.residentValid = ER.residentValid = false;  // the extractors' own results on the device are overwritten below
    // the caller's keypoints/descriptors are authoritative (Frame::mvKeys etc.): put them on the device
    TRY(orbx_check(cudaMemcpyAsync(EL.P.outKp, kpsL, (size_t)nL * 28, cudaMemcpyHostToDevice, s), "H2D kpsL", g_err));
    TRY(orbx_check(cudaMemcpyAsync(EL.P.outDesc, descL, (size_t)nL * 32, cudaMemcpyHostToDevice, s), "H2D descL", g_err));
    TRY(orbx_check(cudaMemcpyAsync(EL.P.outCnt, &nL, 4, cudaMemcpyHostToDevice, s), "H2D nL", g_err));
    if (nR) {
        TRY(orbx_check(cudaMemcpyAsync(ER.P.outKp, kpsR, (size_t)nR * 28, cudaMemcpyHostToDevice, s), "H2D kpsR", g_err));
        TRY(orbx_check(cudaMemcpyAsync(ER.P.outDesc, descR, (size_t)nR * 32, cudaMemcpyHostToDevice, s), "H2D descR", g_err));
    }
    TRY(orbx_check(cudaMemcpyAsync(ER.P.outCnt, &nR, 4, cudaMemcpyHostToDevice, s), "H2D nR", g_err));
    OrbxParams pl = EL.P, pr = ER.P;
    pl.imgs = pr.imgs = 1;
    OrbxStereoParams sp{};
    sp.frames = 1; sp.nlevels = pl.nlevels; sp.rows0 = EL.rows; sp.maxKp = pl.maxKp; sp.mbf = mbf; sp.mb = mb;
    sp.imgStep = 1; sp.imgOfsL = 0; sp.imgOfsR = 0;
    sp.uRight = l->dU; sp.depth = l->dD; sp.sad = l->dSad; sp.nMatched = l->dNm;
    sp.rsorted = l->dSorted; sp.binStart = l->dBin; sp.binR = orbx_stereo_bin_radius(l->cfg.scale.data(), l->cfg.nlevels);
    orbx_launch_stereo(pl, pr, sp, s);
    TRY(orbx_check(cudaGetLastError(), "stereo launch", g_err));
    TRY(orbx_check(cudaMemcpyAsync(uRight, l->dU, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H uRight", g_err));
    TRY(orbx_check(cudaMemcpyAsync(depth, l->dD, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H depth", g_err));
    return EL.sync(g_err);
}

// ComputeStereoMatches on what the two extractors left on the device: nothing goes up.
extern "C" int orbx_stereo_match_resident(orbx_extractor* l, orbx_extractor* r, int nL, int nR, float mbf, float mb, float* uRight,
                                          float* depth)
{
    if (!l || !r || nL < 0 || nR < 0 || (nL && (!uRight || !depth))) return fail(ORBX_ERR_ARG, "bad arguments");
    if (!l->eng || !r->eng || !l->hasRun || !r->hasRun) return fail(ORBX_ERR_STATE, "both extractors must have run");
    Engine &EL = *l->eng, &ER = *r->eng;
    if (EL.rows != ER.rows || EL.cols != ER.cols || EL.P.nlevels != ER.P.nlevels || l->device != r->device)
        return fail(ORBX_ERR_STATE, "left/right extractors differ in image size, levels or device");
    if (!EL.residentValid || !ER.residentValid || nL != *reinterpret_cast<const int*>(EL.hOut) || nR != *reinterpret_cast<const int*>(ER.hOut))
        return fail(ORBX_ERR_STATE, "keypoint counts differ from what the extractors produced last: use orbx_stereo_match");
    for (int i = 0; i < nL; ++i) uRight[i] = depth[i] = -1.f;  // src/Frame.cc:524-525
    if (nL == 0) return ORBX_OK;
    TRY(orbx_check(cudaSetDevice(l->device), "cudaSetDevice", g_err));
    TRY(ensure_stereo_scratch(l, EL));
    if (!l->hOut || l->hOutCap < (size_t)EL.P.maxKp * 8) {
        if (l->hOut) cudaFreeHost(l->hOut);
        l->hOut = nullptr; l->hOutCap = 0;
        TRY(orbx_check(cudaHostAlloc(&l->hOut, (size_t)EL.P.maxKp * 8, cudaHostAllocDefault), "pinned stereo staging", g_err));
        l->hOutCap = (size_t)EL.P.maxKp * 8;
    }
    cudaStream_t s = EL.stream;  // both extracts have synchronised their streams before returning
    OrbxParams pl = EL.P, pr = ER.P;
    pl.imgs = pr.imgs = 1;
    OrbxStereoParams sp{};
    sp.frames = 1; sp.nlevels = pl.nlevels; sp.rows0 = EL.rows; sp.maxKp = pl.maxKp; sp.mbf = mbf; sp.mb = mb;
    sp.imgStep = 1; sp.imgOfsL = 0; sp.imgOfsR = 0;
    sp.uRight = l->dU; sp.depth = l->dD; sp.sad = l->dSad; sp.nMatched = l->dNm;
    sp.rsorted = l->dSorted; sp.binStart = l->dBin; sp.binR = orbx_stereo_bin_radius(l->cfg.scale.data(), l->cfg.nlevels);
    orbx_launch_stereo(pl, pr, sp, s);
    TRY(orbx_check(cudaGetLastError(), "stereo launch", g_err));
    TRY(orbx_check(cudaMemcpyAsync(l->hOut, l->dU, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H uRight", g_err));
    TRY(orbx_check(cudaMemcpyAsync(l->hOut + (size_t)EL.P.maxKp * 4, l->dD, (size_t)nL * 4, cudaMemcpyDeviceToHost, s), "D2H depth", g_err));
    TRY(EL.sync(g_err));
    std::memcpy(uRight, l->hOut, (size_t)nL * 4);
    std::memcpy(depth, l->hOut + (size_t)EL.P.maxKp * 4, (size_t)nL * 4);
    return ORBX_OK;
}

// One stereo frame in one call (Frame::Frame(stereo), src/Frame.cc:113-125: ExtractORB left and right + ComputeStereoMatches).
extern "C" int orbx_stereo_frame(orbx_extractor* l, orbx_extractor* r, const uint8_t* imgL, const uint8_t* imgR, int rows, int cols,
                                 size_t strideL, size_t strideR, float mbf, float mb, orbx_keypoint* kpsL, uint8_t* descL, int* nL,
                                 orbx_keypoint* kpsR, uint8_t* descR, int* nR, int capacity, float* uRight, float* depth)
{
    if (!l || !r || l == r) return fail(ORBX_ERR_ARG, "two distinct handles are needed");
    if (rows == 0 || cols == 0 || !imgL || !imgR) {  // src/ORBextractor.cc:1044-1045 on both sides: nothing extracted, nothing matched
        if (nL) *nL = 0;
        if (nR) *nR = 0;
        return ORBX_OK;
    }
    if (rows < 0 || cols < 0 || strideL < (size_t)cols || strideR < (size_t)cols || !kpsL || !descL || !nL || !kpsR || !descR || !nR ||
        !uRight || !depth || capacity < 0)
        return fail(ORBX_ERR_ARG, "bad arguments");
    if (l->device != r->device || l->cfg.nfeatures != r->cfg.nfeatures || l->cfg.nlevels != r->cfg.nlevels || l->cfg.scaleFactor != r->cfg.scaleFactor ||
        l->cfg.iniTh != r->cfg.iniTh || l->cfg.minTh != r->cfg.minTh)
        return fail(ORBX_ERR_STATE, "orbx_stereo_frame needs two extractors with the same parameters on the same device");
    if (!l->pair || l->pair->rows != rows || l->pair->cols != cols) {
        std::unique_ptr<Engine> ne(new Engine());
        TRY(ne->init(l->cfg, rows, cols, 2, true, l->device, g_err));
        if (l->pairPeer && l->pairPeer->view == l->pair.get()) l->pairPeer->view = nullptr;
        if (l->view == l->pair.get()) l->view = nullptr;
        l->pair = std::move(ne);
    }
    Engine& E = *l->pair;
    const uint8_t* imgs[2] = {imgL, imgR};
    const size_t strides[2] = {strideL, strideR};
    TRY(E.processFrame(imgs, strides, mbf, mb, g_err));
    // the handles' own engines no longer hold this frame: pyramid levels come from the pair engine, and orbx_stereo_match
    // (which reads the handles' own pyramids) needs a new orbx_extract first
    l->hasRun = r->hasRun = false;
    if (l->pairPeer && l->pairPeer != r) {  // an earlier right-hand partner no longer reads this engine
        if (l->pairPeer->view == &E) l->pairPeer->view = nullptr;
        if (l->pairPeer->pairOwner == l) l->pairPeer->pairOwner = nullptr;
    }
    if (r->pairOwner && r->pairOwner != l && r->pairOwner->pairPeer == r) r->pairOwner->pairPeer = nullptr;
    l->view = r->view = &E; l->viewImg = 0; r->viewImg = 1;
    l->pairPeer = r;
    r->pairOwner = l;
    const size_t Pk = (size_t)E.P.maxKp;
    const int cl = reinterpret_cast<const int*>(E.hOut)[0], cr = reinterpret_cast<const int*>(E.hOut)[1];
    if (cl > capacity || cr > capacity) return fail(ORBX_ERR_CAPACITY, "keypoint capacity too small (see orbx_max_keypoints)");
    std::memcpy(kpsL, E.hOut + E.frameOffKp, (size_t)cl * 28);
    std::memcpy(kpsR, E.hOut + E.frameOffKp + Pk * 28, (size_t)cr * 28);
    std::memcpy(descL, E.hOut + E.frameOffDesc, (size_t)cl * 32);
    std::memcpy(descR, E.hOut + E.frameOffDesc + Pk * 32, (size_t)cr * 32);
    std::memcpy(uRight, E.hOut + E.frameOffU, (size_t)cl * 4);
    std::memcpy(depth, E.hOut + E.frameOffD, (size_t)cl * 4);
    *nL = cl;
    *nR = cr;
    return ORBX_OK;
}

// ---- frame batches ---------------------------------------------------------------------------------

// an asynchronous orbx_batch_process_async may still be in flight: every other entry point drains it first
#define ORBX_DRAIN(b) do { if (!(b)->eng.lastBounds.empty()) TRY((b)->eng.waitPipelined(g_err)); } while (0)

// chunks of >= 16 frames keep every kernel's grid well above one wave while the copies of the neighbouring chunks overlap
static int e2e_chunk_frames(int nFrames)
{
    const int forcedChunks = orbx_tune(ORBX_TUNE_E2E_CHUNKS);
    return forcedChunks > 0 ? (nFrames + forcedChunks - 1) / forcedChunks : nFrames >= 64 ? (nFrames + 7) / 8 : nFrames >= 32 ? 16 : nFrames;
}

extern "C" int orbx_batch_create(int nfeatures, float sf, int nlevels, int ini, int mn, int rows, int cols, int maxFrames,
                                 int stereo, int device, orbx_batch** out)
{
    if (!out) return fail(ORBX_ERR_ARG, "out is NULL");
    *out = nullptr;
    TRY(check_cfg(nfeatures, sf, nlevels, ini, mn));
    if (maxFrames < 1) return fail(ORBX_ERR_ARG, "max_frames < 1");
    std::unique_ptr<orbx_batch> b(new orbx_batch());
    b->cfg.init(nfeatures, sf, nlevels, ini, mn);
    b->maxFrames = maxFrames;
    b->stereo = stereo != 0;
    TRY(b->eng.init(b->cfg, rows, cols, maxFrames * (b->stereo ? 2 : 1), b->stereo, device, g_err));
    *out = b.release();
    return ORBX_OK;
}

extern "C" void orbx_batch_destroy(orbx_batch* b) { delete b; }
extern "C" int orbx_batch_max_keypoints(const orbx_batch* b) { return b ? b->eng.P.maxKp : ORBX_ERR_ARG; }
extern "C" void* orbx_batch_stream(orbx_batch* b) { return b ? (void*)b->eng.stream : nullptr; }
extern "C" int orbx_batch_launches_per_run(const orbx_batch* b)
{
    return b ? (b->eng.launchesPerExtract + (b->stereo ? 3 : 0)) * b->eng.lastRunChunks : ORBX_ERR_ARG;
}
extern "C" long long orbx_batch_device_bytes(const orbx_batch* b) { return b ? (long long)b->eng.bytesAllocated : 0; }

extern "C" int orbx_batch_upload(orbx_batch* b, const uint8_t* left, const uint8_t* right, int nFrames)
{
    if (!b) return fail(ORBX_ERR_ARG, "handle is NULL");
    ORBX_DRAIN(b);
    return b->eng.uploadImages(left, right, nFrames, false, g_err);
}

extern "C" int orbx_batch_upload_device(orbx_batch* b, const void* left, const void* right, int nFrames)
{
    if (!b) return fail(ORBX_ERR_ARG, "handle is NULL");
    ORBX_DRAIN(b);
    return b->eng.uploadImages((const uint8_t*)left, (const uint8_t*)right, nFrames, true, g_err);
}

// chunks of a device-resident run (see Engine::runBatch); ORBX_RUN_CHUNKS overrides for experiments
static int run_chunks(int nFrames)
{
    const int forced = orbx_tune(ORBX_TUNE_RUN_CHUNKS);
    if (forced > 0) return forced;
    return nFrames >= 32 ? 2 : 1;
}

extern "C" int orbx_batch_run(orbx_batch* b, int nFrames, float mbf, float mb, int sync)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames) return fail(ORBX_ERR_ARG, "bad frame count");
    ORBX_DRAIN(b);
    TRY(b->eng.runBatch(nFrames, mbf, mb, run_chunks(nFrames), g_err));
    return sync ? b->eng.sync(g_err) : ORBX_OK;
}

// one strided D2H copy: host [n][rowBytes] <- device rows of stride devStride starting at src
static int d2h_rows(Engine& E, void* dst, const void* src, size_t rowBytes, size_t devStride, int n, const char* what)
{
    if (!dst) return ORBX_OK;
    return orbx_check(cudaMemcpy2DAsync(dst, rowBytes, src, devStride, rowBytes, n, cudaMemcpyDeviceToHost, E.stream), what, g_err);
}

extern "C" int orbx_batch_download(orbx_batch* b, int nFrames, int* nLeft, orbx_keypoint* kpsLeft, uint8_t* descLeft, int* nRight,
                                   orbx_keypoint* kpsRight, uint8_t* descRight, float* uRight, float* depth)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames) return fail(ORBX_ERR_ARG, "bad frame count");
    ORBX_DRAIN(b);
    Engine& E = b->eng;
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    const size_t per = b->stereo ? 2 : 1, P = E.P.maxKp;
    const uint8_t* kp = (const uint8_t*)E.P.outKp;
    TRY(d2h_rows(E, nLeft, E.P.outCnt, 4, per * 4, nFrames, "D2H nLeft"));
    TRY(d2h_rows(E, kpsLeft, kp, P * 28, per * P * 28, nFrames, "D2H kpsLeft"));
    TRY(d2h_rows(E, descLeft, E.P.outDesc, P * 32, per * P * 32, nFrames, "D2H descLeft"));
    if (b->stereo) {
        TRY(d2h_rows(E, nRight, E.P.outCnt + 1, 4, per * 4, nFrames, "D2H nRight"));
        TRY(d2h_rows(E, kpsRight, kp + P * 28, P * 28, per * P * 28, nFrames, "D2H kpsRight"));
        TRY(d2h_rows(E, descRight, E.P.outDesc + P * 32, P * 32, per * P * 32, nFrames, "D2H descRight"));
        TRY(d2h_rows(E, uRight, E.SP.uRight, P * 4, P * 4, nFrames, "D2H uRight"));
        TRY(d2h_rows(E, depth, E.SP.depth, P * 4, P * 4, nFrames, "D2H depth"));
    }
    return E.sync(g_err);
}

// Asynchronous orbx_batch_process: returns when the work is enqueued; consecutive calls pipeline (the uploads of call k+1
// overlap the kernels of call k). orbx_batch_wait() drains; every other orbx_batch_* call drains first by itself.
extern "C" int orbx_batch_process_async(orbx_batch* b, const uint8_t* left, const uint8_t* right, int nFrames, float mbf, float mb,
                                        int* nLeft, orbx_keypoint* kpsLeft, uint8_t* descLeft, int* nRight, orbx_keypoint* kpsRight,
                                        uint8_t* descRight, float* uRight, float* depth)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames) return fail(ORBX_ERR_ARG, "bad frame count");
    Engine::HostOutputs o{nLeft, kpsLeft, descLeft, nRight, kpsRight, descRight, uRight, depth};
    // four large chunks: without a ramp to hide, fewer and larger launches run closer to the device-resident rate
    const int forcedChunks = orbx_tune(ORBX_TUNE_E2E_CHUNKS);
    const int chunk = forcedChunks > 0 ? (nFrames + forcedChunks - 1) / forcedChunks : nFrames >= 64 ? (nFrames + 3) / 4 : nFrames >= 32 ? 16 : nFrames;
    return b->eng.processPipelined(left, right, nFrames, mbf, mb, o, chunk, g_err, false);
}

extern "C" int orbx_batch_wait(orbx_batch* b)
{
    if (!b) return fail(ORBX_ERR_ARG, "null batch");
    return b->eng.waitPipelined(g_err);
}

extern "C" int orbx_batch_process(orbx_batch* b, const uint8_t* left, const uint8_t* right, int nFrames, float mbf, float mb,
                                  int* nLeft, orbx_keypoint* kpsLeft, uint8_t* descLeft, int* nRight, orbx_keypoint* kpsRight,
                                  uint8_t* descRight, float* uRight, float* depth)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames) return fail(ORBX_ERR_ARG, "bad frame count");
    Engine::HostOutputs o{nLeft, kpsLeft, descLeft, nRight, kpsRight, descRight, uRight, depth};
    return b->eng.processPipelined(left, right, nFrames, mbf, mb, o, e2e_chunk_frames(nFrames), g_err);
}

extern "C" int orbx_batch_stats(orbx_batch* b, int nFrames, long long* nLeft, long long* nRight, long long* nMatched)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames) return fail(ORBX_ERR_ARG, "bad frame count");
    ORBX_DRAIN(b);
    Engine& E = b->eng;
    const int per = b->stereo ? 2 : 1;
    std::vector<int> cnt((size_t)nFrames * per), nm(nFrames, 0);
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    TRY(orbx_check(cudaMemcpyAsync(cnt.data(), E.P.outCnt, cnt.size() * 4, cudaMemcpyDeviceToHost, E.stream), "D2H counts", g_err));
    if (b->stereo) TRY(orbx_check(cudaMemcpyAsync(nm.data(), E.SP.nMatched, nm.size() * 4, cudaMemcpyDeviceToHost, E.stream), "D2H matched", g_err));
    TRY(E.sync(g_err));
    long long l = 0, r = 0, m = 0;
    for (int f = 0; f < nFrames; ++f) { l += cnt[(size_t)f * per]; if (b->stereo) { r += cnt[(size_t)f * 2 + 1]; m += nm[f]; } }
    if (nLeft) *nLeft = l;
    if (nRight) *nRight = r;
    if (nMatched) *nMatched = m;
    return ORBX_OK;
}

extern "C" int orbx_batch_level_size(const orbx_batch* b, int level, int* rows, int* cols)
{
    if (!b || level < 0 || level >= b->cfg.nlevels || !rows || !cols) return fail(ORBX_ERR_ARG, "bad arguments");
    *rows = b->eng.P.lv[level].h;
    *cols = b->eng.P.lv[level].w;
    return ORBX_OK;
}

extern "C" int orbx_batch_level(orbx_batch* b, int frame, int side, int level, int blurred, uint8_t* dst, size_t dstStride)
{
    if (!b || !dst || level < 0 || level >= b->cfg.nlevels || frame < 0 || frame >= b->maxFrames || side < 0 || side > (b->stereo ? 1 : 0))
        return fail(ORBX_ERR_ARG, "bad arguments");
    ORBX_DRAIN(b);
    return copy_level(b->eng, frame * (b->stereo ? 2 : 1) + side, level, blurred != 0, dst, dstStride);
}

extern "C" int orbx_batch_candidates(orbx_batch* b, int frame, int side, int level, orbx_keypoint* out, int capacity, int* n)
{
    if (!b || !n || level < 0 || level >= b->cfg.nlevels || frame < 0 || frame >= b->maxFrames || side < 0 || side > (b->stereo ? 1 : 0))
        return fail(ORBX_ERR_ARG, "bad arguments");
    ORBX_DRAIN(b);
    Engine& E = b->eng;
    const int img = frame * (b->stereo ? 2 : 1) + side;
    const OrbxLevel& L = E.P.lv[level];
    int cnt = 0;
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    TRY(orbx_check(cudaMemcpy(&cnt, E.P.candCnt + (size_t)img * E.P.nlevels + level, 4, cudaMemcpyDeviceToHost), "D2H candCnt", g_err));
    *n = cnt;
    const int m = cnt < capacity ? cnt : capacity;
    if (m <= 0 || !out) return ORBX_OK;
    std::vector<uint32_t> xy(m);
    std::vector<uint8_t> rs(m);
    TRY(orbx_check(cudaMemcpy(xy.data(), E.P.candXY + (size_t)img * E.P.candTotal + L.candBase, (size_t)m * 4, cudaMemcpyDeviceToHost), "D2H candXY", g_err));
    TRY(orbx_check(cudaMemcpy(rs.data(), E.P.candRs + (size_t)img * E.P.candTotal + L.candBase, (size_t)m, cudaMemcpyDeviceToHost), "D2H candRs", g_err));
    for (int i = 0; i < m; ++i) {
        out[i].x = (float)(xy[i] & 0xFFFF); out[i].y = (float)(xy[i] >> 16); out[i].size = 7.f; out[i].angle = -1.f;
        out[i].response = (float)rs[i]; out[i].octave = 0; out[i].class_id = -1;
    }
    return ORBX_OK;
}

// per-stage CUDA-event timing (stages: pyramid, fast, octree, blur, describe, stereo)
extern "C" int orbx_batch_set_profiling(orbx_batch* b, int on)
{
    if (!b) return fail(ORBX_ERR_ARG, "handle is NULL");
    return b->eng.setProfiling(on != 0, g_err);
}
extern "C" int orbx_batch_stage_count(void) { return Engine::kStages - 1; }
extern "C" const char* orbx_batch_stage_name(int i) { return Engine::stageName(i); }
extern "C" int orbx_batch_stage_times(orbx_batch* b, double* ms, int* runs)
{
    if (!b || !ms || !runs) return fail(ORBX_ERR_ARG, "bad arguments");
    ORBX_DRAIN(b);
    double t[Engine::kStages];
    TRY(b->eng.stageTimes(t, runs, g_err));
    for (int i = 0; i < Engine::kStages - 1; ++i) ms[i] = t[i];
    return ORBX_OK;
}

// pinned host memory for the end-to-end path, placed on the NUMA node the current device hangs off: with one process per
// GPU on a two-socket host, pages first touched on the far socket make every H2D/D2H cross the socket interconnect.
// The node comes from sysfs (the device's PCI function); the placement is a PREFERRED memory policy of the calling
// thread held only around the allocation (the driver populates and pins the pages inside cudaHostAlloc). Where the
// policy call is not permitted (seccomp) the allocation runs on one of the node's cores instead (local allocation is the
// default policy); where the node is unknown the allocation is the plain one. orbx_set_tuning("numa_off", 1) disables.
#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>
static int g_numaNode = -2;   // node of the last orbx_host_alloc: -2 not tried, -1 unknown/unavailable
static int g_numaBound = 0;   // 1: the last allocation ran under the policy, 2: on the node's cores, 0: plain

static int device_numa_node()
{
    int dev = 0;
    char bus[32] = {0};
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) { cudaGetLastError(); return -1; }
    for (char* c = bus; *c; ++c) *c = (char)tolower(*c);
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}

static long set_policy(int mode, const unsigned long* mask, unsigned long maxnode)
{
#ifdef SYS_set_mempolicy
    return syscall(SYS_set_mempolicy, mode, mask, maxnode);
#else
    return -1;
#endif
}

// "0-13,56-69" -> cpu set
static bool node_cpus(int node, cpu_set_t* set)
{
    const std::string path = "/sys/devices/system/node/node" + std::to_string(node) + "/cpulist";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return false;
    CPU_ZERO(set);
    int a, b, n = 0;
    while (fscanf(f, "%d", &a) == 1) {
        b = a;
        int c = fgetc(f);
        if (c == '-') { if (fscanf(f, "%d", &b) != 1) break; c = fgetc(f); }
        for (int i = a; i <= b && i < CPU_SETSIZE; ++i) { CPU_SET(i, set); ++n; }
        if (c != ',') break;
    }
    fclose(f);
    return n > 0;
}

extern "C" void* orbx_host_alloc(size_t bytes)
{
    const int node = orbx_tune(ORBX_TUNE_NUMA_OFF) ? -1 : device_numa_node();
    unsigned long mask[16] = {0};  // up to 1024 nodes
    bool bound = false;
    if (node >= 0 && node < 1024) {
        mask[node / (8 * sizeof(unsigned long))] = 1ul << (node % (8 * sizeof(unsigned long)));
        bound = set_policy(/*MPOL_PREFERRED*/ 1, mask, 1024) == 0;
    }
    // no policy call on this host: run the allocation on one of the node's cores instead (default policy = local node)
    cpu_set_t old, local;
    bool moved = false;
    if (node >= 0 && !bound && sched_getaffinity(0, sizeof old, &old) == 0 && node_cpus(node, &local))
        moved = sched_setaffinity(0, sizeof local, &local) == 0;
    void* p = nullptr;
    const cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
    if (bound) set_policy(/*MPOL_DEFAULT*/ 0, nullptr, 0);
    if (moved) sched_setaffinity(0, sizeof old, &old);
    g_numaNode = node;
    g_numaBound = bound ? 1 : moved ? 2 : 0;
    if (e != cudaSuccess) { g_err = "cudaHostAlloc failed"; cudaGetLastError(); return nullptr; }
    return p;
}
// node of the device seen by the last orbx_host_alloc (-1 unknown, -2 no allocation yet); *bound = the policy was applied
extern "C" int orbx_host_numa_node(int* bound)
{
    if (bound) *bound = g_numaBound;
    return g_numaNode;
}
extern "C" void orbx_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- raw-frame input path (include/orbx_input.h; kernel in input.cu) -----------------------------------------
#include "../../include/orbx_input.h"

extern "C" int orbx_batch_set_rectification(orbx_batch* b, int side, const float* map1, const float* map2, int mapRows, int mapCols,
                                            size_t mapStrideBytes, int cropTop)
{
    if (!b || side < 0 || side > (b->stereo ? 1 : 0)) return fail(ORBX_ERR_ARG, "bad handle or side");
    ORBX_DRAIN(b);
    Engine& E = b->eng;
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    TRY(E.sync(g_err));
    if (E.dMap1[side]) { cudaFree(E.dMap1[side]); cudaFree(E.dMap2[side]); E.dMap1[side] = E.dMap2[side] = nullptr; }
    E.mapRows[side] = E.mapCols[side] = 0;
    E.cropTop[side] = cropTop;
    if (!map1 && !map2) return cropTop >= 0 ? ORBX_OK : fail(ORBX_ERR_ARG, "negative crop");
    if (!map1 || !map2 || mapRows < 1 || mapCols != E.cols || cropTop < 0 || cropTop + E.rows > mapRows || mapStrideBytes < (size_t)mapCols * 4 ||
        (mapStrideBytes & 3))
        return fail(ORBX_ERR_ARG, "maps must be map_rows x (engine cols) CV_32FC1 with crop_top + engine rows <= map_rows");
    const size_t bytes = (size_t)mapRows * mapCols * 4;
    TRY(orbx_check(cudaMalloc(&E.dMap1[side], bytes), "map alloc", g_err));
    TRY(orbx_check(cudaMalloc(&E.dMap2[side], bytes), "map alloc", g_err));
    TRY(orbx_check(cudaMemcpy2D(E.dMap1[side], (size_t)mapCols * 4, map1, mapStrideBytes, (size_t)mapCols * 4, mapRows, cudaMemcpyHostToDevice), "map H2D", g_err));
    TRY(orbx_check(cudaMemcpy2D(E.dMap2[side], (size_t)mapCols * 4, map2, mapStrideBytes, (size_t)mapCols * 4, mapRows, cudaMemcpyHostToDevice), "map H2D", g_err));
    E.mapRows[side] = mapRows;
    E.mapCols[side] = mapCols;
    return ORBX_OK;
}

extern "C" int orbx_batch_upload_raw(orbx_batch* b, const uint8_t* left, const uint8_t* right, int nFrames, int srcRows, int srcCols,
                                     int channels, int rgbOrder)
{
    if (!b || nFrames < 1 || nFrames > b->maxFrames || !left || (b->stereo && !right)) return fail(ORBX_ERR_ARG, "bad upload arguments");
    ORBX_DRAIN(b);
    if (srcRows < 1 || srcCols < 1 || (channels != 1 && channels != 3 && channels != 4)) return fail(ORBX_ERR_ARG, "bad raw frame format");
    Engine& E = b->eng;
    const int per = b->stereo ? 2 : 1;
    for (int s = 0; s < per; ++s)
        if (!E.dMap1[s] && (srcCols != E.cols || E.cropTop[s] + E.rows > srcRows))
            return fail(ORBX_ERR_ARG, "without maps the raw frame must contain the engine's image (same width, crop inside)");
    TRY(orbx_check(cudaSetDevice(E.device), "cudaSetDevice", g_err));
    const size_t frame = (size_t)srcRows * srcCols * channels, need = frame * nFrames * per;
    if (need > E.rawCap) {
        TRY(E.sync(g_err));
        if (E.dRaw) cudaFree(E.dRaw);
        E.dRaw = nullptr; E.rawCap = 0;
        TRY(orbx_check(cudaMalloc(&E.dRaw, frame * b->maxFrames * per), "raw frame buffer", g_err));
        E.rawCap = frame * b->maxFrames * per;
    }
    for (int s = 0; s < per; ++s) {
        uint8_t* d = E.dRaw + (size_t)s * frame * nFrames;
        TRY(orbx_check(cudaMemcpyAsync(d, s ? right : left, frame * nFrames, cudaMemcpyHostToDevice, E.stream), "H2D raw frames", g_err));
        OrbxStageArgs a{};
        a.src = d; a.srcImgStride = frame; a.srcRows = srcRows; a.srcCols = srcCols; a.srcStride = srcCols * channels;
        a.map1 = E.dMap1[s]; a.map2 = E.dMap2[s]; a.mapStride = E.mapCols[s];
        a.cropTop = E.cropTop[s]; a.rows = E.rows; a.cols = E.cols;
        a.dst = E.P.img0; a.dstImgStride = E.P.img0Stride; a.dstPitch = E.P.lv[0].pitch; a.imgStep = per; a.imgOfs = s; a.rgbOrder = rgbOrder;
        orbx_launch_stage(a, channels, nFrames, E.stream);
    }
    return orbx_check(cudaGetLastError(), "stage launch", g_err);
}
