// bow.cu — SURVEY.md §8f N2, declared in include/orbx_bow.h: Frame::ComputeBoW (/root/reference/src/Frame.cc:447-454) =
// TemplatedVocabulary::transform(features, BowVector, FeatureVector, levelsup)
// (/root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1207, 1231-1272), TF_IDF weighting, L1 scoring.
//
//   orbx_bow_descend_kernel  warp per feature: at every tree level lane c takes child c (k <= 32), Hamming distance by
//                            popcount (FORB::distance, FORB.cpp:81-101), warp-min of (distance, child order) = the
//                            reference's first strict minimum; the node at level L - levelsup is remembered.
//   orbx_bow_reduce_kernel   CTA per frame: bitonic sort of (word id, feature index) in shared memory; a word's weights
//                            are added in feature order (BowVector::addWeight), the L1 norm is summed in word order by
//                            one thread (floating-point addition is not associative; BowVector::normalize is a loop
//                            over the map) and every value divided by it; the same sort on (node id, feature index)
//                            gives the FeatureVector in CSR form.
#include <cstring>
#include <string>
#include <vector>

#include "../../include/orbx_bow.h"
#include "engine.h"
#include "matcher.h"
#include "handles.h"

struct orbx_vocabulary {
    orbx_matcher ctx;  // stream + packed call buffers
    int L = 0, nNodes = 0, nWords = 0;
    int* dChildStart = nullptr; int* dChildren = nullptr; uint8_t* dDesc = nullptr; double* dWeight = nullptr; int* dWordId = nullptr;
};

namespace {

struct BowFrame {
    int n, cap;  // features, power-of-two sort capacity is per launch
    const uint8_t* desc;
    int* word; int* nid; double* w;  // per feature
    int* wordIds; double* wordVals; int* fvNodes; int* fvStart; int* fvFeats; int* counts;  // counts[0] = nWords, [1] = nFv
};

struct VocDev {
    const int* childStart; const int* children; const uint8_t* desc; const double* weight; const int* wordId;
    int nidLevel;
};

__global__ void __launch_bounds__(256) orbx_bow_descend_kernel(VocDev V, const BowFrame* frames)
{
    const BowFrame& F = frames[blockIdx.y];
    const int lane = threadIdx.x & 31, i = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= F.n) return;
    const uint4* f4 = reinterpret_cast<const uint4*>(F.desc + 32 * (size_t)i);
    const uint4 a0 = f4[0], a1 = f4[1];
    int node = 0, lvl = 0, nid = 0;
    while (true) {
        ++lvl;
        const int b = V.childStart[node], e = V.childStart[node + 1];
        unsigned key = 0xFFFFFFFFu;
        if (lane < e - b) {
            const int c = V.children[b + lane];
            const uint4* d4 = reinterpret_cast<const uint4*>(V.desc + 32 * (size_t)c);
            const uint4 b0 = d4[0], b1 = d4[1];
            const int dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                             __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
            key = ((unsigned)dist << 8) | (unsigned)lane;
        }
        key = __reduce_min_sync(0xffffffffu, key);  // REDUX.MIN: first minimum in child order
        node = V.children[b + (int)(key & 255u)];
        if (lvl == V.nidLevel) nid = node;
        if (V.childStart[node + 1] == V.childStart[node]) break;  // isLeaf() = no children
    }
    if (lane == 0) {
        const double w = V.weight[node];
        F.word[i] = w > 0 ? V.wordId[node] : -1;  // "not stopped" :1170
        F.nid[i] = nid;
        F.w[i] = w;
    }
}

// ascending bitonic sort of keys[0..cap) in shared memory, cap a power of two, all threads of the CTA call
__device__ void bitonic_sort(unsigned long long* keys, int cap)
{
    for (int k = 2; k <= cap; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < cap; t += blockDim.x) {
                const int p = t ^ j;
                if (p > t) {
                    const unsigned long long a = keys[t], b = keys[p];
                    const bool up = (t & k) == 0;
                    if ((a > b) == up) { keys[t] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
}

// positions of the segment heads among keys[0..nValid): exclusive scan of head flags; returns the number of heads
__device__ int head_positions(const unsigned long long* keys, int nValid, int cap, int* pos, int* wsum)
{
    const int T = blockDim.x, per = (cap + T - 1) / T, b = threadIdx.x * per, e = min(b + per, nValid);
    int s = 0;
    for (int i = b; i < e; ++i) s += (i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32)) ? 1 : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (threadIdx.x == 0) { int run = 0; for (int w = 0; w < T / 32; ++w) { const int t = wsum[w]; wsum[w] = run; run += t; } wsum[T / 32] = run; }
    __syncthreads();
    int run = wsum[warp] + inc - s;
    for (int i = b; i < e; ++i) {
        const bool head = i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32);
        pos[i] = head ? run : -1;
        run += head;
    }
    const int total = wsum[T / 32];
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(1024) orbx_bow_reduce_kernel(const BowFrame* frames, int cap)
{
    extern __shared__ unsigned long long keys[];  // [cap], then int pos[cap]
    __shared__ int wsum[33];
    __shared__ int nValidSh;
    __shared__ double normSh;
    int* pos = reinterpret_cast<int*>(keys + cap);
    const BowFrame& F = frames[blockIdx.x];
    const int tid = threadIdx.x, T = blockDim.x;
    // ---- BowVector
    if (tid == 0) nValidSh = 0;
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < cap; i += T) {
        const bool ok = i < F.n && F.word[i] >= 0;
        keys[i] = ok ? ((unsigned long long)(unsigned)F.word[i] << 32) | (unsigned)i : ~0ull;
        mine += ok;
    }
    if (mine) atomicAdd(&nValidSh, mine);
    __syncthreads();
    const int nValid = nValidSh;
    bitonic_sort(keys, cap);
    const int nWords = head_positions(keys, nValid, cap, pos, wsum);
    for (int i = tid; i < nValid; i += T)
        if (pos[i] >= 0) {
            double v = F.w[(int)(keys[i] & 0xFFFFFFFFu)];  // addWeight: insert, then += in feature order
            for (int j = i + 1; j < nValid && (keys[j] >> 32) == (keys[i] >> 32); ++j) v += F.w[(int)(keys[j] & 0xFFFFFFFFu)];
            F.wordIds[pos[i]] = (int)(keys[i] >> 32);
            F.wordVals[pos[i]] = v;
        }
    __syncthreads();
    if (tid == 0) {  // BowVector::normalize(L1): the reference's summation order
        double norm = 0.0;
        for (int k = 0; k < nWords; ++k) norm += fabs(F.wordVals[k]);
        normSh = norm;
        F.counts[0] = nWords;
    }
    __syncthreads();
    const double norm = normSh;
    if (norm > 0.0)
        for (int k = tid; k < nWords; k += T) F.wordVals[k] = __ddiv_rn(F.wordVals[k], norm);
    __syncthreads();
    // ---- FeatureVector
    for (int i = tid; i < cap; i += T) {
        const bool ok = i < F.n && F.word[i] >= 0;
        keys[i] = ok ? ((unsigned long long)(unsigned)F.nid[i] << 32) | (unsigned)i : ~0ull;
    }
    __syncthreads();
    bitonic_sort(keys, cap);
    const int nFv = head_positions(keys, nValid, cap, pos, wsum);
    for (int i = tid; i < nValid; i += T) {
        F.fvFeats[i] = (int)(keys[i] & 0xFFFFFFFFu);
        if (pos[i] >= 0) { F.fvNodes[pos[i]] = (int)(keys[i] >> 32); F.fvStart[pos[i]] = i; }
    }
    if (tid == 0) { F.fvStart[nFv] = nValid; F.counts[1] = nFv; }
}

}  // namespace

// ======================================= host side / C-ABI ===================================================
extern "C" int orbx_vocabulary_create(int device, int L, int nNodes, const int* parent, const uint8_t* isLeaf, const uint8_t* desc,
                                      const double* weight, orbx_vocabulary** out)
{
    if (!out) return orbx_fail(ORBX_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (L < 1 || nNodes < 1 || !parent || !isLeaf || !desc || !weight) return orbx_fail(ORBX_ERR_ARG, "bad vocabulary arguments");
    const int N = nNodes + 1;  // with the root
    std::vector<int> cnt(N + 1, 0), wordId(N, 0);
    int nWords = 0;
    for (int i = 0; i < nNodes; ++i) {
        if (parent[i] < 0 || parent[i] > i) return orbx_fail(ORBX_ERR_ARG, "a node's parent must precede it (loadFromTextFile order)");
        cnt[parent[i] + 1]++;
        if (isLeaf[i]) wordId[i + 1] = nWords++;
    }
    for (int p = 0; p < N; ++p) {
        if (cnt[p + 1] > 32) return orbx_fail(ORBX_ERR_ARG, "more than 32 children per node");
        if (p > 0 && (cnt[p + 1] == 0) != (isLeaf[p - 1] != 0)) return orbx_fail(ORBX_ERR_ARG, "leaf flag and children disagree");
    }
    if (cnt[1] == 0) return orbx_fail(ORBX_ERR_ARG, "the root has no children");
    std::vector<int> start(N + 1, 0), children(nNodes), cursor;
    for (int p = 0; p < N; ++p) start[p + 1] = start[p] + cnt[p + 1];
    cursor = start;
    for (int i = 0; i < nNodes; ++i) children[cursor[parent[i]]++] = i + 1;  // push_back order
    std::vector<uint8_t> d((size_t)N * 32, 0);
    std::memcpy(d.data() + 32, desc, (size_t)nNodes * 32);
    std::vector<double> w(N, 0.0);
    std::memcpy(w.data() + 1, weight, (size_t)nNodes * 8);

    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(device), "cudaSetDevice", err));
    orbx_vocabulary* v = new orbx_vocabulary();
    v->ctx.device = device;
    v->L = L; v->nNodes = N; v->nWords = nWords;
    auto up = [&](void** dp, const void* src, size_t bytes, const char* what) -> int {
        TRY(orbx_check(cudaMalloc(dp, bytes), what, err));
        return orbx_check(cudaMemcpy(*dp, src, bytes, cudaMemcpyHostToDevice), what, err);
    };
    int rc = orbx_check(cudaStreamCreateWithFlags(&v->ctx.stream, cudaStreamNonBlocking), "cudaStreamCreate", err);
    if (!rc) rc = up((void**)&v->dChildStart, start.data(), (size_t)(N + 1) * 4, "vocabulary childStart");
    if (!rc) rc = up((void**)&v->dChildren, children.data(), (size_t)nNodes * 4, "vocabulary children");
    if (!rc) rc = up((void**)&v->dDesc, d.data(), d.size(), "vocabulary descriptors");
    if (!rc) rc = up((void**)&v->dWeight, w.data(), (size_t)N * 8, "vocabulary weights");
    if (!rc) rc = up((void**)&v->dWordId, wordId.data(), (size_t)N * 4, "vocabulary word ids");
    if (rc) { orbx_vocabulary_destroy(v); return rc; }
    *out = v;
    return ORBX_OK;
}

extern "C" void orbx_vocabulary_destroy(orbx_vocabulary* v)
{
    if (!v) return;
    cudaSetDevice(v->ctx.device);
    for (void* p : {(void*)v->dChildStart, (void*)v->dChildren, (void*)v->dDesc, (void*)v->dWeight, (void*)v->dWordId, (void*)v->ctx.dBuf})
        if (p) cudaFree(p);
    if (v->ctx.hIn) cudaFreeHost(v->ctx.hIn);
    if (v->ctx.hOut) cudaFreeHost(v->ctx.hOut);
    if (v->ctx.stream) cudaStreamDestroy(v->ctx.stream);
    delete v;
}

extern "C" int orbx_vocabulary_words(const orbx_vocabulary* v) { return v ? v->nWords : ORBX_ERR_ARG; }

// desc[f]: host descriptors (resident == NULL) or, with `resident`, device pointers into a batch engine (nothing is uploaded)
static int compute_bow_impl(orbx_vocabulary* v, const uint8_t* const* desc, orbx_batch* resident, const int* n, int nFrames, int levelsup,
                            int* const* wordIds, double* const* wordVals, int* nWords, int* const* fvNodes, int* const* fvStart,
                            int* const* fvFeats, int* nFvNodes)
{
    if (!v || !desc || !n || nFrames < 1 || !wordIds || !wordVals || !nWords || !fvNodes || !fvStart || !fvFeats || !nFvNodes)
        return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    int maxN = 0;
    for (int f = 0; f < nFrames; ++f) {
        if (n[f] < 0 || n[f] > 8192 || !fvStart[f] || (n[f] && (!desc[f] || !wordIds[f] || !wordVals[f] || !fvNodes[f] || !fvFeats[f])))
            return orbx_fail(ORBX_ERR_ARG, "bad frame (0 <= n <= 8192 features, non-NULL buffers)");
        maxN = n[f] > maxN ? n[f] : maxN;
    }
    orbx_matcher* m = &v->ctx;
    std::string& err = orbx_err_slot();
    TRY(orbx_check(cudaSetDevice(m->device), "cudaSetDevice", err));
    struct Off { size_t desc, word, nid, w, ids, vals, fn, fs, ff, cnt; };
    std::vector<Off> off(nFrames);
    OrbxPacker pk;
    const size_t oFr = pk.take(sizeof(BowFrame) * (size_t)nFrames);
    for (int f = 0; f < nFrames; ++f) off[f].desc = pk.take(resident ? 0 : (size_t)n[f] * 32);
    const size_t inBytes = pk.off;
    for (int f = 0; f < nFrames; ++f) { off[f].word = pk.take((size_t)n[f] * 4); off[f].nid = pk.take((size_t)n[f] * 4); off[f].w = pk.take((size_t)n[f] * 8); }
    const size_t outBase = pk.off;
    for (int f = 0; f < nFrames; ++f) {
        off[f].vals = pk.take((size_t)n[f] * 8); off[f].ids = pk.take((size_t)n[f] * 4); off[f].fn = pk.take((size_t)n[f] * 4);
        off[f].fs = pk.take((size_t)n[f] * 4 + 4); off[f].ff = pk.take((size_t)n[f] * 4); off[f].cnt = pk.take(8);
    }
    const size_t outBytes = pk.off - outBase;
    TRY(orbx_matcher_grow(m, pk.off, inBytes, outBytes));
    BowFrame* hf = reinterpret_cast<BowFrame*>(m->hIn + oFr);
    for (int f = 0; f < nFrames; ++f) {
        uint8_t* d = m->dBuf;
        BowFrame F{};
        F.n = n[f];
        F.desc = resident ? desc[f] : d + off[f].desc; F.word = (int*)(d + off[f].word); F.nid = (int*)(d + off[f].nid); F.w = (double*)(d + off[f].w);
        F.wordIds = (int*)(d + off[f].ids); F.wordVals = (double*)(d + off[f].vals); F.fvNodes = (int*)(d + off[f].fn);
        F.fvStart = (int*)(d + off[f].fs); F.fvFeats = (int*)(d + off[f].ff); F.counts = (int*)(d + off[f].cnt);
        hf[f] = F;
        if (n[f] && !resident) std::memcpy(m->hIn + off[f].desc, desc[f], (size_t)n[f] * 32);
    }
    cudaStream_t s = m->stream;
    if (resident) {  // behind whatever the batch engine still has in flight
        Engine& E = resident->eng;
        if (!E.lastBounds.empty()) TRY(E.waitPipelined(err));
        if (!m->orderEvent) TRY(orbx_check(cudaEventCreateWithFlags(&m->orderEvent, cudaEventDisableTiming), "cudaEventCreate", err));
        TRY(orbx_check(cudaEventRecord(m->orderEvent, E.stream), "event", err));
        TRY(orbx_check(cudaStreamWaitEvent(s, m->orderEvent, 0), "wait", err));
    }
    TRY(orbx_check(cudaMemcpyAsync(m->dBuf, m->hIn, inBytes, cudaMemcpyHostToDevice, s), "H2D descriptors", err));
    const BowFrame* df = (const BowFrame*)(m->dBuf + oFr);
    VocDev V{v->dChildStart, v->dChildren, v->dDesc, v->dWeight, v->dWordId, v->L - levelsup};
    if (maxN) orbx_bow_descend_kernel<<<dim3((maxN + 7) / 8, nFrames), 256, 0, s>>>(V, df);
    int cap = 32;
    while (cap < maxN) cap <<= 1;
    const size_t smem = (size_t)cap * 12;
    if (smem > 32 * 1024) cudaFuncSetAttribute(orbx_bow_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    orbx_bow_reduce_kernel<<<nFrames, 1024, smem, s>>>(df, cap);
    TRY(orbx_check(cudaGetLastError(), "bow launch", err));
    TRY(orbx_check(cudaMemcpyAsync(m->hOut, m->dBuf + outBase, outBytes, cudaMemcpyDeviceToHost, s), "D2H bow", err));
    TRY(orbx_check(cudaStreamSynchronize(s), "sync", err));
    for (int f = 0; f < nFrames; ++f) {
        const uint8_t* h = m->hOut - outBase;
        const int* cnt = (const int*)(h + off[f].cnt);
        nWords[f] = cnt[0];
        nFvNodes[f] = cnt[1];
        if (cnt[0]) { std::memcpy(wordIds[f], h + off[f].ids, (size_t)cnt[0] * 4); std::memcpy(wordVals[f], h + off[f].vals, (size_t)cnt[0] * 8); }
        if (cnt[1]) std::memcpy(fvNodes[f], h + off[f].fn, (size_t)cnt[1] * 4);
        std::memcpy(fvStart[f], h + off[f].fs, ((size_t)cnt[1] + 1) * 4);
        const int nv = ((const int*)(h + off[f].fs))[cnt[1]];
        if (nv) std::memcpy(fvFeats[f], h + off[f].ff, (size_t)nv * 4);
    }
    return ORBX_OK;
}

extern "C" int orbx_compute_bow(orbx_vocabulary* v, const uint8_t* const* desc, const int* n, int nFrames, int levelsup, int* const* wordIds,
                                double* const* wordVals, int* nWords, int* const* fvNodes, int* const* fvStart, int* const* fvFeats,
                                int* nFvNodes)
{
    return compute_bow_impl(v, desc, nullptr, n, nFrames, levelsup, wordIds, wordVals, nWords, fvNodes, fvStart, fvFeats, nFvNodes);
}

extern "C" int orbx_compute_bow_resident(orbx_vocabulary* v, orbx_batch* b, int firstFrame, int side, const int* n, int nFrames, int levelsup,
                                         int* const* wordIds, double* const* wordVals, int* nWords, int* const* fvNodes, int* const* fvStart,
                                         int* const* fvFeats, int* nFvNodes)
{
    if (!v || !b || !n || nFrames < 1) return orbx_fail(ORBX_ERR_ARG, "bad arguments");
    const int per = b->stereo ? 2 : 1;
    if (firstFrame < 0 || firstFrame + nFrames > b->maxFrames || side < 0 || side >= per) return orbx_fail(ORBX_ERR_ARG, "frames outside the batch");
    if (b->eng.device != v->ctx.device) return orbx_fail(ORBX_ERR_STATE, "batch and vocabulary live on different devices");
    std::vector<const uint8_t*> dd(nFrames);
    for (int f = 0; f < nFrames; ++f) {
        if (n[f] > b->eng.P.maxKp) return orbx_fail(ORBX_ERR_ARG, "more features than the batch engine can hold");
        dd[f] = b->eng.P.outDesc + ((size_t)(firstFrame + f) * per + side) * b->eng.P.maxKp * 32;
    }
    return compute_bow_impl(v, dd.data(), b, n, nFrames, levelsup, wordIds, wordVals, nWords, fvNodes, fvStart, fvFeats, nFvNodes);
}
