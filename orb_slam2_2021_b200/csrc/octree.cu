// octree.cu — deterministic data-parallel DistributeOctTree: one CTA per (pyramid level, image).
// Replaces ORBextractor::DistributeOctTree and ExtractorNode::DivideNode
// (/root/reference/src/ORBextractor.cc:484-540, 542-766) and the candidate concatenation of
// ComputeKeyPointsOctTree (:823-828).
//
// The reference walks a std::list, splitting nodes and pushing children to the front; the output
// order of a level is the final list order. Here a node IS its list position and every candidate
// carries the position of the node that owns it (label). One pass = for a set of nodes to split,
// given in "creation order" r = 0..S-1:
//     count the 4 quadrant populations (shared-memory atomics over the candidates),
//     c_r = #non-empty children, g_r = c_r - 1, exclusive scan of g  ->  child k of node r is the
//     (scan_r + r + k)-th created child; with T children in total it lands at position T-1-k'
//     (push_front reverses creation order), unsplit nodes keep their relative order behind them.
// Phase 1 (:609-668) splits every multi-key node in list order. Phase 2 (:679-740) sorts the
// children created by the previous pass by (size, creation index) and splits from the largest
// down until the list reaches N: that stop index t* is the first r with m + inclusive_scan(g)_r >= N.
// The reference breaks equal sizes by heap address; the canonical rule is creation
// index, later first == smaller list position first. The intra-node key order only matters for
// the final "first maximum wins" (:747-763), which is reproduced as max over
// (response << 24 | ~candidate index).
#include "orbx_dev.cuh"

#define OCT_THREADS 256
#define OCT_K 8  // candidates per thread kept in registers (levels with up to OCT_THREADS*OCT_K candidates)

struct OctSmem {
    short4* box[2];     // x0,y0,x1,y1 per list position (ping-pong)
    int* cnt[2];        // keys per node
    uint8_t* inV[2];    // node was created by the previous pass and has >1 keys
    int* childCnt;      // [cap*4]
    uint16_t* newPos;   // [cap*4] position of child q of a split node
    uint16_t* movePos;  // [cap] position of an unsplit node
    int* rankOf;        // [cap] creation-order rank of a split candidate, -1 otherwise
    int* byRank;        // [cap] inverse
    int* scan;          // [cap+1]
    int* vlist;         // [cap]
    unsigned* best;     // [cap]
};

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// In-place exclusive scan of a[0..n) in shared memory; a[n] receives the total. All threads call. Two barriers: every warp
// publishes its sum, then every thread adds up the sums of the warps before its own (8 values) instead of waiting for a
// second, single-warp scan.
__device__ void block_excl_scan(int* a, int n, int* wsum /*[OCT_THREADS/32 + 1]*/)
{
    const int T = OCT_THREADS, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk = (n + T - 1) / T;
    const int b = min(tid * chunk, n), e = min(b + chunk, n);
    int s = 0;
    for (int i = b; i < e; ++i) s += a[i];
    const int inc = warp_incl_scan(s, lane);
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < T / 32; ++w) {
        const int v = wsum[w];
        before += w < warp ? v : 0;
        total += v;
    }
    int run = before + inc - s;
    for (int i = b; i < e; ++i) {
        int v = a[i];
        a[i] = run;
        run += v;
    }
    if (tid == 0) a[n] = total;
    __syncthreads();  // (also: nobody overwrites wsum for the next scan before everybody has read it)
}

__device__ __forceinline__ int quadrant(uint32_t xy, short4 b)
{
    // DivideNode :486-487, :515-529; coordinates are small integers, float compare == int compare
    const int x = xy & 0xFFFF, y = xy >> 16;
    const int mx = b.x + ((b.z - b.x + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
    const int my = b.y + ((b.w - b.y + 1) >> 1);
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);
}

__device__ __forceinline__ short4 child_box(short4 b, int q)
{
    const short mx = b.x + ((b.z - b.x + 1) >> 1), my = b.y + ((b.w - b.y + 1) >> 1);
    short4 c;
    c.x = (q & 1) ? mx : b.x;
    c.z = (q & 1) ? b.z : mx;
    c.y = (q & 2) ? my : b.y;
    c.w = (q & 2) ? b.w : my;
    return c;
}

struct __align__(8) OrbxCellRec {  // same layout as OrbxCell in fast.cu
    short x0, y0;
    uint8_t rw, rh, level, pad;
};

#define OCT_CELL_CHUNK 512  // cells gathered per round of step 0 (2 KB of static shared memory; KITTI's largest level has 429)

// SPILL: the node list lives in p.octSpill (global memory, L2-resident) instead of dynamic shared memory -- for per-level quotas
// above ~2600 nodes, which the reference accepts and the shared memory of one CTA cannot hold. Same code, generic addressing.
template <bool SPILL>
__global__ void __launch_bounds__(OCT_THREADS) orbx_octree_kernel(OrbxParams p, const OrbxCellRec* cells)
{
    extern __shared__ __align__(16) uint8_t smraw[];
    __shared__ int wsum[OCT_THREADS / 32 + 1];
    __shared__ int cellScan[OCT_CELL_CHUNK + 1];
    __shared__ int sh_m, sh_tstar, sh_nexp;

    const int level = blockIdx.x, img = blockIdx.y;
    const OrbxLevel& L = p.lv[level];
    const int cap = L.keepCap, N = L.quota;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // carve shared memory (largest alignment first)
    OctSmem S;
    uint8_t* ptr = SPILL ? p.octSpill + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * p.octSpillStride : smraw;
    S.box[0] = (short4*)ptr; ptr += sizeof(short4) * cap;
    S.box[1] = (short4*)ptr; ptr += sizeof(short4) * cap;
    S.cnt[0] = (int*)ptr; ptr += 4 * cap;
    S.cnt[1] = (int*)ptr; ptr += 4 * cap;
    S.childCnt = (int*)ptr; ptr += 16 * cap;
    S.rankOf = (int*)ptr; ptr += 4 * cap;
    S.byRank = (int*)ptr; ptr += 4 * cap;
    S.scan = (int*)ptr; ptr += 4 * (cap + 1);
    S.vlist = (int*)ptr; ptr += 4 * (cap + 1);
    int* tmp = (int*)ptr;            // sizes of the nodes to rank (phase 2, inside the passes) ...
    S.best = (unsigned*)ptr; ptr += 4 * cap;  // ... and, after the passes, the best response per node: never live together
    S.newPos = (uint16_t*)ptr; ptr += 8 * cap;
    S.movePos = (uint16_t*)ptr; ptr += 2 * cap;
    S.inV[0] = ptr; ptr += cap;
    S.inV[1] = ptr; ptr += cap;

    // ping-pong node arrays as plain pointers (indexing S.box[cur] with a run-time `cur` put the struct in local memory)
    short4 *boxC = S.box[0], *boxN = S.box[1];
    int *cntC = S.cnt[0], *cntN = S.cnt[1];
    uint8_t *inVC = S.inV[0], *inVN = S.inV[1];
    uint32_t* candXY = p.candXY + (size_t)img * p.candTotal + L.candBase;
    uint8_t* candRs = p.candRs + (size_t)img * p.candTotal + L.candBase;
    uint16_t* label = p.label + (size_t)img * p.candTotal + L.candBase;

    // ---- step 0: concatenate the cells' keypoints in cell order (:823-828) and bin into ini nodes (:571-575)
    const int nIni = L.nIni;
    const float hX = L.hX;
    for (int i = tid; i < nIni; i += OCT_THREADS) S.childCnt[i] = 0;
    __syncthreads();
    int n = 0;
    {
        const int* cc = p.cellCnt + (size_t)img * p.cellsTotal + L.cellBase;
        const uint32_t* slots = p.cellKp + (size_t)img * p.slotTotal + L.slotBase;
        const OrbxCellRec* ct = cells + L.cellBase;
        for (int c0 = 0; c0 < L.nCells; c0 += OCT_CELL_CHUNK) {
            const int nc = min(OCT_CELL_CHUNK, L.nCells - c0);
            for (int i = tid; i < nc; i += OCT_THREADS) cellScan[i] = cc[c0 + i];
            __syncthreads();
            block_excl_scan(cellScan, nc, wsum);
            // one thread per candidate (its cell by binary search over the scanned counts): every thread has an independent
            // load of its slot in flight; a warp per cell with a lane per slot kept ~7 lanes busy behind one L2 round trip per cell
            const int tot = cellScan[nc];
            for (int q = tid; q < tot; q += OCT_THREADS) {
                int lo = 0, hi = nc;  // cellScan[lo] <= q < cellScan[hi]
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (cellScan[mid] <= q) lo = mid; else hi = mid;
                }
                const int e = q - cellScan[lo];
                const OrbxCellRec cr = ct[c0 + lo];
                const int sx = cr.x0 - (ORBX_EDGE - 3), sy = cr.y0 - (ORBX_EDGE - 3);
                const uint32_t v = slots[(size_t)(c0 + lo) * L.slotCap + e];
                const int x = (v & 255) + sx, y = ((v >> 8) & 255) + sy;
                const int idx = n + q;
                candXY[idx] = (uint32_t)x | ((uint32_t)y << 16);
                candRs[idx] = (uint8_t)(v >> 16);
                int b = (int)orbx_fdiv((float)x, hX);
                b = min(b, nIni - 1);
                label[idx] = (uint16_t)b;
                atomicAdd(&S.childCnt[b], 1);
            }
            n += cellScan[nc];
            __syncthreads();
        }
    }
    if (tid == 0) {
        p.candCnt[img * p.nlevels + level] = n;
        int m = 0;
        for (int i = 0; i < nIni; ++i) {  // :555-588 non-empty ini nodes, in order
            const int k = S.childCnt[i];
            S.newPos[i] = (uint16_t)m;
            if (k > 0) {
                boxC[m] = make_short4((short)(int)orbx_fmul(hX, (float)i), 0, (short)(int)orbx_fmul(hX, (float)(i + 1)),
                                          (short)(L.h - 2 * (ORBX_EDGE - 3)));
                cntC[m] = k;
                inVC[m] = 0;
                ++m;
            }
        }
        sh_m = m;
    }
    __syncthreads();
    // Candidates of this thread (i = tid + k*OCT_THREADS) live in registers for the passes below: position and node label
    // are read and rewritten in every pass, and as global-memory traffic their latency was most of the kernel's time.
    // Levels with more candidates than fit (large images) take the same loops through global memory.
    const bool inReg = n <= OCT_THREADS * OCT_K;
    uint32_t rxy[OCT_K];
    int rlb[OCT_K];
#pragma unroll
    for (int k = 0; k < OCT_K; ++k) {
        const int i = tid + k * OCT_THREADS;
        rxy[k] = (inReg && i < n) ? candXY[i] : 0u;
        rlb[k] = (inReg && i < n) ? (int)S.newPos[label[i]] : 0;
    }
    if (!inReg)
        for (int i = tid; i < n; i += OCT_THREADS) label[i] = S.newPos[label[i]];
// BODY sees the candidate index i, its packed position XY and its label LB (assignable)
#define OCT_FOR_CANDIDATES(BODY)                                                            \
    if (inReg) {                                                                            \
        _Pragma("unroll") for (int k_ = 0; k_ < OCT_K; ++k_) {                              \
            const int i = tid + k_ * OCT_THREADS;                                           \
            if (i < n) { const uint32_t XY = rxy[k_]; int& LB = rlb[k_]; BODY }             \
        }                                                                                   \
    } else {                                                                                \
        for (int i = tid; i < n; i += OCT_THREADS) {                                        \
            const uint32_t XY = candXY[i]; int LB = label[i]; const int lb0_ = LB; BODY     \
            if (LB != lb0_) label[i] = (uint16_t)LB;                                        \
        }                                                                                   \
    }
    int m = sh_m, phase = 1;
    __syncthreads();

    // ---- passes
    while (true) {
        // (1) nodes to split and their creation order
        int Sn;
        if (phase == 1) {
            for (int i = tid; i < m; i += OCT_THREADS) S.scan[i] = cntC[i] > 1;
            __syncthreads();
            block_excl_scan(S.scan, m, wsum);
            Sn = S.scan[m];
            for (int i = tid; i < m; i += OCT_THREADS) {
                const bool f = cntC[i] > 1;
                S.rankOf[i] = f ? S.scan[i] : -1;
                if (f) S.byRank[S.scan[i]] = i;
            }
        } else {
            for (int i = tid; i < m; i += OCT_THREADS) S.scan[i] = inVC[i];
            __syncthreads();
            block_excl_scan(S.scan, m, wsum);
            Sn = S.scan[m];
            for (int i = tid; i < m; i += OCT_THREADS) {
                S.rankOf[i] = -1;
                if (inVC[i]) {
                    S.vlist[S.scan[i]] = i;
                    tmp[S.scan[i]] = cntC[i];
                }
            }
            __syncthreads();
            // sort descending by (size, creation index); creation index descends with list position
            for (int i = tid; i < Sn; i += OCT_THREADS) {
                const int ci = tmp[i];
                int r = 0;
                for (int j = 0; j < Sn; ++j) {
                    const int cj = tmp[j];
                    r += (cj > ci) || (cj == ci && j < i);
                }
                S.rankOf[S.vlist[i]] = r;
                S.byRank[r] = S.vlist[i];
            }
        }
        __syncthreads();
        if (Sn == 0) break;  // nothing can be split: list unchanged (:672 / :737)

        // (2) quadrant populations of the nodes that may split
        for (int i = tid; i < Sn; i += OCT_THREADS) {
            const int pos = S.byRank[i];
            *(int4*)&S.childCnt[pos * 4] = make_int4(0, 0, 0, 0);
        }
        __syncthreads();
        OCT_FOR_CANDIDATES(if (S.rankOf[LB] >= 0) atomicAdd(&S.childCnt[LB * 4 + quadrant(XY, boxC[LB])], 1);)
        __syncthreads();

        // (3) growth per node in creation order, stop index t*
        for (int r = tid; r < Sn; r += OCT_THREADS) {
            const int4 c = *(int4*)&S.childCnt[S.byRank[r] * 4];
            S.scan[r] = (c.x > 0) + (c.y > 0) + (c.z > 0) + (c.w > 0) - 1;
        }
        if (tid == 0) { sh_tstar = Sn; sh_nexp = 0; }
        __syncthreads();
        block_excl_scan(S.scan, Sn, wsum);
        if (phase == 2) {
            for (int r = tid; r < Sn; r += OCT_THREADS)
                if (m + S.scan[r + 1] >= N) atomicMin(&sh_tstar, r + 1);  // scan[r+1] = inclusive growth (:733)
            __syncthreads();
        }
        const int tstar = sh_tstar;
        const int T = S.scan[tstar] + tstar;

        // (4) new list: reversed children first, then the unsplit nodes in their old order
        for (int r = tid; r < tstar; r += OCT_THREADS) {
            const int pos = S.byRank[r];
            const int4 c = *(int4*)&S.childCnt[pos * 4];
            const int cc[4] = {c.x, c.y, c.z, c.w};
            const short4 pb = boxC[pos];
            int k = S.scan[r] + r, multi = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (cc[q] > 0) {
                    const int np = T - 1 - k;
                    S.newPos[pos * 4 + q] = (uint16_t)np;
                    boxN[np] = child_box(pb, q);
                    cntN[np] = cc[q];
                    inVN[np] = cc[q] > 1;
                    multi += cc[q] > 1;
                    ++k;
                }
            }
            if (multi) atomicAdd(&sh_nexp, multi);
        }
        for (int i = tid; i < m; i += OCT_THREADS) {
            const int r = S.rankOf[i];
            S.vlist[i] = !(r >= 0 && r < tstar);
        }
        __syncthreads();
        block_excl_scan(S.vlist, m, wsum);
        for (int i = tid; i < m; i += OCT_THREADS) {
            const int r = S.rankOf[i];
            if (!(r >= 0 && r < tstar)) {
                const int np = T + S.vlist[i];
                S.movePos[i] = (uint16_t)np;
                boxN[np] = boxC[i];
                cntN[np] = cntC[i];
                inVN[np] = 0;
            }
        }
        __syncthreads();
        // (5) relabel the candidates
        OCT_FOR_CANDIDATES(const int r = S.rankOf[LB];
                           LB = (r >= 0 && r < tstar) ? S.newPos[LB * 4 + quadrant(XY, boxC[LB])] : S.movePos[LB];)
        const int prev = m;
        m = T + (m - tstar);
        const int nexp = sh_nexp;
        { short4* tb = boxC; boxC = boxN; boxN = tb; int* tc = cntC; cntC = cntN; cntN = tc; uint8_t* tv = inVC; inVC = inVN; inVN = tv; }
        __syncthreads();
        if (m >= N || m == prev) break;                    // :672, :737
        if (phase == 1 && m + 3 * nexp > N) phase = 2;      // :676
    }

    // ---- best response per node, first candidate wins ties (:744-763); output in list order
    for (int i = tid; i < m; i += OCT_THREADS) S.best[i] = 0;
    __syncthreads();
    OCT_FOR_CANDIDATES((void)XY; atomicMax(&S.best[LB], ((unsigned)candRs[i] << 24) | (0xFFFFFFu - (unsigned)i));)
#undef OCT_FOR_CANDIDATES
    __syncthreads();
    uint2* keep = p.keep + (size_t)img * p.keepTotal + L.keepBase;
    for (int i = tid; i < m; i += OCT_THREADS) {
        const unsigned b = S.best[i];
        const int idx = (int)(0xFFFFFFu - (b & 0xFFFFFFu));
        keep[i] = make_uint2(candXY[idx], b >> 24);
    }
    if (tid == 0) p.keepCnt[img * p.nlevels + level] = m;
}

static size_t octree_smem_bytes(int cap) { return (size_t)cap * 80 + 64; }
size_t orbx_octree_list_bytes(int cap) { return octree_smem_bytes(cap); }
// dynamic + static shared memory of one quadtree CTA for a per-level list capacity (Engine::plan rejects what the SM cannot hold)
size_t orbx_octree_smem_total(int cap) { return octree_smem_bytes(cap) + sizeof(int) * (OCT_THREADS / 32 + 1 + OCT_CELL_CHUNK + 1 + 3) + 64; }

void orbx_launch_octree(const OrbxParams& p, const void* cellTable, cudaStream_t s)
{
    int cap = 0;
    for (int l = 0; l < p.nlevels; ++l) cap = p.lv[l].keepCap > cap ? p.lv[l].keepCap : cap;
    const size_t smem = octree_smem_bytes(cap);
    dim3 grid(p.nlevels, p.imgs);
    if (p.octSpillStride) {  // Engine::plan chose the global-memory list (the quota does not fit a CTA's shared memory)
        orbx_octree_kernel<true><<<grid, OCT_THREADS, 0, s>>>(p, (const OrbxCellRec*)cellTable);
        return;
    }
    if (smem > 32 * 1024)  // static + dynamic above 48 KB needs the opt-in; per-device attribute, set on every such launch
        if (cudaFuncSetAttribute(orbx_octree_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return;  // the error stays pending for the caller's cudaGetLastError; no launch on unwritten state
    orbx_octree_kernel<false><<<grid, OCT_THREADS, smem, s>>>(p, (const OrbxCellRec*)cellTable);
}
