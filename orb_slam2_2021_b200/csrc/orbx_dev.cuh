// orbx_dev.cuh — device-side data layout shared by all kernels and the host engine.
//
// HBM layout for a chunk of `imgs` images (stereo: image 2f = left of frame f, 2f+1 = right):
//   img0   [imgs][H*pitch0]       u8  level 0 = the uploaded image, row pitch = align16(W) (TMA needs 16-B strides)
//   pyr    [imgs][pyrStride]      u8  levels 1..L-1 packed, row pitch = align16(w_l)
//   blur   [imgs][blurStride]     u8  7x7 Gaussian of levels 0..L-1, row pitch = align16(w_l)
//   cellCnt[imgs][cellsTotal]     i32 FAST keypoints per 30-px cell
//   cellKp [imgs][slotTotal]      u32 per-cell slots (x | y<<8 | score<<16), ROI coordinates
//   candXY [imgs][candTotal]      u32 level candidates in reference order (x | y<<16, rel. minBorder)
//   candRs [imgs][candTotal]      u8  FAST score
//   label  [imgs][candTotal]      u16 quadtree node (list position) of each candidate
//   keep   [imgs][keepTotal]      u32x2 kept keypoints per level in final list order (xy, score)
//   keepCnt[imgs][L]              i32
//   outKp  [imgs][maxKp] 28 B, outDesc [imgs][maxKp][32], outCnt [imgs]
//   uRight/depth [frames][maxKp]  f32, matchSad [frames][maxKp] i32 (scratch for the median cut)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "orbx_math.h"

#define ORBX_MAX_LEVELS 12
#ifndef ORBX_RS_TMA_ROWS
#define ORBX_RS_TMA_ROWS 16  // destination rows per thread of the TMA-staged resize (tile = 128 columns x 4 x this many rows); measured: 8 -> 0.476, 16 -> 0.445, 32 -> 0.459 ms per 512 KITTI images
#endif
#define ORBX_EDGE 19
#define ORBX_HALF_PATCH 15
#define ORBX_BLUR_TW 128  // blur output tile (a 160 x 70 TMA box: 16 px left and right, 3 px above/below)
#define ORBX_BLUR_TH 64
#define ORBX_DESC_BOX0_W 48  // describe: TMA box of the radius-15 disc ((cx-15)&15 + 31 <= 46 columns)
#define ORBX_DESC_BOX0_H 31
#define ORBX_DESC_BOX1_W 64  // describe: TMA box of the rotated pattern's reach, radius 18 ((cx-18)&15 + 37 <= 52 columns)
#define ORBX_DESC_BOX1_H 37

struct OrbxLevel {
    int w, h, pitch;      // level size; row pitch = align16(w) (level 0 lives in img0, levels >= 1 in pyr)
    int bpitch;           // row pitch in blur
    unsigned off;         // byte offset inside one image's pyr block (levels >= 1)
    unsigned boff;        // byte offset inside one image's blur block
    int nCols, nRows, wCell, hCell;  // FAST cell grid (src/ORBextractor.cc:784-790)
    int cellBase, nCells;            // index of the first cell of this level in cellCnt
    int slotCap;                     // per-cell slot capacity = ceil(wCell/2)*ceil(hCell/2)
    unsigned slotBase;               // first slot of this level in cellKp
    unsigned candBase;               // first candidate of this level in candXY/candRs/label
    int candCap;
    int quota, nIni, keepCap, keepBase;
    float scale, invScale;
    int patchSize;                   // (int)(31*scale) as float size of the keypoint
    float hX;                        // (float)(maxBX-minBX)/nIni
};

struct OrbxParams {
    int nlevels, iniTh, minTh;
    int rows, cols;
    int imgs;  // images in this launch
    int imgBase;  // index of image 0 of this launch inside the engine (buffer pointers are already advanced; TMA z needs it)
    OrbxLevel lv[ORBX_MAX_LEVELS];
    // per-image strides (elements of the respective type)
    size_t img0Stride, pyrStride, blurStride;
    int cellsTotal;
    unsigned slotTotal, candTotal;
    int keepTotal, maxKp;
    // buffers
    uint8_t* img0;
    uint8_t* pyr;
    uint8_t* blur;
    int* cellCnt;
    uint32_t* cellKp;
    uint32_t* candXY;
    uint8_t* candRs;
    uint16_t* label;
    uint2* keep;
    int* keepCnt;
    int* candCnt;  // [imgs][L] number of candidates per level (written by the quadtree kernel)
    int* keepOfs;  // [imgs][L] exclusive prefix of keepCnt = first output index of a level (written by the count kernel)
    int* chainState;   // counters of the one-launch pyramid (orbx_launch_pyramid_chain): chainInts per image, advanced by slice()
    int chainInts;
    uint8_t* octSpill;      // quadtree node lists in global memory, octSpillStride bytes per (image, level); only for quotas whose
    size_t octSpillStride;  // list does not fit the shared memory of a CTA (0 otherwise), advanced by slice()
    void* outKp;   // orbx_keypoint[imgs][maxKp]
    uint8_t* outDesc;
    int* outCnt;
};

__device__ __forceinline__ const uint8_t* orbx_level_ptr(const OrbxParams& p, int img, int l, int* pitch)
{
    if (l == 0) {
        *pitch = p.lv[0].pitch;
        return p.img0 + (size_t)img * p.img0Stride;
    }
    *pitch = p.lv[l].pitch;
    return p.pyr + (size_t)img * p.pyrStride + p.lv[l].off;
}

// Stereo stage parameters: left/right images may live in two different engines (two handles).
struct OrbxStereoParams {
    int frames, nlevels, rows0;
    int maxKp;
    float mbf, mb;
    // per-side image geometry comes from the side's OrbxParams; image index of frame f on side s is
    // f*imgStep + imgOfs[s] (batch engine: step 2, ofs {0,1}; two handles: step 1, ofs {0,0}).
    int imgStep, imgOfsL, imgOfsR;
    float* uRight;  // [frames][maxKp]
    float* depth;
    int* sad;       // [frames][maxKp] best SAD of accepted matches, -1 otherwise
    int* nMatched;  // [frames]
    // right keypoints of each frame counting-sorted by (int)y (the row-band table of src/Frame.cc:532-549 in
    // data-parallel form): rsorted[3j] = {x bits, minr | maxr<<16, octave, iR}, rsorted[3j+1..3j+2] = descriptor; bin b = binStart[b]..binStart[b+1]
    uint4* rsorted;  // [frames][maxKp][3]
    int* binStart;   // [frames][rows0 + 1]
    int binR;        // a left keypoint on row v can only match right keypoints binned in [v-binR, v+binR]
};

// Launch wrappers (one per kernel group), defined in the .cu files.
// tight [frame][rows][cols] images of one side -> pitched level 0 of image frame*per + side
// input.cu: fused rectification (remap) + crop + gray conversion into level 0
struct OrbxStageArgs {
    const uint8_t* src; size_t srcImgStride; int srcRows, srcCols, srcStride;  // raw frames, bytes
    const float* map1; const float* map2; int mapStride;                      // floats per map row; NULL = identity
    int cropTop, rows, cols;                                                   // output geometry
    uint8_t* dst; size_t dstImgStride; int dstPitch;
    int imgStep, imgOfs;  // output image index = imgStep * blockIdx.z + imgOfs (stereo engines interleave L/R)
    int rgbOrder;
};

void orbx_launch_stage(const OrbxStageArgs& a, int channels, int nImgs, cudaStream_t s);
void orbx_launch_ingest(const OrbxParams& p, const uint8_t* src, int nFrames, int per, int side, cudaStream_t s);
void orbx_launch_pyramid(const OrbxParams& p, const int* const* tabs, cudaStream_t s);
void orbx_launch_pyramid_tma(const OrbxParams& p, const int* const* tabs, const void* srcMaps, const int* boxW, const int* boxH, cudaStream_t s);
int orbx_pyramid_chain_state_ints(const OrbxParams& p);
void orbx_launch_pyramid_chain(const OrbxParams& p, const int* const* tabs, int* state, cudaStream_t s);
void orbx_fast_geometry(const OrbxParams& p, int* boxW, int* rhMax);  // TMA box of the FAST kernel
void orbx_launch_fast(const OrbxParams& p, const void* cellTable, const void* levelMaps, cudaStream_t s);
void orbx_launch_octree(const OrbxParams& p, const void* cellTable, cudaStream_t s);
size_t orbx_octree_smem_total(int cap);  // shared memory of one quadtree CTA with a per-level list capacity of cap
size_t orbx_octree_list_bytes(int cap);  // bytes of one node list (the dynamic shared memory, or the spill slot in global memory)
void orbx_launch_blur(const OrbxParams& p, const void* tileTable, int nTiles, const void* levelMaps, cudaStream_t s);
void orbx_launch_describe(const OrbxParams& p, const void* mapsPlain, const void* mapsBlur, cudaStream_t s);
void orbx_launch_stereo(const OrbxParams& pl, const OrbxParams& pr, const OrbxStereoParams& sp, cudaStream_t s);
