// tma_test.cu — minimal check of the 3-D u8 TMA box load used by blur.cu (descriptor in global memory vs __grid_constant__).
#include <cstdio>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../orb_slam2_2021_b200/csrc/orbx_tma.cuh"

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__global__ void kstage(int stage, const uint8_t* g, uint8_t* out)
{
    __shared__ __align__(128) uint8_t src[70 * 128];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { orbx_mbar_init(&bar, 1); if (stage >= 1) orbx_fence_mbar_init(); }
    __syncthreads();
    if (stage >= 2) {
        if (threadIdx.x == 0) {
            orbx_mbar_expect_tx(&bar, 1024);
            if (stage >= 3)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(orbx_smem_addr(src)), "l"(g), "r"(1024), "r"(orbx_smem_addr(&bar)) : "memory");
        }
        if (stage >= 3) orbx_mbar_wait(&bar, 0);
    }
    out[threadIdx.x] = src[threadIdx.x];
}

template <bool GC>
__global__ void k(const CUtensorMap* gmap, const __grid_constant__ CUtensorMap cmap, int x, int y, int z, uint8_t* out)
{
    __shared__ __align__(128) uint8_t src[70 * 128];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { orbx_mbar_init(&bar, 1); orbx_fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        orbx_mbar_expect_tx(&bar, 70 * 128);
        orbx_tma_load_3d(src, GC ? &cmap : gmap, x, y, z, &bar);
    }
    orbx_mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < 70 * 128; i += blockDim.x) out[i] = src[i];
}

int main(int argc, char** argv)
{
    const int w = 1241, h = 376, pitch = 1248, imgs = 3;
    std::vector<uint8_t> host((size_t)pitch * h * imgs);
    for (size_t i = 0; i < host.size(); ++i) host[i] = (uint8_t)(i * 7 + (i >> 8));
    uint8_t *d, *out;
    cudaMalloc(&d, host.size()); cudaMalloc(&out, 70 * 128);
    cudaMemcpy(d, host.data(), host.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap m;
    const cuuint64_t dims[3] = {w, h, imgs}; const cuuint64_t strides[2] = {pitch, (cuuint64_t)pitch * h};
    const cuuint32_t box[3] = {128, 70, 1}; const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    CUtensorMap* dm; cudaMalloc(&dm, sizeof(m)); cudaMemcpy(dm, &m, sizeof(m), cudaMemcpyHostToDevice);
    if (argc > 1) {
        const int stage = atoi(argv[1]);
        kstage<<<1, 256>>>(stage, d, out);
        printf("stage %d: %s\n", stage, cudaGetErrorString(cudaDeviceSynchronize()));
        return 0;
    }
    std::vector<uint8_t> res(70 * 128);
    for (int variant = 0; variant < 2; ++variant) {
        const int x = 116, y = 61, z = 1;
        if (variant == 0) k<true><<<1, 256>>>(dm, m, x, y, z, out); else k<false><<<1, 256>>>(dm, m, x, y, z, out);
        cudaError_t e = cudaDeviceSynchronize();
        printf("variant %s: %s\n", variant == 0 ? "grid_constant" : "global", cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        cudaMemcpy(res.data(), out, res.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r2 = 0; r2 < 70; ++r2) for (int c = 0; c < 128; ++c) {
            const int gx = x + c, gy = y + r2;
            const uint8_t want = (gx >= 0 && gx < w && gy >= 0 && gy < h) ? host[((size_t)z * h + gy) * pitch + gx] : 0;
            bad += res[r2 * 128 + c] != want;
        }
        printf("  mismatches: %d\n", bad);
    }
    // negative coordinates
    k<true><<<1, 256>>>(dm, m, -4, -3, 0, out);
    printf("negative coords: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    cudaMemcpy(res.data(), out, res.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r2 = 0; r2 < 70; ++r2) for (int c = 0; c < 128; ++c) {
        const int gx = -4 + c, gy = -3 + r2;
        const uint8_t want = (gx >= 0 && gx < w && gy >= 0 && gy < h) ? host[((size_t)0 * h + gy) * pitch + gx] : 0;
        bad += res[r2 * 128 + c] != want;
    }
    printf("  mismatches: %d\n", bad);
    return 0;
}
