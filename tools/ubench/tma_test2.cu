#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../orb_slam2_2021_b200/csrc/orbx_tma.cuh"
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap cmap, int rank, int bytes, int x, int y, int z, uint8_t* out)
{
#ifdef STATIC_SMEM
    __shared__ __align__(128) uint8_t src[70 * 128];
#else
    extern __shared__ __align__(1024) uint8_t src[];
#endif
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { orbx_mbar_init(&bar, 1); orbx_fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        orbx_mbar_expect_tx(&bar, bytes);
        if (rank == 3) orbx_tma_load_3d(src, &cmap, x, y, z, &bar);
        else asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(orbx_smem_addr(src)), "l"(&cmap), "r"(orbx_smem_addr(&bar)), "r"(x), "r"(y) : "memory");
    }
    orbx_mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = src[i];
}
int main(int argc, char** argv)
{
    const int rank = atoi(argv[1]), bw = atoi(argv[2]), bh = atoi(argv[3]), l2 = atoi(argv[4]);
    const int w = 1241, h = 376, pitch = 1248, imgs = 3;
    std::vector<uint8_t> host((size_t)pitch * h * imgs);
    for (size_t i = 0; i < host.size(); ++i) host[i] = (uint8_t)(i * 7 + (i >> 8));
    uint8_t *d, *out;
    cudaMalloc(&d, host.size()); cudaMalloc(&out, 65536);
    cudaMemcpy(d, host.data(), host.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap m;
    const cuuint64_t dims[3] = {w, h, imgs}; const cuuint64_t strides[2] = {pitch, (cuuint64_t)pitch * h};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}; const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const int bytes = bw * bh;
    k<<<1, 256, bytes>>>(m, rank, bytes, CX, CY, 1, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("rank %d box %dx%d l2 %d: encode %d, run: %s\n", rank, bw, bh, l2, (int)r, cudaGetErrorString(e));
    if (e) return 1;
    std::vector<uint8_t> res(bytes);
    cudaMemcpy(res.data(), out, bytes, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r2 = 0; r2 < bh; ++r2) for (int c = 0; c < bw; ++c) {
        const int gx = CX + c, gy = CY + r2, z = rank == 3 ? 1 : 0;
        bad += res[r2 * bw + c] != ((gx >= 0 && gx < w && gy >= 0 && gy < h) ? host[((size_t)z * h + gy) * pitch + gx] : 0);
    }
    printf("  mismatches %d\n", bad);
    return 0;
}
