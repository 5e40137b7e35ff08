// h2d2d.cu — pinned H2D: one flat copy vs a 2-D copy of 1241-byte rows into a 1248-byte pitch (KITTI level 0).
// Build: nvcc -O3 -o h2d2d h2d2d.cu ; run on a B200.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
int main()
{
    const size_t W = 1241, H = 376, P = 1248, N = 512;
    uint8_t *h, *d;
    cudaMallocHost(&h, W * H * N);
    cudaMalloc(&d, P * H * N);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 4; ++mode) {
        float best = 1e9f;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0, s);
            if (mode == 0) cudaMemcpyAsync(d, h, W * H * N, cudaMemcpyHostToDevice, s);
            if (mode == 1) cudaMemcpy2DAsync(d, P, h, W, W, H * N, cudaMemcpyHostToDevice, s);
            if (mode == 2) for (int c = 0; c < 8; ++c) cudaMemcpy2DAsync(d + c * (N / 8) * H * P, P, h + c * (N / 8) * H * W, W, W, H * N / 8, cudaMemcpyHostToDevice, s);
            if (mode == 3) for (int c = 0; c < 8; ++c) cudaMemcpyAsync(d + c * (N / 8) * H * W, h + c * (N / 8) * H * W, W * H * N / 8, cudaMemcpyHostToDevice, s);
            cudaEventRecord(e1, s);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const char* names[] = {"flat 1 copy", "2-D 1 copy", "2-D 8 copies", "flat 8 copies"};
        printf("%-14s %.3f ms  %.1f GB/s  (%s)\n", names[mode], best, W * H * N / best / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
