// pipes.cu — issue-rate microbenchmark for the integer/half min-max instructions the FAST kernel could use.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on a B200.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 4096
#define ILP 8

template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned* out, unsigned seed, long long* cyc)
{
    unsigned a[ILP], b[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) { a[i] = seed * (threadIdx.x + i + 1); b[i] = seed ^ (0x01010101u * (i + 3)); }
    unsigned c = seed | 0x00100010u;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (MODE == 0) a[i] = __vimin3_u16x2(a[i], b[i], c);
            if (MODE == 1) { __half2 x = *(__half2*)&a[i], y = *(__half2*)&b[i]; x = __hmin2(x, y); a[i] = *(unsigned*)&x; }
            if (MODE == 2) { if (i & 1) a[i] = __vimax3_u16x2(a[i], b[i], c); else { __half2 x = *(__half2*)&a[i], y = *(__half2*)&b[i]; x = __hmax2(x, y); a[i] = *(unsigned*)&x; } }
            if (MODE == 3) a[i] = a[i] * 3u + b[i];                           // IMAD
            if (MODE == 4) a[i] = __byte_perm(a[i], b[i], 0x5140);            // PRMT
            if (MODE == 5) { if (i & 1) a[i] = __vimin3_u16x2(a[i], b[i], c); else a[i] = a[i] * 3u + b[i]; }
            if (MODE == 6) { __half2 x = *(__half2*)&a[i], y = *(__half2*)&b[i]; x = __hsub2(x, y); a[i] = *(unsigned*)&x; }
            if (MODE == 7) { if (i & 1) a[i] = __vimin3_u16x2(a[i], b[i], c); else { __half2 x = *(__half2*)&a[i], y = *(__half2*)&b[i]; x = __hsub2(x, y); a[i] = *(unsigned*)&x; } }
            if (MODE == 8) a[i] = min(a[i], b[i]);                            // IMNMX 32-bit
            if (MODE == 9) a[i] = __vmins2(a[i], b[i]);                  // 2-input packed
            if (MODE == 10) { float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(__uint_as_float(a[i])), "f"(__uint_as_float(b[i])), "f"(__uint_as_float(c))); a[i] = __float_as_uint(r); }
            if (MODE == 11) { if (i % 3 == 0) a[i] = __vimin3_u16x2(a[i], b[i], c); else { __half2 x = *(__half2*)&a[i], y = *(__half2*)&b[i]; x = __hmin2(x, y); a[i] = *(unsigned*)&x; } }
            if (MODE == 12) a[i] = __vabsdiffu4(a[i], b[i]);
            if (MODE == 13) a[i] = __vmaxu4(a[i], b[i]);
            if (MODE == 14) a[i] = __vcmpgtu4(a[i], b[i]);
            if (MODE == 15) a[i] = __vsub2(a[i], b[i]);
            if (MODE == 16) a[i] = __dp4a(b[i], 0x03020100u, a[i]);
            if (MODE == 17) { if (i & 1) a[i] = __vimin3_u16x2(a[i], b[i], c); else a[i] = __dp4a(b[i], 0x03020100u, a[i]); }
            if (MODE == 18) { if (i & 1) a[i] = a[i] * 3u + b[i]; else a[i] = __dp4a(b[i], 0x03020100u, a[i]); }
            if (MODE == 19) a[i] = __float_as_uint(__uint_as_float(a[i]) + 12582912.0f);
            if (MODE == 20) a[i] = (unsigned)__float2int_rn(__uint_as_float(a[i] & 0x3fffffffu));
            if (MODE == 21) a[i] = __vsadu4(a[i], b[i]);
            if (MODE == 22) a[i] = __funnelshift_r(a[i], b[i], c);
        }
    }
    long long t1 = clock64();
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, unsigned* out, long long* dcyc)
{
    // 8 warps per SMSP-set: 256 threads x 4 CTAs per SM
    k<MODE><<<148 * 4, 256>>>(out, 12345u, dcyc);
    cudaDeviceSynchronize();
    k<MODE><<<148 * 4, 256>>>(out, 12345u, dcyc);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, dcyc, 8, cudaMemcpyDeviceToHost);
    // per SMSP: 8 warps (32 warps per SM / 4) each issuing ITERS*ILP instr
    double per = (double)c / ((double)ITERS * ILP * 8);
    printf("%-28s cycles/warp-instr/SMSP = %.3f  (%s)\n", name, per, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    unsigned* out; long long* dcyc;
    cudaMalloc(&out, 148 * 4 * 256 * 4); cudaMalloc(&dcyc, 8);
    run<0>("vimin3_u16x2", out, dcyc);
    run<1>("hmin2 (2-in)", out, dcyc);
    run<2>("vimax3_u16x2 + hmax2 mix", out, dcyc);
    run<3>("imad", out, dcyc);
    run<4>("prmt", out, dcyc);
    run<5>("vimin3_u16x2 + imad mix", out, dcyc);
    run<6>("hsub2", out, dcyc);
    run<7>("vimin3_u16x2 + hsub2 mix", out, dcyc);
    run<8>("imnmx u32", out, dcyc);
    run<9>("vimin_s16x2 (2-in)", out, dcyc);
    run<10>("min.f32 3-in", out, dcyc);
    run<11>("vimin3_u16x2 : hmin2 = 1:2", out, dcyc);
    run<12>("vabsdiffu4", out, dcyc);
    run<13>("vmaxu4", out, dcyc);
    run<14>("vcmpgtu4", out, dcyc);
    run<15>("vsub2", out, dcyc);
    run<16>("dp4a", out, dcyc);
    run<17>("dp4a + vimin3 mix", out, dcyc);
    run<18>("dp4a + imad mix", out, dcyc);
    run<19>("fadd", out, dcyc);
    run<20>("f2i.rn", out, dcyc);
    run<21>("vsadu4", out, dcyc);
    run<22>("funnelshift", out, dcyc);
    return 0;
}
