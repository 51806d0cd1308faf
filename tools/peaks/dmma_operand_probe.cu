// Probe: DMMA.8x8x4 throughput of one SM when the A / B fragments are (0) loop-invariant registers, (1) fresh from shared
// memory for every instruction, (2) fresh A, B shared by two consecutive instructions, (3) fresh registers (no LDS:
// values rotated in registers).  8 independent accumulator chains per warp, WARPS warps on ONE SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_operand_probe dmma_operand_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void probe(double* out, long long* clk, int iters) {
    __shared__ double sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = 1e-3 * (i % 7);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    double c[8][2];
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = 0.0;
    double ra[8], rb[8];
    for (int i = 0; i < 8; i++) { ra[i] = sm[lane + 32 * i]; rb[i] = sm[lane + 32 * (i + 8)]; }
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        const double* p = sm + (it & 7) * 32 * 8 + lane;      // a different slice every iteration
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 8; i++) dmma(c[i], ra[0], rb[0]);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; i++) dmma(c[i], p[32 * i], p[32 * i + 16 * 64]);
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; i++) dmma(c[i], p[32 * i], p[32 * (i >> 1) + 16 * 64]);
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) dmma(c[i], ra[i], rb[i]);
        }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) clk[0] = t1 - t0;
}

int main() {
    double* out; long long* clk;
    cudaMalloc(&out, 1024 * 8); cudaMalloc(&clk, 8);
    const int iters = 2000;
    for (int warps = 4; warps <= 16; warps *= 2)
        for (int mode = 0; mode < 4; mode++) {
            if (mode == 0) probe<0><<<1, 32 * warps>>>(out, clk, iters);
            if (mode == 1) probe<1><<<1, 32 * warps>>>(out, clk, iters);
            if (mode == 2) probe<2><<<1, 32 * warps>>>(out, clk, iters);
            if (mode == 3) probe<3><<<1, 32 * warps>>>(out, clk, iters);
            cudaDeviceSynchronize();
            long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
            printf("warps %2d mode %d: %.1f clocks per DMMA per sub-partition\n", warps, mode, (double)h / (iters * 8.0 * warps / 4.0));
        }
    return 0;
}
