// Micro-benchmarks for the pipes the GP hot path is bounded by on B200 (sm_100a):
//   FP64 DFMA (SIMT), FP64 DMMA m8n8k4 (mma.sync; SASS DMMA.8x8x4), FP32 FFMA, MUFU ex2.
// Prints one JSON object.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peaks peaks.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); return 1;}}while(0)

template<int ILP>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double s) {
    double acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x * 1e-9 + i;
    double m = 1.0 + s, a = s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) acc[i] = fma(acc[i], m, a);
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += acc[i];
    if (r == 12345.678) out[0] = r;
}

template<int ILP>
__global__ void __launch_bounds__(256) k_ffma(float* out, int iters, float s) {
    float acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x * 1e-9f + i;
    float m = 1.0f + s, a = s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) acc[i] = fmaf(acc[i], m, a);
    }
    float r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += acc[i];
    if (r == 12345.678f) out[0] = r;
}

template<int ILP>
__global__ void __launch_bounds__(256) k_ex2(float* out, int iters, float s) {
    float acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x * 1e-3f + i * 0.01f + s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) { float t; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(acc[i])); acc[i] = t; }
    }
    float r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += acc[i];
    if (r == 12345.678f) out[0] = r;
}

template<int ILP>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters, double s) {
    double c[ILP][2];
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = 0; c[i][1] = 0; }
    double a = s + threadIdx.x * 1e-9, b = s * 0.5;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) r += c[i][0] + c[i][1];
    if (r == 12345.678) out[0] = r;
}

// DMMA and DFMA issued from the same warps: do the two FP64 paths overlap?
template<int NM, int NF>
__global__ void __launch_bounds__(256) k_mix(double* out, int iters, double s) {
    double c[NM][2], f[NF];
#pragma unroll
    for (int i = 0; i < NM; i++) { c[i][0] = 0; c[i][1] = 0; }
#pragma unroll
    for (int i = 0; i < NF; i++) f[i] = threadIdx.x * 1e-9 + i;
    double a = s + threadIdx.x * 1e-9, b = s * 0.5, m = 1.0 + s;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NM; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
#pragma unroll
        for (int i = 0; i < NF; i++) f[i] = fma(f[i], m, s);
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < NM; i++) r += c[i][0] + c[i][1];
#pragma unroll
    for (int i = 0; i < NF; i++) r += f[i];
    if (r == 12345.678) out[0] = r;
}

// numerical check of the m8n8k4 fragment layout: C(8x8) = A(8x4) * B(4x8)
__global__ void k_dmma_check(const double* A, const double* B, double* C) {
    int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    double a = A[g * 4 + t];       // A row-major 8x4: (row g, k t)
    double b = B[t * 8 + g];       // B row-major 4x8: (k t, col g)
    double c0 = 0, c1 = 0;
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
    C[g * 8 + 2 * t] = c0; C[g * 8 + 2 * t + 1] = c1;
}

template<typename F>
float timeit(F f, int reps) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    double* dout; CK(cudaMalloc(&dout, 1024));
    float* fout = (float*)dout;
    printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
    const int iters = 20000;
    for (int bps = 1; bps <= 8; bps *= 2) {
        int grid = sms * bps;
        float ms = timeit([&] { k_dfma<8><<<grid, 256>>>(dout, iters, 1e-12); }, 5);
        double tf = 2.0 * 8 * iters * 256.0 * grid / (ms * 1e-3) / 1e12;
        printf(", \"dfma_tflops_bps%d\": %.2f", bps, tf);
    }
    for (int bps = 1; bps <= 8; bps *= 2) {
        int grid = sms * bps;
        float ms = timeit([&] { k_dmma<8><<<grid, 256>>>(dout, iters / 4, 1e-12); }, 5);
        double tf = 2.0 * 256 * 8 * (iters / 4) * 8.0 * grid / (ms * 1e-3) / 1e12;   // 256 MAC per warp-mma, 8 warps
        printf(", \"dmma_tflops_bps%d\": %.2f", bps, tf);
    }
    {
        int grid = sms * 4;
        float ms = timeit([&] { k_dmma<2><<<grid, 256>>>(dout, iters, 1e-12); }, 5);
        double tf = 2.0 * 256 * 2 * iters * 8.0 * grid / (ms * 1e-3) / 1e12;
        printf(", \"dmma_ilp2_tflops\": %.2f", tf);
        ms = timeit([&] { k_dmma<4><<<grid, 256>>>(dout, iters / 2, 1e-12); }, 5);
        tf = 2.0 * 256 * 4 * (iters / 2) * 8.0 * grid / (ms * 1e-3) / 1e12;
        printf(", \"dmma_ilp4_tflops\": %.2f", tf);
        // single warp, dependent chain: latency per DMMA in ns
        ms = timeit([&] { k_dmma<1><<<1, 32>>>(dout, iters, 1e-12); }, 3);
        printf(", \"dmma_dep_latency_ns\": %.2f", ms * 1e6 / iters);
        ms = timeit([&] { k_dfma<1><<<1, 32>>>(dout, iters, 1e-12); }, 3);
        printf(", \"dfma_dep_latency_ns\": %.2f", ms * 1e6 / iters);
    }
    for (int bps = 2; bps <= 8; bps *= 2) {
        int grid = sms * bps;
        float ms = timeit([&] { k_ffma<8><<<grid, 256>>>(fout, iters * 4, 1e-7f); }, 5);
        double tf = 2.0 * 8 * iters * 4 * 256.0 * grid / (ms * 1e-3) / 1e12;
        printf(", \"ffma_tflops_bps%d\": %.2f", bps, tf);
    }
    {
        int grid = sms * 8;
        float ms = timeit([&] { k_ex2<8><<<grid, 256>>>(fout, iters, -0.5f); }, 5);
        double g = 8.0 * iters * 256.0 * grid / (ms * 1e-3) / 1e12;
        printf(", \"ex2_tera_per_s\": %.3f", g);
    }
    {   // mixed DMMA + DFMA: NM DMMA (512 flop/lane-group... 2*256 per warp) and NF DFMA (2*32 per warp) per iteration
        int grid = sms * 4;
        float ms = timeit([&] { k_mix<8, 8><<<grid, 256>>>(dout, iters / 4, 1e-12); }, 5);
        double n = (double)(iters / 4) * 8.0 * grid;   // warp-iterations
        printf(", \"mix8_8_dmma_tflops\": %.2f, \"mix8_8_dfma_tflops\": %.2f", 2.0 * 256 * 8 * n / (ms * 1e-3) / 1e12, 2.0 * 32 * 8 * n / (ms * 1e-3) / 1e12);
        ms = timeit([&] { k_mix<8, 32><<<grid, 256>>>(dout, iters / 4, 1e-12); }, 5);
        printf(", \"mix8_32_dmma_tflops\": %.2f, \"mix8_32_dfma_tflops\": %.2f", 2.0 * 256 * 8 * n / (ms * 1e-3) / 1e12, 2.0 * 32 * 32 * n / (ms * 1e-3) / 1e12);
        ms = timeit([&] { k_mix<4, 32><<<grid, 256>>>(dout, iters / 4, 1e-12); }, 5);
        printf(", \"mix4_32_dmma_tflops\": %.2f, \"mix4_32_dfma_tflops\": %.2f", 2.0 * 256 * 4 * n / (ms * 1e-3) / 1e12, 2.0 * 32 * 32 * n / (ms * 1e-3) / 1e12);
    }
    {   // layout check
        double hA[32], hB[32], hC[64], ref[64];
        for (int i = 0; i < 32; i++) { hA[i] = (i * 37 % 11) - 5 + 0.25 * i; hB[i] = (i * 13 % 7) - 3 + 0.5 * i; }
        for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) { double s = 0; for (int k = 0; k < 4; k++) s += hA[i * 4 + k] * hB[k * 8 + j]; ref[i * 8 + j] = s; }
        double *dA, *dB, *dC; CK(cudaMalloc(&dA, 256)); CK(cudaMalloc(&dB, 256)); CK(cudaMalloc(&dC, 512));
        CK(cudaMemcpy(dA, hA, 256, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB, 256, cudaMemcpyHostToDevice));
        k_dmma_check<<<1, 32>>>(dA, dB, dC); CK(cudaMemcpy(hC, dC, 512, cudaMemcpyDeviceToHost));
        double err = 0; for (int i = 0; i < 64; i++) { double e = hC[i] - ref[i]; if (e < 0) e = -e; if (e > err) err = e; }
        printf(", \"dmma_layout_max_err\": %.3e", err);
    }
    printf("}\n");
    CK(cudaDeviceSynchronize());
    return 0;
}
