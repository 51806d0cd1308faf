// Microbenchmark of tcgen05.mma kind::i8 issue/throughput on B200: cycles per instruction for M = 128 and
// N in {64,...,256}, operands resident in shared memory (no TMA traffic), one CTA per SM, and the same loop
// with a concurrent 1-D bulk-TMA stream into other shared-memory buffers (smem write interference).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../gpmpc_b200/csrc -o umma_bench umma_bench.cu
#include "ozaki.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
using namespace gpk;
namespace gpk { thread_local char g_err[512]; std::atomic<long> g_launches{0}; int fail(int c, const char*, ...) { return c; } }

// mode 0: `reps` x one MMA of width N.  mode 1: the Ozaki k-step mix for S digits (concatenated widths).
// tma_bytes > 0: warp 2 streams tma_bytes per MMA batch from global into a separate smem ring.
__device__ __forceinline__ uint64_t mkdesc(uint32_t addr, int layout) {
    // layout 0: no swizzle (LBO 128, SBO 512); 2/4/6: 128B/64B/32B swizzle, SBO = 8 rows
    if (layout == 0) return umma_smem_desc(addr, 128, 512);
    const uint32_t sbo = layout == 2 ? 1024 : layout == 4 ? 512 : 256;
    return umma_smem_desc(addr, 16, sbo) | ((uint64_t)layout << 61);
}

__global__ void __launch_bounds__(128, 1) bench(int N, int mode, int S, int reps, int tma_bytes, const uint8_t* src,
                                               long long* cycles, int layout) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t done_bar, tma_bar[2];
    __shared__ uint32_t tmem_base_smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { mbar_init(&done_bar, 1); mbar_init(&tma_bar[0], 1); mbar_init(&tma_bar[1], 1); mbar_fence_init(); }
    if (warp == 1) tmem_alloc(&tmem_base_smem, 512);
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_smem;
    const uint32_t a_base = smem_u32(smem), b_base = a_base + 48 * 1024;
    if (warp == 1) {
        const uint32_t leader = elect_one() ? 1u : 0u;
        const long long t0 = clock64();
        if (mode == 0) {
            const uint64_t ad = mkdesc(a_base, layout), bd = mkdesc(b_base, layout);
            const uint32_t id = umma_idesc_i8(N);
#pragma unroll 1
            for (int r = 0; r < reps; r += 8) {
#pragma unroll
                for (int u = 0; u < 8; u++) umma_i8_pred(tmem + u * 32, ad + u * 256, bd, id, 1, leader);
            }
        } else if (mode == 1) {
#pragma unroll 1
            for (int r = 0; r < reps; r++) {
#pragma unroll
                for (int t = 0; t < 7; t++)
#pragma unroll
                    for (int s0 = 0; s0 < 7; s0 += 4) {
                        if (t < S && s0 < S - t) {
                            const int ns = (S - t - s0) < 4 ? (S - t - s0) : 4;
                            umma_i8_pred(tmem + (t + s0) * 64, umma_smem_desc(a_base + t * 8192, 128, 512),
                                         umma_smem_desc(b_base + s0 * 4096, 128, 512), umma_idesc_i8(ns * 64), 1, leader);
                        }
                    }
            }
        }
        if (leader) tc_commit(&done_bar);
        __syncwarp();
        mbar_wait(&done_bar, 0);
        if (leader) cycles[blockIdx.x] = clock64() - t0;
    } else if (warp == 2 && lane == 0 && tma_bytes > 0) {
        uint8_t* ring = smem + 96 * 1024;
        uint32_t ph[2] = {0, 0};
        for (int r = 0; r < reps; r++) {
            const int b = r & 1;
            if (r >= 2) { mbar_wait(&tma_bar[b], ph[b]); ph[b] ^= 1; }
            mbar_expect_tx(&tma_bar[b], tma_bytes);
            tma_bulk_g2s(ring + b * 60 * 1024, src + ((size_t)(blockIdx.x * 131 + r) % 512) * 65536, tma_bytes, &tma_bar[b]);
        }
        for (int b = 0; b < 2; b++) if (reps > b) mbar_wait(&tma_bar[b], ph[b]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main(int argc, char** argv) {
    const int smem = 220 * 1024;
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    long long* cyc; cudaMallocManaged(&cyc, 148 * sizeof(long long));
    uint8_t* src; cudaMalloc(&src, 512 * 65536); cudaMemset(src, 1, 512 * 65536);
    int layout = 0;
    auto run = [&](int N, int mode, int S, int reps, int tma, int grid) {
        bench<<<grid, 128, smem>>>(N, mode, S, reps, tma, src, cyc, layout);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
        long long mx = 0; for (int i = 0; i < grid; i++) mx = cyc[i] > mx ? cyc[i] : mx;
        return (double)mx / reps;
    };
    run(64, 0, 6, 100, 0, 148);
    if (argc > 1 && !strcmp(argv[1], "--json")) {
        // measured dense int8 rate of the tcgen05 pipe: M=128, N=256, K=32 instructions back to back on every SM,
        // operands resident in shared memory, wall time by CUDA events (so the clock the chip actually holds under
        // this load is included).  burst = one 13 ms launch; sustained = 1.5 s of such launches.
        int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
        const int reps = 200000;
        const double ops = 2.0 * 128 * 256 * 32 * (double)reps * sms;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        double best = 0, clk = 0;
        for (int it = 0; it < 5; it++) {
            cudaEventRecord(e0);
            bench<<<sms, 128, smem>>>(256, 0, 6, reps, 0, src, cyc, 0);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ops / ms * 1e-9 > best) { best = ops / ms * 1e-9; clk = (double)cyc[0] / reps; }
        }
        cudaEventRecord(e0);
        int n = 0; float ms = 0;
        do {
            for (int k = 0; k < 8; k++) bench<<<sms, 128, smem>>>(256, 0, 6, reps, 0, src, cyc, 0);
            n += 8; cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        } while (ms < 1500.f);
        printf("{\"int8_tops_burst\": %.1f, \"int8_tops_sustained\": %.1f, \"clk_per_n256_instr\": %.2f, \"sms\": %d, "
               "\"how\": \"tcgen05.mma kind::i8 M128 N256 K32 back to back, smem-resident operands, CUDA events\"}\n",
               best, ops * n / ms * 1e-9, clk, sms);
        return 0;
    }
    for (int lay : {0, 6, 4, 2}) {
        layout = lay;
        const int grid = 148;
        for (int N : {16, 64, 128, 256})
            printf("layout %d  single MMA  N=%3d : %7.1f clk/instr (floor %d)\n", lay, N, run(N, 0, 6, 4000, 0, grid), N / 2);
        for (int S : {5, 6}) {
            const int units = S * (S + 1) / 2;
            printf("layout %d  ozaki k-step S=%d  : %7.1f clk/k-step (floor %d)\n", lay, S, run(0, 1, S, 2000, 0, grid), units * 32);
            printf("layout %d  ozaki k-step S=%d + TMA 36864 B/k-step: %7.1f clk/k-step\n", lay, S, run(0, 1, S, 2000, 36864, grid));
        }
    }
    return 0;
}
