// Probe: under which launch attributes does a small kernel B become co-resident with a persistent kernel A that
// fills every SM?  A mimics oz_var_kernel's footprint (384 threads, 120 registers + setmaxnreg, 182 KB of shared memory,
// optional cluster of 2), B mimics oz_front_co_kernel (128 threads, <= 152 registers, 28 KB static shared memory).
// Each CTA of B records the global timer when it starts; A records when it starts and ends.  B is co-resident when its
// CTAs start while A is still running.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o coresidency_probe coresidency_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

template <bool SETMAX>
__device__ __forceinline__ void a_body(unsigned long long* t, long long spin_ns) {
    extern __shared__ unsigned char dsm[];
    const int warp = threadIdx.x >> 5;
    if (SETMAX) {
        if (warp >= 8) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 160;\n");
    }
    const unsigned long long t0 = gtimer();
    if (threadIdx.x == 0) { t[2 * blockIdx.x] = t0; dsm[0] = 1; }
    while ((long long)(gtimer() - t0) < spin_ns) { }
    if (threadIdx.x == 0) t[2 * blockIdx.x + 1] = gtimer();
}
template <bool SETMAX>
__global__ void __maxnreg__(120) kernel_a(unsigned long long* t, long long spin_ns) { a_body<SETMAX>(t, spin_ns); }
template <bool SETMAX>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(120) kernel_a_cl(unsigned long long* t, long long spin_ns) { a_body<SETMAX>(t, spin_ns); }

template <int LIVE>
__global__ void __maxnreg__(152) kernel_b(unsigned long long* t, float* sink, int n) {
    __shared__ double pad[28672 / 8];
    const unsigned long long t0 = gtimer();
    if (threadIdx.x == 0) t[blockIdx.x] = t0;
    // keep ~140 registers live
    float acc[LIVE];
#pragma unroll
    for (int i = 0; i < LIVE; i++) acc[i] = sink[(i % 120) * 128 + threadIdx.x];
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int i = 0; i < LIVE; i++) acc[i] = acc[i] * acc[(i + 7) % LIVE] + 1.0f;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < LIVE; i++) s += acc[i];
    pad[threadIdx.x] = s;
    __syncthreads();
    sink[threadIdx.x] = (float)pad[(threadIdx.x + 1) & 127];
}

int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;       // bit 0: cluster, bit 1: setmaxnreg, bit 2: A max carveout, bit 3: B max carveout
    const int dsmem = argc > 2 ? atoi(argv[2]) : 184320;
    const bool cl = variant & 1, sm = variant & 2, ca = variant & 4, cb = variant & 8;
    unsigned long long *ta, *tb;
    float* sink;
    cudaMalloc(&ta, 2 * 148 * 8); cudaMalloc(&tb, 148 * 8); cudaMalloc(&sink, 128 * 128 * 4);
    cudaMemset(ta, 0, 2 * 148 * 8); cudaMemset(tb, 0, 148 * 8); cudaMemset(sink, 0, 128 * 128 * 4);
    cudaStream_t s1, s2;
    cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    const void* fa = cl ? (sm ? (const void*)kernel_a_cl<true> : (const void*)kernel_a_cl<false>)
                        : (sm ? (const void*)kernel_a<true> : (const void*)kernel_a<false>);
    cudaFuncSetAttribute(fa, cudaFuncAttributeMaxDynamicSharedMemorySize, dsmem);
    if (ca) cudaFuncSetAttribute(fa, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    const int live = argc > 3 ? atoi(argv[3]) : 120;
    const void* fb = live <= 104 ? (const void*)kernel_b<104> : live <= 120 ? (const void*)kernel_b<120> : live <= 132 ? (const void*)kernel_b<132> : (const void*)kernel_b<142>;
    if (cb) cudaFuncSetAttribute(fb, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    cudaFuncAttributes fattr; cudaFuncGetAttributes(&fattr, fb);
    long long spin = 2000000;   // 2 ms
    void* args[] = {&ta, &spin};
    cudaLaunchKernel(fa, dim3(148), dim3(384), args, dsmem, s1);
    // give A time to occupy the SMs, then launch B
    cudaEvent_t e; cudaEventCreate(&e);
    int n = 100;
    void* bargs[] = {&tb, &sink, &n};
    cudaLaunchKernel(fb, dim3(148), dim3(128), bargs, 0, s2);
    cudaError_t err = cudaDeviceSynchronize();
    unsigned long long ha[2 * 148], hb[148];
    cudaMemcpy(ha, ta, sizeof(ha), cudaMemcpyDeviceToHost); cudaMemcpy(hb, tb, sizeof(hb), cudaMemcpyDeviceToHost);
    unsigned long long a0 = ~0ull, a1 = 0;
    for (int i = 0; i < 148; i++) { if (ha[2 * i] < a0) a0 = ha[2 * i]; if (ha[2 * i + 1] > a1) a1 = ha[2 * i + 1]; }
    int during = 0;
    for (int i = 0; i < 148; i++) if (hb[i] > a0 && hb[i] + 500000 < a1) during++;
    printf("B regs %d: ", fattr.numRegs);
    printf("variant %d (cluster %d setmaxnreg %d carveoutA %d carveoutB %d, A dyn smem %d): %s; A ran %.3f ms; B CTAs started >0.5 ms before A ended: %d / 148\n",
           variant, cl, sm, ca, cb, dsmem, cudaGetErrorString(err), (a1 - a0) * 1e-6, during);
    return 0;
}
