// Probe: which setmaxnreg arrangements complete on this part?  (build: nvcc -gencode arch=compute_100a,code=sm_100a -O3)
// variant 0: warps 8 and 9 execute DIFFERENT dec instructions (one per role branch), 320 threads
// variant 1: warps 8 and 9 execute the SAME dec instruction, 320 threads
// variant 2: 384 threads (three full warpgroups), warps 8..11 execute the same dec instruction
// variant 3: as 1, but dec comes first in program order for everyone via a CTA barrier before inc
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template <int V>
__global__ void __maxnreg__(144) probe(float* p, int n) {
    const int warp = threadIdx.x >> 5;
    if (V == 0) {
        if (warp == 8) { asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n"); p[threadIdx.x] = 1.f; return; }
        if (warp == 9) { asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n"); p[threadIdx.x] = 2.f; return; }
    } else {
        if (warp >= 8) { asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n"); p[threadIdx.x] = (float)warp; if (V == 3) __syncthreads(); return; }
    }
    if (V == 3) __syncthreads();
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
    float acc[150];
#pragma unroll
    for (int i = 0; i < 150; i++) acc[i] = p[i * 512 + threadIdx.x];
#pragma unroll 1
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int i = 0; i < 150; i++) acc[i] = acc[i] * acc[(i + 1) % 150] + 1.0f;
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 150; i++) s += acc[i];
    p[threadIdx.x] = s;
}

int main(int argc, char** argv) {
    const int v = argc > 1 ? atoi(argv[1]) : 0;
    float* p;
    cudaMalloc(&p, 512 * 160 * 4);
    cudaMemset(p, 0, 512 * 160 * 4);
    if (v == 0) probe<0><<<148, 320>>>(p, 10);
    if (v == 1) probe<1><<<148, 320>>>(p, 10);
    if (v == 2) probe<2><<<148, 384>>>(p, 10);
    if (v == 3) probe<3><<<148, 320>>>(p, 10);
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d: %s\n", v, cudaGetErrorString(e));
    return e != cudaSuccess;
}
