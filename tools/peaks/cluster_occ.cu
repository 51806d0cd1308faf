// How many thread-block clusters of a 1-CTA/SM kernel (320 threads, 180 KB dynamic smem) can be co-resident on this GPU
// for cluster sizes 1, 2, 4, 8 -- i.e. how many SMs a persistent clustered kernel can actually use.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cluster_occ cluster_occ.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    const int smem = 180 * 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("{\"sms\": %d", sms);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(sms / cs * cs); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension; at.val.clusterDim.x = cs; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
        printf(", \"clusters_of_%d\": %d, \"sms_used_%d\": %d", cs, e == cudaSuccess ? n : -1, cs, e == cudaSuccess ? n * cs : -1);
    }
    printf("}\n");
    return 0;
}
