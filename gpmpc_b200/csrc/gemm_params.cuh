// Argument block shared by the DMMA GEMM (dgemm_nt.cuh) and the tcgen05 int8 digit GEMM (ozgemm.cu).
#pragma once
#include "common.cuh"

namespace gpk {

struct GemmParams {
    const double* A;
    const double* B;
    double* C;
    long lda, ldb, ldc;
    long strideA, strideB, strideC;   // batch strides (elements); grid.z = batch
    int M, N, K;
    double alpha, beta;
    int a_tri;       // 0 full, 1 lower, 2 upper
    int c_lower;     // 1: skip tiles entirely above the diagonal
    int c_inplace;   // 1: C aliases A (one CTA owns all columns of its rows)
    int tiles_m, tiles_n;
    int group_m, group_n;   // super-tile for L2 reuse across the ~148 concurrently resident CTAs
    int order, nbatch;      // tile enumeration (see tile_coords)
    double* colsq;          // epilogue 1 output: [tiles_m][ld_colsq]
    long ld_colsq, stride_colsq;
};

// int8 digit path for the large fit GEMMs (ozgemm.cu): 1 = handled, 0 = not eligible, < 0 = error
int ozg_try(const GemmParams& p, int batch, cudaStream_t st);
// which half of the registered scratch the following GEMMs of this host thread may use (0 = all, 1 / 2 = a half)
void ozg_set_partition(int part);

}  // namespace gpk
