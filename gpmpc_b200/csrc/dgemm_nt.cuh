// FP64 "NT" GEMM core on the DMMA pipe (mma.sync m8n8k4 f64 -> SASS DMMA.8x8x4; measured 37.1 TFLOP/s
// on B200, profiles/peaks_micro_r01.json).  tcgen05 has no f64 kind, so this is the tensor path for
// every fp64 contraction of the GP hot path:
//     C[i,j] = beta * C[i,j] + alpha * sum_{k in range(i)} A[i,k] * B[j,k]
// A: M x K row-major, B: N x K row-major (both K-contiguous), C: M x N row-major.
// All dimensions are multiples of the 128 tile (the host pads every matrix), so there is no edge code.
//
// One CTA = one 128x128 tile of C, 8 warps as 2 (M) x 4 (N), warp tile 64x32 = 8x4 DMMA fragments,
// 64 fp64 accumulators per thread.  K is streamed in blocks of 16 through a 4-stage cp.async ring
// (40 KB per stage, rows padded to 160 B so the 8-byte fragment loads are bank-conflict free).
//
// Triangular operands (used by potrf/trtri/K^-1/variance) only shorten the k-range per tile row:
//     a_tri == 1 (A lower triangular, diagonal at A[0][0]):  k <  (tile_m + 1) * 128
//     a_tri == 2 (A upper triangular):                       k >= tile_m * 128
// c_lower == 1 computes only tiles on or below the diagonal (SYRK-like updates).
// Epilogue 0 stores C; epilogue 1 (posterior variance) never stores C and instead writes the
// per-tile column sums of squares  colsq[tile_m][j] = sum_{i in tile} C[i,j]^2.
#pragma once
#include "common.cuh"
#include "gemm_params.cuh"

namespace gpk {

constexpr int G_BM = 128, G_BN = 128, G_BK = 16, G_STAGES = 4, G_THREADS = 256;
constexpr int G_ROW = G_BK + 4;                         // padded smem row, doubles (160 B)
constexpr int G_STAGE_ELEMS = (G_BM + G_BN) * G_ROW;    // doubles per stage
constexpr int G_SMEM_BYTES = G_STAGES * G_STAGE_ELEMS * 8;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// linear CTA id -> (tile_m, tile_n, batch).
// order 0: super-tiles of group_m x group_n tiles enumerated row-major (L2 reuse of both operand panels);
//          the batch index comes from blockIdx.z.
// order 1 (lower-triangular A, work per tile grows with tile_m): the batch is flattened into the columns,
//          columns are cut into chunks of group_n tiles whose B panels stay L2-resident, and inside a
//          chunk the heaviest tile rows go first (longest-processing-time order, so the grid ends on the
//          cheapest tiles and the tail is short).
__device__ __forceinline__ void tile_coords(const GemmParams& p, int id, int& tm, int& tn, int& batch) {
    if (p.order == 1) {
        const int cols = p.tiles_n * p.nbatch;
        const int per_chunk = p.tiles_m * p.group_n;
        const int chunk = id / per_chunk;
        const int rem = id - chunk * per_chunk;
        const int gn = min(p.group_n, cols - chunk * p.group_n);
        tm = p.tiles_m - 1 - rem / gn;
        const int col = chunk * p.group_n + rem % gn;
        batch = col / p.tiles_n;
        tn = col - batch * p.tiles_n;
        return;
    }
    batch = blockIdx.z;
    const int per_srow = p.group_m * p.tiles_n;
    const int sr = id / per_srow;
    const int rem = id - sr * per_srow;
    const int gm = min(p.group_m, p.tiles_m - sr * p.group_m);
    const int per_st = gm * p.group_n;
    const int sc = rem / per_st;
    const int rem2 = rem - sc * per_st;
    tm = sr * p.group_m + rem2 % gm;
    tn = sc * p.group_n + rem2 / gm;
}

// Tile configurations: BM x BN CTA tile, WM x WN warps, warp tile (BM/WM) x (BN/WN) = FM x FN DMMA fragments.
//   128 x 128 (2 x 4 warps, 64 x 32 warp tiles): the throughput configuration (variance GEMM, the largest fit GEMMs)
//    64 x  64 (2 x 2 warps, 32 x 32): mid-size fit GEMMs (up to ~9000 128-tiles): their triangular k-ranges leave a
//              128-tile grid waiting for its longest tile; 4x the CTAs balance better
//    32 x  32 (2 x 2 warps, 16 x 16): every GEMM with fewer than ~200 128-tiles, i.e. the many small GEMMs of the
//              recursive factorisations -- one SM only sustains ~0.25 TFLOP/s of DMMA, so a GEMM with a handful of
//              tiles is bound by that and wants as many CTAs as possible (measured at N=2048: MLL iteration
//              3.65 ms with 64 x 64 tiles, 3.17 ms with 32 x 32)
//    16 x 128 (1 x 4 warps, 16 x 32): for the in-place TRSM leaves (C aliases A; a CTA must own all 128
//              columns of its rows), 8x the CTAs along M
template <int EPI, int BM, int BN, int WM, int WN>
__device__ __forceinline__ void dgemm_nt_body(const GemmParams& p) {
    extern __shared__ __align__(16) double smem[];
    constexpr int THREADS = 32 * WM * WN;
    constexpr int FM = BM / WM / 8, FN = BN / WN / 8;
    constexpr int STAGE_ELEMS = (BM + BN) * G_ROW;
    constexpr int NCH = (BM + BN) * 8 / THREADS;         // 16-byte chunks per thread per k-block
    static_assert((BM + BN) * 8 % THREADS == 0, "chunk count must divide");

    int tm, tn, batch_i;
    tile_coords(p, blockIdx.x, tm, tn, batch_i);
    if (p.c_lower && (tm + 1) * BM <= tn * BN) return;   // tile entirely above the diagonal

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp / WN, wn = warp % WN;

    const long batch = batch_i;
    const double* __restrict__ A = p.A + batch * p.strideA + (long)tm * BM * p.lda;
    const double* __restrict__ B = p.B + batch * p.strideB + (long)tn * BN * p.ldb;

    int kbeg = 0, kend = p.K;
    if (p.a_tri == 1) kend = min(p.K, (tm + 1) * BM);
    if (p.a_tri == 2) kbeg = min(p.K, tm * BM) / G_BK * G_BK;
    kend = (kend + G_BK - 1) / G_BK * G_BK;
    const int nk = (kend - kbeg) / G_BK;

    // chunk c -> row c/8 of the stacked [A rows | B rows] stage, k-chunk c%8.  The copies of one k-block are
    // issued in 4 parts, one per k-step, so the LDGSTS never queue up in front of the fragment LDS of the
    // compute loop (MIO is in-order).
    auto load_part = [&](int stage, int kb, int part) {
        double* As = smem + stage * STAGE_ELEMS;
        const long k0 = kbeg + (long)kb * G_BK;
#pragma unroll
        for (int i = 0; i < NCH; i++) {
            if ((i & 3) != part) continue;
            const int c = tid + i * THREADS;
            const int row = c >> 3, kc = (c & 7) * 2;
            const double* src = row < BM ? A + (long)row * p.lda + k0 + kc : B + (long)(row - BM) * p.ldb + k0 + kc;
            cp_async16(As + row * G_ROW + kc, src);
        }
    };
    auto load_stage = [&](int stage, int kb) {
#pragma unroll
        for (int i = 0; i < 4; i++) load_part(stage, kb, i);
    };

    double acc[FM][FN][2];
#pragma unroll
    for (int i = 0; i < FM; i++)
#pragma unroll
        for (int j = 0; j < FN; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    pdl_wait();                                          // the operands may come from the previous kernel in the stream

#pragma unroll
    for (int s = 0; s < G_STAGES - 1; s++) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }

    const int a_off = (wm * (BM / WM) + g) * G_ROW + t;
    const int b_off = BM * G_ROW + (wn * (BN / WN) + g) * G_ROW + t;

    for (int kb = 0; kb < nk; kb++) {
        cp_async_wait<G_STAGES - 2>();
        __syncthreads();
        // the stage consumed in the previous iteration is refilled piecewise inside the k-step loop
        const int nxt = kb + G_STAGES - 1;
        const bool refill = nxt < nk;
        const int nstage = nxt % G_STAGES;
        const double* st = smem + (kb % G_STAGES) * STAGE_ELEMS;
#pragma unroll
        for (int ks = 0; ks < G_BK / 4; ks++) {
            double a[FM], b[FN];
#pragma unroll
            for (int f = 0; f < FM; f++) a[f] = st[a_off + f * 8 * G_ROW + ks * 4];
#pragma unroll
            for (int f = 0; f < FN; f++) b[f] = st[b_off + f * 8 * G_ROW + ks * 4];
            if (refill) load_part(nstage, nxt, ks);
#pragma unroll
            for (int i = 0; i < FM; i++)
#pragma unroll
                for (int j = 0; j < FN; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        cp_async_commit();
    }
    cp_async_wait<0>();
    pdl_launch_dependents();                             // the next kernel's CTAs may be scheduled during the epilogue (an
                                                         // earlier trigger, at the top, measured no better)

    if constexpr (EPI == 0) {
        if (p.c_inplace) __syncthreads();                // every warp has read its A rows before C (= A) is written
        double* __restrict__ C = p.C + batch * p.strideC + ((long)tm * BM + wm * (BM / WM) + g) * p.ldc +
                                 (long)tn * BN + wn * (BN / WN) + 2 * t;
        const double alpha = p.alpha, beta = p.beta;
#pragma unroll
        for (int i = 0; i < FM; i++)
#pragma unroll
            for (int j = 0; j < FN; j++) {
                double2* ptr = reinterpret_cast<double2*>(C + (long)i * 8 * p.ldc + j * 8);
                double2 v;
                v.x = alpha * acc[i][j][0];
                v.y = alpha * acc[i][j][1];
                if (beta != 0.0) {
                    const double2 old = *ptr;
                    v.x += beta * old.x;
                    v.y += beta * old.y;
                }
                *ptr = v;
            }
    } else {
        static_assert(EPI == 0 || (BM == 128 && BN == 128 && WM == 2 && WN == 4), "column-norm epilogue: 128 x 128 only");
        // column sums of squares over the 128 rows of this tile
        __syncthreads();                       // all warps are done with the pipeline smem
        double* red = smem;                    // [2 (wm)][128 columns]
        const double alpha2 = p.alpha * p.alpha;
#pragma unroll
        for (int j = 0; j < FN; j++) {
            double s0 = 0.0, s1 = 0.0;
#pragma unroll
            for (int i = 0; i < FM; i++) {
                s0 = fma(acc[i][j][0], acc[i][j][0], s0);
                s1 = fma(acc[i][j][1], acc[i][j][1], s1);
            }
            // reduce over g (lane bits 2..4)
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                s0 += __shfl_xor_sync(0xffffffffu, s0, o);
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            }
            if (g == 0) {
                red[wm * BN + wn * 32 + j * 8 + 2 * t] = s0;
                red[wm * BN + wn * 32 + j * 8 + 2 * t + 1] = s1;
            }
        }
        __syncthreads();
        if (tid < BN) {
            double* out = p.colsq + batch * p.stride_colsq + (long)tm * p.ld_colsq + (long)tn * BN;
            out[tid] = alpha2 * (red[tid] + red[BN + tid]);
        }
    }
}

// EPI 0: C = beta C + alpha A B^T (potrf / trtri / K^-1 building block)
__global__ void __launch_bounds__(G_THREADS, 1) dgemm_nt_kernel(const GemmParams p) { dgemm_nt_body<0, 128, 128, 2, 4>(p); }
__global__ void __launch_bounds__(128, 2) dgemm_nt64_kernel(const GemmParams p) { dgemm_nt_body<0, 64, 64, 2, 2>(p); }
__global__ void __launch_bounds__(128, 4) dgemm_nt32_kernel(const GemmParams p) { dgemm_nt_body<0, 32, 32, 2, 2>(p); }
__global__ void __launch_bounds__(128, 2) dgemm_nt16x128_kernel(const GemmParams p) { dgemm_nt_body<0, 16, 128, 1, 4>(p); }
// EPI 1: posterior-variance contraction, column sums of squares of W K_ZX^T (DMMA variance path)
__global__ void __launch_bounds__(G_THREADS, 1) var_colsq_kernel(const GemmParams p) { dgemm_nt_body<1, 128, 128, 2, 4>(p); }

// host-side launcher (defined in linalg.cu)
int launch_dgemm_nt(GemmParams p, int epilogue, int batch, cudaStream_t st);

}  // namespace gpk
