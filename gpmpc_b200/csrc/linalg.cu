// Blocked fp64 Cholesky / triangular inverse / K^-1 / alpha for the N x N training kernel matrix.
//
// Replaces what the reference reaches through gpytorch -> torch LAPACK at gpmpc/gp.py:62 (potrf + solves
// inside ExactMarginalLogLikelihood) and gpmpc/gp.py:46 (torch.linalg.inv), see SURVEY.md section 2.
// Everything O(N^3) is expressed through ONE tensor-pipe kernel, the DMMA "NT" GEMM of dgemm_nt.cuh:
//   potrf   : recursive right-looking Cholesky; leaves are 128x128 single-CTA factorisations that also
//             emit the inverse of the diagonal block, so every triangular solve becomes a GEMM
//   trtri   : W = L^-1 (lower) and U = W^T (upper), recursive, two triangular GEMMs per level
//   kinv    : K^-1 = U U^T (lower tiles on the tensor pipe, mirrored afterwards)
//   alpha   : t = W y, alpha = U t (two triangular GEMVs, HBM-bound), quad = |t|^2, logdet = 2 sum log L_ii
// All matrices are row-major, padded to a multiple of 128 with an identity block (see engine.py).
#include <cstdlib>
#include "dgemm_nt.cuh"
#include <type_traits>
#include <vector>

namespace gpk {

thread_local char g_err[512] = {0};
std::atomic<long> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// ------------------------------------------------------------------------------------------------
// GEMM launcher
// ------------------------------------------------------------------------------------------------
int launch_dgemm_nt(GemmParams p, int epilogue, int batch, cudaStream_t st) {
    // tile thresholds in units of 128 x 128 output tiles, from interleaved A/B runs of the MLL iteration
    // (N=2048: 2.77 -> 2.58 ms, N=4096: 7.58 -> 7.12 ms, N=8192: 28.8 -> 28.3 ms against 100 / 0)
    constexpr long GPK_T32 = 200, GPK_T64 = 9000;
    GPK_ARG(p.M > 0 && p.N > 0 && p.K > 0, "empty GEMM");
    GPK_ARG(p.M % G_BM == 0 && p.N % G_BN == 0 && p.K % G_BK == 0, "GEMM dims must be multiples of 128/128/16");
    GPK_ARG(p.lda % 2 == 0 && p.ldb % 2 == 0, "lda/ldb must be even");
    GPK_ARG(((uintptr_t)p.A % 16) == 0 && ((uintptr_t)p.B % 16) == 0, "A/B must be 16-byte aligned");
    if (epilogue == 0) GPK_ARG(p.ldc % 2 == 0 && ((uintptr_t)p.C % 16) == 0, "C must be 16-byte aligned with even ldc");
    if (epilogue == 1) GPK_ARG(p.colsq != nullptr, "colsq output missing");
    p.nbatch = batch;
    p.c_inplace = (epilogue == 0 && p.C == p.A) ? 1 : 0;
    if (epilogue == 0 && !p.c_inplace) {
        // large GEMMs go to the tcgen05 int8 digit path when the caller registered its scratch (ozgemm.cu)
        const int r = ozg_try(p, batch, st);
        if (r != 0) return r < 0 ? r : 0;
    }
    // tile configuration: the 128 x 128 tiles need ~one wave of CTAs to pay off (see dgemm_nt.cuh)
    const long tiles128 = (long)(p.M / G_BM) * (p.N / G_BN) * batch / (p.c_lower ? 2 : 1);
    int bm = G_BM, bn = G_BN;
    if (epilogue == 0 && tiles128 < GPK_T32) {
        if (p.c_inplace) { bm = 16; bn = 128; }
        else { bm = 32; bn = 32; }
    } else if (epilogue == 0 && !p.c_inplace && tiles128 < GPK_T64) {
        bm = 64; bn = 64;
    }
    GPK_ARG(!p.c_inplace || p.N == bn, "in-place GEMM needs a single tile column");
    p.tiles_m = p.M / bm;
    p.tiles_n = p.N / bn;
    dim3 grid((unsigned)(p.tiles_m * p.tiles_n), 1, (unsigned)batch);
    if (epilogue == 1 && p.a_tri == 1 && !p.c_lower) {
        // LPT order inside L2-sized column chunks (B panel of one column tile = 128 x K doubles)
        p.order = 1;
        const long panel = (long)G_BN * p.K * 8;
        long gn = (48L << 20) / panel;
        const long min_gn = (2 * 148 + p.tiles_m - 1) / p.tiles_m;
        if (gn < min_gn) gn = min_gn;
        p.group_n = (int)gn;
        p.group_m = p.tiles_m;
        grid = dim3((unsigned)(p.tiles_m * p.tiles_n * batch), 1, 1);
    } else {
        p.order = 0;
        p.group_m = p.tiles_m <= 18 ? p.tiles_m : 12;
        p.group_n = 148 / p.group_m > 0 ? 148 / p.group_m : 1;
    }
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(dgemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_BYTES));
        GPK_CUDA(cudaFuncSetAttribute(var_colsq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_BYTES));
        GPK_CUDA(cudaFuncSetAttribute(dgemm_nt32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_BYTES));
        GPK_CUDA(cudaFuncSetAttribute(dgemm_nt64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_BYTES));
        GPK_CUDA(cudaFuncSetAttribute(dgemm_nt16x128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_BYTES));
        attr_done = true;
    }
    const int smem = G_STAGES * (bm + bn) * G_ROW * 8;
    if (epilogue == 1)
        var_colsq_kernel<<<grid, G_THREADS, smem, st>>>(p);
    else if (bm == 128)
        GPK_CUDA(launch_pdl(dgemm_nt_kernel, grid, dim3(G_THREADS), smem, st, p));
    else if (bm == 64)
        GPK_CUDA(launch_pdl(dgemm_nt64_kernel, grid, dim3(128), smem, st, p));
    else if (bn == 32)
        GPK_CUDA(launch_pdl(dgemm_nt32_kernel, grid, dim3(128), smem, st, p));
    else
        GPK_CUDA(launch_pdl(dgemm_nt16x128_kernel, grid, dim3(128), smem, st, p));
    GPK_LAUNCHED();
    return 0;
}

static GemmParams gemm_args(const double* A, long lda, const double* B, long ldb, double* C, long ldc, long M, long N,
                            long K, double alpha, double beta, int a_tri, int c_lower) {
    GemmParams p{};
    p.A = A; p.B = B; p.C = C;
    p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.M = (int)M; p.N = (int)N; p.K = (int)K;
    p.alpha = alpha; p.beta = beta;
    p.a_tri = a_tri; p.c_lower = c_lower;
    return p;
}

// ------------------------------------------------------------------------------------------------
// 128x128 leaf: Cholesky of a diagonal block (one CTA) + inverse of its factor (16 CTAs)
// ------------------------------------------------------------------------------------------------
constexpr int P_N = 128, P_LD = 129, P_B = 8, P_NB = P_N / P_B;

// Cholesky of a 128 x 128 leaf, blocked right-looking with block size 8 (16 block steps), rank-8 updates on the FP64
// tensor pipe (DMMA m8n8k4).  Round 1's leaf kept one 8 x 8 block per THREAD (SIMT DFMAs): a clock64 trace
// (profiles/pipeline_r02.md) showed 2200 clocks per rank-8 update and 2270 per panel step -- operand-fresh DFMAs issue
// every ~4 clocks per sub-partition and from block step 8 on most lanes idle -- 47 us per leaf.  Here a
// WARP owns whole 8 x 8 blocks in the DMMA accumulator layout (lane l holds [l/4][2(l%4)], [l/4][2(l%4)+1]), so the
// update of a block uses all 32 lanes and every decision is warp-uniform.  Nine warps:
//   warp 8 (diagonal warp) owns the 16 diagonal blocks.  Step kb: it parks block kb in shared memory, lane 0 factors
//       it (8 x 8 Cholesky in registers, 1/sqrt by MUFU + two Newton steps), publishes L_kk and 1 / l_jj and ARRIVES
//       on barrier 1; it then brings its later diagonal blocks up to date with the PREVIOUS panel (lazily: nobody
//       waits for that), waits for the panel of this step (barrier 2) and updates block kb + 1 first;
//   warps 0-7 (workers) own the 120 off-diagonal blocks, dealt round-robin in descending block-column order (block
//       i -> warp i % 8, slot i / 8: the blocks a step touches are a prefix, every warp has at most two panel blocks).
//       Step kb: park the panel blocks (block column kb) in the panel buffer, wait for L_kk (barrier 1), 16 lanes take
//       one panel row each and do the forward substitution L_ik = A_ik L_kk^-T (36-deep FMA chain, all rows in
//       parallel), write it to the panel [block][row][k] (row stride 12 doubles: conflict-free fragment reads) and to
//       HBM, barrier 2, then A_ic -= L_ik L_ck^T = 2 DMMAs per trailing block, four blocks per basic block.
// So the serial part of a step is max(diagonal factor, trailing updates) + substitution, not their sum; the panel
// buffer alternates between two copies because the diagonal warp reads the previous one a step late.
constexpr int PM_WORKERS = 8, PM_THREADS = 32 * (PM_WORKERS + 1), PM_SLOTS = 16, PM_LD = 12;
__device__ __forceinline__ void pm_slot_block(int w, int s, int& ty, int& tx) {
    const int i = (s < 15 ? s : 14) * PM_WORKERS + w;   // < 120, off-diagonal blocks ordered by tx descending (slot 15 unused)
    int m = 0;
    while ((m + 1) * (m + 2) / 2 <= i) m++;   // tx = 14 - m has m + 1 blocks
    tx = 14 - m;
    ty = tx + 1 + (i - m * (m + 1) / 2);
}
__device__ __forceinline__ void pm_dmma(double (&c)[2], double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
// 1 / sqrt(d) for a normal positive d: MUFU.RSQ64H seed (~2^-22) and two Newton steps y += y/2 (1 - d y^2)
__device__ __forceinline__ double pm_rsqrt(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;\n" : "=d"(y) : "d"(d));
#pragma unroll
    for (int it = 0; it < 2; it++) {
        const double e = fma(-d * y, y, 1.0);
        y = fma(0.5 * y, e, y);
    }
    return y;
}
__device__ __forceinline__ void pm_bar_sync(int id) { asm volatile("bar.sync %0, %1;\n" ::"r"(id), "n"(PM_THREADS) : "memory"); }
__device__ __forceinline__ void pm_bar_arrive(int id) { asm volatile("bar.arrive %0, %1;\n" ::"r"(id), "n"(PM_THREADS) : "memory"); }

__global__ void __launch_bounds__(PM_THREADS, 1)
potf2_mma_kernel(double* __restrict__ A, long lda, int* __restrict__ info, int row_offset) {
    __shared__ __align__(16) double Pn2[2][P_NB * 8 * PM_LD]; // panel (two buffers, by step parity): [block][row][k]
    __shared__ __align__(16) double Dg[8 * PM_LD];            // the diagonal block, then its factor L_kk
    __shared__ double Rv[8];                                  // 1 / l_jj of the current diagonal block
    __shared__ __align__(16) double Cd[P_NB * 8 * PM_LD];     // the 16 diagonal blocks (diagonal warp only)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int fr = lane >> 2, fq = lane & 3;                  // fragment row, position in the quad
    const int fo = fr * PM_LD + fq;                           // fragment element offset inside a panel block
    pdl_wait();
    if (warp == PM_WORKERS) {
        // ================= diagonal warp: its 16 blocks live in shared memory (run-time block index) =================
        for (int j = 0; j < P_NB; j++)
            *reinterpret_cast<double2*>(Cd + j * 8 * PM_LD + fr * PM_LD + 2 * fq) =
                *reinterpret_cast<const double2*>(A + (long)(j * 8 + fr) * lda + j * 8 + 2 * fq);
        __syncwarp();
        auto update = [&](int j, const double* Pp) {          // C_jj -= L_jk L_jk^T with the panel Pp
            double* cj = Cd + j * 8 * PM_LD + fr * PM_LD + 2 * fq;
            const double* pr = Pp + j * 8 * PM_LD + fo;
            const double v0 = pr[0], v1 = pr[4];
            const double2 cv = *reinterpret_cast<const double2*>(cj);
            double d[2] = {cv.x, cv.y};
            pm_dmma(d, -v0, v0);
            pm_dmma(d, -v1, v1);
            *reinterpret_cast<double2*>(cj) = make_double2(d[0], d[1]);
        };
#pragma unroll 1
        for (int kb = 0; kb < P_NB; kb++) {
            double* Ck = Cd + kb * 8 * PM_LD;
            if (lane == 0) {
                double f[8][8];
#pragma unroll
                for (int r = 0; r < 8; r++)
#pragma unroll
                    for (int cc = 0; cc <= r; cc++) f[r][cc] = Ck[r * PM_LD + cc];
                // 1-based index of the first bad pivot, branch-free: the whole factorisation stays one basic block, so
                // the next pivot's MUFU issues as soon as its operand is ready
                int bad = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    double d = f[j][j];
                    const bool ok = d >= 1e-290 && d <= 1.0e300;   // not positive (also NaN), too small to invert, or inf
                    bad = (!ok && bad == 0) ? j + 1 : bad;
                    d = ok ? d : 1.0;
                    const double rinv = pm_rsqrt(d);
                    double rt = d * rinv;
                    rt = fma(fma(-rt, rt, d), 0.5 * rinv, rt);
                    f[j][j] = rt;
                    Rv[j] = rinv;
#pragma unroll
                    for (int r = j + 1; r < 8; r++) f[r][j] *= rinv;
#pragma unroll
                    for (int cc = j + 1; cc < 8; cc++)
#pragma unroll
                        for (int r = cc; r < 8; r++) f[r][cc] = fma(-f[r][j], f[cc][j], f[r][cc]);
                }
#pragma unroll
                for (int r = 0; r < 8; r++)                    // lower triangle only: nobody reads above the diagonal
#pragma unroll
                    for (int cc = 0; cc <= r; cc++) Dg[r * PM_LD + cc] = f[r][cc];
                if (bad) atomicCAS(info, 0, row_offset + kb * 8 + bad);
            }
            __syncwarp();
            __threadfence_block();
            pm_bar_arrive(1);                                  // L_kk and 1 / l_jj are published
            // the factored block goes back to global memory (lower triangle only, as before)
            {
                const double2 v = *reinterpret_cast<const double2*>(Dg + fr * PM_LD + 2 * fq);
                double* dst = A + (long)(kb * 8 + fr) * lda + kb * 8 + 2 * fq;
                if (2 * fq <= fr) dst[0] = v.x;
                if (2 * fq + 1 <= fr) dst[1] = v.y;
            }
            pm_bar_sync(2);                                    // the panel of step kb is final
            // block kb + 1 is the next one to be factored: this warp updates it itself, right away (the workers
            // bring the later diagonal blocks up to date during their trailing updates)
            if (kb + 1 < P_NB) update(kb + 1, Pn2[kb & 1]);
            __syncwarp();
        }
        return;
    }
    double c[PM_SLOTS][2];
    int sty[PM_SLOTS], stx[PM_SLOTS];
#pragma unroll
    for (int s = 0; s < PM_SLOTS; s++) {
        pm_slot_block(warp, s, sty[s], stx[s]);
        const double2 v = *reinterpret_cast<const double2*>(A + (long)(sty[s] * 8 + fr) * lda + stx[s] * 8 + 2 * fq);
        c[s][0] = v.x; c[s][1] = v.y;
    }
    // ================= workers =================
#pragma unroll 1
    for (int kb = 0; kb < P_NB; kb++) {
        double* Pn = Pn2[kb & 1];
        // ---- (b) panel: this warp's (at most two) blocks of block column kb go to the panel buffer as they are ----
        int pty0 = -1, pty1 = -1;
#pragma unroll
        for (int s = 0; s < 15; s++) {
            if (stx[s] != kb) continue;
            *reinterpret_cast<double2*>(Pn + (sty[s] * 8 + fr) * PM_LD + 2 * fq) = make_double2(c[s][0], c[s][1]);
            if (pty0 < 0) pty0 = sty[s]; else pty1 = sty[s];
        }
        __syncwarp();
        pm_bar_sync(1);                                        // L_kk and 1 / l_jj are published
        {
            // one row per lane: l[c] = (a[c] - sum_{m<c} l[m] L_kk[c][m]) / L_kk[c][c]
            const int pty = (lane >> 3) == 0 ? pty0 : ((lane >> 3) == 1 ? pty1 : -1);
            if (pty >= 0) {
                double* row = Pn + (pty * 8 + (lane & 7)) * PM_LD;
                double l[8];
#pragma unroll
                for (int cc = 0; cc < 8; cc += 2) {
                    const double2 v = *reinterpret_cast<const double2*>(row + cc);
                    l[cc] = v.x; l[cc + 1] = v.y;
                }
                // right-looking: as soon as l[m] is final it leaves every later column, so the serial chain is one FMA and
                // one multiply per column (the factor's entries are fetched up front)
                double lk[8][8], rv[8];
#pragma unroll
                for (int cc = 0; cc < 8; cc++) {
                    rv[cc] = Rv[cc];
#pragma unroll
                    for (int m = 0; m < cc; m++) lk[cc][m] = Dg[cc * PM_LD + m];
                }
#pragma unroll
                for (int m = 0; m < 8; m++) {
                    l[m] *= rv[m];
#pragma unroll
                    for (int cc = m + 1; cc < 8; cc++) l[cc] = fma(-l[m], lk[cc][m], l[cc]);
                }
                double* gdst = A + (long)(pty * 8 + (lane & 7)) * lda + kb * 8;
#pragma unroll
                for (int cc = 0; cc < 8; cc += 2) {
                    *reinterpret_cast<double2*>(row + cc) = make_double2(l[cc], l[cc + 1]);
                    *reinterpret_cast<double2*>(gdst + cc) = make_double2(l[cc], l[cc + 1]);
                }
            }
        }
        pm_bar_sync(2);                                        // the panel of step kb is final
        // ---- (c) trailing update: A_ic -= L_ik L_ck^T; the active slots are the prefix s < nact ----
        {
            const int nblk = (14 - kb) * (15 - kb) / 2;       // off-diagonal blocks with tx > kb
            const int nact = nblk > warp ? (nblk - warp + PM_WORKERS - 1) / PM_WORKERS : 0;
            {
                // diagonal blocks j >= kb + 2 (shared memory; j % 8 == this warp): C_jj -= L_jk L_jk^T
                const int j0 = warp + (warp >= ((kb + 2) & 7) ? 0 : 8) + ((kb + 2) & ~7);   // first j >= kb + 2 with j % 8 == warp
                const bool on0 = j0 < P_NB, on1 = j0 + 8 < P_NB;
                const int ja = on0 ? j0 : 0, jb = on1 ? j0 + 8 : 0;
                double* ca = Cd + ja * 8 * PM_LD + fr * PM_LD + 2 * fq;
                double* cb = Cd + jb * 8 * PM_LD + fr * PM_LD + 2 * fq;
                const double a0 = on0 ? Pn[ja * 8 * PM_LD + fo] : 0.0, a1 = on0 ? Pn[ja * 8 * PM_LD + fo + 4] : 0.0;
                const double b0 = on1 ? Pn[jb * 8 * PM_LD + fo] : 0.0, b1 = on1 ? Pn[jb * 8 * PM_LD + fo + 4] : 0.0;
                if (on0) {                                     // warp-uniform
                    const double2 va = *reinterpret_cast<const double2*>(ca);
                    const double2 vb = *reinterpret_cast<const double2*>(cb);
                    double da[2] = {va.x, va.y}, db[2] = {vb.x, vb.y};
                    pm_dmma(da, -a0, a0);
                    pm_dmma(db, -b0, b0);
                    pm_dmma(da, -a1, a1);
                    pm_dmma(db, -b1, b1);
                    *reinterpret_cast<double2*>(ca) = make_double2(da[0], da[1]);
                    if (on1) *reinterpret_cast<double2*>(cb) = make_double2(db[0], db[1]);
                }
            }
#pragma unroll
            for (int g4 = 0; g4 < 16; g4 += 4) {
                if (g4 >= nact) break;
                double fa[4][2], fb[4][2];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const bool on = g4 + u < nact;
                    const double* pr = Pn + sty[g4 + u] * 8 * PM_LD + fo;
                    const double* pc = Pn + stx[g4 + u] * 8 * PM_LD + fo;
                    fa[u][0] = on ? -pr[0] : 0.0; fa[u][1] = on ? -pr[4] : 0.0;
                    fb[u][0] = on ? pc[0] : 0.0;  fb[u][1] = on ? pc[4] : 0.0;
                }
#pragma unroll
                for (int u = 0; u < 4; u++) pm_dmma(c[g4 + u], fa[u][0], fb[u][0]);
#pragma unroll
                for (int u = 0; u < 4; u++) pm_dmma(c[g4 + u], fa[u][1], fb[u][1]);
            }
        }
    }
}

// X = L^-1 of a factored 128 x 128 leaf, one CTA per 8-wide block column k: the 16 block columns of the inverse are
// independent forward substitutions,  X_kk = L_kk^-1,  X_jk = -L_jj^-1 sum_{k <= m < j} L_jm X_mk  for j > k.
// 256 threads = 4 groups x (8 x 8 entries); group g takes the terms m = k + g, k + g + 4, ... of the sum.
// This used to be the second half of the leaf kernel (one SM, ~35 us); as 16 CTAs it is ~6 us of the critical path.
constexpr int PI_THREADS = 256;
constexpr int PI_SMEM = (P_N * P_LD + 2 * P_NB * 64 + 4 * 64 + P_N) * 8;
__global__ void __launch_bounds__(PI_THREADS)
leaf_inv_kernel(const double* __restrict__ L, long ldl, double* __restrict__ Dinv) {
    extern __shared__ __align__(16) double S[];
    double* D8 = S + P_N * P_LD;              // [16][8][8] inverses of the diagonal 8 x 8 factors
    double* Xs = D8 + P_NB * 64;              // [16][8][8] block column k of X
    double* part = Xs + P_NB * 64;            // [4][64] partial sums of the groups
    double* dinv = part + 4 * 64;             // [128] 1 / l_ii
    const int tid = threadIdx.x, k = blockIdx.x;
    const int g = tid >> 6, e = tid & 63, r = e >> 3, c = e & 7;
    const int row_lo = k * 8;
    pdl_wait();
    // rows >= 8k of L (columns 8k .. row) into shared memory
    for (int idx = tid; idx < (P_N - row_lo) * (P_N / 2); idx += PI_THREADS) {
        const int row = row_lo + idx / (P_N / 2), cp = (idx % (P_N / 2)) * 2;
        if (cp + 1 >= row_lo && cp <= row) {
            const double2 v = *reinterpret_cast<const double2*>(L + (long)row * ldl + cp);
            S[row * P_LD + cp] = v.x;
            S[row * P_LD + cp + 1] = v.y;
        }
    }
    __syncthreads();
    if (tid < P_N - row_lo) dinv[row_lo + tid] = 1.0 / S[(row_lo + tid) * P_LD + row_lo + tid];
    __syncthreads();
    if (tid < P_NB - k) {
        // inverse of the diagonal 8 x 8 factor j, column by column
        const int j = k + tid;
        const double* Lk = S + (j * 8) * P_LD + j * 8;
        double x[8][8];
#pragma unroll
        for (int cc = 0; cc < 8; cc++) {
#pragma unroll
            for (int rr = 0; rr < 8; rr++) {
                if (rr < cc) x[rr][cc] = 0.0;
                else if (rr == cc) x[rr][cc] = dinv[j * 8 + cc];
                else {
                    double sacc = 0.0;
#pragma unroll
                    for (int m = cc; m < rr; m++) sacc = fma(Lk[rr * P_LD + m], x[m][cc], sacc);
                    x[rr][cc] = -dinv[j * 8 + rr] * sacc;
                }
            }
        }
#pragma unroll
        for (int rr = 0; rr < 8; rr++)
#pragma unroll
            for (int cc = 0; cc < 8; cc++) D8[j * 64 + rr * 8 + cc] = x[rr][cc];
    }
    __syncthreads();
    if (tid < 64) Xs[k * 64 + e] = D8[k * 64 + e];
    __syncthreads();
    for (int j = k + 1; j < P_NB; j++) {
        double acc = 0.0;
        for (int m = k + g; m < j; m += 4) {
            const double* lrow = S + (j * 8 + r) * P_LD + m * 8;
            const double* xm = Xs + m * 64 + c;
#pragma unroll
            for (int i = 0; i < 8; i++) acc = fma(lrow[i], xm[i * 8], acc);
        }
        part[g * 64 + e] = acc;
        __syncthreads();
        if (tid < 64) {
            const double* dd = D8 + j * 64 + r * 8;
            double x = 0.0;
#pragma unroll
            for (int i = 0; i <= 7; i++) {
                const double tot = (part[i * 8 + c] + part[64 + i * 8 + c]) + (part[128 + i * 8 + c] + part[192 + i * 8 + c]);
                x = fma(dd[i], tot, x);          // dd[i] = 0 for i > r
            }
            Xs[j * 64 + e] = -x;
        }
        __syncthreads();
    }
    // block column k of X: zero above the diagonal block
    for (int idx = tid; idx < P_N * 8; idx += PI_THREADS) {
        const int row = idx >> 3, cc = idx & 7, jb = row >> 3;
        Dinv[row * P_N + k * 8 + cc] = jb < k ? 0.0 : Xs[jb * 64 + (row & 7) * 8 + cc];
    }
}

// B (m x 128, in place) <- B L^-T for a factored 128 x 128 leaf L: the TRSM leaf of the Cholesky recursion, by blocked
// substitution on the FP64 tensor pipe.  It needs only L itself, so the 128 x 128 inverse of the leaf (leaf_inv_kernel,
// 19 us) is no longer on the factorisation's critical path -- it moves to the inverse factor's side stream.
// One warp owns 8 rows of B as 16 accumulator fragments (8 x 8 blocks, lane l holds [l/4][2(l%4)], [l/4][2(l%4)+1]);
// right-looking over the 16 block columns: X_m = R_m L_mm^-T (2 DMMAs against the 8 x 8 inverse of the diagonal block,
// computed per CTA by 128 threads, one column each), then R_k -= X_m L_km^T for every k > m (2 DMMAs per block, B
// fragments straight from the shared-memory copy of L, row stride 132 doubles: conflict-free).  Fully unrolled:
// 272 DMMAs per warp, ~150 clocks of latency per block column.
constexpr int TS_WARPS = 4, TS_THREADS = 32 * TS_WARPS, TS_LD = 132;
constexpr int TS_SMEM = (P_N * TS_LD + P_NB * 8 * PM_LD) * 8;
__global__ void __launch_bounds__(TS_THREADS, 1)
trsm128_kernel(double* __restrict__ B, long ldb, int m, const double* __restrict__ L, long ldl) {
    extern __shared__ __align__(16) double tsm[];
    double* Ls = tsm;                                     // [128][TS_LD], lower triangle of L
    double* Li = tsm + P_N * TS_LD;                       // [16][8][PM_LD] inverses of the diagonal 8 x 8 blocks
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int fr = lane >> 2, fq = lane & 3;
    pdl_wait();
    for (int idx = tid; idx < P_N * (P_N / 2); idx += TS_THREADS) {
        const int row = idx / (P_N / 2), cp = (idx % (P_N / 2)) * 2;
        if (cp <= row) cp_async16(Ls + row * TS_LD + cp, L + (long)row * ldl + cp);   // all copies in flight at once
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();
    {
        // column cc of the inverse of diagonal block j:  x_cc = 1 / l_cc,  x_i = -(sum_{cc <= q < i} l_iq x_q) / l_ii
        const int j = tid >> 3, cc = tid & 7;
        const double* Lj = Ls + (j * 8) * TS_LD + j * 8;
        double x[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < i; q++) acc = fma(Lj[i * TS_LD + q], x[q], acc);
            const double di = 1.0 / Lj[i * TS_LD + i];
            x[i] = i < cc ? 0.0 : (i == cc ? di : -acc * di);
            Li[(j * 8 + i) * PM_LD + cc] = x[i];
        }
    }
    __syncthreads();
    const long r0 = ((long)blockIdx.x * TS_WARPS + warp) * 8;
    if (r0 >= m) return;
    double* Bw = B + (r0 + fr) * ldb + 2 * fq;
    double c[P_NB][2];
#pragma unroll
    for (int k = 0; k < P_NB; k++) {
        const double2 v = *reinterpret_cast<const double2*>(Bw + 8 * k);
        c[k][0] = v.x; c[k][1] = v.y;
    }
    const int src0 = (lane & ~3) + (fq >> 1), src1 = src0 + 2;
    const bool odd = fq & 1;
#pragma unroll
    for (int mb = 0; mb < P_NB; mb++) {
        // accumulator layout -> A fragments of R_m, then X_m = R_m L_mm^-T
        double a0, a1;
        {
            const double x00 = __shfl_sync(0xffffffffu, c[mb][0], src0), x01 = __shfl_sync(0xffffffffu, c[mb][1], src0);
            const double x10 = __shfl_sync(0xffffffffu, c[mb][0], src1), x11 = __shfl_sync(0xffffffffu, c[mb][1], src1);
            a0 = odd ? x01 : x00; a1 = odd ? x11 : x10;
        }
        double x[2] = {0.0, 0.0};
        pm_dmma(x, a0, Li[(mb * 8 + fr) * PM_LD + fq]);
        pm_dmma(x, a1, Li[(mb * 8 + fr) * PM_LD + 4 + fq]);
        c[mb][0] = x[0]; c[mb][1] = x[1];
        if (mb + 1 < P_NB) {
            const double x00 = __shfl_sync(0xffffffffu, x[0], src0), x01 = __shfl_sync(0xffffffffu, x[1], src0);
            const double x10 = __shfl_sync(0xffffffffu, x[0], src1), x11 = __shfl_sync(0xffffffffu, x[1], src1);
            a0 = -(odd ? x01 : x00); a1 = -(odd ? x11 : x10);
            // consecutive DMMAs share their A fragment (operand reuse), and the two updates of a block are 15 - mb
            // instructions apart instead of back to back
#pragma unroll
            for (int k = mb + 1; k < P_NB; k++) pm_dmma(c[k], a0, Ls[(k * 8 + fr) * TS_LD + mb * 8 + fq]);
#pragma unroll
            for (int k = mb + 1; k < P_NB; k++) pm_dmma(c[k], a1, Ls[(k * 8 + fr) * TS_LD + mb * 8 + 4 + fq]);
        }
    }
    pdl_launch_dependents();
#pragma unroll
    for (int k = 0; k < P_NB; k++) *reinterpret_cast<double2*>(Bw + 8 * k) = make_double2(c[k][0], c[k][1]);
}

static int launch_trsm128(double* B, long ldb, long m, const double* L, long ldl, cudaStream_t st) {
    static bool attr_dev[64] = {};
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(trsm128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM));
        attr_done = true;
    }
    const unsigned grid = (unsigned)((m + 8 * TS_WARPS - 1) / (8 * TS_WARPS));
    GPK_CUDA(launch_pdl(trsm128_kernel, dim3(grid), dim3(TS_THREADS), TS_SMEM, st, B, ldb, (int)m, L, ldl));
    GPK_LAUNCHED();
    return 0;
}

// the leaf's factorisation, and (separately: it is off the factorisation's critical path) the inverse of its factor
static int launch_potf2(double* A, long lda, int* info, int row_offset, cudaStream_t st) {
    GPK_CUDA(launch_pdl(potf2_mma_kernel, dim3(1), dim3(PM_THREADS), 0, st, A, lda, info, row_offset));
    GPK_LAUNCHED();
    return 0;
}
static int launch_leaf_inv(const double* A, long lda, double* Dinv, cudaStream_t st) {
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(leaf_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PI_SMEM));
        attr_done = true;
    }
    GPK_CUDA(launch_pdl(leaf_inv_kernel, dim3(P_NB), dim3(PI_THREADS), PI_SMEM, st, A, lda, Dinv));
    GPK_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// small memory-bound helpers
// ------------------------------------------------------------------------------------------------
// dst[c][r] = src[r][c]  for a rows x cols block
__global__ void transpose_kernel(const double* __restrict__ src, long lds, double* __restrict__ dst, long ldd, int rows,
                                 int cols) {
    __shared__ double tile[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    pdl_wait();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = by + i, c = bx + threadIdx.x;
        if (r < rows && c < cols) tile[i][threadIdx.x] = src[(long)r * lds + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = bx + i, r = by + threadIdx.x;
        if (r < rows && c < cols) dst[(long)c * ldd + r] = tile[threadIdx.x][i];
    }
}

static int launch_transpose(const double* src, long lds, double* dst, long ldd, long rows, long cols, cudaStream_t st) {
    dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
    GPK_CUDA(launch_pdl(transpose_kernel, grid, dim3(32, 8), 0, st, src, lds, dst, ldd, (int)rows, (int)cols));
    GPK_LAUNCHED();
    return 0;
}

// upper[r][c] = lower[c][r] for c > r (mirror the computed lower triangle of a symmetric matrix)
__global__ void mirror_lower_kernel(double* __restrict__ A, long lda, int n) {
    __shared__ double tile[32][33];
    const int bx = blockIdx.x, by = blockIdx.y;   // tile (by, bx) with by >= bx is the source
    if (by < bx) return;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = by * 32 + i, c = bx * 32 + threadIdx.x;
        tile[i][threadIdx.x] = (r < n && c < n) ? A[(long)r * lda + c] : 0.0;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = bx * 32 + i, c = by * 32 + threadIdx.x;      // destination element (r, c) = source (c, r)
        if (r < n && c < n && c > r) A[(long)r * lda + c] = tile[threadIdx.x][i];
    }
}

// y[i] = sum_{k<=i} A[i][k] x[k]  (mode 0, lower)   or   sum_{k>=i} A[i][k] x[k]  (mode 1, upper); one warp per row
__global__ void trmv_rows_kernel(const double* __restrict__ A, long lda, int n, const double* __restrict__ x,
                                 double* __restrict__ y, int mode) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const int k0 = mode == 0 ? 0 : (row & ~1);
    const int k1 = mode == 0 ? ((row + 2) & ~1) : n;      // exclusive, even; the other triangle holds exact zeros
    const double* a = A + (long)row * lda;
    double s = 0.0;
    for (int k = k0 + 2 * lane; k < k1; k += 64) {
        const double2 av = *reinterpret_cast<const double2*>(a + k);
        const double2 xv = *reinterpret_cast<const double2*>(x + k);
        s = fma(av.x, xv.x, s);
        s = fma(av.y, xv.y, s);
    }
    s = warp_sum(s);
    if (lane == 0) y[row] = s;
}

// out[0] = sum t_i^2 ; out[1] = 2 sum log L_ii   (single block)
__global__ void fit_scalars_kernel(const double* __restrict__ t, const double* __restrict__ L, long ldl, int n,
                                   double* __restrict__ out) {
    __shared__ double sq[32], sl[32];
    double q = 0.0, l = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        q = fma(t[i], t[i], q);
        l += log(L[(long)i * ldl + i]);
    }
    q = warp_sum(q);
    l = warp_sum(l);
    if ((threadIdx.x & 31) == 0) { sq[threadIdx.x >> 5] = q; sl[threadIdx.x >> 5] = l; }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int nw = blockDim.x >> 5;
        q = threadIdx.x < nw ? sq[threadIdx.x] : 0.0;
        l = threadIdx.x < nw ? sl[threadIdx.x] : 0.0;
        q = warp_sum(q);
        l = warp_sum(l);
        if (threadIdx.x == 0) { out[0] = q; out[1] = 2.0 * l; }
    }
}

// ------------------------------------------------------------------------------------------------
// recursions (host side; every step is enqueued on `st`, no synchronisation)
// ------------------------------------------------------------------------------------------------
struct Rec {
    cudaStream_t st;
    double* A;      // matrix being factored (lower triangle), lda
    long lda;
    double* Dinv;   // (n/128) inverted diagonal blocks, 128x128 each
    int* info;
};

static long split(long n) { return (n / TILE / 2) * TILE; }   // n >= 256, multiple of 128

// B (m x n at rows r0.., cols c0..) <- B * L^-T, L = A[c0:c0+n, c0:c0+n]
static int trsm_rec(const Rec& r, long r0, long m, long c0, long n) {
    double* Bp = r.A + r0 * r.lda + c0;
    if (n == TILE)    // in place, by substitution with the factored leaf itself (its inverse is not needed here)
        return launch_trsm128(Bp, r.lda, m, r.A + c0 * r.lda + c0, r.lda, r.st);
    const long n1 = split(n), n2 = n - n1;
    GPK_TRY(trsm_rec(r, r0, m, c0, n1));
    // B2 -= B1 * Lb^T,  Lb = A[c0+n1 : c0+n, c0 : c0+n1]
    GPK_TRY(launch_dgemm_nt(gemm_args(Bp, r.lda, r.A + (c0 + n1) * r.lda + c0, r.lda, Bp + n1, r.lda, m, n2, n1, -1.0,
                                      1.0, 0, 0), 0, 1, r.st));
    return trsm_rec(r, r0, m, c0 + n1, n2);
}

static int potrf_rec(const Rec& r, long off, long n) {
    if (n == TILE) {
        GPK_TRY(launch_potf2(r.A + off * r.lda + off, r.lda, r.info, (int)off, r.st));
        return launch_leaf_inv(r.A + off * r.lda + off, r.lda, r.Dinv + (off / TILE) * TILE * TILE, r.st);
    }
    const long n1 = split(n), n2 = n - n1;
    GPK_TRY(potrf_rec(r, off, n1));
    GPK_TRY(trsm_rec(r, off + n1, n2, off, n1));
    double* A21 = r.A + (off + n1) * r.lda + off;
    double* A22 = r.A + (off + n1) * r.lda + off + n1;
    GPK_TRY(launch_dgemm_nt(gemm_args(A21, r.lda, A21, r.lda, A22, r.lda, n2, n2, n1, -1.0, 1.0, 0, 1), 0, 1, r.st));
    return potrf_rec(r, off + n1, n2);
}

struct RecInv {
    cudaStream_t st;
    const double* L; long ldl;
    const double* Dinv;
    double* W; long ldw;
    double* U; long ldu;
    double* T;      // scratch, at least (n/2 rounded) x (n/2 rounded) doubles
};

static int trtri_rec(const RecInv& r, long off, long n) {
    if (n == TILE) {
        const double* Di = r.Dinv + (off / TILE) * TILE * TILE;
        GPK_CUDA(cudaMemcpy2DAsync(r.W + off * r.ldw + off, r.ldw * 8, Di, TILE * 8, TILE * 8, TILE,
                                   cudaMemcpyDeviceToDevice, r.st));
        return launch_transpose(Di, TILE, r.U + off * r.ldu + off, r.ldu, TILE, TILE, r.st);
    }
    const long n1 = split(n), n2 = n - n1;
    GPK_TRY(trtri_rec(r, off, n1));
    GPK_TRY(trtri_rec(r, off + n1, n2));
    // T (n1 x n2) = U11 * L21^T     (= (L21 W11)^T; U11 upper triangular)
    const double* U11 = r.U + off * r.ldu + off;
    const double* L21 = r.L + (off + n1) * r.ldl + off;
    GPK_TRY(launch_dgemm_nt(gemm_args(U11, r.ldu, L21, r.ldl, r.T, n2, n1, n2, n1, 1.0, 0.0, 2, 0), 0, 1, r.st));
    // W21 (n2 x n1) = - W22 * T^T   (W22 lower triangular)
    const double* W22 = r.W + (off + n1) * r.ldw + off + n1;
    double* W21 = r.W + (off + n1) * r.ldw + off;
    GPK_TRY(launch_dgemm_nt(gemm_args(W22, r.ldw, r.T, n2, W21, r.ldw, n2, n1, n2, -1.0, 0.0, 1, 0), 0, 1, r.st));
    // U12 = W21^T
    return launch_transpose(W21, r.ldw, r.U + off * r.ldu + off + n1, r.ldu, n2, n1, r.st);
}

// ------------------------------------------------------------------------------------------------
// Factorisation with the inverse factor computed BESIDE it (gpk_potrf_trtri).  The Cholesky recursion is a chain of
// small dependent kernels at its bottom levels (a 128-leaf is ONE CTA on 148 SMs); W = L^-1 of an already factored
// diagonal block only needs that block, so all of trtri's work goes to a side stream in the recursion's post-order
// and fills the idle SMs.  Events carry the two dependencies main -> side: "leaf factored + its inverse emitted"
// and "L21 of this node final"; the side stream's own order gives W11, W22 before a node's combine.
// ------------------------------------------------------------------------------------------------
struct FitSide {
    cudaStream_t side = nullptr;
    std::vector<cudaEvent_t> pool;
    size_t next = 0;
    int get(cudaEvent_t* e) {
        if (next == pool.size()) {
            cudaEvent_t ev;
            GPK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            pool.push_back(ev);
        }
        *e = pool[next++];
        return 0;
    }
};
static FitSide g_fit_side[64];

// side stream + `events` pooled events of the current device (gpk_init); gpk_potrf_trtri creates them lazily otherwise
int fit_side_prepare(int events) {
    FitSide& fs = g_fit_side[current_device()];
    if (!fs.side) GPK_CUDA(cudaStreamCreateWithFlags(&fs.side, cudaStreamNonBlocking));
    while ((int)fs.pool.size() < events) {
        cudaEvent_t ev;
        GPK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        fs.pool.push_back(ev);
    }
    return 0;
}
void fit_side_release() {
    FitSide& fs = g_fit_side[current_device()];
    for (cudaEvent_t ev : fs.pool) cudaEventDestroy(ev);
    fs.pool.clear();
    fs.next = 0;
    if (fs.side) cudaStreamDestroy(fs.side);
    fs.side = nullptr;
}

struct RecBoth {
    Rec f;             // factorisation, on f.st (the caller's stream)
    RecInv v;          // inverse factor, on v.st (the side stream)
    FitSide* fs;
};

static int potrf_trtri_rec(const RecBoth& r, long off, long n) {
    cudaEvent_t ev;
    if (n == TILE) {
        ozg_set_partition(1);
        GPK_TRY(launch_potf2(r.f.A + off * r.f.lda + off, r.f.lda, r.f.info, (int)off, r.f.st));
        GPK_TRY(r.fs->get(&ev));
        GPK_CUDA(cudaEventRecord(ev, r.f.st));
        GPK_CUDA(cudaStreamWaitEvent(r.v.st, ev, 0));
        // the leaf's inverse only feeds the inverse factor: side stream (the TRSM leaves substitute with L itself)
        GPK_TRY(launch_leaf_inv(r.f.A + off * r.f.lda + off, r.f.lda, r.f.Dinv + (off / TILE) * TILE * TILE, r.v.st));
        ozg_set_partition(2);
        return trtri_rec(r.v, off, TILE);                       // leaf: copy + transpose of the emitted inverse
    }
    const long n1 = split(n), n2 = n - n1;
    GPK_TRY(potrf_trtri_rec(r, off, n1));
    ozg_set_partition(1);
    GPK_TRY(trsm_rec(r.f, off + n1, n2, off, n1));
    GPK_TRY(r.fs->get(&ev));
    GPK_CUDA(cudaEventRecord(ev, r.f.st));                       // L21 final
    double* A21 = r.f.A + (off + n1) * r.f.lda + off;
    double* A22 = r.f.A + (off + n1) * r.f.lda + off + n1;
    GPK_TRY(launch_dgemm_nt(gemm_args(A21, r.f.lda, A21, r.f.lda, A22, r.f.lda, n2, n2, n1, -1.0, 1.0, 0, 1), 0, 1, r.f.st));
    GPK_TRY(potrf_trtri_rec(r, off + n1, n2));
    // combine on the side stream (W11 / U11 and W22 are already ordered there)
    ozg_set_partition(2);
    GPK_CUDA(cudaStreamWaitEvent(r.v.st, ev, 0));
    const RecInv& v = r.v;
    const double* U11 = v.U + off * v.ldu + off;
    const double* L21 = v.L + (off + n1) * v.ldl + off;
    GPK_TRY(launch_dgemm_nt(gemm_args(U11, v.ldu, L21, v.ldl, v.T, n2, n1, n2, n1, 1.0, 0.0, 2, 0), 0, 1, v.st));
    const double* W22 = v.W + (off + n1) * v.ldw + off + n1;
    double* W21 = v.W + (off + n1) * v.ldw + off;
    GPK_TRY(launch_dgemm_nt(gemm_args(W22, v.ldw, v.T, n2, W21, v.ldw, n2, n1, n2, -1.0, 0.0, 1, 0), 0, 1, v.st));
    return launch_transpose(W21, v.ldw, v.U + off * v.ldu + off + n1, v.ldu, n2, n1, v.st);
}

}  // namespace gpk

bool gpk::pdl_enabled() {
    static const bool on = !(getenv("GPK_PDL") && atoi(getenv("GPK_PDL")) == 0);
    return on;
}

// ================================================================================================
// C ABI
// ================================================================================================
using namespace gpk;

extern "C" {

const char* gpk_last_error(void) { return g_err; }
long gpk_launch_count(void) { return g_launches.load(); }
void gpk_reset_launch_count(void) { g_launches.store(0); }
int gpk_abi_version(void) { return 2; }

// Per-device library state (what the compute entries would otherwise create on first use): the exp table of the
// prediction front end (8 KB, one cudaMalloc + copy + synchronisation of `stream`), the side stream of
// gpk_potrf_trtri and a pool of events for it.  Idempotent; call once per device before timing-sensitive work.
int gpk_init(void* stream) {
    GPK_ARG(exp_table((cudaStream_t)stream) != nullptr, "could not allocate the exp table");
    GPK_TRY(oz_side_prepare());
    return fit_side_prepare(1024);
}
// Frees the current device's library state: exp table, side stream, events, cached int8-GEMM item lists, and forgets
// the registered int8-GEMM workspace (which stays the caller's).  The library can be used again afterwards.
int gpk_shutdown(void) {
    GPK_CUDA(cudaDeviceSynchronize());
    exp_table_release();
    fit_side_release();
    oz_side_release();
    ozg_items_release();
    return 0;
}

int gpk_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc, long M, long N, long K,
                 double alpha, double beta, int a_tri, int c_lower, void* stream) {
    return launch_dgemm_nt(gemm_args(A, lda, B, ldb, C, ldc, M, N, K, alpha, beta, a_tri, c_lower), 0, 1,
                           (cudaStream_t)stream);
}

int gpk_potrf(double* A, long lda, long npad, double* dinv, int* info_dev, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    GPK_ARG(lda >= npad && lda % 2 == 0, "lda");
    GPK_ARG(A && dinv && info_dev, "null pointer");
    Rec r{(cudaStream_t)stream, A, lda, dinv, info_dev};
    GPK_CUDA(cudaMemsetAsync(info_dev, 0, sizeof(int), r.st));
    return potrf_rec(r, 0, npad);
}

int gpk_trtri(const double* L, long ldl, long npad, const double* dinv, double* W, long ldw, double* U, long ldu,
              double* tmp, long tmp_elems, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    GPK_ARG(L && dinv && W && U, "null pointer");
    GPK_ARG(ldw >= npad && ldu >= npad && ldw % 2 == 0 && ldu % 2 == 0, "ldw/ldu");
    if (npad > TILE) {
        const long n1 = split(npad), n2 = npad - n1;
        GPK_ARG(tmp && tmp_elems >= n1 * n2, "scratch too small (need split(n) * (n - split(n)) doubles)");
    }
    cudaStream_t st = (cudaStream_t)stream;
    GPK_CUDA(cudaMemset2DAsync(W, ldw * 8, 0, npad * 8, npad, st));
    GPK_CUDA(cudaMemset2DAsync(U, ldu * 8, 0, npad * 8, npad, st));
    RecInv r{st, L, ldl, dinv, W, ldw, U, ldu, tmp};
    return trtri_rec(r, 0, npad);
}

// gpk_potrf followed by gpk_trtri, with the inverse factor computed on an internal side stream beside the
// factorisation (fork after the caller's prior work, join before returning control of `stream`): same results,
// same kernels, shorter critical path.
int gpk_potrf_trtri(double* A, long lda, long npad, double* dinv, int* info_dev, double* W, long ldw, double* U, long ldu,
                    double* tmp, long tmp_elems, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    GPK_ARG(lda >= npad && lda % 2 == 0, "lda");
    GPK_ARG(A && dinv && info_dev && W && U, "null pointer");
    GPK_ARG(ldw >= npad && ldu >= npad && ldw % 2 == 0 && ldu % 2 == 0, "ldw/ldu");
    if (npad > TILE) {
        const long n1 = split(npad), n2 = npad - n1;
        GPK_ARG(tmp && tmp_elems >= n1 * n2, "scratch too small (need split(n) * (n - split(n)) doubles)");
    }
    int dev = 0;
    cudaGetDevice(&dev);
    GPK_ARG(dev >= 0 && dev < 64, "device index");
    FitSide& fs = g_fit_side[dev];
    if (!fs.side) GPK_CUDA(cudaStreamCreateWithFlags(&fs.side, cudaStreamNonBlocking));
    fs.next = 0;
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t ev;
    GPK_TRY(fs.get(&ev));
    GPK_CUDA(cudaEventRecord(ev, st));                            // fork: the side stream starts after the caller's prior work
    GPK_CUDA(cudaStreamWaitEvent(fs.side, ev, 0));
    GPK_CUDA(cudaMemsetAsync(info_dev, 0, sizeof(int), st));
    GPK_CUDA(cudaMemset2DAsync(W, ldw * 8, 0, npad * 8, npad, fs.side));
    GPK_CUDA(cudaMemset2DAsync(U, ldu * 8, 0, npad * 8, npad, fs.side));
    RecBoth r{Rec{st, A, lda, dinv, info_dev}, RecInv{fs.side, A, lda, dinv, W, ldw, U, ldu, tmp}, &fs};
    const int rc = potrf_trtri_rec(r, 0, npad);
    ozg_set_partition(0);
    GPK_TRY(rc);
    GPK_TRY(fs.get(&ev));
    GPK_CUDA(cudaEventRecord(ev, fs.side));                       // join
    GPK_CUDA(cudaStreamWaitEvent(st, ev, 0));
    return 0;
}

long gpk_trtri_scratch_elems(long npad) {
    if (npad <= TILE) return 0;
    const long n1 = split(npad);
    return n1 * (npad - n1);
}

int gpk_kinv(const double* U, long ldu, long npad, double* Kinv, long ldk, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    // lower tiles of U U^T on the tensor pipe, then mirror
    GPK_TRY(launch_dgemm_nt(gemm_args(U, ldu, U, ldu, Kinv, ldk, npad, npad, npad, 1.0, 0.0, 2, 1), 0, 1, st));
    dim3 grid((unsigned)(npad / 32), (unsigned)(npad / 32));
    mirror_lower_kernel<<<grid, dim3(32, 8), 0, st>>>(Kinv, ldk, (int)npad);
    GPK_LAUNCHED();
    return 0;
}

// rows [r0, r1) of K^-1 = U U^T, columns [0, ncols) (row-split MLL gradient: every rank holds U and forms only what its
// share of the trace reads -- all columns, or with the symmetric trace only ncols = r1; no mirror step).  Row i only needs k >= i: upper-triangular A window.
int gpk_kinv_rows(const double* U, long ldu, long npad, long r0, long r1, long ncols, double* Kinv, long ldk, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    GPK_ARG(r0 >= 0 && r1 > r0 && r1 <= npad && r0 % TILE == 0 && r1 % TILE == 0, "row range must be 128-aligned inside [0, npad]");
    GPK_ARG(ncols > 0 && ncols <= npad && ncols % TILE == 0, "ncols must be a multiple of 128 in (0, npad]");
    GPK_ARG(U && Kinv && ldu >= npad && ldk >= npad, "null pointer / leading dimensions");
    return launch_dgemm_nt(gemm_args(U + r0 * ldu + r0, ldu, U + r0, ldu, Kinv + r0 * ldk, ldk, r1 - r0, ncols, npad - r0,
                                     1.0, 0.0, 2, 0), 0, 1, (cudaStream_t)stream);
}

// t = W y ; alpha = U t ; scal[0] = y^T K^-1 y ; scal[1] = log det K
int gpk_solve_alpha(const double* W, long ldw, const double* U, long ldu, const double* L, long ldl, long npad,
                    const double* y, double* t_work, double* alpha, double* scal, void* stream) {
    GPK_ARG(npad > 0 && npad % TILE == 0, "npad must be a positive multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    const int wpb = 8;
    trmv_rows_kernel<<<(unsigned)((npad + wpb - 1) / wpb), wpb * 32, 0, st>>>(W, ldw, (int)npad, y, t_work, 0);
    GPK_LAUNCHED();
    trmv_rows_kernel<<<(unsigned)((npad + wpb - 1) / wpb), wpb * 32, 0, st>>>(U, ldu, (int)npad, t_work, alpha, 1);
    GPK_LAUNCHED();
    fit_scalars_kernel<<<1, 1024, 0, st>>>(t_work, L, ldl, (int)npad, scal);
    GPK_LAUNCHED();
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// posterior variance  var_f(z) = sf2 - || W k(X,z) ||^2 ,  W = L^-1   (gpmpc/gpmpc.py:443-445 keeps only
// this diagonal of the posterior covariance).  Per chunk of queries: (1) K_ZX written once to HBM by the
// SIMT cross kernel, (2) triangular DMMA GEMM V = W K_ZX^T whose epilogue reduces V to per-tile column
// sums of squares (V is never stored, the Q x Q covariance is never formed), (3) finalise.
// ------------------------------------------------------------------------------------------------
extern "C" int gpk_cross_kernel(int dtype_a, const void* A, long lda, long na, long strideA, const double* B, long ldb,
                                long nb, long strideB, int d, const double* hyp, double* out, long ldo, long strideO,
                                long rows_pad, long cols_pad, int mode, int batch, void* stream);
extern "C" int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                                const double* hyp, void* var, long stride_var, int batch, void* stream);

// V = W K_ZX^T reduced to per-tile column sums of squares (the dominant kernel of the variance path),
// exposed on its own so that tests and bench.py can drive / time it in isolation.
extern "C" int gpk_var_colsq(const double* W, long ldw, long strideW, const double* kzx, long ldk, long strideK,
                             long npad, long qpad, double* colsq, long ld_colsq, long stride_colsq, int batch,
                             void* stream) {
    GPK_ARG(W && kzx && colsq, "null pointer");
    GPK_ARG(npad % TILE == 0 && qpad % TILE == 0 && ld_colsq >= qpad, "npad / qpad / ld_colsq");
    GemmParams p = gemm_args(W, ldw, kzx, ldk, nullptr, 0, npad, qpad, npad, 1.0, 0.0, 1, 0);
    p.strideA = strideW;
    p.strideB = strideK;
    p.colsq = colsq;
    p.ld_colsq = ld_colsq;
    p.stride_colsq = stride_colsq;
    return launch_dgemm_nt(p, 1, batch, (cudaStream_t)stream);
}

extern "C" long gpk_predict_var_work_bytes(long q_chunk, long npad, int batch) {
    const long qc = (q_chunk + TILE - 1) / TILE * TILE;
    return (qc * npad + (npad / TILE) * qc) * 8 * batch;
}

extern "C" int gpk_predict_var(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                               const double* X, long ldx, long n, long strideX, const double* W, long ldw,
                               long strideW, long npad, void* var, long stride_var, void* work, long work_bytes,
                               int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && X && W && var && work, "null pointer");
    GPK_ARG(npad % TILE == 0 && npad >= n && n > 0 && batch >= 1, "n / npad / batch");
    if (q == 0) return 0;
    const int tiles_m = (int)(npad / TILE);
    // largest query chunk (multiple of 128) whose K_ZX block and column partials fit the scratch
    long qc = work_bytes / ((npad + tiles_m) * 8 * batch);
    qc = qc / TILE * TILE;
    GPK_ARG(qc >= TILE, "scratch too small (see gpk_predict_var_work_bytes)");
    const long q_round = (q + TILE - 1) / TILE * TILE;
    if (qc > q_round) qc = q_round;
    double* kzx = (double*)work;                       // [batch][qc][npad]
    double* colsq = kzx + (long)batch * qc * npad;     // [batch][tiles_m][qc]
    const int esz = dtype == 0 ? 4 : 8;
    cudaStream_t st = (cudaStream_t)stream;
    for (long q0 = 0; q0 < q; q0 += qc) {
        const long qn = (q - q0 < qc) ? (q - q0) : qc;             // valid queries in this chunk
        const long qp = (qn + TILE - 1) / TILE * TILE;             // padded
        const void* Zc = (const char*)Z + q0 * ldz * esz;
        GPK_TRY(gpk_cross_kernel(dtype, Zc, ldz, qn, strideZ, X, ldx, n, strideX, d, hyp, kzx, npad, qc * npad, qp,
                                 npad, 0, batch, st));
        GPK_TRY(gpk_var_colsq(W, ldw, strideW, kzx, npad, qc * npad, npad, qp, colsq, qc, (long)tiles_m * qc, batch, st));
        void* vout = (char*)var + q0 * esz;
        GPK_TRY(gpk_var_finalize(dtype, colsq, qc, (long)tiles_m * qc, tiles_m, qn, hyp, vout, stride_var, batch, st));
    }
    return 0;
}
