// Posterior variance on the tcgen05 int8 tensor pipe (see ozaki.cuh for the scheme and the operand images).
//
//   var_f(z) = sf2 - || W k(X,z) ||^2 ,  W = L^-1          gpmpc/gpmpc.py:443-445 keeps this diagonal of
//                                                           gp.likelihood(gp(z)).covariance_matrix
// Kernels in this file:
//   oz_row_scale_kernel   2^ew_i > max_k |W[i][k]|  (per row of W, fit time)
//   oz_slice_w_kernel     W (fp64) -> int8 digit images + row scales (fit time)
//   oz_cross_kernel       queries -> E = exp(-r^2/2) -> int8 digit images (per prediction step; K_ZX in fp64
//                         never touches HBM)
//   oz_var_kernel<S>      persistent, warp-specialised: 1-D bulk-TMA producer warp, one-thread tcgen05.mma
//                         issuer, 4 epilogue warps (TMEM -> fp64 Horner recombination -> column sums of squares)
#include "ozaki.cuh"
#include <cmath>
#include <cstdlib>
#include <type_traits>

namespace gpk {

// ------------------------------------------------------------------------------------------------
// W row scales: scale[i] = 2^e with 2^e > max_k |W[i][k]| (frexp exponent), 1 for an all-zero row
// ------------------------------------------------------------------------------------------------
__global__ void oz_row_scale_kernel(const double* __restrict__ W, long ldw, int npad, double* __restrict__ scale) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= npad) return;
    const double* w = W + (long)row * ldw;
    double m = 0.0;
    for (int k = lane; k <= (row | 127) && k < npad; k += 32) m = fmax(m, fabs(w[k]));   // lower triangular + padding
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) {
        int e = 0;
        if (m > 0.0) frexp(m, &e);                       // m = f * 2^e, f in [0.5, 1)
        scale[row] = m > 0.0 ? ldexp(1.0, e) : 1.0;
    }
}

// ------------------------------------------------------------------------------------------------
// W -> digit image.  One CTA per (kb, ib): 64 rows x 64 k; thread = (row, 16-wide k chunk).
// Blocks above the diagonal of the block PAIR (kb > ib | 1) are never read by the GEMM and are skipped.
// ------------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(256)
oz_slice_w_kernel(const double* __restrict__ W, long ldw, int npad, const double* __restrict__ scale,
                  int8_t* __restrict__ img) {
    const int kb = blockIdx.x, ib = blockIdx.y;
    if (kb > (ib | 1)) return;           // the even block of a pair also carries the (all-zero) block kb = ib + 1
    const int kb_total = npad / OZ_BK;
    int8_t* out = img + ((long)ib * kb_total + kb) * S * OZ_B_SLICE;
    const int kc = threadIdx.x & 3, r = threadIdx.x >> 2;    // 4 threads of one row read 64 contiguous doubles
    const long grow = (long)ib * OZ_WB + r;
    const double inv = 1.0 / scale[grow];                    // power of two: exact
    const double* src = W + grow * ldw + (long)kb * OZ_BK + kc * 16;
    int8_t dig[16][S];
#pragma unroll
    for (int k = 0; k < 16; k++) signed_digits<S>(to_fixed<S>(src[k] * inv), dig[k]);
#pragma unroll
    for (int s = 0; s < S; s++) {
        uint32_t w4[4];
#pragma unroll
        for (int v = 0; v < 4; v++)
            w4[v] = (uint32_t)(uint8_t)dig[4 * v][s] | ((uint32_t)(uint8_t)dig[4 * v + 1][s] << 8) |
                    ((uint32_t)(uint8_t)dig[4 * v + 2][s] << 16) | ((uint32_t)(uint8_t)dig[4 * v + 3][s] << 24);
        *reinterpret_cast<uint4*>(out + (long)s * OZ_B_SLICE + oz_tile_off(r, kc * 16)) =
            make_uint4(w4[0], w4[1], w4[2], w4[3]);
    }
}

// ------------------------------------------------------------------------------------------------
// Fused prediction front end (north-star subsystem a + the operand of b):  for a tile of 128 queries and a
// range of training points, E = exp(-1/2 |z~ - x~|^2) is evaluated ONCE in fp64 and used for
//   (1) the int8 digit image of E (operand of the tcgen05 variance contraction), and
//   (2) the posterior mean and input-Jacobian partial sums  m = sum_i a_i E_i,  S_j = sum_i a_i E_i x~_ij
//       (a_i = sf2 alpha_i;  mean = m,  jac_j = (S_j - z~_j m) / ell_j;  gpmpc/gp.py:85 and the CasADi AD of
//       gpmpc/gp.py:12-14 inside acados, gpmpc/gpmpc.py:189-211).
// thread = (query row, every other 16-wide k chunk); digits come from ONE fused multiply-add per value
// (E 2^(8S-2) + 1.5 2^52 leaves the fixed-point integer in the mantissa), a bias of 128 per digit that turns
// the balanced base-256 digits into the plain BYTES of that integer (two's complement = byte ^ 0x80), and
// byte permutes -- the FP64 pipe only sees the GEMMs, exp, and the mean accumulation.
// ------------------------------------------------------------------------------------------------
#ifndef OZ_FRONT_UNROLL
#define OZ_FRONT_UNROLL 1
#endif
struct OzFrontArgs {
    const void* Z; long ldz, q, strideZ;
    const double* X; long ldx, n, strideX;
    const double* alpha; long stride_alpha;  // [gp][>= n]  (only when MJ)
    int d;
    const double* hyp;
    int8_t* img; long strideE;               // digit image of E, [gp][qtile][kb][S][128 x 64]
    int kb_total, kb_per_cta;
    double* mj_part; long stride_mj;         // [gp][split][q][DP + 1] partial (m, S_j)
    const double* exp_tab;                   // 2^(j/1024), j < 1024 (device, built once per process)
    unsigned long long* tdbg;                // experiments only: per-CTA (start, end) global timer of the co-resident kernel
};

__device__ __forceinline__ unsigned long long oz_gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void oz_dmma(double (&c)[2], double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// Both GEMM-shaped parts run on the FP64 tensor pipe (DMMA m8n8k4), one warp per 16 queries:
//   GEMM 1   -1/2 r^2 [8 queries x 8 points] = (-1/2|z~|^2 - 1/2|x~|^2) + z~ [8 x d] x~^T [d x 8]
//   GEMM 2   S [8 queries x d] += (a E) [8 x 4 points] x~ [4 x d]
// The point <-> fragment-column mapping is chosen so that the accumulator elements a thread holds after GEMM 1
// are (i) exactly the A-fragment elements it must supply to GEMM 2 and (ii) four CONSECUTIVE training points
// k = 4t .. 4t+3 of a 16-point unit, i.e. one aligned 32-bit word of each digit row: a warp's digit store is
// one contiguous 128-byte core matrix.  Column c = 2t+u of block b  <->  point 4t + 2b + u.
template <int DP, bool MJ>
struct OzFrontSmem {
    static constexpr int DPB = DP < 8 ? 8 : DP;
    static constexpr int LDA = ((DP + 15) / 16) * 16 + 4;       // = 4 mod 16: conflict-free B fragments of GEMM 1
    static constexpr int LDB = DPB + 1;                         // = 1 mod 4 : conflict-free B fragments of GEMM 2
    alignas(16) double XsA[OZ_BK * LDA];                         // rows in fragment-column order (see above)
    alignas(16) double XsB[MJ ? OZ_BK * LDB : 1];                // rows in natural order
    alignas(16) double ex[OZ_BK];
    alignas(16) double as[OZ_BK];
    double exp_tab[EXP_TAB];
};

// One work item = (GP, 16 NW queries starting at row q_off of query tile qt, split of the training range), NW warps.
template <typename TZ, int S, int DP, bool MJ, int NW, bool IMG>
__device__ __forceinline__ void oz_front_item(const OzFrontArgs& a, OzFrontSmem<DP, MJ>& sm, long gp, int qt, int q_off, int split) {
    constexpr int KS = DP / 4;                           // GEMM 1 k-steps
    constexpr int DPB = OzFrontSmem<DP, MJ>::DPB, NB = DPB / 8;   // GEMM 2 column blocks
    constexpr int LDA = OzFrontSmem<DP, MJ>::LDA, LDB = OzFrontSmem<DP, MJ>::LDB;
    constexpr int NT = 32 * NW;                          // threads of the CTA
    constexpr int UNIT_UNROLL = OZ_FRONT_UNROLL;
    constexpr int PASSES = 4 * OZ_BK / NT;               // staging: 4 threads per training point, NT / 4 points per pass
    double* XsA = sm.XsA; double* XsB = sm.XsB; double* ex = sm.ex; double* as = sm.as;
    const double* exp_tab = sm.exp_tab;
    const double* hyp = a.hyp + gp * HYP_STRIDE;
    const TZ* Z = reinterpret_cast<const TZ*>(a.Z) + gp * a.strideZ;
    const double* X = a.X + gp * a.strideX;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;

    // A fragments of GEMM 1 (scaled queries) for the two 8-query blocks of this warp
    double za[2][KS], ezq[2];
    bool qv[2];
    int qrow[2];
#pragma unroll
    for (int mb = 0; mb < 2; mb++) {
        qrow[mb] = q_off + warp * 16 + mb * 8 + g;
        const long gq = (long)qt * OZ_QT + qrow[mb];
        qv[mb] = gq < a.q;
        double sq = 0.0;
#pragma unroll
        for (int s_ = 0; s_ < KS; s_++) {
            const int j = t + 4 * s_;
            double v = 0.0;
            if (j < a.d && qv[mb]) v = ((double)Z[gq * a.ldz + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
            za[mb][s_] = v;
            sq = fma(v, v, sq);
        }
        sq += __shfl_xor_sync(0xffffffffu, sq, 1);
        sq += __shfl_xor_sync(0xffffffffu, sq, 2);
        ezq[mb] = qv[mb] ? -0.5 * sq : -1.0e300;        // queries beyond q: exp -> 0, like the padding points
    }
    double S2[2][NB][2], msum[2] = {0.0, 0.0};
#pragma unroll
    for (int mb = 0; mb < 2; mb++)
#pragma unroll
        for (int nb = 0; nb < NB; nb++) S2[mb][nb][0] = S2[mb][nb][1] = 0.0;

    const int kb0 = split * a.kb_per_cta;
    const int kb1 = min(a.kb_total, kb0 + a.kb_per_cta);
    // staging role: 4 threads per training point, thread `part` owns dimensions part, part + 4, ...
    const int part = tid & 3;
    constexpr int NPF = (DPB + 3) / 4;
    double pf[PASSES][NPF], pf_alpha[PASSES];
    auto prefetch = [&](int kb) {                        // raw values of the next k block into registers
#pragma unroll
        for (int ps = 0; ps < PASSES; ps++) {
            const int sp = (tid >> 2) + ps * (NT / 4);
            const long gi = (long)kb * OZ_BK + sp;
#pragma unroll
            for (int u = 0; u < NPF; u++) {
                const int j = part + 4 * u;
                pf[ps][u] = (gi < a.n && j < a.d) ? X[gi * a.ldx + j] : 0.0;
            }
            pf_alpha[ps] = 0.0;
            if (MJ && part == 0) pf_alpha[ps] = gi < a.n ? a.alpha[gp * a.stride_alpha + gi] : 0.0;
        }
    };
    if (kb0 < kb1) prefetch(kb0);
    for (int kb = kb0; kb < kb1; kb++) {
        __syncthreads();
#pragma unroll
        for (int ps = 0; ps < PASSES; ps++) {
            const int sp = (tid >> 2) + ps * (NT / 4);
            const int kk = sp & 15;
            const int slotA = (sp & ~15) + ((kk >> 1) & 1) * 8 + 2 * (kk >> 2) + (kk & 1);
            const long gi = (long)kb * OZ_BK + sp;
            double sq = 0.0;
#pragma unroll
            for (int u = 0; u < NPF; u++) {
                const int j = part + 4 * u;
                const double v = (gi < a.n && j < a.d) ? (pf[ps][u] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j] : 0.0;
                if (j < DP) XsA[slotA * LDA + j] = v;
                if (MJ) XsB[sp * LDB + j] = v;
                sq = fma(v, v, sq);
            }
            sq += __shfl_xor_sync(0xffffffffu, sq, 1);
            sq += __shfl_xor_sync(0xffffffffu, sq, 2);
            if (part == 0) {
                ex[sp] = gi < a.n ? -0.5 * sq : -1.0e300;  // padding points: exp -> 0
                as[sp] = MJ ? hyp[HYP_SF2] * pf_alpha[ps] : 0.0;
            }
        }
        __syncthreads();
        if (kb + 1 < kb1) prefetch(kb + 1);              // in flight during the compute below
        int8_t* out = a.img + gp * a.strideE + ((long)qt * a.kb_total + kb) * S * OZ_A_SLICE;
#pragma unroll UNIT_UNROLL
        for (int unit = 0; unit < OZ_BK / 16; unit++) {
            const int p0 = unit * 16 + 4 * t;            // this thread's four consecutive training points
            double b1[2][KS];
#pragma unroll
            for (int b = 0; b < 2; b++)
#pragma unroll
                for (int s_ = 0; s_ < KS; s_++) b1[b][s_] = XsA[(unit * 16 + b * 8 + g) * LDA + t + 4 * s_];
            const double2 ex01 = *reinterpret_cast<const double2*>(&ex[p0]);
            const double2 ex23 = *reinterpret_cast<const double2*>(&ex[p0 + 2]);
            double b2[4][NB], as4[4];
            if (MJ) {
                const double2 as01 = *reinterpret_cast<const double2*>(&as[p0]);
                const double2 as23 = *reinterpret_cast<const double2*>(&as[p0 + 2]);
                as4[0] = as01.x; as4[1] = as01.y; as4[2] = as23.x; as4[3] = as23.y;
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int nb = 0; nb < NB; nb++) b2[i][nb] = XsB[(p0 + i) * LDB + g + 8 * nb];
            }
            // straight-line code for both 8-query blocks (no branch inside a unit: one basic block, so that the
            // independent chains of the two blocks interleave)
            double c[2][2][2];
#pragma unroll
            for (int mb = 0; mb < 2; mb++) {
                c[mb][0][0] = ezq[mb] + ex01.x; c[mb][0][1] = ezq[mb] + ex01.y;
                c[mb][1][0] = ezq[mb] + ex23.x; c[mb][1][1] = ezq[mb] + ex23.y;
            }
#pragma unroll
            for (int s_ = 0; s_ < KS; s_++)
#pragma unroll
                for (int mb = 0; mb < 2; mb++) {
                    oz_dmma(c[mb][0], za[mb][s_], b1[0][s_]);
                    oz_dmma(c[mb][1], za[mb][s_], b1[1][s_]);
                }
            double e[2][4];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int mb = 0; mb < 2; mb++) e[mb][i] = exp_neg_tab(c[mb][i >> 1][i & 1], exp_tab);   // invalid query: ezq = -1e300 -> 0
            if (IMG) {
#pragma unroll
                for (int mb = 0; mb < 2; mb++) {
                    uint32_t w[S];
                    oz_pack_digits4<S>(e[mb], w);
                    int8_t* dst = out + oz_tile_off(qrow[mb], unit * 16 + 4 * t);
#pragma unroll
                    for (int s_ = 0; s_ < S; s_++) *reinterpret_cast<uint32_t*>(dst + (long)s_ * OZ_A_SLICE) = w[s_];
                }
            }
            if (MJ) {
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int mb = 0; mb < 2; mb++) {
                        const double wgt = e[mb][i] * as4[i];
                        msum[mb] += wgt;
#pragma unroll
                        for (int nb = 0; nb < NB; nb++) oz_dmma(S2[mb][nb], wgt, b2[i][nb]);
                    }
            }
        }
    }
    if (MJ) {
#pragma unroll
        for (int mb = 0; mb < 2; mb++) {
            double m = msum[mb];
            m += __shfl_xor_sync(0xffffffffu, m, 1);
            m += __shfl_xor_sync(0xffffffffu, m, 2);
            if (!qv[mb]) continue;
            const long gq = (long)qt * OZ_QT + qrow[mb];
            double* dst = a.mj_part + gp * a.stride_mj + ((long)split * a.q + gq) * (DP + 1);
            if (t == 0) dst[0] = m;
#pragma unroll
            for (int nb = 0; nb < NB; nb++)
#pragma unroll
                for (int cc = 0; cc < 2; cc++) {
                    const int j = 8 * nb + 2 * t + cc;
                    if (j < DP) dst[1 + j] = S2[mb][nb][cc];
                }
        }
    }
}

// stand-alone front end: one CTA (8 warps) per (split, query tile, GP), two CTAs per SM
template <typename TZ, int S, int DP, bool MJ, bool IMG>
__global__ void __launch_bounds__(256, 2) oz_front_kernel(const OzFrontArgs a) {
    __shared__ OzFrontSmem<DP, MJ> sm;
    for (int i = threadIdx.x; i < EXP_TAB; i += 256) sm.exp_tab[i] = a.exp_tab[i];   // visible after the item's first barrier
    oz_front_item<TZ, S, DP, MJ, 8, IMG>(a, sm, blockIdx.z, blockIdx.y, 0, blockIdx.x);
}

// co-resident front end: persistent, at most one 4-warp CTA per SM (152 registers x 128 threads and 29 KB of shared
// memory are what a resident oz_var_kernel CTA leaves free), so that the E image of the NEXT query sub-chunk is
// produced on the FP64 pipe while the tensor pipe contracts the current one.  Items = (GP, query tile, half tile,
// split), static round-robin.
template <typename TZ, int S, int DP, bool MJ>
__global__ void __maxnreg__(152) oz_front_co_kernel(const OzFrontArgs a, int n_split, int qtiles, int batch) {
    __shared__ OzFrontSmem<DP, MJ> sm;
    if (a.tdbg && threadIdx.x == 0) a.tdbg[2 * blockIdx.x] = oz_gtimer();
    for (int i = threadIdx.x; i < EXP_TAB; i += 128) sm.exp_tab[i] = a.exp_tab[i];
    const long total = (long)batch * qtiles * 2 * n_split;
    for (long item = blockIdx.x; item < total; item += gridDim.x) {
        const int split = (int)(item % n_split);
        long r = item / n_split;
        const int half = (int)(r & 1);
        r >>= 1;
        const int qt = (int)(r % qtiles);
        const long gp = r / qtiles;
        oz_front_item<TZ, S, DP, MJ, 4, true>(a, sm, gp, qt, half * 64, split);
    }
    if (a.tdbg && threadIdx.x == 0) a.tdbg[2 * blockIdx.x + 1] = oz_gtimer();
}

// mean = sum of the split partials; jac_j = (S_j - z~_j mean) / ell_j
template <typename TZ, int DP>
__global__ void oz_mj_finalize_kernel(const OzFrontArgs a, int n_split, TZ* __restrict__ mean, long stride_mean,
                                      TZ* __restrict__ jac, long ldj, long stride_jac) {
    const long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long gp = blockIdx.z;
    if (q >= a.q) return;
    const double* hyp = a.hyp + gp * HYP_STRIDE;
    const TZ* Z = reinterpret_cast<const TZ*>(a.Z) + gp * a.strideZ;
    const double* part = a.mj_part + gp * a.stride_mj;
    double acc[DP + 1];
#pragma unroll
    for (int j = 0; j <= DP; j++) acc[j] = 0.0;
    for (int s = 0; s < n_split; s++) {
        const double* w = part + ((long)s * a.q + q) * (DP + 1);
#pragma unroll
        for (int j = 0; j <= DP; j++) acc[j] += w[j];
    }
    mean[gp * stride_mean + q] = (TZ)acc[0];
#pragma unroll
    for (int j = 0; j < DP; j++)
        if (j < a.d) {
            const double zs = ((double)Z[q * a.ldz + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
            jac[gp * stride_jac + q * ldj + j] = (TZ)((acc[1 + j] - zs * acc[0]) * hyp[HYP_INV_ELL + j]);
        }
}

// ------------------------------------------------------------------------------------------------
// the GEMM: persistent CTAs (one per SM), 10 warps
//   warp 8      producer: per stage two 1-D bulk TMA copies (all S digit slices of the E and of the W k-block)
//   warp 9      TMEM allocation + MMA issue (one thread).  For E digit t one instruction multiplies against the
//               CONCATENATION of W digits 0..S-1-t (N = 64 (S-t), split at 256) and accumulates into digit
//               groups t..S-1, which sit side by side in TMEM: S+2 instructions per k-step instead of
//               S(S+1)/2, and each E slice is read from shared memory once per k-step
//   warps 0..7  epilogue: tcgen05.ld the S int32 accumulators, Horner in fp64, row scale, square, and add into
//               the per-query running sum (one query per TMEM lane = per thread; two warps per lane quarter
//               split the 64 columns)
// CTA pairs (cluster of 2): both CTAs work on the SAME query tile and on W row blocks 2j and 2j+1, so each
// loads half of the E k-block and multicasts it into both shared memories (the kernel is bound by L2 -> SM
// bandwidth without this: 72 KB per stage per SM); a stage is released by a multicast tcgen05.commit of both
// MMA issuers.  CTA 0 runs one k-block past its diagonal (zero digits) to stay in lockstep with CTA 1.
// Work item of a pair = (query tile, piece of the W row-block pairs); the pieces of a query tile have equal
// (triangular) work; partial sums part[gp][2 piece + rank][q] are summed by var_finalize_kernel (deterministic).
// ------------------------------------------------------------------------------------------------
struct OzArgs {
    const int8_t* Eimg; long strideE;        // bytes per GP
    const int8_t* Wimg; long strideW;
    long w_kb_bytes;                         // bytes per (row block, k block) of the W image = its digit count * 4 KB
    const double* wscale; long stride_ws;    // [gp][npad]
    const double* hyp;                       // [gp][HYP_STRIDE]
    double* part; long ld_part, stride_part; // [gp][piece][ld_part]
    int qtiles, nbatch, kb_total, n_pieces, total_items;
    int piece_begin[OZ_MAX_PIECES + 1];      // W row-block PAIRS [piece_begin[p], piece_begin[p+1]) of piece p
    int32_t* dbg;                            // optional raw accumulators [gp*qtiles + qt][ib][S][128][64]
    int flags;                               // profiling only: 1 = epilogue skips the TMEM reads and the math
    unsigned long long* tdbg;                // experiments only: per-CTA (start, end) global timer
};

constexpr int OZ_EPI_WARPS = 8;
// 12 warps = three FULL warpgroups: setmaxnreg only completes when every warp of a 4-warp group executes the same
// instruction (a 10-warp CTA hangs on this part: tools/peaks/setmaxnreg_probe.cu); warps 10 and 11 only hand back registers
constexpr int OZ_THREADS = 128 + 32 * OZ_EPI_WARPS;
// the two single-thread roles get the HIGHEST warp ids: the SM's issue arbiter prefers higher warp ids, and the
// MMA issuer must never queue behind the epilogue warps that share its scheduler
constexpr int OZ_PRODUCER_WARP = OZ_EPI_WARPS + 2, OZ_MMA_WARP = OZ_EPI_WARPS + 3;

__device__ __forceinline__ void oz_item_coords(const OzArgs& a, int item, int& qt, int& gp, int& jp0, int& jp1) {
    const int col = item / a.n_pieces;            // consecutive items share the query tile (its E image stays in L2)
    const int piece = a.n_pieces - 1 - (item - col * a.n_pieces);
    gp = col / a.qtiles;
    qt = col - gp * a.qtiles;
    jp0 = a.piece_begin[piece];
    jp1 = a.piece_begin[piece + 1];
}

// Register budget: the kernel is launched with OZ_LAUNCH_REGS registers per thread (384 x 120 = 46,080 of the SM's
// 65,536), the warpgroup of the two single-thread roles hands its registers back (setmaxnreg.dec) and the epilogue
// warps take theirs (8 x 32 x OZ_EPI_REGS + 4 x 32 x OZ_AUX_REGS = 46,080): what is left (19,456 registers, ~45 KB of
// shared memory at S <= 5) holds one 4-warp CTA of oz_front_co_kernel, which produces the next sub-chunk's E image on
// the otherwise idle FP64 pipe.
constexpr int OZ_LAUNCH_REGS = 120;
#ifndef OZ_EPI_REGS
#define OZ_EPI_REGS 160
#define OZ_AUX_REGS 40
#endif
template <int S>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(OZ_LAUNCH_REGS) oz_var_kernel(const OzArgs a) {
    constexpr int NST = oz_stages(S);
    constexpr uint32_t STAGE_BYTES = oz_stage_bytes(S);
    constexpr uint32_t A_BYTES = S * OZ_A_SLICE, B_BYTES = S * OZ_B_SLICE;
    // TMEM: a tile's S digit groups take 64 S of the 512 columns.  Two buffers at columns 0 and SPARE = 512 - 64 S
    // alternate between tiles; they share NOV = 2 S - 8 groups (the last ones of buffer 0 = the first ones of
    // buffer 1).  The epilogue pulls those shared groups into registers first and releases the MMA issuer, so the
    // next tile's main loop overlaps with the rest of the epilogue.  With more than 2 shared groups (S = 6) the
    // registers do not suffice and one buffer is used.
    constexpr int NOV = 2 * S - 8;
    constexpr bool DB = NOV <= 2;
    constexpr uint32_t SPARE = 512 - 64 * S;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[NST], empty_bar[NST], tmem_full_bar, tmem_empty_bar;
    __shared__ uint32_t tmem_base_smem;
    __shared__ double half_sum[OZ_QT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NST; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 2); }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&tmem_empty_bar, 32 * OZ_EPI_WARPS);
        mbar_fence_init();
    }
    if (warp == OZ_MMA_WARP) tmem_alloc(&tmem_base_smem, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // the peer's barriers are initialised before anything is multicast
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    if (a.tdbg && threadIdx.x == 0) a.tdbg[2 * blockIdx.x] = oz_gtimer();

    if (warp >= OZ_EPI_WARPS) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(OZ_AUX_REGS));
    if (warp == OZ_PRODUCER_WARP) {
        // ===== producer (warp-uniform, elected lane issues) =====
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            int qt, gp, jp0, jp1;
            oz_item_coords(a, item, qt, gp, jp0, jp1);
            const int8_t* Ag = a.Eimg + gp * a.strideE + (long)qt * a.kb_total * A_BYTES + rank * (A_BYTES / 2);
            for (int jp = jp1 - 1; jp >= jp0; jp--) {
                // the image may carry more digits than this kernel multiplies (float32 queries use the leading 4)
                const int8_t* Bg = a.Wimg + gp * a.strideW + (long)(2 * jp + rank) * a.kb_total * a.w_kb_bytes;
                for (int kb = 0; kb <= 2 * jp + 1; kb++) {
                    mbar_wait(&empty_bar[st], ph ^ 1);          // both CTAs are done reading this stage
                    mbar_expect_tx_pred(&full_bar[st], STAGE_BYTES, leader);
                    const uint32_t dst = smem_u32(smem) + (uint32_t)st * STAGE_BYTES;
                    tma_bulk_g2s_mcast_pred(dst + rank * (A_BYTES / 2), Ag + (long)kb * A_BYTES, A_BYTES / 2, &full_bar[st],
                                            (uint16_t)3, leader);
                    tma_bulk_g2s_pred(dst + A_BYTES, Bg + (long)kb * a.w_kb_bytes, B_BYTES, &full_bar[st], leader);
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == OZ_MMA_WARP) {
        // ===== MMA issuer: the whole warp runs the loops (warp-uniform), the elected lane issues =====
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0, tph = 0, par = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            int qt, gp, jp0, jp1;
            oz_item_coords(a, item, qt, gp, jp0, jp1);
            for (int jp = jp1 - 1; jp >= jp0; jp--) {
                // the epilogue of the previous tile has drained the columns this tile's buffer shares with it
                mbar_wait(&tmem_empty_bar, tph ^ 1);
                tc_fence_after();
                const uint32_t dbase = tmem_base + (DB ? par * SPARE : 0u);
                par ^= 1;
                for (int kb = 0; kb <= 2 * jp + 1; kb++) {
                    mbar_wait(&full_bar[st], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(smem) + (uint32_t)st * STAGE_BYTES;
                    const uint32_t b_base = a_base + A_BYTES;
#pragma unroll
                    for (int j = 0; j < OZ_BK / 32; j++) {
#pragma unroll
                        for (int t = 0; t < S; t++) {
                            // E digit t against the concatenated W digits 0..S-1-t: N = 64 (S - t) columns landing
                            // in digit groups t..S-1; wider than 256 -> two equal halves (multiples of 16)
                            const uint64_t adesc = umma_smem_desc(a_base + t * OZ_A_SLICE + j * 256, 128, OZ_BK * 8);
                            const uint32_t acc = (kb > 0 || j > 0 || t > 0) ? 1u : 0u;
                            constexpr int dummy = 0; (void)dummy;
                            const int ntot = OZ_WB * (S - t);
                            const int nparts = ntot > 256 ? 2 : 1;
                            const int npart = ntot / nparts;
#pragma unroll
                            for (int h = 0; h < nparts; h++) {
                                // rows [h npart, (h+1) npart) of the concatenation: byte offset = rows * 64
                                const uint64_t bdesc = umma_smem_desc(b_base + h * npart * OZ_BK + j * 256, 128, OZ_BK * 8);
                                umma_i8_pred(dbase + (uint32_t)(t * OZ_WB + h * npart), adesc, bdesc, umma_idesc_i8(npart), acc, leader);
                            }
                        }
                    }
                    tc_commit_mcast_pred(&empty_bar[st], (uint16_t)3, leader);   // frees the stage in BOTH CTAs
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
                tc_commit_pred(&tmem_full_bar, leader);
                tph ^= 1;
            }
        }
    } else if (warp < OZ_EPI_WARPS) {
        // ===== epilogue: warps 0..7; lane quarter = warp % 4, column half = warp / 4 =====
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(OZ_EPI_REGS));
        const int quarter = warp & 3;
        const int half = warp >> 2;
        const int row = quarter * 32 + lane;                   // query within the tile
        const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 32);
        uint32_t tph = 0, par = 0;
        // one tile through the double-buffered path; PAR = which TMEM buffer holds it
        auto tile_db = [&](auto par_tag, const double* wsb, int32_t* dp, double& sum) {
            constexpr int PAR = decltype(par_tag)::value;
            constexpr int NREST = S - NOV;
            constexpr int OV0 = PAR == 0 ? S - NOV : 0;           // first shared group (in this buffer's numbering)
            constexpr int RS0 = PAR == 0 ? 0 : NOV;               // first non-shared group
            const uint32_t tb = tbase + (PAR ? SPARE : 0u);
            int32_t vo[NOV > 0 ? NOV : 1][4][8];
            int32_t v[2][NREST][8];
#pragma unroll
            for (int g = 0; g < NOV; g++)
#pragma unroll
                for (int c = 0; c < 4; c++) tmem_ld8(tb + (uint32_t)((OV0 + g) * OZ_WB + c * 8), vo[g][c]);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&tmem_empty_bar);                          // the next tile may start accumulating
#pragma unroll
            for (int g = 0; g < NREST; g++) tmem_ld8(tb + (uint32_t)((RS0 + g) * OZ_WB), v[0][g]);
#pragma unroll
            for (int c = 0; c < 4; c++) {
                tmem_ld_wait();
                if (c < 3) {
#pragma unroll
                    for (int g = 0; g < NREST; g++) tmem_ld8(tb + (uint32_t)((RS0 + g) * OZ_WB + (c + 1) * 8), v[(c + 1) & 1][g]);
                }
                int32_t vv[S][8];
#pragma unroll
                for (int g = 0; g < S; g++)
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const bool shared = g >= OV0 && g < OV0 + NOV;
                        vv[g][j] = shared ? vo[shared ? g - OV0 : 0][c][j] : v[c & 1][shared ? 0 : g - RS0][j];
                    }
                if (dp) {
#pragma unroll
                    for (int g = 0; g < S; g++)
#pragma unroll
                        for (int j = 0; j < 8; j++) dp[(long)g * OZ_QT * OZ_WB + c * 8 + j] = vv[g][j];
                }
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    double acc = oz_recombine<S>(vv, j);
                    acc *= __ldg(wsb + c * 8 + j);
                    sum = fma(acc, acc, sum);
                }
            }
        };
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            int qt, gp, jp0, jp1;
            oz_item_coords(a, item, qt, gp, jp0, jp1);
            const double sf2s = a.hyp[(long)gp * HYP_STRIDE + HYP_SF2] * (1.0 / 4096.0);
            const double* ws = a.wscale + gp * a.stride_ws;
            double sum = 0.0;
            for (int jp = jp1 - 1; jp >= jp0; jp--) {
                const int ib = 2 * jp + rank;
                mbar_wait(&tmem_full_bar, tph);
                tph ^= 1;
                tc_fence_after();
                const double* wsb = ws + (long)ib * OZ_WB + half * 32;
                if (a.flags & 1) {
                    tc_fence_before();
                    mbar_arrive(&tmem_empty_bar);
                    continue;
                }
                int32_t* dp = a.dbg ? a.dbg + ((((long)(gp * a.qtiles + qt) * a.kb_total + ib) * S) * OZ_QT + row) * OZ_WB + half * 32
                                    : nullptr;
                if constexpr (DB) {
                    if (par) tile_db(std::integral_constant<int, 1>{}, wsb, dp, sum);
                    else tile_db(std::integral_constant<int, 0>{}, wsb, dp, sum);
                    par ^= 1;
                    continue;
                }
                int32_t v[2][S][8];
#pragma unroll
                for (int g = 0; g < S; g++) tmem_ld8(tbase + (uint32_t)(g * OZ_WB), v[0][g]);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    tmem_ld_wait();
                    if (c < 3) {
#pragma unroll
                        for (int g = 0; g < S; g++) tmem_ld8(tbase + (uint32_t)(g * OZ_WB + (c + 1) * 8), v[(c + 1) & 1][g]);
                    } else {                                   // all TMEM reads of this tile are in registers
                        tc_fence_before();
                        mbar_arrive(&tmem_empty_bar);
                    }
                    if (dp) {
#pragma unroll
                        for (int g = 0; g < S; g++)
#pragma unroll
                            for (int j = 0; j < 8; j++) dp[(long)g * OZ_QT * OZ_WB + c * 8 + j] = v[c & 1][g][j];
                    }
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        double acc = oz_recombine<S>(v[c & 1], j);
                        acc *= __ldg(wsb + c * 8 + j);
                        sum = fma(acc, acc, sum);
                    }
                }
            }
            sum *= sf2s * sf2s;
            // combine the two column halves of each query and write the partial of this piece
            const int piece = a.n_pieces - 1 - (item % a.n_pieces);
            if (half == 1) half_sum[row] = sum;
            asm volatile("bar.sync 1, %0;\n" ::"n"(32 * OZ_EPI_WARPS) : "memory");
            if (half == 0)
                a.part[gp * a.stride_part + (long)(2 * piece + rank) * a.ld_part + (long)qt * OZ_QT + row] = sum + half_sum[row];
            asm volatile("bar.sync 1, %0;\n" ::"n"(32 * OZ_EPI_WARPS) : "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.tdbg && threadIdx.x == 0) a.tdbg[2 * blockIdx.x + 1] = oz_gtimer();
    cluster_sync_all();                      // the peer may still be multicasting into / arriving on this CTA
    if (warp == OZ_MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace gpk

// ================================================================================================
// C ABI
// ================================================================================================
using namespace gpk;

static int oz_sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

// 2^(j/1024) table of the front kernel's exp: computed on the host in long double, rounded to double.  Built by
// gpk_init (or lazily by the first prediction on a device: one cudaMalloc + one stream synchronisation), freed by
// gpk_shutdown.
static double* g_exp_tab[64] = {nullptr};
const double* gpk::exp_table(cudaStream_t st) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return nullptr;
    if (!g_exp_tab[dev]) {
        static double host_tab[EXP_TAB];
        for (int j = 0; j < EXP_TAB; j++) host_tab[j] = (double)exp2l((long double)j / EXP_TAB);
        double* p = nullptr;
        if (cudaMalloc(&p, sizeof(host_tab)) != cudaSuccess) return nullptr;
        if (cudaMemcpyAsync(p, host_tab, sizeof(host_tab), cudaMemcpyHostToDevice, st) != cudaSuccess) return nullptr;
        cudaStreamSynchronize(st);
        g_exp_tab[dev] = p;
    }
    return g_exp_tab[dev];
}
void gpk::exp_table_release() {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !g_exp_tab[dev]) return;
    cudaFree(g_exp_tab[dev]);
    g_exp_tab[dev] = nullptr;
}

static int oz_pad_dim(int d) { return d <= 4 ? 4 : d <= 8 ? 8 : d <= 16 ? 16 : 32; }

// pieces per query tile: the count (<= 8, <= number of W row-block pairs) that wastes the least in the last
// round of the static round-robin over the CTA pairs
static int oz_num_pieces(long qtiles_total, long npairs_w) {
    const long hw_pairs = oz_sm_count() / 2;
    int best = 1;
    double best_waste = 1e30;
    for (int p = 1; p <= OZ_MAX_PIECES && p <= npairs_w; p++) {
        const long items = qtiles_total * p;
        const long rounds = (items + hw_pairs - 1) / hw_pairs;
        const double waste = (double)(rounds * hw_pairs) / (double)items + 0.002 * p;   // mild preference for few pieces
        if (waste < best_waste) { best_waste = waste; best = p; }
    }
    return best;
}

template <int S>
static int oz_launch_var(const OzArgs& a, cudaStream_t st) {
    constexpr int smem = oz_stages(S) * oz_stage_bytes(S);
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(oz_var_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        // the full carveout, not the smallest that fits this kernel: the co-resident front CTA lives in the rest
        GPK_CUDA(cudaFuncSetAttribute(oz_var_kernel<S>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                      (int)cudaSharedmemCarveoutMaxShared));
        attr_done = true;
    }
    const int hw_pairs = oz_sm_count() / 2;
    const int grid = 2 * (a.total_items < hw_pairs ? a.total_items : hw_pairs);   // CTA pairs (static cluster dims 2)
    oz_var_kernel<S><<<grid, OZ_THREADS, smem, st>>>(a);
    GPK_LAUNCHED();
    return 0;
}

// splits of the training range per query tile (<= 16, <= k blocks).  Two 256-thread CTAs are resident per SM, so the
// grid runs in waves of 2 x SMs CTAs; a CTA costs its k blocks plus a fixed part (exp table into shared memory, query
// fragments, partial-sum stores: about 1.5 k blocks of time).  Picks the count that minimises waves x CTA cost:
// few large CTAs when the tiles alone fill the waves, more splits only when they buy a fuller last wave.
static int oz_front_splits(long qtiles_total, int kb_total) {
    if (const char* e = getenv("GPK_OZ_FRONT_SPLITS")) {
        const int ns = atoi(e);
        if (ns >= 1 && ns <= 16 && ns <= kb_total) return ns;
    }
    const long slots = 2L * oz_sm_count();
    int best = 1;
    double best_cost = 1e30;
    for (int ns = 1; ns <= 16 && ns <= kb_total; ns++) {
        const long ctas = qtiles_total * ns;
        const long waves = (ctas + slots - 1) / slots;
        const int kb_per_cta = (kb_total + ns - 1) / ns;
        const double cost = (double)waves * (kb_per_cta + 1.5);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = ns; }
    }
    return best;
}

// splits of the co-resident front end: persistent CTAs (one per SM), items = columns x 2 half tiles x splits
static int oz_front_co_splits(long cols, int kb_total) {
    const long slots = oz_sm_count();
    int best = 1;
    double best_cost = 1e30;
    for (int ns = 1; ns <= 16 && ns <= kb_total; ns++) {
        const long items = cols * 2 * ns;
        const long rounds = (items + slots - 1) / slots;
        const int kb_per_cta = (kb_total + ns - 1) / ns;
        const double cost = (double)rounds * (kb_per_cta + 0.5);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = ns; }
    }
    return best;
}

template <typename TZ, int S, bool MJ, bool IMG = true>
static int oz_launch_front(const OzFrontArgs& a, bool co, int n_split, long qtiles, int batch, TZ* mean, long stride_mean, TZ* jac,
                           long ldj, long stride_jac, cudaStream_t st) {
    const int dp = oz_pad_dim(a.d);
    if (co && IMG) {
        const long items = (long)batch * qtiles * 2 * n_split;
        const unsigned grid = (unsigned)(items < oz_sm_count() ? items : oz_sm_count());
        // the shared-memory carveout is an SM-wide setting: a kernel that prefers a smaller one than the resident
        // oz_var_kernel (which needs the maximum) is not scheduled onto that SM until it drains
        static bool attr_dev[64][4] = {};
#define OZ_FRONT_CASE(DPV, I) case DPV: \
            if (!attr_dev[current_device()][I]) { \
                GPK_CUDA(cudaFuncSetAttribute(oz_front_co_kernel<TZ, S, DPV, MJ>, cudaFuncAttributePreferredSharedMemoryCarveout, \
                                              (int)cudaSharedmemCarveoutMaxShared)); \
                attr_dev[current_device()][I] = true; \
            } \
            oz_front_co_kernel<TZ, S, DPV, MJ><<<grid, 128, 0, st>>>(a, n_split, (int)qtiles, batch); break;
        switch (dp) { OZ_FRONT_CASE(4, 0) OZ_FRONT_CASE(8, 1) OZ_FRONT_CASE(16, 2) OZ_FRONT_CASE(32, 3) }
#undef OZ_FRONT_CASE
    } else {
        dim3 grid((unsigned)n_split, (unsigned)qtiles, (unsigned)batch);
#define OZ_FRONT_CASE(DPV) case DPV: oz_front_kernel<TZ, S, DPV, MJ, IMG><<<grid, 256, 0, st>>>(a); break;
        switch (dp) { OZ_FRONT_CASE(4) OZ_FRONT_CASE(8) OZ_FRONT_CASE(16) OZ_FRONT_CASE(32) }
#undef OZ_FRONT_CASE
    }
    GPK_LAUNCHED();
    if (MJ) {
        dim3 g2((unsigned)((a.q + 127) / 128), 1, (unsigned)batch);
#define OZ_FIN_CASE(DPV) case DPV: oz_mj_finalize_kernel<TZ, DPV><<<g2, 128, 0, st>>>(a, n_split, mean, stride_mean, jac, ldj, stride_jac); break;
        switch (dp) { OZ_FIN_CASE(4) OZ_FIN_CASE(8) OZ_FIN_CASE(16) OZ_FIN_CASE(32) }
#undef OZ_FIN_CASE
        GPK_LAUNCHED();
    }
    return 0;
}

// co = the persistent 4-warp variant that shares the SMs with a running oz_var_kernel (gpk_predict_i8's pipeline)
template <typename TZ>
static int oz_front_dispatch(const OzFrontArgs& a_in, int S, bool mj, bool co, long qtiles, int batch, void* mean,
                             long stride_mean, void* jac, long ldj, long stride_jac, cudaStream_t st) {
    OzFrontArgs a = a_in;
    if (a.img != nullptr && getenv("GPK_OZ_FRONT_CO")) co = atoi(getenv("GPK_OZ_FRONT_CO")) != 0;   // experiments only
    a.exp_tab = exp_table(st);
    GPK_ARG(a.exp_tab != nullptr, "could not allocate the exp table");
    const int n_split = co ? oz_front_co_splits(qtiles * batch, a.kb_total) : oz_front_splits(qtiles * batch, a.kb_total);
    a.kb_per_cta = (a.kb_total + n_split - 1) / n_split;
    // no image wanted (mean + Jacobian alone): the image-free instantiation
    if (a.img == nullptr) {
        if (!mj || co) return fail(-1, "front end without an image needs mean/Jacobian outputs");
        return oz_launch_front<TZ, OZ_MIN_S, true, false>(a, false, n_split, qtiles, batch, (TZ*)mean, stride_mean, (TZ*)jac, ldj, stride_jac, st);
    }
#define OZ_F(SV) \
    if (S == SV) return mj ? oz_launch_front<TZ, SV, true>(a, co, n_split, qtiles, batch, (TZ*)mean, stride_mean, (TZ*)jac, ldj, stride_jac, st) \
                           : oz_launch_front<TZ, SV, false>(a, co, n_split, qtiles, batch, nullptr, 0, nullptr, 0, 0, st);
    OZ_F(4) OZ_F(5) OZ_F(6)
#undef OZ_F
    return fail(-1, "digits must be 4, 5 or 6");
}

// ---- side stream of the prediction pipeline (per device; created by gpk_init or lazily) -----------------------
struct OzSide {
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_front[2] = {nullptr, nullptr}, ev_gemm[2] = {nullptr, nullptr};
};
static OzSide g_oz_side[64];
static unsigned long long* g_oz_tdbg_gemm = nullptr;   // experiments only (GPK_OZ_PIPE_TRACE)
static unsigned long long* g_oz_tdbg_front = nullptr;
int gpk::oz_side_prepare() {
    OzSide& sd = g_oz_side[current_device()];
    if (sd.side) return 0;
    GPK_CUDA(cudaStreamCreateWithFlags(&sd.side, cudaStreamNonBlocking));
    cudaEvent_t* evs[5] = {&sd.ev_fork, &sd.ev_front[0], &sd.ev_front[1], &sd.ev_gemm[0], &sd.ev_gemm[1]};
    for (cudaEvent_t* e : evs) GPK_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    return 0;
}
void gpk::oz_side_release() {
    OzSide& sd = g_oz_side[current_device()];
    if (!sd.side) return;
    cudaEvent_t evs[5] = {sd.ev_fork, sd.ev_front[0], sd.ev_front[1], sd.ev_gemm[0], sd.ev_gemm[1]};
    for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    cudaStreamDestroy(sd.side);
    sd = OzSide{};
}

// Sub-chunk plan of gpk_predict_i8.  Cost model in units of one column's (GP x 128-query tile) contraction time:
// the stand-alone front end costs f per column (f = 0.42 at npad = 2048, ~ 1/npad), the co-resident one about
// OZ_CO_SLOWDOWN x that (one 4-warp CTA per SM instead of two 8-warp CTAs) and hides behind the previous sub-chunk's
// contraction; a contraction launch of c columns runs ceil(c p / pairs) rounds of items of 1/p column each and costs a
// fixed L per launch (ramp, tail, cross-stream event).  Not pipelined: every front end is exposed.
struct OzPlan { long tsub; bool pipelined; };
static double oz_gemm_cost(long cols, long npairs_w) {
    const long hw_pairs = oz_sm_count() / 2;
    double best = 1e300;
    for (int p = 1; p <= OZ_MAX_PIECES && p <= npairs_w; p++) {
        const long rounds = (cols * p + hw_pairs - 1) / hw_pairs;
        const double c = (double)rounds * hw_pairs / p * (1.0 + 0.002 * p);
        if (c < best) best = c;
    }
    return best;
}
static OzPlan oz_plan(long tiles_total, int batch, long npad, int S, long tcap) {
    const long npw = npad / (2 * OZ_WB);
    const double f = 0.42 * 2048.0 / (double)npad;
    double L = 3.0 * (2048.0 / (double)npad) * (2048.0 / (double)npad);
    double co_slow = 4.5;
    // Default OFF: measured on B200 (profiles/pipeline_r02.md) the co-resident front CTAs start with the contraction
    // but are starved until it ends -- oz_var_kernel's tensor-core operand fetches take ~115 of the SM's 128 B/clk of
    // shared-memory bandwidth, which the front end's fragment loads need too -- so the pipelined step is slower
    // (config 2: 0.77 vs 0.72 ms) or equal (config 3).  The path stays for experiments and is covered by tests.
    int mode = 0;                                        // 0 = never pipeline, 1 = by the model, 2 = always (tests)
    if (const char* e = getenv("GPK_OZ_PIPELINE")) mode = atoi(e);
    if (const char* e = getenv("GPK_OZ_PIPE_L")) L = atof(e);
    if (const char* e = getenv("GPK_OZ_PIPE_CO")) co_slow = atof(e);
    long force_t = 0;
    if (const char* e = getenv("GPK_OZ_PIPE_TSUB")) force_t = atol(e);
    const long t1 = tiles_total < tcap ? tiles_total : tcap;
    // not pipelined: chunks of t1 tiles
    double cost1 = 0.0;
    for (long done = 0; done < tiles_total; done += t1) {
        const long t = tiles_total - done < t1 ? tiles_total - done : t1;
        cost1 += f * t * batch + oz_gemm_cost(t * batch, npw) + L;
    }
    OzPlan plan{t1, false};
    // the co-resident front CTA needs the shared memory a contraction CTA leaves free: S <= 5 (see OZ_LAUNCH_REGS)
    const long tcap2 = tcap / 2;
    if (mode == 0 || S > 5 || tcap2 < 1 || tiles_total < 2) return plan;
    double best = mode == 2 ? 1e300 : cost1;
    for (long t = 1; t <= tcap2 && t < tiles_total; t++) {
        if (force_t && t != force_t) continue;
        double cost = f * t * batch;
        for (long done = 0; done < tiles_total; done += t) {
            const long tc = tiles_total - done < t ? tiles_total - done : t;
            const long tn = tiles_total - done - tc < t ? tiles_total - done - tc : t;     // next sub-chunk's tiles
            const double g = oz_gemm_cost(tc * batch, npw), fn = co_slow * f * tn * batch;
            cost += (g > fn ? g : fn) + L;
        }
        if (cost < best) { best = cost; plan = OzPlan{t, true}; }
    }
    return plan;
}

constexpr int OZ_SLOTS = 2 * OZ_MAX_PIECES;
constexpr int OZ_MAX_SPLITS = 16;

extern "C" {

int gpk_oz_default_digits(void) { return 5; }

// Largest training-set size whose int32 digit-group accumulators provably cannot overflow.  The last group (g = S-1)
// sums S digit products per k: (a_0, b_g) and (a_g, b_0) with a leading digit (|a_0| <= 64, 0 <= b_0 <= 64 because
// E = exp(.) lies in (0, 1]) and S-2 products of trailing digits in [-128, 127].  Most negative sum per k:
// -(64*128 + 128*64 + (S-2)*128*127) = -(16384 + (S-2)*16256) >= -2^31 / N; the positive side is smaller because
// a_g * b_0 > 0 needs a_g > 0 (<= 127).  S = 5: N <= 32961 (config 4's N = 32768 fits), S = 6: 26379, S = 4: 43919.
long gpk_oz_max_npad(int S) {
    if (S < OZ_MIN_S || S > OZ_MAX_S) return -1;
    const long per_k = 16384 + (long)(S - 2) * 16256;
    return ((1L << 31) / per_k) / TILE * TILE;
}
int gpk_oz_query_tile(void) { return OZ_QT; }

long gpk_oz_w_image_bytes(long npad, int S) { return (npad % TILE || S < OZ_MIN_S || S > OZ_MAX_S) ? -1 : (long)S * npad * npad; }
long gpk_oz_e_image_bytes(long qpad, long npad, int S) {
    return (qpad % OZ_QT || npad % TILE || S < OZ_MIN_S || S > OZ_MAX_S) ? -1 : (long)S * qpad * npad;
}
int gpk_oz_pieces(long qtiles_total, long npad) { return oz_num_pieces(qtiles_total, npad / (2 * OZ_WB)); }

// W (npad x npad fp64, lower triangular, identity padding) -> int8 digit image + per-row power-of-two scales
int gpk_oz_slice_w(const double* W, long ldw, long npad, int S, void* img, double* wscale, void* stream) {
    GPK_ARG(W && img && wscale, "null pointer");
    GPK_ARG(npad > 0 && npad % TILE == 0 && ldw >= npad, "npad / ldw");
    GPK_ARG(S >= OZ_MIN_S && S <= OZ_MAX_S, "digits must be 4, 5 or 6");
    GPK_ARG(((uintptr_t)img % 16) == 0, "image must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    oz_row_scale_kernel<<<(unsigned)((npad + 7) / 8), 256, 0, st>>>(W, ldw, (int)npad, wscale);
    GPK_LAUNCHED();
    dim3 grid((unsigned)(npad / OZ_BK), (unsigned)(npad / OZ_WB));
    if (S == 4) oz_slice_w_kernel<4><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else if (S == 5) oz_slice_w_kernel<5><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else oz_slice_w_kernel<6><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    GPK_LAUNCHED();
    return 0;
}

// queries -> digit image of E = exp(-r^2/2) for `batch` GPs; qpad = q rounded up to 128
int gpk_oz_cross(int dtype, const void* Z, long ldz, long q, long strideZ, const double* X, long ldx, long n,
                 long strideX, int d, const double* hyp, void* img, long strideE, long npad, long qpad, int S,
                 int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && X && hyp && img, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0 && batch >= 1, "n / d / batch");
    GPK_ARG(npad % TILE == 0 && npad >= n && qpad % OZ_QT == 0 && qpad >= q && q > 0, "npad / qpad");
    GPK_ARG(S >= OZ_MIN_S && S <= OZ_MAX_S, "digits must be 4, 5 or 6");
    GPK_ARG(((uintptr_t)img % 16) == 0 && strideE % 16 == 0, "image alignment");
    OzFrontArgs a{};
    a.Z = Z; a.ldz = ldz; a.q = q; a.strideZ = strideZ;
    a.X = X; a.ldx = ldx; a.n = n; a.strideX = strideX;
    a.d = d; a.hyp = hyp;
    a.img = (int8_t*)img; a.strideE = strideE;
    a.kb_total = (int)(npad / OZ_BK);
    const long qtiles = qpad / OZ_QT;
    if (dtype == 0) return oz_front_dispatch<float>(a, S, false, false, qtiles, batch, nullptr, 0, nullptr, 0, 0, (cudaStream_t)stream);
    return oz_front_dispatch<double>(a, S, false, false, qtiles, batch, nullptr, 0, nullptr, 0, 0, (cudaStream_t)stream);
}

// mean + Jacobian alone on the DMMA front kernel (no digit image, no variance): the want_var = false path for fp64
long gpk_mean_jac_dmma_work_bytes(long q_chunk, int d, int batch) {
    if (q_chunk <= 0 || d < 1 || d > HYP_MAXD || batch < 1) return -1;
    const long qc = (q_chunk + OZ_QT - 1) / OZ_QT * OZ_QT;
    return (long)OZ_MAX_SPLITS * qc * (oz_pad_dim(d) + 1) * 8 * batch;
}

int gpk_predict_mean_jac_dmma(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                              const double* X, long ldx, long n, long strideX, const double* alpha, long stride_alpha,
                              void* mean, long stride_mean, void* jac, long ldj, long stride_jac, void* work,
                              long work_bytes, int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && X && alpha && mean && jac && work, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0 && batch >= 1 && ldj >= d, "n / d / batch / ldj");
    if (q == 0) return 0;
    const int dp = oz_pad_dim(d);
    const long per_q = (long)OZ_MAX_SPLITS * (dp + 1) * 8 * batch;
    long qc = work_bytes / per_q / OZ_QT * OZ_QT;
    GPK_ARG(qc >= OZ_QT, "scratch too small (see gpk_mean_jac_dmma_work_bytes)");
    const long q_round = (q + OZ_QT - 1) / OZ_QT * OZ_QT;
    if (qc > q_round) qc = q_round;
    const int esz = dtype == 0 ? 4 : 8;
    cudaStream_t st = (cudaStream_t)stream;
    for (long q0 = 0; q0 < q; q0 += qc) {
        const long qn = (q - q0 < qc) ? (q - q0) : qc;
        const long qtiles = (qn + OZ_QT - 1) / OZ_QT;
        OzFrontArgs a{};
        a.Z = (const char*)Z + q0 * ldz * esz; a.ldz = ldz; a.q = qn; a.strideZ = strideZ;
        a.X = X; a.ldx = ldx; a.n = n; a.strideX = strideX;
        a.alpha = alpha; a.stride_alpha = stride_alpha;
        a.d = d; a.hyp = hyp;
        a.img = nullptr; a.strideE = 0;
        a.kb_total = (int)((n + OZ_BK - 1) / OZ_BK);
        a.mj_part = (double*)work; a.stride_mj = (long)OZ_MAX_SPLITS * qc * (dp + 1);
        void* mout = (char*)mean + q0 * esz;
        void* jout = (char*)jac + q0 * ldj * esz;
        if (dtype == 0) GPK_TRY(oz_front_dispatch<float>(a, OZ_MIN_S, true, false, qtiles, batch, mout, stride_mean, jout, ldj, stride_jac, st));
        else GPK_TRY(oz_front_dispatch<double>(a, OZ_MIN_S, true, false, qtiles, batch, mout, stride_mean, jout, ldj, stride_jac, st));
    }
    return 0;
}

// part[gp][slot][q], slot < 2 n_pieces: sums of V[i][q]^2 over disjoint sets of W rows (V = sf2 W E^T);
// their sum over the slots is || W k(X, z_q) ||^2.  n_pieces = gpk_oz_pieces()
static int oz_var_partials(const void* Eimg, long strideE, const void* Wimg, long strideW, int S_img, const double* wscale,
                           long stride_ws, const double* hyp, long npad, long qpad, int S, double* part, long ld_part,
                           long stride_part, int n_pieces, int batch, int* dbg, void* stream) {
    GPK_ARG(Wimg && wscale && Eimg && hyp && part, "null pointer");
    GPK_ARG(S_img >= S && S_img <= OZ_MAX_S, "the W image must carry at least the digits that are multiplied");
    GPK_ARG(npad > 0 && npad % TILE == 0 && qpad > 0 && qpad % OZ_QT == 0 && ld_part >= qpad && batch >= 1, "sizes");
    GPK_ARG(S >= OZ_MIN_S && S <= OZ_MAX_S, "digits must be 4, 5 or 6");
    GPK_ARG(((uintptr_t)Wimg % 16) == 0 && ((uintptr_t)Eimg % 16) == 0 && strideW % 16 == 0 && strideE % 16 == 0, "alignment");
    GPK_ARG(npad <= gpk_oz_max_npad(S), "npad too large for exact int32 accumulation with this many digits");
    const long nib = npad / (2 * OZ_WB);     // W row-block pairs
    GPK_ARG(n_pieces >= 1 && n_pieces <= OZ_MAX_PIECES && n_pieces <= nib, "n_pieces");
    OzArgs a{};
    a.Eimg = (const int8_t*)Eimg; a.strideE = strideE;
    a.Wimg = (const int8_t*)Wimg; a.strideW = strideW;
    a.w_kb_bytes = (long)S_img * OZ_B_SLICE;
    a.wscale = wscale; a.stride_ws = stride_ws;
    a.hyp = hyp;
    a.part = part; a.ld_part = ld_part; a.stride_part = stride_part;
    a.qtiles = (int)(qpad / OZ_QT);
    a.nbatch = batch;
    a.kb_total = (int)(npad / OZ_BK);
    a.n_pieces = n_pieces;
    a.total_items = a.qtiles * batch * n_pieces;
    // equal triangular work: sum_{ib < b} (ib + 1) ~ b^2 / 2  ->  b_p = nib sqrt(p / P)
    a.piece_begin[0] = 0;
    for (int p = 1; p < n_pieces; p++) {
        int b = (int)llround((double)nib * sqrt((double)p / n_pieces));
        if (b <= a.piece_begin[p - 1]) b = a.piece_begin[p - 1] + 1;
        if (b > nib - (n_pieces - p)) b = (int)nib - (n_pieces - p);
        a.piece_begin[p] = b;
    }
    a.piece_begin[n_pieces] = (int)nib;
    a.dbg = dbg;
    a.tdbg = g_oz_tdbg_gemm;
    a.flags = getenv("GPK_OZ_FLAGS") ? atoi(getenv("GPK_OZ_FLAGS")) : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (S == 4) return oz_launch_var<4>(a, st);
    if (S == 5) return oz_launch_var<5>(a, st);
    return oz_launch_var<6>(a, st);
}

int gpk_oz_var_partials(const void* Eimg, long strideE, const void* Wimg, long strideW, const double* wscale,
                        long stride_ws, const double* hyp, long npad, long qpad, int S, double* part, long ld_part,
                        long stride_part, int n_pieces, int batch, int* dbg, void* stream) {
    return oz_var_partials(Eimg, strideE, Wimg, strideW, S, wscale, stride_ws, hyp, npad, qpad, S, part, ld_part, stride_part,
                           n_pieces, batch, dbg, stream);
}

// ---- (6+7) fused posterior mean + Jacobian + variance ----------------------------------------------------
// One pass over the (query, training point) pairs evaluates the kernel once: it feeds the mean / Jacobian sums
// and the int8 digit image for the tcgen05 variance contraction  var_f(z) = sf2 - || W k(X,z) ||^2.
// mean / jac may be NULL (variance only).  Queries go through in chunks that fit `work`.  mj_done_event
// (cudaEvent_t or NULL) is recorded once mean and jac are complete.

long gpk_predict_i8_work_bytes(long q_chunk, long npad, int d, int S, int batch) {
    const long qc = (q_chunk + OZ_QT - 1) / OZ_QT * OZ_QT;
    return ((long)S * qc * npad + OZ_SLOTS * qc * 8 + (long)OZ_MAX_SPLITS * qc * (oz_pad_dim(d) + 1) * 8) * batch;
}

int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                     const double* hyp, void* var, long stride_var, int batch, void* stream);

int gpk_predict_i8(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                   const double* X, long ldx, long n, long strideX, const double* alpha, long stride_alpha,
                   const void* Wimg, long strideW, const double* wscale, long stride_ws, long npad, int S,
                   void* mean, long stride_mean, void* jac, long ldj, long stride_jac, void* var, long stride_var,
                   void* work, long work_bytes, int batch, int mul_digits, void* mj_done_event, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && X && Wimg && wscale && var && work, "null pointer");
    GPK_ARG(mul_digits == 0 || (mul_digits >= OZ_MIN_S && mul_digits <= S), "mul_digits must be 0 or in [4, digits]");
    GPK_ARG((mean == nullptr) == (jac == nullptr), "mean and jac must both be given or both be NULL");
    GPK_ARG(mean == nullptr || (alpha != nullptr && ldj >= d), "alpha / ldj");
    GPK_ARG(d >= 1 && d <= HYP_MAXD, "d must be in [1, 32]");
    GPK_ARG(npad % TILE == 0 && npad >= n && n > 0 && batch >= 1, "n / npad / batch");
    GPK_ARG(S >= OZ_MIN_S && S <= OZ_MAX_S, "digits must be 4, 5 or 6");
    GPK_ARG(((uintptr_t)work % 16) == 0, "work must be 16-byte aligned");
    if (q == 0) return 0;
    // mul_digits < digits multiplies only the leading mul_digits digits of both operands, whatever the W image carries
    // (the digits are balanced, so dropping the trailing ones rounds): the caller's accuracy policy picks it, e.g. 4
    // digits (10 digit products) for float32 queries held to 1e-4 when the fit's conditioning allows
    // (engine.select_digits); there is no implicit cut by dtype
    const int S_img = S;
    if (mul_digits) S = mul_digits;
    const bool mj = mean != nullptr;
    const int dp = oz_pad_dim(d);
    const long per_tile = ((long)S * npad + OZ_SLOTS * 8 + (long)OZ_MAX_SPLITS * (dp + 1) * 8) * batch * OZ_QT;
    const long tiles_total = (q + OZ_QT - 1) / OZ_QT;
    const long tcap = work_bytes / per_tile;
    GPK_ARG(tcap >= 1, "scratch too small (see gpk_predict_i8_work_bytes)");
    // sub-chunks of `tsub` query tiles; pipelined = the front end of sub-chunk c + 1 runs on the library's side stream
    // (co-resident 4-warp CTAs, FP64 pipe) while the tensor pipe contracts sub-chunk c, scratch split in two halves
    const OzPlan plan = oz_plan(tiles_total, batch, npad, S, tcap);
    const long qc = plan.tsub * OZ_QT;
    const int nbuf = plan.pipelined ? 2 : 1;
    int8_t* eimg[2]; double* part[2]; double* mjp[2];
    {
        char* w = (char*)work;
        for (int b = 0; b < nbuf; b++) { eimg[b] = (int8_t*)w; w += (long)batch * S * qc * npad; }
        for (int b = 0; b < nbuf; b++) { part[b] = (double*)w; w += (long)batch * OZ_SLOTS * qc * 8; }
        for (int b = 0; b < nbuf; b++) { mjp[b] = (double*)w; w += (long)batch * OZ_MAX_SPLITS * qc * (dp + 1) * 8; }
    }
    const int esz = dtype == 0 ? 4 : 8;
    cudaStream_t st = (cudaStream_t)stream;
    OzSide* sd = nullptr;
    if (plan.pipelined) {
        GPK_TRY(oz_side_prepare());
        sd = &g_oz_side[current_device()];
    }
    const long nchunks = (tiles_total + plan.tsub - 1) / plan.tsub;
    auto front = [&](long c, bool co, cudaStream_t s_) -> int {
        const long q0 = c * qc;
        const long qn = (q - q0 < qc) ? (q - q0) : qc;
        const long qtiles = (qn + OZ_QT - 1) / OZ_QT;
        OzFrontArgs a{};
        a.Z = (const char*)Z + q0 * ldz * esz; a.ldz = ldz; a.q = qn; a.strideZ = strideZ;
        a.X = X; a.ldx = ldx; a.n = n; a.strideX = strideX;
        a.alpha = alpha; a.stride_alpha = stride_alpha;
        a.d = d; a.hyp = hyp;
        a.img = eimg[c % nbuf]; a.strideE = (long)S * qc * npad;
        a.kb_total = (int)(npad / OZ_BK);
        a.mj_part = mjp[c % nbuf]; a.stride_mj = (long)OZ_MAX_SPLITS * qc * (dp + 1);
        a.tdbg = (co && c == 1) ? g_oz_tdbg_front : nullptr;
        void* mout = mj ? (char*)mean + q0 * esz : nullptr;
        void* jout = mj ? (char*)jac + q0 * ldj * esz : nullptr;
        if (dtype == 0) GPK_TRY(oz_front_dispatch<float>(a, S, mj, co, qtiles, batch, mout, stride_mean, jout, ldj, stride_jac, s_));
        else GPK_TRY(oz_front_dispatch<double>(a, S, mj, co, qtiles, batch, mout, stride_mean, jout, ldj, stride_jac, s_));
        // mean and Jacobian are final after the last sub-chunk's front end: a caller that passed an event can start
        // copying them out while the variance contraction of the last sub-chunk runs
        if (mj && mj_done_event && c + 1 == nchunks) GPK_CUDA(cudaEventRecord((cudaEvent_t)mj_done_event, s_));
        return 0;
    };
    // experiments only: per-kernel times of the first sub-chunks on both streams (synchronises the stream)
    const bool trace = getenv("GPK_OZ_PIPE_TRACE") != nullptr;
    cudaEvent_t tev[8][4] = {};
    if (trace) {
        for (auto& r : tev) for (auto& e : r) cudaEventCreate(&e);
        cudaMalloc(&g_oz_tdbg_gemm, 2 * 160 * 8); cudaMalloc(&g_oz_tdbg_front, 2 * 160 * 8);
        cudaMemset(g_oz_tdbg_gemm, 0, 2 * 160 * 8); cudaMemset(g_oz_tdbg_front, 0, 2 * 160 * 8);
    }
    unsigned long long* const tdbg_gemm = g_oz_tdbg_gemm;
    for (long c = 0; c < nchunks; c++) {
        // first sub-chunk, or not pipelined: nothing to overlap the front end with -> the stand-alone 8-warp kernel
        if (c == 0 || !plan.pipelined) GPK_TRY(front(c, false, st));
        const long q0 = c * qc;
        const long qn = (q - q0 < qc) ? (q - q0) : qc;
        const long qp = (qn + OZ_QT - 1) / OZ_QT * OZ_QT;
        const int pieces = oz_num_pieces(qp / OZ_QT * batch, npad / (2 * OZ_WB));
        if (plan.pipelined && c + 1 < nchunks) {
            // sub-chunk c + 1's front end: its buffers were last read by the contraction of sub-chunk c - 1
            // (c = 0: ordered after the first front end, so the two front kernels do not compete)
            cudaEvent_t gate = c == 0 ? sd->ev_fork : sd->ev_gemm[(c - 1) & 1];
            if (c == 0) GPK_CUDA(cudaEventRecord(gate, st));
            GPK_CUDA(cudaStreamWaitEvent(sd->side, gate, 0));
            if (trace && c < 8) cudaEventRecord(tev[c][0], sd->side);
            GPK_TRY(front(c + 1, true, sd->side));
            if (trace && c < 8) cudaEventRecord(tev[c][1], sd->side);
            GPK_CUDA(cudaEventRecord(sd->ev_front[(c + 1) & 1], sd->side));
        }
        if (plan.pipelined && c > 0) GPK_CUDA(cudaStreamWaitEvent(st, sd->ev_front[c & 1], 0));
        if (trace && c < 8) cudaEventRecord(tev[c][2], st);
        g_oz_tdbg_gemm = c == 0 ? tdbg_gemm : nullptr;      // per-CTA timers of the first contraction only
        GPK_TRY(oz_var_partials(eimg[c % nbuf], (long)S * qc * npad, Wimg, strideW, S_img, wscale, stride_ws, hyp, npad, qp, S,
                                part[c % nbuf], qc, (long)OZ_SLOTS * qc, pieces, batch, nullptr, stream));
        if (trace && c < 8) cudaEventRecord(tev[c][3], st);
        if (plan.pipelined && c + 2 < nchunks) GPK_CUDA(cudaEventRecord(sd->ev_gemm[c & 1], st));
        void* vout = (char*)var + q0 * esz;
        GPK_TRY(gpk_var_finalize(dtype, part[c % nbuf], qc, (long)OZ_SLOTS * qc, 2 * pieces, qn, hyp, vout, stride_var, batch, stream));
    }
    if (trace) {
        cudaStreamSynchronize(st);
        for (long c = 0; c < nchunks && c < 8; c++) {
            float tf = -1.f, tg = -1.f, lag = -1.f;
            if (plan.pipelined && c + 1 < nchunks) { cudaEventElapsedTime(&tf, tev[c][0], tev[c][1]); cudaEventElapsedTime(&lag, tev[c][2], tev[c][1]); }
            cudaEventElapsedTime(&tg, tev[c][2], tev[c][3]);
            fprintf(stderr, "[oz trace] chunk %ld: contraction %.4f ms, co front of chunk %ld %.4f ms (ends %.4f ms after the contraction starts)\n",
                    c, tg, c + 1, tf, lag);
        }
        for (auto& r : tev) for (auto& e : r) cudaEventDestroy(e);
        unsigned long long hg[320], hf[320];
        cudaMemcpy(hg, tdbg_gemm, sizeof(hg), cudaMemcpyDeviceToHost);
        cudaMemcpy(hf, g_oz_tdbg_front, sizeof(hf), cudaMemcpyDeviceToHost);
        unsigned long long g0 = ~0ull, g1 = 0;
        for (int i = 0; i < 160; i++) if (hg[2 * i]) { if (hg[2 * i] < g0) g0 = hg[2 * i]; if (hg[2 * i + 1] > g1) g1 = hg[2 * i + 1]; }
        fprintf(stderr, "[oz trace] contraction 0 CTAs: first start 0, last end %.1f us\n", (g1 - g0) * 1e-3);
        int nf = 0, early = 0; double smin = 1e30, smax = -1e30, emin = 1e30, emax = -1e30, dsum = 0;
        for (int i = 0; i < 160; i++) if (hf[2 * i]) {
            const double s0 = ((double)hf[2 * i] - (double)g0) * 1e-3, e0 = ((double)hf[2 * i + 1] - (double)g0) * 1e-3;
            nf++; if (s0 < (g1 - g0) * 0.5e-3) early++;
            if (s0 < smin) smin = s0; if (s0 > smax) smax = s0; if (e0 < emin) emin = e0; if (e0 > emax) emax = e0; dsum += e0 - s0;
        }
        fprintf(stderr, "[oz trace] co front 1: %d CTAs, %d started in the contraction's first half; start %.1f .. %.1f us, end %.1f .. %.1f us, mean duration %.1f us\n",
                nf, early, smin, smax, emin, emax, nf ? dsum / nf : 0.0);
        cudaFree(tdbg_gemm); cudaFree(g_oz_tdbg_front);
        g_oz_tdbg_gemm = g_oz_tdbg_front = nullptr;
    }
    return 0;
}

}  // extern "C"
