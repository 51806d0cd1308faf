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
#include <cstdlib>

namespace gpk {

// ------------------------------------------------------------------------------------------------
// signed base-128 digits of a non-negative / signed fixed-point integer (most significant digit first)
//   I = sum_s dig[s] * 128^(S-1-s),  dig[s] in [-64, 64]
// ------------------------------------------------------------------------------------------------
template <int S>
__device__ __forceinline__ void signed_digits(long long I, int8_t (&dig)[S]) {
#pragma unroll
    for (int s = S - 1; s > 0; s--) {
        const int low = (int)(I & 127);
        const int dgt = low >= 64 ? low - 128 : low;     // in [-64, 63]
        dig[s] = (int8_t)dgt;
        I = (I - dgt) >> 7;
    }
    dig[0] = (int8_t)I;                                  // |I| <= 64 by construction of the scaling
}

// x in [-1, 1] -> nearest integer of x * 2^(6 + 7 (S-1)) (exact for S <= 7: fits the 52-bit mantissa)
template <int S>
__device__ __forceinline__ long long to_fixed(double x) {
    constexpr double SCALE = (double)(1ull << (6 + 7 * (S - 1)));
    return __double2ll_rn(x * SCALE);
}

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
// queries -> E = exp(-1/2 |z~ - x~|^2) -> digit image.  One CTA per (kb group, qtile, gp): 128 queries,
// thread = (query row, two 16-wide k chunks); the CTA walks over OZ_CK_KB k blocks.
// Kernel formula gpmpc/gp.py:12-14 (sf2 is applied in the GEMM epilogue).
// ------------------------------------------------------------------------------------------------
constexpr int OZ_CK_KB = 4;          // k blocks (of 64 training points) per CTA
constexpr int OZ_LD = HYP_MAXD + 1;

template <typename TZ, int S, int DP>
__global__ void __launch_bounds__(256)
oz_cross_kernel(const TZ* __restrict__ Z, long ldz, long q, long strideZ, const double* __restrict__ X, long ldx,
                long n, long strideX, int d, const double* __restrict__ hyp_all, int8_t* __restrict__ img, long strideE,
                int kb_total) {
    __shared__ double Xs[OZ_BK * OZ_LD], ex[OZ_BK];
    const long gp = blockIdx.z;
    const int qt = blockIdx.y;
    const double* hyp = hyp_all + gp * HYP_STRIDE;
    Z += gp * strideZ;
    X += gp * strideX;
    const int tid = threadIdx.x;
    const int r = tid & 127, kh = tid >> 7;              // 8 consecutive threads -> one 128-byte core matrix
    const long gq = (long)qt * OZ_QT + r;

    double z[DP];
    double ez = 0.0;
#pragma unroll
    for (int j = 0; j < DP; j++) {
        double v = 0.0;
        if (j < d && gq < q) v = ((double)Z[gq * ldz + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
        z[j] = v;
        ez = fma(v, v, ez);
    }
    ez *= -0.5;

    for (int kbi = 0; kbi < OZ_CK_KB; kbi++) {
        const int kb = blockIdx.x * OZ_CK_KB + kbi;
        if (kb >= kb_total) break;
        __syncthreads();
        {   // stage 64 scaled training points: 4 threads per point
            const int p = tid >> 2, part = tid & 3;
            const long gi = (long)kb * OZ_BK + p;
            double s = 0.0;
            for (int j = part; j < DP; j += 4) {
                const double v = (gi < n && j < d) ? (X[gi * ldx + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j] : 0.0;
                Xs[p * OZ_LD + j] = v;
                s = fma(v, v, s);
            }
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            if (part == 0) ex[p] = -0.5 * s;
        }
        __syncthreads();
        int8_t* out = img + gp * strideE + ((long)qt * kb_total + kb) * S * OZ_A_SLICE;
#pragma unroll 1
        for (int kc = kh; kc < OZ_BK / 16; kc += 2) {
            int8_t dig[16][S];
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const int p = kc * 16 + k;
                double a = ez + ex[p];
#pragma unroll
                for (int j = 0; j < DP; j++) a = fma(z[j], Xs[p * OZ_LD + j], a);
                const long gi = (long)kb * OZ_BK + p;
                const double e = (gi < n && gq < q) ? exp(fmin(a, 0.0)) : 0.0;
                signed_digits<S>(to_fixed<S>(e), dig[k]);
            }
#pragma unroll
            for (int s = 0; s < S; s++) {
                uint32_t w4[4];
#pragma unroll
                for (int v = 0; v < 4; v++)
                    w4[v] = (uint32_t)(uint8_t)dig[4 * v][s] | ((uint32_t)(uint8_t)dig[4 * v + 1][s] << 8) |
                            ((uint32_t)(uint8_t)dig[4 * v + 2][s] << 16) | ((uint32_t)(uint8_t)dig[4 * v + 3][s] << 24);
                *reinterpret_cast<uint4*>(out + (long)s * OZ_A_SLICE + oz_tile_off(r, kc * 16)) =
                    make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// the GEMM: persistent CTAs (one per SM), 10 warps
//   warp 0      producer: per stage two 1-D bulk TMA copies (all S digit slices of the E and of the W k-block)
//   warp 1      TMEM allocation + MMA issue (one thread).  For E digit t one instruction multiplies against the
//               CONCATENATION of W digits 0..S-1-t (N = 64 (S-t), split at 256) and accumulates into digit
//               groups t..S-1, which sit side by side in TMEM: S+2 instructions per k-step instead of
//               S(S+1)/2, and each E slice is read from shared memory once per k-step
//   warps 2..9  epilogue: tcgen05.ld the S int32 accumulators, Horner in fp64, row scale, square, and add into
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
    const double* wscale; long stride_ws;    // [gp][npad]
    const double* hyp;                       // [gp][HYP_STRIDE]
    double* part; long ld_part, stride_part; // [gp][piece][ld_part]
    int qtiles, nbatch, kb_total, n_pieces, total_items;
    int piece_begin[OZ_MAX_PIECES + 1];      // W row-block PAIRS [piece_begin[p], piece_begin[p+1]) of piece p
    int32_t* dbg;                            // optional raw accumulators [gp*qtiles + qt][ib][S][128][64]
    int flags;                               // profiling only: 1 = epilogue skips the TMEM reads and the math
};

constexpr int OZ_EPI_WARPS = 8;
constexpr int OZ_THREADS = 64 + 32 * OZ_EPI_WARPS;

__device__ __forceinline__ void oz_item_coords(const OzArgs& a, int item, int& qt, int& gp, int& jp0, int& jp1) {
    const int col = item / a.n_pieces;            // consecutive items share the query tile (its E image stays in L2)
    const int piece = a.n_pieces - 1 - (item - col * a.n_pieces);
    gp = col / a.qtiles;
    qt = col - gp * a.qtiles;
    jp0 = a.piece_begin[piece];
    jp1 = a.piece_begin[piece + 1];
}

__device__ __forceinline__ double oz_ll2d(long long v) {
    // exact int64 -> double for |v| < 2^51 without the conversion pipe: bit pattern of 1.5 * 2^52 + v, minus the bias
    return __longlong_as_double(0x4338000000000000LL + v) - 6755399441055744.0;
}
// R = sum_g 128^-g D_g: digit groups are combined three at a time in exact int64 arithmetic (integer pipe),
// so the FP64 pipe only sees one conversion add per triple and the Horner steps between triples
template <int S>
__device__ __forceinline__ double oz_recombine(const int32_t (&v)[S][8], int j) {
    constexpr int NT = (S + 2) / 3;
    double acc = 0.0;
#pragma unroll
    for (int tr = NT - 1; tr >= 0; tr--) {
        const int g0 = 3 * tr;
        const int cnt = (S - g0) < 3 ? (S - g0) : 3;
        long long t = v[g0][j];
#pragma unroll
        for (int u = 1; u < cnt; u++) t = t * 128 + v[g0 + u][j];
        // t = sum_u D_{g0+u} 128^(cnt-1-u); its weight relative to the triple's leading digit is 128^-(cnt-1)
        const double dv = oz_ll2d(t) * (cnt == 3 ? 1.0 / 16384.0 : cnt == 2 ? 1.0 / 128.0 : 1.0);
        acc = (tr == NT - 1) ? dv : fma(acc, 1.0 / 2097152.0, dv);
    }
    return acc;
}

template <int S>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(OZ_THREADS, 1) oz_var_kernel(const OzArgs a) {
    constexpr int NST = oz_stages(S);
    constexpr uint32_t STAGE_BYTES = oz_stage_bytes(S);
    constexpr uint32_t A_BYTES = S * OZ_A_SLICE, B_BYTES = S * OZ_B_SLICE;
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
    if (warp == 1) tmem_alloc(&tmem_base_smem, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // the peer's barriers are initialised before anything is multicast
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ===== producer (warp-uniform, elected lane issues) =====
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            int qt, gp, jp0, jp1;
            oz_item_coords(a, item, qt, gp, jp0, jp1);
            const int8_t* Ag = a.Eimg + gp * a.strideE + (long)qt * a.kb_total * A_BYTES + rank * (A_BYTES / 2);
            for (int jp = jp1 - 1; jp >= jp0; jp--) {
                const int8_t* Bg = a.Wimg + gp * a.strideW + (long)(2 * jp + rank) * a.kb_total * B_BYTES;
                for (int kb = 0; kb <= 2 * jp + 1; kb++) {
                    mbar_wait(&empty_bar[st], ph ^ 1);          // both CTAs are done reading this stage
                    mbar_expect_tx_pred(&full_bar[st], STAGE_BYTES, leader);
                    const uint32_t dst = smem_u32(smem) + (uint32_t)st * STAGE_BYTES;
                    tma_bulk_g2s_mcast_pred(dst + rank * (A_BYTES / 2), Ag + (long)kb * A_BYTES, A_BYTES / 2, &full_bar[st],
                                            (uint16_t)3, leader);
                    tma_bulk_g2s_pred(dst + A_BYTES, Bg + (long)kb * B_BYTES, B_BYTES, &full_bar[st], leader);
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp runs the loops (warp-uniform), the elected lane issues =====
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0, tph = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            int qt, gp, jp0, jp1;
            oz_item_coords(a, item, qt, gp, jp0, jp1);
            for (int jp = jp1 - 1; jp >= jp0; jp--) {
                mbar_wait(&tmem_empty_bar, tph ^ 1);          // epilogue has drained the accumulators
                tc_fence_after();
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
                                umma_i8_pred(tmem_base + (uint32_t)(t * OZ_WB + h * npart), adesc, bdesc, umma_idesc_i8(npart), acc, leader);
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
    } else {
        // ===== epilogue: warps 2..9; lane quarter = warp % 4, column half = (warp - 2) / 4 =====
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = quarter * 32 + lane;                   // query within the tile
        const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 32);
        uint32_t tph = 0;
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
                    if (a.dbg) {
                        int32_t* dp = a.dbg + ((((long)(gp * a.qtiles + qt) * a.kb_total + ib) * S) * OZ_QT + row) * OZ_WB + half * 32 + c * 8;
#pragma unroll
                        for (int g = 0; g < S; g++)
#pragma unroll
                            for (int j = 0; j < 8; j++) dp[(long)g * OZ_QT * OZ_WB + j] = v[c & 1][g][j];
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
    cluster_sync_all();                      // the peer may still be multicasting into / arriving on this CTA
    if (warp == 1) {
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
    static bool attr_done = false;
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(oz_var_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    const int hw_pairs = oz_sm_count() / 2;
    const int grid = 2 * (a.total_items < hw_pairs ? a.total_items : hw_pairs);   // CTA pairs (static cluster dims 2)
    oz_var_kernel<S><<<grid, OZ_THREADS, smem, st>>>(a);
    GPK_LAUNCHED();
    return 0;
}

template <typename TZ, int S>
static int oz_launch_cross(const TZ* Z, long ldz, long q, long strideZ, const double* X, long ldx, long n, long strideX,
                           int d, const double* hyp, int8_t* img, long strideE, int kb_total, long qtiles, int batch,
                           cudaStream_t st) {
    dim3 grid((unsigned)((kb_total + OZ_CK_KB - 1) / OZ_CK_KB), (unsigned)qtiles, (unsigned)batch);
#define OZ_CROSS_CASE(DPV) case DPV: oz_cross_kernel<TZ, S, DPV><<<grid, 256, 0, st>>>(Z, ldz, q, strideZ, X, ldx, n, strideX, d, hyp, img, strideE, kb_total); break;
    switch (oz_pad_dim(d)) {
        OZ_CROSS_CASE(4) OZ_CROSS_CASE(8) OZ_CROSS_CASE(16) OZ_CROSS_CASE(32)
    }
#undef OZ_CROSS_CASE
    GPK_LAUNCHED();
    return 0;
}

extern "C" {

int gpk_oz_default_digits(void) { return 6; }
int gpk_oz_query_tile(void) { return OZ_QT; }

long gpk_oz_w_image_bytes(long npad, int S) { return (npad % TILE || S < 5 || S > OZ_MAX_S) ? -1 : (long)S * npad * npad; }
long gpk_oz_e_image_bytes(long qpad, long npad, int S) {
    return (qpad % OZ_QT || npad % TILE || S < 5 || S > OZ_MAX_S) ? -1 : (long)S * qpad * npad;
}
int gpk_oz_pieces(long qtiles_total, long npad) { return oz_num_pieces(qtiles_total, npad / (2 * OZ_WB)); }

// W (npad x npad fp64, lower triangular, identity padding) -> int8 digit image + per-row power-of-two scales
int gpk_oz_slice_w(const double* W, long ldw, long npad, int S, void* img, double* wscale, void* stream) {
    GPK_ARG(W && img && wscale, "null pointer");
    GPK_ARG(npad > 0 && npad % TILE == 0 && ldw >= npad, "npad / ldw");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)img % 16) == 0, "image must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    oz_row_scale_kernel<<<(unsigned)((npad + 7) / 8), 256, 0, st>>>(W, ldw, (int)npad, wscale);
    GPK_LAUNCHED();
    dim3 grid((unsigned)(npad / OZ_BK), (unsigned)(npad / OZ_WB));
    if (S == 5) oz_slice_w_kernel<5><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else if (S == 6) oz_slice_w_kernel<6><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else oz_slice_w_kernel<7><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
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
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)img % 16) == 0 && strideE % 16 == 0, "image alignment");
    cudaStream_t st = (cudaStream_t)stream;
    const int kb_total = (int)(npad / OZ_BK);
    const long qtiles = qpad / OZ_QT;
#define OZ_X(TZ, SV) return oz_launch_cross<TZ, SV>((const TZ*)Z, ldz, q, strideZ, X, ldx, n, strideX, d, hyp, (int8_t*)img, strideE, kb_total, qtiles, batch, st)
    if (dtype == 0) { if (S == 5) OZ_X(float, 5); if (S == 6) OZ_X(float, 6); OZ_X(float, 7); }
    if (S == 5) OZ_X(double, 5);
    if (S == 6) OZ_X(double, 6);
    OZ_X(double, 7);
#undef OZ_X
}

// part[gp][slot][q], slot < 2 n_pieces: sums of V[i][q]^2 over disjoint sets of W rows (V = sf2 W E^T);
// their sum over the slots is || W k(X, z_q) ||^2.  n_pieces = gpk_oz_pieces()
int gpk_oz_var_partials(const void* Eimg, long strideE, const void* Wimg, long strideW, const double* wscale,
                        long stride_ws, const double* hyp, long npad, long qpad, int S, double* part, long ld_part,
                        long stride_part, int n_pieces, int batch, int* dbg, void* stream) {
    GPK_ARG(Wimg && wscale && Eimg && hyp && part, "null pointer");
    GPK_ARG(npad > 0 && npad % TILE == 0 && qpad > 0 && qpad % OZ_QT == 0 && ld_part >= qpad && batch >= 1, "sizes");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)Wimg % 16) == 0 && ((uintptr_t)Eimg % 16) == 0 && strideW % 16 == 0 && strideE % 16 == 0, "alignment");
    GPK_ARG(npad <= 65536, "npad too large for the int32 accumulators");
    const long nib = npad / (2 * OZ_WB);     // W row-block pairs
    GPK_ARG(n_pieces >= 1 && n_pieces <= OZ_MAX_PIECES && n_pieces <= nib, "n_pieces");
    OzArgs a{};
    a.Eimg = (const int8_t*)Eimg; a.strideE = strideE;
    a.Wimg = (const int8_t*)Wimg; a.strideW = strideW;
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
    a.flags = getenv("GPK_OZ_FLAGS") ? atoi(getenv("GPK_OZ_FLAGS")) : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (S == 5) return oz_launch_var<5>(a, st);
    if (S == 6) return oz_launch_var<6>(a, st);
    return oz_launch_var<7>(a, st);
}

// ---- (7') posterior variance on the int8 tensor pipe ------------------------------------------------
// var_f(z) = sf2 - || W k(X,z) ||^2 for q queries of `batch` GPs; queries go through in chunks that fit `work`
// (digit image of E + per-piece partial sums).  Same contract as gpk_predict_var, W given as digit image.
constexpr int OZ_SLOTS = 2 * OZ_MAX_PIECES;
long gpk_predict_var_i8_work_bytes(long q_chunk, long npad, int S, int batch) {
    const long qc = (q_chunk + OZ_QT - 1) / OZ_QT * OZ_QT;
    return ((long)S * qc * npad + OZ_SLOTS * qc * 8) * batch;
}

int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                     const double* hyp, void* var, long stride_var, int batch, void* stream);

int gpk_predict_var_i8(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                       const double* X, long ldx, long n, long strideX, const void* Wimg, long strideW,
                       const double* wscale, long stride_ws, long npad, int S, void* var, long stride_var,
                       void* work, long work_bytes, int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && X && Wimg && wscale && var && work, "null pointer");
    GPK_ARG(npad % TILE == 0 && npad >= n && n > 0 && batch >= 1, "n / npad / batch");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)work % 16) == 0, "work must be 16-byte aligned");
    if (q == 0) return 0;
    long qc = work_bytes / (((long)S * npad + OZ_SLOTS * 8) * batch);
    qc = qc / OZ_QT * OZ_QT;
    GPK_ARG(qc >= OZ_QT, "scratch too small (see gpk_predict_var_i8_work_bytes)");
    const long q_round = (q + OZ_QT - 1) / OZ_QT * OZ_QT;
    if (qc > q_round) qc = q_round;
    int8_t* eimg = (int8_t*)work;                                         // [batch][S qc npad]
    double* part = (double*)(eimg + (long)batch * S * qc * npad);         // [batch][pieces][qc]
    const int esz = dtype == 0 ? 4 : 8;
    for (long q0 = 0; q0 < q; q0 += qc) {
        const long qn = (q - q0 < qc) ? (q - q0) : qc;
        const long qp = (qn + OZ_QT - 1) / OZ_QT * OZ_QT;
        const int pieces = oz_num_pieces(qp / OZ_QT * batch, npad / (2 * OZ_WB));
        const void* Zc = (const char*)Z + q0 * ldz * esz;
        GPK_TRY(gpk_oz_cross(dtype, Zc, ldz, qn, strideZ, X, ldx, n, strideX, d, hyp, eimg, (long)S * qc * npad, npad, qp,
                             S, batch, stream));
        GPK_TRY(gpk_oz_var_partials(eimg, (long)S * qc * npad, Wimg, strideW, wscale, stride_ws, hyp, npad, qp, S, part,
                                    qc, (long)OZ_SLOTS * qc, pieces, batch, nullptr, stream));
        void* vout = (char*)var + q0 * esz;
        GPK_TRY(gpk_var_finalize(dtype, part, qc, (long)OZ_SLOTS * qc, 2 * pieces, qn, hyp, vout, stride_var, batch, stream));
    }
    return 0;
}

}  // extern "C"
