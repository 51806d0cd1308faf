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
// W -> digit image.  One CTA per (tile_m, kb): 128 rows x 64 k; thread = (row, 16-wide k chunk) pairs.
// Blocks above the diagonal (kb*64 >= (tile_m+1)*128) are never read by the GEMM and are skipped.
// ------------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(256)
oz_slice_w_kernel(const double* __restrict__ W, long ldw, int npad, const double* __restrict__ scale,
                  int8_t* __restrict__ img) {
    const int kb = blockIdx.x, tm = blockIdx.y;
    if (kb * OZ_BK >= (tm + 1) * OZ_BM) return;
    const int kb_total = npad / OZ_BK;
    int8_t* out = img + ((long)tm * kb_total + kb) * S * OZ_A_SLICE;
    for (int item = threadIdx.x; item < OZ_BM * (OZ_BK / 16); item += 256) {
        const int kc = item & 3, r = item >> 2;          // 4 threads of one row read 64 contiguous doubles
        const long grow = (long)tm * OZ_BM + r;
        const double inv = 1.0 / scale[grow];            // power of two: exact
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
            *reinterpret_cast<uint4*>(out + (long)s * OZ_A_SLICE + oz_tile_off(r, kc * 16)) =
                make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// queries -> E = exp(-1/2 |z~ - x~|^2) -> digit image.  One CTA per (kb group, qtile, gp): 64 queries,
// thread = (query row, 16-wide k chunk); the CTA walks over OZ_CK_KB k blocks.
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
    const int r = tid & 63, kc = tid >> 6;               // 8 consecutive threads -> one 128-byte core matrix
    const long gq = (long)qt * OZ_BN + r;

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
        int8_t* out = img + gp * strideE + ((long)qt * kb_total + kb) * S * OZ_B_SLICE;
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
}


// ------------------------------------------------------------------------------------------------
// the GEMM: persistent CTAs (one per SM), static LPT tile schedule, 6 warps
//   warp 0      producer: per stage two 1-D bulk TMA copies (all S digit slices of the A and of the B k-block)
//   warp 1      TMEM allocation + MMA issue (one thread).  For A digit s one instruction multiplies against the
//               CONCATENATION of B digits 0..S-1-s (N = 64 (S-s), split at 256) and accumulates into digit
//               groups s..S-1, which sit side by side in TMEM: 2S+O(1) instructions per k-step instead of
//               S(S+1)/2, and each A slice is read from shared memory once per k-step
//   warps 2..5  epilogue: tcgen05.ld the S int32 accumulators, Horner in fp64, scale, square, column sums
// ------------------------------------------------------------------------------------------------
struct OzArgs {
    const int8_t* Wimg; long strideW;        // bytes per GP
    const int8_t* Eimg; long strideE;
    const double* wscale; long stride_ws;    // [gp][npad]
    const double* hyp;                       // [gp][HYP_STRIDE]
    double* colsq; long ld_colsq, stride_colsq;   // [gp][tiles_m][ld_colsq]
    int tiles_m, qtiles, nbatch, group_n, total_tiles, kb_total;
    int32_t* dbg;                            // optional raw accumulators [tile][S][128][64]
};

__device__ __forceinline__ void oz_tile_coords(const OzArgs& a, int id, int& tm, int& qt, int& gp) {
    // LPT inside L2-sized column chunks (heaviest tile rows first), as in dgemm_nt.cuh order 1
    const int cols = a.qtiles * a.nbatch;
    const int per_chunk = a.tiles_m * a.group_n;
    const int chunk = id / per_chunk;
    const int rem = id - chunk * per_chunk;
    const int gn = min(a.group_n, cols - chunk * a.group_n);
    tm = a.tiles_m - 1 - rem / gn;
    const int col = chunk * a.group_n + rem % gn;
    gp = col / a.qtiles;
    qt = col - gp * a.qtiles;
}

constexpr int OZ_THREADS = 192;

template <int S>
__global__ void __launch_bounds__(OZ_THREADS, 1) oz_var_kernel(const OzArgs a) {
    constexpr int NST = oz_stages(S);
    constexpr uint32_t STAGE_BYTES = oz_stage_bytes(S);
    constexpr uint32_t A_BYTES = S * OZ_A_SLICE, B_BYTES = S * OZ_B_SLICE;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[NST], empty_bar[NST], tmem_full_bar, tmem_empty_bar;
    __shared__ uint32_t tmem_base_smem;
    __shared__ double red[2][4][OZ_BN];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NST; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&tmem_empty_bar, 128);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_base_smem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            int st = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
                int tm, qt, gp;
                oz_tile_coords(a, tile, tm, qt, gp);
                const int nkb = (tm + 1) * (OZ_BM / OZ_BK);
                const int8_t* Ag = a.Wimg + gp * a.strideW + (long)tm * a.kb_total * A_BYTES;
                const int8_t* Bg = a.Eimg + gp * a.strideE + (long)qt * a.kb_total * B_BYTES;
                for (int kb = 0; kb < nkb; kb++) {
                    mbar_wait(&empty_bar[st], ph ^ 1);
                    mbar_expect_tx(&full_bar[st], STAGE_BYTES);
                    uint8_t* dst = smem + (long)st * STAGE_BYTES;
                    tma_bulk_g2s(dst, Ag + (long)kb * A_BYTES, A_BYTES, &full_bar[st]);
                    tma_bulk_g2s(dst + A_BYTES, Bg + (long)kb * B_BYTES, B_BYTES, &full_bar[st]);
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            int st = 0; uint32_t ph = 0, tph = 0;
            for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
                int tm, qt, gp;
                oz_tile_coords(a, tile, tm, qt, gp);
                const int nkb = (tm + 1) * (OZ_BM / OZ_BK);
                mbar_wait(&tmem_empty_bar, tph ^ 1);          // epilogue has drained the accumulators
                tc_fence_after();
                for (int kb = 0; kb < nkb; kb++) {
                    mbar_wait(&full_bar[st], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(smem + (long)st * STAGE_BYTES);
                    const uint32_t b_base = a_base + A_BYTES;
#pragma unroll
                    for (int j = 0; j < OZ_BK / 32; j++) {
#pragma unroll
                        for (int s = 0; s < S; s++) {
                            const uint64_t adesc = umma_smem_desc(a_base + s * OZ_A_SLICE + j * 256, 128, OZ_BK * 8);
                            const uint32_t acc = (kb > 0 || j > 0 || s > 0) ? 1u : 0u;
#pragma unroll
                            for (int t0 = 0; t0 < S - s; t0 += 4) {
                                const int nt = (S - s - t0) < 4 ? (S - s - t0) : 4;
                                const uint64_t bdesc = umma_smem_desc(b_base + t0 * OZ_B_SLICE + j * 256, 128, OZ_BK * 8);
                                umma_i8(tmem_base + (uint32_t)((s + t0) * OZ_BN), adesc, bdesc, umma_idesc_i8(nt * OZ_BN), acc);
                            }
                        }
                    }
                    tc_commit(&empty_bar[st]);                 // stage reusable once these MMAs have read it
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
                tc_commit(&tmem_full_bar);
                tph ^= 1;
            }
        }
    } else {
        // ===== epilogue (warps 2..5 -> TMEM lane quarters 2,3,0,1) =====
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        uint32_t tph = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, it++) {
            int tm, qt, gp;
            oz_tile_coords(a, tile, tm, qt, gp);
            const double scale = a.hyp[(long)gp * HYP_STRIDE + HYP_SF2] *
                                 a.wscale[gp * a.stride_ws + (long)tm * OZ_BM + row] * (1.0 / 4096.0);
            mbar_wait(&tmem_full_bar, tph);
            tph ^= 1;
            tc_fence_after();
            const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16);
            double* rbuf = red[it & 1][quarter];
#pragma unroll 1
            for (int c = 0; c < OZ_BN / 16; c++) {
                int32_t v[S][16];
#pragma unroll
                for (int g = 0; g < S; g++) tmem_ld16(tbase + (uint32_t)(g * OZ_BN + c * 16), v[g]);
                tmem_ld_wait();
                if (c == OZ_BN / 16 - 1) {                     // all TMEM reads of this tile are done
                    tc_fence_before();
                    mbar_arrive(&tmem_empty_bar);
                }
                if (a.dbg) {
#pragma unroll
                    for (int g = 0; g < S; g++)
#pragma unroll
                        for (int j = 0; j < 16; j++)
                            a.dbg[(((long)tile * S + g) * OZ_BM + row) * OZ_BN + c * 16 + j] = v[g][j];
                }
                double x[16];
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    double acc = (double)v[S - 1][j];
#pragma unroll
                    for (int g = S - 2; g >= 0; g--) acc = fma(acc, 1.0 / 128.0, (double)v[g][j]);
                    acc *= scale;
                    x[j] = acc * acc;
                }
                // transpose-reduce over the 32 lanes: afterwards lane l holds column (l >> 1) of this 16-wide chunk
#pragma unroll
                for (int w = 8, m = 16; w >= 1; w >>= 1, m >>= 1) {
                    const bool up = (lane & m) != 0;
#pragma unroll
                    for (int i = 0; i < w; i++) {
                        const double keep = up ? x[i + w] : x[i];
                        const double send = up ? x[i] : x[i + w];
                        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
                    }
                }
                x[0] += __shfl_xor_sync(0xffffffffu, x[0], 1);
                if ((lane & 1) == 0) rbuf[c * 16 + (lane >> 1)] = x[0];
            }
            asm volatile("bar.sync 1, 128;\n" ::: "memory");   // the 4 epilogue warps
            const int et = threadIdx.x - 64;
            if (et < OZ_BN) {
                const double s = red[it & 1][0][et] + red[it & 1][1][et] + red[it & 1][2][et] + red[it & 1][3][et];
                a.colsq[gp * a.stride_colsq + (long)tm * a.ld_colsq + (long)qt * OZ_BN + et] = s;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
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

template <int S>
static int oz_launch_var(const OzArgs& a, cudaStream_t st) {
    constexpr int smem = oz_stages(S) * oz_stage_bytes(S);
    static bool attr_done = false;
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(oz_var_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    const int grid = a.total_tiles < oz_sm_count() ? a.total_tiles : oz_sm_count();
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

long gpk_oz_w_image_bytes(long npad, int S) { return (npad % OZ_BM || S < 5 || S > OZ_MAX_S) ? -1 : (long)S * npad * npad; }
long gpk_oz_e_image_bytes(long qpad, long npad, int S) {
    return (qpad % OZ_BN || npad % OZ_BM || S < 5 || S > OZ_MAX_S) ? -1 : (long)S * qpad * npad;
}

// W (npad x npad fp64, lower triangular, identity padding) -> int8 digit image + per-row power-of-two scales
int gpk_oz_slice_w(const double* W, long ldw, long npad, int S, void* img, double* wscale, void* stream) {
    GPK_ARG(W && img && wscale, "null pointer");
    GPK_ARG(npad > 0 && npad % OZ_BM == 0 && ldw >= npad, "npad / ldw");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)img % 16) == 0, "image must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    oz_row_scale_kernel<<<(unsigned)((npad + 7) / 8), 256, 0, st>>>(W, ldw, (int)npad, wscale);
    GPK_LAUNCHED();
    dim3 grid((unsigned)(npad / OZ_BK), (unsigned)(npad / OZ_BM));
    if (S == 5) oz_slice_w_kernel<5><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else if (S == 6) oz_slice_w_kernel<6><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    else oz_slice_w_kernel<7><<<grid, 256, 0, st>>>(W, ldw, (int)npad, wscale, (int8_t*)img);
    GPK_LAUNCHED();
    return 0;
}

// queries -> digit image of E = exp(-r^2/2) for `batch` GPs; qpad = q rounded up to 64
int gpk_oz_cross(int dtype, const void* Z, long ldz, long q, long strideZ, const double* X, long ldx, long n,
                 long strideX, int d, const double* hyp, void* img, long strideE, long npad, long qpad, int S,
                 int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && X && hyp && img, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0 && batch >= 1, "n / d / batch");
    GPK_ARG(npad % OZ_BM == 0 && npad >= n && qpad % OZ_BN == 0 && qpad >= q && q > 0, "npad / qpad");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)img % 16) == 0 && strideE % 16 == 0, "image alignment");
    cudaStream_t st = (cudaStream_t)stream;
    const int kb_total = (int)(npad / OZ_BK);
    const long qtiles = qpad / OZ_BN;
#define OZ_X(TZ, SV) return oz_launch_cross<TZ, SV>((const TZ*)Z, ldz, q, strideZ, X, ldx, n, strideX, d, hyp, (int8_t*)img, strideE, kb_total, qtiles, batch, st)
    if (dtype == 0) { if (S == 5) OZ_X(float, 5); if (S == 6) OZ_X(float, 6); OZ_X(float, 7); }
    if (S == 5) OZ_X(double, 5);
    if (S == 6) OZ_X(double, 6);
    OZ_X(double, 7);
#undef OZ_X
}

// colsq[gp][tile_m][q] = sum over the 128 rows of tile_m of (sf2 W K_ZX^T / sf2 ...)^2, i.e. of V[i][q]^2
int gpk_oz_var_colsq(const void* Wimg, long strideW, const double* wscale, long stride_ws, const void* Eimg,
                     long strideE, const double* hyp, long npad, long qpad, int S, double* colsq, long ld_colsq,
                     long stride_colsq, int batch, int* dbg, void* stream) {
    GPK_ARG(Wimg && wscale && Eimg && hyp && colsq, "null pointer");
    GPK_ARG(npad > 0 && npad % OZ_BM == 0 && qpad > 0 && qpad % OZ_BN == 0 && ld_colsq >= qpad && batch >= 1, "sizes");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)Wimg % 16) == 0 && ((uintptr_t)Eimg % 16) == 0 && strideW % 16 == 0 && strideE % 16 == 0, "alignment");
    GPK_ARG((npad / OZ_BK) * (long)(2 * 4096) < (1L << 31), "npad too large for the int32 accumulators");
    OzArgs a{};
    a.Wimg = (const int8_t*)Wimg; a.strideW = strideW;
    a.Eimg = (const int8_t*)Eimg; a.strideE = strideE;
    a.wscale = wscale; a.stride_ws = stride_ws;
    a.hyp = hyp;
    a.colsq = colsq; a.ld_colsq = ld_colsq; a.stride_colsq = stride_colsq;
    a.tiles_m = (int)(npad / OZ_BM);
    a.qtiles = (int)(qpad / OZ_BN);
    a.nbatch = batch;
    a.kb_total = (int)(npad / OZ_BK);
    a.total_tiles = a.tiles_m * a.qtiles * batch;
    const long panel = (long)S * OZ_BN * npad;            // E image bytes of one query tile
    long gn = (48L << 20) / panel;
    const long min_gn = (2L * oz_sm_count() + a.tiles_m - 1) / a.tiles_m;
    if (gn < min_gn) gn = min_gn;
    a.group_n = (int)gn;
    a.dbg = dbg;
    cudaStream_t st = (cudaStream_t)stream;
    if (S == 5) return oz_launch_var<5>(a, st);
    if (S == 6) return oz_launch_var<6>(a, st);
    return oz_launch_var<7>(a, st);
}


// ---- (7') posterior variance on the int8 tensor pipe ------------------------------------------------
// var_f(z) = sf2 - || W k(X,z) ||^2 for q queries of `batch` GPs; queries go through in chunks that fit `work`
// (digit image of E + per-tile column partials).  Same contract as gpk_predict_var, W given as digit image.
long gpk_predict_var_i8_work_bytes(long q_chunk, long npad, int S, int batch) {
    const long qc = (q_chunk + OZ_BN - 1) / OZ_BN * OZ_BN;
    return ((long)S * qc * npad + (npad / OZ_BM) * qc * 8) * batch;
}

int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                     const double* hyp, void* var, long stride_var, int batch, void* stream);

int gpk_predict_var_i8(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                       const double* X, long ldx, long n, long strideX, const void* Wimg, long strideW,
                       const double* wscale, long stride_ws, long npad, int S, void* var, long stride_var,
                       void* work, long work_bytes, int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && X && Wimg && wscale && var && work, "null pointer");
    GPK_ARG(npad % OZ_BM == 0 && npad >= n && n > 0 && batch >= 1, "n / npad / batch");
    GPK_ARG(S >= 5 && S <= OZ_MAX_S, "digits must be 5, 6 or 7");
    GPK_ARG(((uintptr_t)work % 16) == 0, "work must be 16-byte aligned");
    if (q == 0) return 0;
    const int tiles_m = (int)(npad / OZ_BM);
    long qc = work_bytes / (((long)S * npad + tiles_m * 8) * batch);
    qc = qc / OZ_BN * OZ_BN;
    GPK_ARG(qc >= OZ_BN, "scratch too small (see gpk_predict_var_i8_work_bytes)");
    const long q_round = (q + OZ_BN - 1) / OZ_BN * OZ_BN;
    if (qc > q_round) qc = q_round;
    int8_t* eimg = (int8_t*)work;                                         // [batch][S qc npad]
    double* colsq = (double*)(eimg + (long)batch * S * qc * npad);        // [batch][tiles_m][qc]
    const int esz = dtype == 0 ? 4 : 8;
    for (long q0 = 0; q0 < q; q0 += qc) {
        const long qn = (q - q0 < qc) ? (q - q0) : qc;
        const long qp = (qn + OZ_BN - 1) / OZ_BN * OZ_BN;
        const void* Zc = (const char*)Z + q0 * ldz * esz;
        GPK_TRY(gpk_oz_cross(dtype, Zc, ldz, qn, strideZ, X, ldx, n, strideX, d, hyp, eimg, (long)S * qc * npad, npad, qp,
                             S, batch, stream));
        GPK_TRY(gpk_oz_var_colsq(Wimg, strideW, wscale, stride_ws, eimg, (long)S * qc * npad, hyp, npad, qp, S, colsq, qc,
                                 (long)tiles_m * qc, batch, nullptr, stream));
        void* vout = (char*)var + q0 * esz;
        GPK_TRY(gpk_var_finalize(dtype, colsq, qc, (long)tiles_m * qc, tiles_m, qn, hyp, vout, stride_var, batch, stream));
    }
    return 0;
}

}  // extern "C"
