// SIMT kernels of the GP hot path: ARD squared-exponential kernel assembly, the fused posterior
// mean + input-Jacobian kernel (north-star subsystem a), the variance finalisation and the MLL gradient
// trace reduction (subsystem c).
//
// Math restated from the reference (file:line under /root/reference):
//   k(a,b) = sf2 * exp(-1/2 sum_j (a_j - b_j)^2 / ell_j^2)                      gpmpc/gp.py:12-14, 17-21
//   K      = k(X,X) + noise I                                                   gpmpc/gp.py:45
//   mean   = k(z,X) K^-1 y  = sum_i alpha_i k(z,x_i)                            gpmpc/gp.py:85
//   jac    = d mean / d z   = sum_i alpha_i k(z,x_i) (x_i - z) / ell^2          CasADi AD of gp.py:12-14 in acados
// Points are centred by `center` and divided by ell on the fly (what gpytorch's RBFKernel does before its
// distance computation); r^2 is then expanded as |a|^2 + |b|^2 - 2 a.b so the inner loop is one FMA per
// dimension.
#include "common.cuh"
#include <mutex>

namespace gpk {

// ------------------------------------------------------------------------------------------------
// cross kernel / training-matrix assembly:  out[r][c] = sf2 exp(-1/2 |a~_r - b~_c|^2)
//   mode 0: cross kernel K(A, B); rows >= na and columns >= nb are written as 0 (padding)
//   mode 1: training matrix K(A, A) + noise I, exact diagonal sf2 + noise, identity on the padding
// ------------------------------------------------------------------------------------------------
constexpr int CK_T = 64, CK_LD = HYP_MAXD + 1;

template <typename TP_>
__device__ __forceinline__ void stage_scaled_rows(const TP_* __restrict__ P, long ldp, long n_valid, long row0,
                                                  int d, int dfill, const double* __restrict__ hyp, double* sm,
                                                  double* se) {
    // 64 rows, scaled and centred, into sm[64][CK_LD] (columns d..dfill-1 zeroed); se[r] = -1/2 |row|^2.
    // 256 threads: 4 threads per row.
    const int r = threadIdx.x >> 2, q = threadIdx.x & 3;
    const long gr = row0 + r;
    double s = 0.0;
    for (int j = q; j < dfill; j += 4) {
        double v = 0.0;
        if (gr < n_valid && j < d) v = ((double)P[gr * ldp + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
        sm[r * CK_LD + j] = v;
        s = fma(v, v, s);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (q == 0) se[r] = -0.5 * s;
}

template <typename TA>
__global__ void __launch_bounds__(256)
cross_kernel(const TA* __restrict__ A, long lda, long na, long strideA, const double* __restrict__ B, long ldb,
             long nb, long strideB, int d, const double* __restrict__ hyp_all, double* __restrict__ out, long ldo,
             long strideO, int mode) {
    __shared__ double As[CK_T * CK_LD], Bs[CK_T * CK_LD], ea[CK_T], eb[CK_T];
    const long batch = blockIdx.z;
    const double* hyp = hyp_all + batch * HYP_STRIDE;
    A += batch * strideA;
    B += batch * strideB;
    out += batch * strideO;
    const long row0 = (long)blockIdx.y * CK_T, col0 = (long)blockIdx.x * CK_T;
    stage_scaled_rows(A, lda, na, row0, d, d, hyp, As, ea);
    stage_scaled_rows(B, ldb, nb, col0, d, d, hyp, Bs, eb);
    __syncthreads();
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[r][c] = ea[ty * 4 + r] + eb[tx + 16 * c];
#pragma unroll 4
    for (int j = 0; j < d; j++) {
        double a[4], b[4];
#pragma unroll
        for (int r = 0; r < 4; r++) a[r] = As[(ty * 4 + r) * CK_LD + j];
#pragma unroll
        for (int c = 0; c < 4; c++) b[c] = Bs[(tx + 16 * c) * CK_LD + j];
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc[r][c] = fma(a[r], b[c], acc[r][c]);
    }
    const double sf2 = hyp[HYP_SF2], noise = hyp[HYP_NOISE];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const long gr = row0 + ty * 4 + r, gc = col0 + tx + 16 * c;
            double v;
            if (gr < na && gc < nb) {
                v = sf2 * exp(fmin(acc[r][c], 0.0));
                if (mode == 1 && gr == gc) v = sf2 + noise;
            } else {
                v = (mode == 1 && gr == gc) ? 1.0 : 0.0;
            }
            out[gr * ldo + gc] = v;
        }
}

// ------------------------------------------------------------------------------------------------
// train pack: per tile of TP points  [x~ (TP x DP)] [e (TP)] [a (TP)]  in the prediction dtype
//   fp64: e = -1/2 |x~|^2          (natural exponent, exp())
//   fp32: e = -1/2 |x~|^2 log2(e)  (base-2 exponent, ex2.approx)
//   a = sf2 * alpha (0 on the padding)
// ------------------------------------------------------------------------------------------------
constexpr int TP = 256;
constexpr double LOG2E = 1.4426950408889634074;

template <typename T>
__global__ void pack_train_kernel(const double* __restrict__ X, long ldx, long n, int d, int dp,
                                  const double* __restrict__ alpha, const double* __restrict__ hyp,
                                  T* __restrict__ pack, long n_tiles) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles * TP) return;
    const long tile = i / TP;
    const int r = (int)(i - tile * TP);
    T* base = pack + tile * (long)TP * (dp + 2);
    double s = 0.0;
    for (int j = 0; j < dp; j++) {
        double v = 0.0;
        if (i < n && j < d) v = (X[i * ldx + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
        const T vt = (T)v;
        base[r * dp + j] = vt;
        s = fma((double)vt, (double)vt, s);      // norm of the value actually used by the kernel
    }
    const double scale = sizeof(T) == 4 ? LOG2E : 1.0;
    base[(long)TP * dp + r] = (T)(-0.5 * s * scale);
    base[(long)TP * dp + TP + r] = (T)(i < n ? hyp[HYP_SF2] * alpha[i] : 0.0);
}

// ------------------------------------------------------------------------------------------------
// fused posterior mean + input-Jacobian (north-star subsystem a)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float kexp(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ double kexp(double x) { return exp(x); }
// fp64 with few input dimensions: the exponential dominates the pair cost and the table version (common.cuh) is
// ~10 % faster end to end; from d = 16 on its random shared-memory lookups cost more than the library polynomial
// (measured: 142 vs 165 G pairs/s), so the library exp stays there
template <bool TAB>
__device__ __forceinline__ double kexp_sel(double x, const double* tab) { return TAB ? exp_neg_tab(fmin(x, 0.0), tab) : exp(x); }
template <bool TAB>
__device__ __forceinline__ float kexp_sel(float x, const double*) { return kexp(x); }

template <typename T, int DP>
__device__ __forceinline__ void load_row(const T* __restrict__ p, T (&x)[DP]) {
    if constexpr ((DP * sizeof(T)) % 16 == 0) {
        constexpr int V = 16 / sizeof(T);
        using VT = typename std::conditional<sizeof(T) == 4, float4, double2>::type;
#pragma unroll
        for (int v = 0; v < DP / V; v++) {
            const VT t = reinterpret_cast<const VT*>(p)[v];
            if constexpr (sizeof(T) == 4) {
                x[v * 4 + 0] = ((const float4&)t).x; x[v * 4 + 1] = ((const float4&)t).y;
                x[v * 4 + 2] = ((const float4&)t).z; x[v * 4 + 3] = ((const float4&)t).w;
            } else {
                x[v * 2 + 0] = ((const double2&)t).x; x[v * 2 + 1] = ((const double2&)t).y;
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < DP; j++) x[j] = p[j];
    }
}

struct MeanJacArgs {
    const void* Z; long ldz; long q; long strideZ;   // raw queries (dtype T), shared between GPs when strideZ == 0
    int d;
    const double* hyp;                               // [batch][HYP_STRIDE]
    const void* pack; long stride_pack;              // [batch][n_tiles * TP * (DP + 2)]
    int n_tiles, tiles_per_split, n_split;
    void* mean; long stride_mean;                    // [batch][q]
    void* jac; long ldj; long stride_jac;            // [batch][q][ldj]
    void* work; long stride_work;                    // [batch][n_split][q][DP + 1] partial sums when n_split > 1
    const double* exp_tab;                           // 2^(j/1024) table of the fp64 exponential (common.cuh)
};

constexpr int MJ_THREADS = 128;
// tuning hooks (A/B builds) of the float32 d = 16 kernel: resident CTAs per SM it is compiled for, queries per thread,
// inner unroll.  Measured on B200 (tools/time_mj_variants.py, N = 2048): (1 CTA, 2 queries, unroll 2) 422 / 601 G pairs/s
// at Q = 8192 / 65536, (4, 1, 4) 460 / 640: one query per thread at 120 registers keeps 16 warps per SM resident
#ifndef MJ_F32_MINB
#define MJ_F32_MINB 4
#endif
#ifndef MJ_F32_RQ
#define MJ_F32_RQ 1
#endif
#ifndef MJ_UNROLL
#define MJ_UNROLL 4
#endif

template <typename T, int DP, int RQ>
__global__ void __launch_bounds__(MJ_THREADS, (sizeof(T) == 4 && DP == 16) ? MJ_F32_MINB : 1)
mean_jac_kernel(const MeanJacArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int TILE_ELEMS = TP * (DP + 2);
    constexpr uint32_t TILE_BYTES = TILE_ELEMS * sizeof(T);
    T* buf = reinterpret_cast<T*>(smem_raw);                        // 2 tiles
    __shared__ __align__(8) uint64_t bar[2];
    constexpr bool TAB = sizeof(T) == 8 && DP <= 8;
    __shared__ double etab[TAB ? EXP_TAB : 1];                      // table of exp_neg_tab

    const int tid = threadIdx.x;
    const long batch = blockIdx.z;
    const int split = blockIdx.y;
    if constexpr (TAB)
        for (int i = tid; i < EXP_TAB; i += MJ_THREADS) etab[i] = a.exp_tab[i];   // visible after the barrier below
    const double* hyp = a.hyp + batch * HYP_STRIDE;
    const T* __restrict__ Z = reinterpret_cast<const T*>(a.Z) + batch * a.strideZ;
    const T* __restrict__ pack = reinterpret_cast<const T*>(a.pack) + batch * a.stride_pack;

    const int t_begin = split * a.tiles_per_split;
    const int t_end = min(a.n_tiles, t_begin + a.tiles_per_split);

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0 && t_begin < t_end) {
        mbar_expect_tx(&bar[0], TILE_BYTES);
        tma_bulk_g2s(buf, pack + (long)t_begin * TILE_ELEMS, TILE_BYTES, &bar[0]);
    }

    // this thread's RQ queries, centred / scaled; for fp32 the exponent runs in base 2
    constexpr T ESCALE = sizeof(T) == 4 ? (T)LOG2E : (T)1;
    T zs[RQ][DP], zl[RQ][DP], ez[RQ], m[RQ], S[RQ][DP];
    long qi[RQ];
#pragma unroll
    for (int r = 0; r < RQ; r++) {
        qi[r] = ((long)blockIdx.x * RQ + r) * MJ_THREADS + tid;
        T s = 0;
#pragma unroll
        for (int j = 0; j < DP; j++) {
            T v = 0;
            if (j < a.d && qi[r] < a.q) v = (T)(((double)Z[qi[r] * a.ldz + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j]);
            zs[r][j] = v;
            zl[r][j] = v * ESCALE;
            s = fma(v, v, s);
            S[r][j] = 0;
        }
        ez[r] = (T)(-0.5) * s * ESCALE;
        m[r] = 0;
    }
    constexpr bool PACK = sizeof(T) == 4 && DP % 2 == 0;
    float2 zl2[RQ][PACK ? DP / 2 : 1], S2[RQ][PACK ? DP / 2 : 1];
    if constexpr (PACK) {
#pragma unroll
        for (int r = 0; r < RQ; r++)
#pragma unroll
            for (int jj = 0; jj < DP / 2; jj++) {
                zl2[r][jj] = make_float2((float)zl[r][2 * jj], (float)zl[r][2 * jj + 1]);
                S2[r][jj] = make_float2(0.f, 0.f);
            }
    }

    for (int t = t_begin; t < t_end; t++) {
        const int it = t - t_begin, cur = it & 1;
        if (tid == 0 && t + 1 < t_end) {   // the other buffer was released by the __syncthreads below
            mbar_expect_tx(&bar[cur ^ 1], TILE_BYTES);
            tma_bulk_g2s(buf + (cur ^ 1) * TILE_ELEMS, pack + (long)(t + 1) * TILE_ELEMS, TILE_BYTES, &bar[cur ^ 1]);
        }
        mbar_wait(&bar[cur], (it >> 1) & 1);
        const T* xs = buf + cur * TILE_ELEMS;
        const T* es = xs + TP * DP;
        const T* as = es + TP;
        constexpr int UNR = (sizeof(T) == 4 && DP == 16) ? MJ_UNROLL : 2;
#pragma unroll UNR
        for (int i = 0; i < TP; i++) {
            T x[DP];
            load_row<T, DP>(xs + i * DP, x);
            const T e = es[i], al = as[i];
            if constexpr (PACK) {
                // float32: both GEMM-shaped parts as packed FFMA2 (fma.rn.f32x2) over pairs of input dimensions --
                // Blackwell's FP32 pipe only reaches its 128 FMA/clk/SM through the packed form
                float2 x2[DP / 2];
#pragma unroll
                for (int jj = 0; jj < DP / 2; jj++) x2[jj] = make_float2((float)x[2 * jj], (float)x[2 * jj + 1]);
#pragma unroll
                for (int r = 0; r < RQ; r++) {
                    float2 s2 = make_float2((float)(e + ez[r]), 0.f);
#pragma unroll
                    for (int jj = 0; jj < DP / 2; jj++) s2 = __ffma2_rn(zl2[r][jj], x2[jj], s2);
                    const float w = kexp(s2.x + s2.y) * (float)al;
                    m[r] += (T)w;
                    const float2 w2 = make_float2(w, w);
#pragma unroll
                    for (int jj = 0; jj < DP / 2; jj++) S2[r][jj] = __ffma2_rn(w2, x2[jj], S2[r][jj]);
                }
            } else {
#pragma unroll
                for (int r = 0; r < RQ; r++) {
                    T s = e + ez[r];
#pragma unroll
                    for (int j = 0; j < DP; j++) s = fma(zl[r][j], x[j], s);
                    const T w = kexp_sel<TAB>(s, etab) * al;
                    m[r] += w;
#pragma unroll
                    for (int j = 0; j < DP; j++) S[r][j] = fma(w, x[j], S[r][j]);
                }
            }
        }
        __syncthreads();
    }
    if constexpr (PACK) {
#pragma unroll
        for (int r = 0; r < RQ; r++)
#pragma unroll
            for (int jj = 0; jj < DP / 2; jj++) {
                S[r][2 * jj] = (T)S2[r][jj].x;
                S[r][2 * jj + 1] = (T)S2[r][jj].y;
            }
    }

    if (a.n_split == 1) {
        T* mean = reinterpret_cast<T*>(a.mean) + batch * a.stride_mean;
        T* jac = reinterpret_cast<T*>(a.jac) + batch * a.stride_jac;
#pragma unroll
        for (int r = 0; r < RQ; r++) {
            if (qi[r] >= a.q) continue;
            mean[qi[r]] = m[r];
#pragma unroll
            for (int j = 0; j < DP; j++)
                if (j < a.d) jac[qi[r] * a.ldj + j] = (T)((double)(S[r][j] - zs[r][j] * m[r]) * hyp[HYP_INV_ELL + j]);
        }
    } else {
        T* work = reinterpret_cast<T*>(a.work) + batch * a.stride_work + (long)split * a.q * (DP + 1);
#pragma unroll
        for (int r = 0; r < RQ; r++) {
            if (qi[r] >= a.q) continue;
            T* w = work + qi[r] * (DP + 1);
            w[0] = m[r];
#pragma unroll
            for (int j = 0; j < DP; j++) w[1 + j] = S[r][j];
        }
    }
}

// sums the split partials and applies jac_j = (S_j - z~_j mean) / ell_j
template <typename T, int DP>
__global__ void mean_jac_finalize_kernel(const MeanJacArgs a) {
    const long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long batch = blockIdx.z;
    if (q >= a.q) return;
    const double* hyp = a.hyp + batch * HYP_STRIDE;
    const T* Z = reinterpret_cast<const T*>(a.Z) + batch * a.strideZ;
    const T* work = reinterpret_cast<const T*>(a.work) + batch * a.stride_work;
    double acc[DP + 1];
#pragma unroll
    for (int j = 0; j <= DP; j++) acc[j] = 0.0;
    for (int s = 0; s < a.n_split; s++) {
        const T* w = work + ((long)s * a.q + q) * (DP + 1);
#pragma unroll
        for (int j = 0; j <= DP; j++) acc[j] += (double)w[j];
    }
    reinterpret_cast<T*>(a.mean)[batch * a.stride_mean + q] = (T)acc[0];
    T* jac = reinterpret_cast<T*>(a.jac) + batch * a.stride_jac + q * a.ldj;
#pragma unroll
    for (int j = 0; j < DP; j++)
        if (j < a.d) {
            const T zsj = (T)(((double)Z[q * a.ldz + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j]);
            jac[j] = (T)((acc[1 + j] - (double)zsj * acc[0]) * hyp[HYP_INV_ELL + j]);
        }
}

// Split of the training tiles over CTAs.  The grid runs in waves of (SMs x resident CTAs) -- the resident count comes
// from the occupancy API for this instantiation -- and a CTA costs its tiles plus about half a tile of set-up (queries,
// first TMA round trip, partial-sum stores): pick the split count that minimises waves x CTA cost.  (Config 2 in
// float32: 192 query blocks, 592 slots, 8 tiles -> 3 splits of 3 tiles fill ONE wave; the old "about two waves of
// CTAs" rule took 4 splits = 1.3 waves.)
template <typename T, int DP, int RQ>
static int mj_pick_split(long qblocks, int n_tiles, long q, int batch, const void* work, long work_bytes) {
    constexpr int smem = 2 * TP * (DP + 2) * (int)sizeof(T);
    static int resident_dev[64] = {};
    int& resident = resident_dev[current_device()];
    if (resident == 0) {
        int nb = 0, sms = 148;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, mean_jac_kernel<T, DP, RQ>, MJ_THREADS, smem);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, current_device());
        resident = (nb > 0 ? nb : 1) * (sms > 0 ? sms : 148);
    }
    int best = 1;
    double best_cost = 1e30;
    for (int ns = 1; ns <= n_tiles && ns <= 64; ns++) {
        const int tps = (n_tiles + ns - 1) / ns;
        const int real_ns = (n_tiles + tps - 1) / tps;
        if (real_ns != ns) continue;                               // same split as a smaller count
        const long need = (long)ns * q * (DP + 1) * (long)sizeof(T) * batch;
        if (ns > 1 && (work == nullptr || work_bytes < need)) break;
        const long waves = (qblocks * ns + resident - 1) / resident;
        const double cost = (double)waves * (tps + 0.5);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = ns; }
    }
    return best;
}

template <typename T, int DP, int RQ>
static int launch_mean_jac_t(MeanJacArgs a, int batch, long work_bytes, cudaStream_t st) {
    constexpr int smem = 2 * TP * (DP + 2) * (int)sizeof(T);
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(mean_jac_kernel<T, DP, RQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    const long per_block = (long)MJ_THREADS * RQ;
    const long qblocks_1 = (a.q + per_block - 1) / per_block;
    const int n_split = mj_pick_split<T, DP, RQ>(qblocks_1 * batch, a.n_tiles, a.q, batch, a.work, work_bytes);
    a.tiles_per_split = (a.n_tiles + n_split - 1) / n_split;
    a.n_split = (a.n_tiles + a.tiles_per_split - 1) / a.tiles_per_split;
    a.stride_work = (long)a.n_split * a.q * (DP + 1);
    dim3 grid((unsigned)qblocks_1, (unsigned)a.n_split, (unsigned)batch);
    mean_jac_kernel<T, DP, RQ><<<grid, MJ_THREADS, smem, st>>>(a);
    GPK_LAUNCHED();
    if (a.n_split > 1) {
        dim3 g2((unsigned)((a.q + 127) / 128), 1, (unsigned)batch);
        mean_jac_finalize_kernel<T, DP><<<g2, 128, 0, st>>>(a);
        GPK_LAUNCHED();
    }
    return 0;
}

template <typename T>
static int launch_mean_jac(const MeanJacArgs& a, int dp, int batch, long work_bytes, cudaStream_t st) {
    constexpr int RQ = sizeof(T) == 4 ? 2 : 1;
    switch (dp) {
        case 1: return launch_mean_jac_t<T, 1, RQ>(a, batch, work_bytes, st);
        case 2: return launch_mean_jac_t<T, 2, RQ>(a, batch, work_bytes, st);
        case 3: return launch_mean_jac_t<T, 3, RQ>(a, batch, work_bytes, st);
        case 4: return launch_mean_jac_t<T, 4, RQ>(a, batch, work_bytes, st);
        case 8: return launch_mean_jac_t<T, 8, RQ>(a, batch, work_bytes, st);
        case 16: return launch_mean_jac_t<T, 16, sizeof(T) == 4 ? MJ_F32_RQ : 1>(a, batch, work_bytes, st);
        case 32: return launch_mean_jac_t<T, 32, 1>(a, batch, work_bytes, st);
        default: return fail(-1, "mean_jac: unsupported padded dimension %d", dp);
    }
}

// ------------------------------------------------------------------------------------------------
// variance finalisation:  var[q] = sf2 - sum_tm colsq[tm][q]
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void var_finalize_kernel(const double* __restrict__ colsq, long ld, long stride_colsq, int tiles_m, long q,
                                    const double* __restrict__ hyp_all, T* __restrict__ var, long stride_var) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long batch = blockIdx.z;
    if (i >= q) return;
    const double* c = colsq + batch * stride_colsq + i;
    double s = 0.0;
    for (int t = 0; t < tiles_m; t++) s += c[(long)t * ld];
    var[batch * stride_var + i] = (T)(hyp_all[batch * HYP_STRIDE + HYP_SF2] - s);
}

// ------------------------------------------------------------------------------------------------
// MLL gradient trace reduction (north-star subsystem c).  With G = alpha alpha^T - K^-1 and K_f the
// noise-free kernel, accumulates over rows a in [r0, r1), all columns b < n:
//   out[0] = sum_a G_aa ;  out[1] = sum_ab G_ab K_f,ab ;  out[2 + j] = sum_ab G_ab K_f,ab (x~_aj - x~_bj)^2
// K_f is recomputed from X (never read from HBM); K^-1 is streamed once: 8 N^2 bytes per full pass.
// ------------------------------------------------------------------------------------------------
template <int DP>
__global__ void __launch_bounds__(256)
grad_trace_kernel(const double* __restrict__ X, long ldx, long n, int d, const double* __restrict__ hyp,
                  const double* __restrict__ Kinv, long ldk, const double* __restrict__ alpha, long r0, long r1,
                  double* __restrict__ partial, int sym) {
    __shared__ double As[CK_T * CK_LD], Bs[CK_T * CK_LD], ea[CK_T], eb[CK_T], red[8][DP + 2];
    const long row0 = r0 + (long)blockIdx.y * CK_T, col0 = (long)blockIdx.x * CK_T;
    // sym: the summand is symmetric in (a, b), so only b <= a is visited (b < a counted twice): half the K^-1
    // traffic and exponentials, and only the lower triangle of K^-1 is read
    if (sym && col0 > row0 + CK_T - 1) {
        if (threadIdx.x < DP + 2) partial[((long)blockIdx.y * gridDim.x + blockIdx.x) * (DP + 2) + threadIdx.x] = 0.0;
        return;
    }
    stage_scaled_rows(X, ldx, n, row0, d, DP, hyp, As, ea);
    stage_scaled_rows(X, ldx, n, col0, d, DP, hyp, Bs, eb);
    __syncthreads();
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const double sf2 = hyp[HYP_SF2];
    double w[4][4];
    double acc_tr = 0.0, acc_k = 0.0;
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) w[r][c] = ea[ty * 4 + r] + eb[tx + 16 * c];
#pragma unroll
    for (int j = 0; j < DP; j++) {
        double av[4], bv[4];
#pragma unroll
        for (int r = 0; r < 4; r++) av[r] = As[(ty * 4 + r) * CK_LD + j];
#pragma unroll
        for (int c = 0; c < 4; c++) bv[c] = Bs[(tx + 16 * c) * CK_LD + j];
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) w[r][c] = fma(av[r], bv[c], w[r][c]);
    }
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const long gr = row0 + ty * 4 + r, gc = col0 + tx + 16 * c;
            double v = 0.0;
            if (gr < r1 && gr < n && gc < n && (!sym || gc <= gr)) {
                const double g = alpha[gr] * alpha[gc] - Kinv[gr * ldk + gc];
                const double kf = gr == gc ? sf2 : sf2 * exp(fmin(w[r][c], 0.0));
                v = g * kf * ((sym && gc < gr) ? 2.0 : 1.0);
                if (gr == gc) acc_tr += g;
            }
            w[r][c] = v;
            acc_k += v;
        }
    // per-dimension sums are reduced across the warp right away so they never live in registers together
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 2
    for (int j = 0; j < DP; j++) {
        double av[4], bv[4];
#pragma unroll
        for (int r = 0; r < 4; r++) av[r] = As[(ty * 4 + r) * CK_LD + j];
#pragma unroll
        for (int c = 0; c < 4; c++) bv[c] = Bs[(tx + 16 * c) * CK_LD + j];
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const double df = av[r] - bv[c];
                s = fma(w[r][c], df * df, s);
            }
        s = warp_sum(s);
        if (lane == 0) red[warp][2 + j] = s;
    }
    acc_tr = warp_sum(acc_tr);
    acc_k = warp_sum(acc_k);
    if (lane == 0) {
        red[warp][0] = acc_tr;
        red[warp][1] = acc_k;
    }
    __syncthreads();
    if (threadIdx.x < DP + 2) {
        double s = 0.0;
#pragma unroll
        for (int wi = 0; wi < 8; wi++) s += red[wi][threadIdx.x];
        partial[((long)blockIdx.y * gridDim.x + blockIdx.x) * (DP + 2) + threadIdx.x] = s;
    }
}

// out[v] = sum_b partial[b][v]   (one block per value v; deterministic order)
__global__ void reduce_partials_kernel(const double* __restrict__ partial, long n_blocks, int width,
                                       double* __restrict__ out) {
    __shared__ double sh[32];
    const int v = blockIdx.x;
    double s = 0.0;
    for (long b = threadIdx.x; b < n_blocks; b += blockDim.x) s += partial[b * width + v];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0;
        s = warp_sum(s);
        if (threadIdx.x == 0) out[v] = s;
    }
}

static int pad_dim(int d) {
    if (d <= 4) return d;
    if (d <= 8) return 8;
    if (d <= 16) return 16;
    return 32;
}

}  // namespace gpk

// ================================================================================================
// C ABI
// ================================================================================================
using namespace gpk;

extern "C" {

int gpk_pad_dim(int d) { return (d >= 1 && d <= HYP_MAXD) ? pad_dim(d) : -1; }
int gpk_hyp_stride(void) { return HYP_STRIDE; }
int gpk_train_tile(void) { return TP; }

long gpk_pack_train_elems(long n, int d) {
    if (n <= 0 || d < 1 || d > HYP_MAXD) return -1;
    const long tiles = (n + TP - 1) / TP;
    return tiles * TP * (pad_dim(d) + 2);
}

// K(A,B) [mode 0] or K(A,A) + noise I with identity padding [mode 1] into out (rows_pad x cols_pad, multiples of 64)
int gpk_cross_kernel(int dtype_a, const void* A, long lda, long na, long strideA, const double* B, long ldb, long nb,
                     long strideB, int d, const double* hyp, double* out, long ldo, long strideO, long rows_pad,
                     long cols_pad, int mode, int batch, void* stream) {
    GPK_ARG(dtype_a == 0 || dtype_a == 1, "dtype_a must be 0 (f32) or 1 (f64)");
    GPK_ARG(A && B && hyp && out, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD, "d must be in [1, 32]");
    GPK_ARG(rows_pad % CK_T == 0 && cols_pad % CK_T == 0 && rows_pad >= na && cols_pad >= nb, "padded sizes");
    GPK_ARG(ldo >= cols_pad && batch >= 1, "ldo / batch");
    GPK_ARG(mode == 0 || mode == 1, "mode");
    dim3 grid((unsigned)(cols_pad / CK_T), (unsigned)(rows_pad / CK_T), (unsigned)batch);
    if (dtype_a == 0)
        cross_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)A, lda, na, strideA, B, ldb, nb, strideB,
                                                                    d, hyp, out, ldo, strideO, mode);
    else
        cross_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>((const double*)A, lda, na, strideA, B, ldb, nb,
                                                                     strideB, d, hyp, out, ldo, strideO, mode);
    GPK_LAUNCHED();
    return 0;
}

int gpk_pack_train(int dtype, const double* X, long ldx, long n, int d, const double* alpha, const double* hyp,
                   void* pack, long pack_elems, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(X && alpha && hyp && pack, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0, "n / d");
    GPK_ARG(pack_elems >= gpk_pack_train_elems(n, d), "pack buffer too small");
    const long tiles = (n + TP - 1) / TP;
    const int dp = pad_dim(d);
    const unsigned blocks = (unsigned)((tiles * TP + 255) / 256);
    if (dtype == 0)
        pack_train_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(X, ldx, n, d, dp, alpha, hyp, (float*)pack, tiles);
    else
        pack_train_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>(X, ldx, n, d, dp, alpha, hyp, (double*)pack, tiles);
    GPK_LAUNCHED();
    return 0;
}

long gpk_mean_jac_work_bytes(int dtype, long q, long n, int d, int batch) {
    // upper bound: every train tile its own split
    const long tiles = (n + TP - 1) / TP;
    return tiles * q * (pad_dim(d) + 1) * (dtype == 0 ? 4 : 8) * batch;
}

// fused posterior mean + Jacobian for `batch` GPs of identical shape.
int gpk_predict_mean_jac(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                         const void* pack, long stride_pack, long n, void* mean, long stride_mean, void* jac,
                         long ldj, long stride_jac, void* work, long work_bytes, int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(Z && hyp && pack && mean && jac, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0 && batch >= 1, "n / d / batch");
    if (q == 0) return 0;
    GPK_ARG(q > 0 && ldz >= d && ldj >= d, "q / ldz / ldj");
    GPK_ARG(((uintptr_t)pack % 16) == 0, "pack must be 16-byte aligned");
    const int dp = pad_dim(d);
    const int esz = dtype == 0 ? 4 : 8;
    MeanJacArgs a{};
    a.Z = Z; a.ldz = ldz; a.q = q; a.strideZ = strideZ; a.d = d; a.hyp = hyp;
    a.pack = pack; a.stride_pack = stride_pack;
    a.exp_tab = nullptr;
    if (dtype == 1) {
        a.exp_tab = exp_table((cudaStream_t)stream);
        GPK_ARG(a.exp_tab != nullptr, "could not allocate the exp table");
    }
    a.n_tiles = (int)((n + TP - 1) / TP);
    a.mean = mean; a.stride_mean = stride_mean; a.jac = jac; a.ldj = ldj; a.stride_jac = stride_jac;
    a.work = work;       // the split of the training tiles over CTAs is chosen per kernel instantiation (mj_pick_split)
    if (dtype == 0) return launch_mean_jac<float>(a, dp, batch, work_bytes, (cudaStream_t)stream);
    return launch_mean_jac<double>(a, dp, batch, work_bytes, (cudaStream_t)stream);
}

int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                     const double* hyp, void* var, long stride_var, int batch, void* stream) {
    GPK_ARG(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    GPK_ARG(colsq && hyp && var && tiles_m >= 1 && q > 0, "arguments");
    dim3 grid((unsigned)((q + 255) / 256), 1, (unsigned)batch);
    if (dtype == 0)
        var_finalize_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(colsq, ld, stride_colsq, tiles_m, q, hyp, (float*)var, stride_var);
    else
        var_finalize_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(colsq, ld, stride_colsq, tiles_m, q, hyp, (double*)var, stride_var);
    GPK_LAUNCHED();
    return 0;
}

long gpk_grad_trace_work_elems(long n, int d, long r0, long r1) {
    const long by = (r1 - r0 + CK_T - 1) / CK_T, bx = (n + CK_T - 1) / CK_T;
    return by * bx * (pad_dim(d) + 2);
}

// out[0] = tr G over rows [r0,r1), out[1] = sum G*K_f, out[2+j] = sum G*K_f*(x~_aj - x~_bj)^2, j < d
static int mll_grad_reduce(const double* X, long ldx, long n, int d, const double* hyp, const double* Kinv, long ldk,
                           const double* alpha, long r0, long r1, double* work, long work_elems, double* out, int sym,
                           void* stream) {
    GPK_ARG(X && hyp && Kinv && alpha && work && out, "null pointer");
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0, "n / d");
    GPK_ARG(0 <= r0 && r0 < r1 && r1 <= n, "row range");
    GPK_ARG(work_elems >= gpk_grad_trace_work_elems(n, d, r0, r1), "scratch too small");
    const int dp = pad_dim(d);
    dim3 grid((unsigned)((n + CK_T - 1) / CK_T), (unsigned)((r1 - r0 + CK_T - 1) / CK_T));
    cudaStream_t st = (cudaStream_t)stream;
#define GT_CASE(DPV) case DPV: grad_trace_kernel<DPV><<<grid, 256, 0, st>>>(X, ldx, n, d, hyp, Kinv, ldk, alpha, r0, r1, work, sym); break;
    switch (dp) {
        GT_CASE(1) GT_CASE(2) GT_CASE(3) GT_CASE(4) GT_CASE(8) GT_CASE(16) GT_CASE(32)
        default: return fail(-1, "grad_trace: unsupported padded dimension %d", dp);
    }
#undef GT_CASE
    GPK_LAUNCHED();
    reduce_partials_kernel<<<dp + 2, 256, 0, st>>>(work, (long)grid.x * grid.y, dp + 2, out);
    GPK_LAUNCHED();
    return 0;
}

int gpk_mll_grad_reduce(const double* X, long ldx, long n, int d, const double* hyp, const double* Kinv, long ldk,
                        const double* alpha, long r0, long r1, double* work, long work_elems, double* out,
                        void* stream) {
    return mll_grad_reduce(X, ldx, n, d, hyp, Kinv, ldk, alpha, r0, r1, work, work_elems, out, 0, stream);
}

int gpk_mll_grad_reduce_sym(const double* X, long ldx, long n, int d, const double* hyp, const double* Kinv, long ldk,
                            const double* alpha, long r0, long r1, double* work, long work_elems, double* out,
                            void* stream) {
    return mll_grad_reduce(X, ldx, n, d, hyp, Kinv, ldk, alpha, r0, r1, work, work_elems, out, 1, stream);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of the gradient trace: the per-CTA partial sums live in a file-scope device array
// instead of a caller-supplied workspace, so the call needs no cudaMalloc anywhere (callers under CUDA-graph
// capture, serving loops, harnesses that forbid allocation).  Every context that loads the library holds the
// array (GT_STATIC_ELEMS doubles = 1.06 MiB) until exit.
// In HOST code `g_gt_static_partials` is only a stand-in for the symbol, not a device address: its address is
// looked up once per device with cudaGetSymbolAddress, with that device current, and that pointer is what the
// kernels get.
// ------------------------------------------------------------------------------------------------
namespace gpk {
constexpr long GT_STATIC_MAX_BLOCKS = 4096;                        // (n / 64)^2 CTAs of a full pass at n = 4096
constexpr long GT_STATIC_ELEMS = GT_STATIC_MAX_BLOCKS * (HYP_MAXD + 2);
static __device__ __align__(256) double g_gt_static_partials[GT_STATIC_ELEMS];

static int gt_static_scratch(double** out_ptr) {
    static std::mutex mu;
    static double* ptr_dev[64] = {};
    int dev = -1;
    GPK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(-1, "grad_trace_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr_dev[dev]) {
        void* p = nullptr;
        GPK_CUDA(cudaGetSymbolAddress(&p, g_gt_static_partials));   // of the CURRENT device's copy of the array
        if (!p) return fail(-1, "grad_trace_static: cudaGetSymbolAddress returned a null address");
        ptr_dev[dev] = (double*)p;
    }
    *out_ptr = ptr_dev[dev];
    return 0;
}
}  // namespace gpk

extern "C" {

long gpk_mll_grad_reduce_static_max_elems(void) { return GT_STATIC_ELEMS; }

int gpk_mll_grad_reduce_static(const double* X, long ldx, long n, int d, const double* hyp, const double* Kinv,
                               long ldk, const double* alpha, long r0, long r1, int sym, double* out, void* stream) {
    GPK_ARG(d >= 1 && d <= HYP_MAXD && n > 0, "n / d");
    GPK_ARG(0 <= r0 && r0 < r1 && r1 <= n, "row range");
    GPK_ARG(sym == 0 || sym == 1, "sym must be 0 or 1");
    const long need = gpk_grad_trace_work_elems(n, d, r0, r1);
    if (need > GT_STATIC_ELEMS)
        return fail(-1, "gpk_mll_grad_reduce_static: n = %ld, rows [%ld, %ld), d = %d need %ld scratch doubles, the "
                        "built-in scratch holds %ld (a full pass up to n = 4096 at d = 32); use gpk_mll_grad_reduce with a "
                        "caller-supplied workspace", n, r0, r1, d, need, (long)GT_STATIC_ELEMS);
    double* scratch = nullptr;
    if (int rc = gt_static_scratch(&scratch)) return rc;
    // every CTA of the trace kernel (also the ones the symmetric form skips) WRITES its slot and the reduction reads
    // exactly the slots of this launch, so nothing of an earlier call's contents is read: no clear is needed
    return mll_grad_reduce(X, ldx, n, d, hyp, Kinv, ldk, alpha, r0, r1, scratch, GT_STATIC_ELEMS, out, sym, stream);
}

}  // extern "C"
