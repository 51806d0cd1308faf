// One-launch MLL value + gradient for SMALL exact GPs (n <= 128): the sizes the reference itself reaches
// (scripts/gp_mpc_config.yaml:24,30 -> N <= 45; three GPs fitted one after the other, gpmpc/gpmpc.py:159-162).
//
// At these sizes the blocked path of linalg.cu is a chain of ~20 dependent launches of which each is one CTA on 148
// SMs, plus three host round trips per Adam iteration -- launch latency, not arithmetic.  Here ONE CTA per GP does the
// whole iteration of gpmpc/gp.py:59-63 out of shared memory:
//   K = K_f + noise I  ->  L (Cholesky, in place, packed lower triangle)  ->  W = L^-1 (second packed triangle)
//   t = W y, alpha = W^T t, quad = t.t = y^T K^-1 y, logdet = 2 sum log L_ii
//   K^-1_ab = sum_{i >= max(a,b)} W_ia W_ib formed on the fly inside the gradient trace
//   G = alpha alpha^T - K^-1:  out = [tr G, sum G K_f, sum G K_f ((x_aj - x_bj) / ell_j)^2 ...]   (as gpk_mll_grad_reduce)
// `batch` GPs of identical (n, d) run concurrently as `batch` CTAs (the reference's thrust / roll / pitch GPs with
// their inputs zero-padded to a common d).
#include "common.cuh"
#include <cmath>

namespace gpk {

constexpr int SM_MAXN = 128;
constexpr int SM_THREADS = 256;
constexpr int SM_OUT = 40;                     // doubles per GP: info, quad, logdet, pad, partials[<= 34]

__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }   // packed lower triangle, j <= i

struct SmallArgs {
    const double* X; long ldx, strideX;       // [batch][n][ldx]
    const double* y; long stride_y;            // [batch][n]
    const double* hyp;                         // [batch][HYP_STRIDE]
    int n, d, dp;
    double* out;                               // [batch][SM_OUT]
    double* alpha; long stride_alpha;          // optional [batch][n]
    double* K; double* Kinv; long ldk, strideK; // optional dense [batch][n][ldk] outputs: K_f + noise I and its inverse
};

__global__ void __launch_bounds__(SM_THREADS, 1) mll_small_kernel(const SmallArgs a) {
    extern __shared__ __align__(16) double sm[];
    const int n = a.n, d = a.d, dp = a.dp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long gp = blockIdx.x;
    const double* hyp = a.hyp + gp * HYP_STRIDE;
    const int ntri = n * (n + 1) / 2;
    double* Lp = sm;                           // packed L  (K first)
    double* Wp = Lp + ntri;                    // packed W = L^-1
    double* Xs = Wp + ntri;                    // [n][d] centred / scaled inputs
    double* ys = Xs + n * d;                   // y, later alpha
    double* ts = ys + n;                       // t = W y
    double* red = ts + n;                      // [8][dp + 2] warp partials
    __shared__ int s_info;
    __shared__ double s_piv;
    if (tid == 0) s_info = 0;
    const double sf2 = hyp[HYP_SF2], noise = hyp[HYP_NOISE];
    for (int e = tid; e < n * d; e += SM_THREADS) {
        const int i = e / d, j = e - i * d;
        Xs[e] = (a.X[gp * a.strideX + (long)i * a.ldx + j] - hyp[HYP_CENTER + j]) * hyp[HYP_INV_ELL + j];
    }
    for (int i = tid; i < n; i += SM_THREADS) ys[i] = a.y[gp * a.stride_y + i];
    __syncthreads();
    // ---- K = sf2 exp(-1/2 r^2) + noise I, lower triangle (gpmpc/gp.py:12-14, 45) -----------------
    constexpr int NW = SM_THREADS / 32;
    for (int i = warp; i < n; i += NW)                     // row i by warp, columns j <= i by lane
        for (int j = lane; j <= i; j += 32) {
            double r2 = 0.0;
            for (int k = 0; k < d; k++) {
                const double df = Xs[i * d + k] - Xs[j * d + k];
                r2 = fma(df, df, r2);
            }
            const double kv = i == j ? sf2 + noise : sf2 * exp(-0.5 * r2);
            Lp[tri(i, j)] = kv;
            if (a.K) {
                a.K[gp * a.strideK + (long)i * a.ldk + j] = kv;
                a.K[gp * a.strideK + (long)j * a.ldk + i] = kv;
            }
        }
    __syncthreads();
    // ---- Cholesky, right-looking, in place; one column per step --------------------------------------
    double logdet = 0.0;
    for (int j = 0; j < n; j++) {
        if (tid == 0) {
            const double p = Lp[tri(j, j)];
            if (!(p > 0.0) && s_info == 0) s_info = j + 1;         // LAPACK-style: first non-positive pivot
            s_piv = p > 0.0 ? sqrt(p) : 1.0;
        }
        __syncthreads();
        const double piv = s_piv, inv = 1.0 / piv;
        logdet += log(piv);
        for (int i = j + tid; i < n; i += SM_THREADS) Lp[tri(i, j)] = i == j ? piv : Lp[tri(i, j)] * inv;
        __syncthreads();
        // trailing update of the lower triangle of rows > j: A_ik -= L_ij L_kj, j < k <= i
        for (int i = j + 1 + warp; i < n; i += NW) {
            const double lij = Lp[tri(i, j)];
            for (int k = j + 1 + lane; k <= i; k += 32) Lp[tri(i, k)] = fma(-lij, Lp[tri(k, j)], Lp[tri(i, k)]);
        }
        __syncthreads();
    }
    // ---- W = L^-1: thread c owns column c, forward substitution down the rows (no barriers) ----------------
    for (int c = tid; c < n; c += SM_THREADS) {
        Wp[tri(c, c)] = 1.0 / Lp[tri(c, c)];
        for (int i = c + 1; i < n; i++) {
            double s0 = 0.0, s1 = 0.0;
            int k = c;
            for (; k + 1 < i; k += 2) {
                s0 = fma(Lp[tri(i, k)], Wp[tri(k, c)], s0);
                s1 = fma(Lp[tri(i, k + 1)], Wp[tri(k + 1, c)], s1);
            }
            if (k < i) s0 = fma(Lp[tri(i, k)], Wp[tri(k, c)], s0);
            Wp[tri(i, c)] = -(s0 + s1) / Lp[tri(i, i)];
        }
    }
    __syncthreads();
    // ---- t = W y ; alpha = W^T t ; quad = t . t ------------------------------------------------------
    for (int i = tid; i < n; i += SM_THREADS) {
        double s = 0.0;
        for (int k = 0; k <= i; k++) s = fma(Wp[tri(i, k)], ys[k], s);
        ts[i] = s;
    }
    __syncthreads();
    double alpha_i = 0.0;
    for (int i = tid; i < n; i += SM_THREADS) {
        double s = 0.0;
        for (int k = i; k < n; k++) s = fma(Wp[tri(k, i)], ts[k], s);
        alpha_i = s;                                                 // n <= 128 < SM_THREADS: at most one per thread
    }
    __syncthreads();
    if (tid < n) {
        ys[tid] = alpha_i;
        if (a.alpha) a.alpha[gp * a.stride_alpha + tid] = alpha_i;
    }
    __syncthreads();
    // ---- gradient trace over the lower triangle (b < a counted twice) -----------------------------------------
    double acc[2 + HYP_MAXD];
    for (int v = 0; v < dp + 2; v++) acc[v] = 0.0;
    for (int i = warp; i < n; i += NW)                     // pair (a = i, b = j), j <= i
        for (int j = lane; j <= i; j += 32) {
            double kinv = 0.0;
            for (int r = i; r < n; r++) kinv = fma(Wp[tri(r, i)], Wp[tri(r, j)], kinv);
            if (a.Kinv) {
                a.Kinv[gp * a.strideK + (long)i * a.ldk + j] = kinv;
                a.Kinv[gp * a.strideK + (long)j * a.ldk + i] = kinv;
            }
            const double g = ys[i] * ys[j] - kinv;
            double r2 = 0.0;
            for (int k = 0; k < d; k++) {
                const double df = Xs[i * d + k] - Xs[j * d + k];
                r2 = fma(df, df, r2);
            }
            const double kf = i == j ? sf2 : sf2 * exp(-0.5 * r2);
            const double v = g * kf * (i == j ? 1.0 : 2.0);
            if (i == j) acc[0] += g;
            acc[1] += v;
            for (int k = 0; k < d; k++) {
                const double df = Xs[i * d + k] - Xs[j * d + k];
                acc[2 + k] = fma(v, df * df, acc[2 + k]);
            }
        }
    double q = 0.0;
    for (int i = tid; i < n; i += SM_THREADS) q = fma(ts[i], ts[i], q);
    q = warp_sum(q);
    for (int v = 0; v < dp + 2; v++) {
        const double s = warp_sum(acc[v]);
        if (lane == 0) red[warp * (dp + 2) + v] = s;
    }
    __shared__ double s_quad[SM_THREADS / 32];
    if (lane == 0) s_quad[warp] = q;
    __syncthreads();
    double* out = a.out + gp * SM_OUT;
    if (tid < dp + 2) {
        double s = 0.0;
        for (int w = 0; w < SM_THREADS / 32; w++) s += red[w * (dp + 2) + tid];
        out[4 + tid] = s;
    }
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < SM_THREADS / 32; w++) s += s_quad[w];
        out[0] = (double)s_info;
        out[1] = s;
        out[2] = 2.0 * logdet;
        out[3] = 0.0;
    }
}

static int small_pad_dim(int d) { return d <= 4 ? d : d <= 8 ? 8 : d <= 16 ? 16 : 32; }

}  // namespace gpk

using namespace gpk;

extern "C" {

int gpk_mll_small_max_n(void) { return SM_MAXN; }
int gpk_mll_small_out_stride(void) { return SM_OUT; }

// out[gp][0] = potrf info (first non-positive pivot, 1-based, 0 = ok), [1] = y^T K^-1 y, [2] = log det K,
// [4 + v] = the d + 2 gradient sums of gpk_mll_grad_reduce (padded to gpk_pad_dim(d) + 2 entries).  alpha (optional)
// receives K^-1 y; K / Kinv (optional, both or neither) receive the dense n x n matrices K_f + noise I and its inverse
// (gp.K, gp.K_inv of gpmpc/gp.py:43-46), row stride ldk, GP stride strideK.  One CTA per GP; n <= gpk_mll_small_max_n().
int gpk_mll_small(const double* X, long ldx, long strideX, const double* y, long stride_y, long n, int d, const double* hyp,
                  double* out, double* alpha, long stride_alpha, double* K, double* Kinv, long ldk, long strideK,
                  int batch, void* stream) {
    GPK_ARG(X && y && hyp && out, "null pointer");
    GPK_ARG(n >= 1 && n <= SM_MAXN && d >= 1 && d <= HYP_MAXD && ldx >= d && batch >= 1, "n / d / ldx / batch");
    GPK_ARG((K == nullptr) == (Kinv == nullptr) && (K == nullptr || ldk >= n), "K and Kinv go together, ldk >= n");
    SmallArgs a{};
    a.K = K; a.Kinv = Kinv; a.ldk = ldk; a.strideK = strideK;
    a.X = X; a.ldx = ldx; a.strideX = strideX; a.y = y; a.stride_y = stride_y; a.hyp = hyp;
    a.n = (int)n; a.d = d; a.dp = small_pad_dim(d); a.out = out; a.alpha = alpha; a.stride_alpha = stride_alpha;
    const int ntri = (int)(n * (n + 1) / 2);
    const int smem = (2 * ntri + (int)n * d + 2 * (int)n + 8 * (a.dp + 2)) * (int)sizeof(double);
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(mll_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done = true;
    }
    mll_small_kernel<<<batch, SM_THREADS, smem, (cudaStream_t)stream>>>(a);
    GPK_LAUNCHED();
    return 0;
}

}  // extern "C"
