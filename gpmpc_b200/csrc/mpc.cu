// Chance-constraint tightening of the GP-MPC: the covariance recursion of
// GPMPC.propagate_constraint_limits (gpmpc/gpmpc.py:425-498) batched over rollouts, consuming the GP variance
// output of gpk_predict_i8 / gpk_predict_var directly in device memory.
//
// Per rollout r and stage i (reference lines in brackets):
//   c_i      = dt^2 (var_i + noise) .* [cos(phi) sin(theta)^2, sin(phi)^2, (cos(phi) cos(theta))^2, 1, 1]   [447-470]
//              with var = [T, T, T, R, P] the three GPs' predictive variances (thrust GP on the first three rows)
//              and the likelihood noise added once more                                                [483-485]
//   state_constraint[:, i] = -inverse_cdf |A_x| sqrt(diag P_i),   A_x = [-I; I]                        [472-476, 487]
//   input_constraint[:, i] = -inverse_cdf |A_u| sqrt(diag(K P_i K^T))                                  [480, 488]
//   P_{i+1}  = A P A^T + A P K^T B^T + B K P A^T + B K P K^T B^T + Bd diag(c_i) Bd^T                    [489-495]
//            = (A + B K) P (A + B K)^T + Bd diag(c_i) Bd^T         (the same sum, factored; P_0 = 0)   [436]
// Sixteen threads per rollout (a 4x4 grid of BS x BS register blocks of P, two rollouts per warp, warp-level
// synchronisation only); the stage loop is sequential (T steps of two small products), the batch of rollouts fills
// the machine.  Limits are produced a chunk of stages at a time so that sqrt and the stores along the stage axis stay
// off the recursion's critical path.
#include "common.cuh"

namespace gpk {

constexpr int TC_ND = 5;        // uncertain directions: dx, dy, dz from the thrust GP, dphi, dtheta (gpmpc.py:68)
constexpr int TC_NGP = 3;       // thrust, roll, pitch GPs (gpmpc.py:59)
constexpr int TC_MAX_NX = 16, TC_MAX_NU = 8;
constexpr int TC_CHUNK = 16;    // stages per chunk: disturbances evaluated ahead, limits written behind
constexpr int TC_WARPS = 4;     // warps per CTA
constexpr int TC_RPC = 2 * TC_WARPS;   // rollouts per CTA

struct TightenArgs {
    const double* var; long stride_var;    // [3][R*T] GP variances
    const double* Z; long ldz;             // [R*T][ldz] stage inputs z = (x, u); columns idx_phi / idx_theta are read
    int idx_phi, idx_theta;
    const double* hyp;                     // 3 hyper-parameter blocks (likelihood noise)
    const double *A, *B, *K, *Bd;          // row-major nx x nx, nx x nu, nu x nx, nx x 5
    int nx, nu, T;
    long R;
    double dt, inv_cdf;
    int var_has_noise;                     // 1: var is the predictive variance (latent + noise); 0: latent var_f
    double* state_c;                       // [R][2 nx][T + 1]
    double* input_c;                       // [R][2 nu][T]
};

// a variance that rounding pushed below zero is 0 (numpy's sqrt would give NaN there); NaN inputs stay NaN
__device__ __forceinline__ double clamp0(double v) { return v < 0.0 ? 0.0 : v; }

template <int BS>
struct TightenSmem {
    static constexpr int NX = 4 * BS, LD = NX + 2;          // even leading dimension: 16-byte row pairs
    double Ms[NX * LD];                                      // closed loop M = A + B K
    double Ks[TC_MAX_NU * LD];
    double Ps[TC_RPC][NX * LD];                              // P_i, rows
    double Ts[TC_RPC][NX * LD];                              // (P M^T)^T
    double Bds[NX * TC_ND];
    double cs[TC_RPC][TC_CHUNK * TC_ND];                     // stage disturbances of the current chunk
    double hx[TC_RPC][TC_CHUNK * NX];                        // diag P_i of the current chunk
    double hu[TC_RPC][TC_CHUNK * TC_MAX_NU];                 // diag(K P_i K^T) of the current chunk
};

template <int BS>
__global__ void __launch_bounds__(TC_WARPS * 32) tighten_kernel(TightenArgs a) {
    using SM = TightenSmem<BS>;
    constexpr int NX = SM::NX, LD = SM::LD;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SM& sm = *reinterpret_cast<SM*>(smem_raw);
    const int t = threadIdx.x, lane = t & 31, hl = lane & 15, bi = hl >> 2, bj = hl & 3;
    const int slot = (t >> 5) * 2 + (lane >> 4);
    const int nx = a.nx, nu = a.nu, T = a.T;
    const long r_true = (long)blockIdx.x * TC_RPC + slot;
    const bool active = r_true < a.R;
    const long r = active ? r_true : a.R - 1;               // idle half-warps shadow the last rollout, without stores

    for (int e = t; e < NX * NX; e += TC_WARPS * 32) {
        const int ra = e / NX, cb = e % NX;
        double m = 0.0;
        if (ra < nx && cb < nx) {
            m = a.A[ra * nx + cb];
            for (int j = 0; j < nu; j++) m = fma(a.B[ra * nu + j], a.K[j * nx + cb], m);
        }
        sm.Ms[ra * LD + cb] = m;
        if (ra < TC_MAX_NU) sm.Ks[ra * LD + cb] = (ra < nu && cb < nx) ? a.K[ra * nx + cb] : 0.0;
        if (cb < TC_ND) sm.Bds[ra * TC_ND + cb] = (ra < nx) ? a.Bd[ra * TC_ND + cb] : 0.0;
    }
    for (int e = t; e < TC_RPC * NX * LD; e += TC_WARPS * 32) (&sm.Ps[0][0])[e] = 0.0;
    __syncthreads();

    double* Ps = sm.Ps[slot];
    double* Ts = sm.Ts[slot];
    double* cs = sm.cs[slot];
    double* hx = sm.hx[slot];
    double* hu = sm.hu[slot];
    const double noise_mult = a.var_has_noise ? 1.0 : 2.0;
    const double dt2 = a.dt * a.dt;
    double* sc = a.state_c + r * 2 * nx * (long)(T + 1);
    double* ic = a.input_c + r * 2 * nu * (long)T;
    const int a0 = BS * bi, b0 = BS * bj;
    double p_own[BS][BS];       // this thread's block of P_i
#pragma unroll
    for (int e = 0; e < BS; e++)
#pragma unroll
        for (int f = 0; f < BS; f++) p_own[e][f] = 0.0;

    for (int i = 0; i <= T; i++) {
        const int ic_ = i % TC_CHUNK;
        if (ic_ == 0 && i < T) {
            // stage disturbances c_i (5 per stage) of the next TC_CHUNK stages, one stage per thread
            const int j = hl;
            if (i + j < T) {
                const long zi = r * T + i + j;
                double sp, cp, sth, cth;
                sincos(a.Z[zi * a.ldz + a.idx_phi], &sp, &cp);
                sincos(a.Z[zi * a.ldz + a.idx_theta], &sth, &cth);
                double v[TC_NGP];
#pragma unroll
                for (int g = 0; g < TC_NGP; g++)
                    v[g] = (a.var[g * a.stride_var + zi] + noise_mult * a.hyp[g * HYP_STRIDE + HYP_NOISE]) * dt2;
                cs[j * TC_ND + 0] = v[0] * (cp * sth * sth);
                cs[j * TC_ND + 1] = v[0] * (sp * sp);
                cs[j * TC_ND + 2] = v[0] * ((cp * cth) * (cp * cth));
                cs[j * TC_ND + 3] = v[1];
                cs[j * TC_ND + 4] = v[2];
            }
        }
        if (bi == bj) {
#pragma unroll
            for (int e = 0; e < BS; e++) hx[ic_ * NX + a0 + e] = p_own[e][e];     // diag P_i
        }
        if (i < T) {
            // diag(K P_i K^T): each thread's block contributes k_a^T P_block k_b, summed over the 16 threads
            for (int j = 0; j < nu; j++) {
                double s = 0.0;
#pragma unroll
                for (int e = 0; e < BS; e++) {
                    double te = 0.0;
#pragma unroll
                    for (int f = 0; f < BS; f++) te = fma(p_own[e][f], sm.Ks[j * LD + b0 + f], te);
                    s = fma(sm.Ks[j * LD + a0 + e], te, s);
                }
#pragma unroll
                for (int o = 1; o < 16; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (hl == j) hu[ic_ * TC_MAX_NU + j] = s;
            }
            double acc[BS][BS];
            // phase 1 (reads P): T1 = P M^T, stored transposed
#pragma unroll
            for (int e = 0; e < BS; e++)
#pragma unroll
                for (int f = 0; f < BS; f++) acc[e][f] = 0.0;
#pragma unroll
            for (int k = 0; k < NX; k += 2) {
                double2 pa[BS], mb[BS];
#pragma unroll
                for (int e = 0; e < BS; e++) pa[e] = *reinterpret_cast<const double2*>(&Ps[(a0 + e) * LD + k]);
#pragma unroll
                for (int f = 0; f < BS; f++) mb[f] = *reinterpret_cast<const double2*>(&sm.Ms[(b0 + f) * LD + k]);
#pragma unroll
                for (int e = 0; e < BS; e++)
#pragma unroll
                    for (int f = 0; f < BS; f++) acc[e][f] = fma(pa[e].y, mb[f].y, fma(pa[e].x, mb[f].x, acc[e][f]));
            }
#pragma unroll
            for (int e = 0; e < BS; e++)
#pragma unroll
                for (int f = 0; f < BS; f++) Ts[(b0 + f) * LD + a0 + e] = acc[e][f];
            __syncwarp();
            // phase 2 (reads T1^T, c; writes P): P' = M T1 + Bd diag(c) Bd^T
            const double* c = cs + ic_ * TC_ND;
#pragma unroll
            for (int e = 0; e < BS; e++)
#pragma unroll
                for (int f = 0; f < BS; f++) acc[e][f] = 0.0;
#pragma unroll
            for (int m = 0; m < TC_ND; m++) {
                const double cm = c[m];
                double ba[BS], bb[BS];
#pragma unroll
                for (int e = 0; e < BS; e++) ba[e] = sm.Bds[(a0 + e) * TC_ND + m];
#pragma unroll
                for (int f = 0; f < BS; f++) bb[f] = cm * sm.Bds[(b0 + f) * TC_ND + m];
#pragma unroll
                for (int e = 0; e < BS; e++)
#pragma unroll
                    for (int f = 0; f < BS; f++) acc[e][f] = fma(ba[e], bb[f], acc[e][f]);
            }
#pragma unroll
            for (int k = 0; k < NX; k += 2) {
                double2 ma[BS], tb[BS];
#pragma unroll
                for (int e = 0; e < BS; e++) ma[e] = *reinterpret_cast<const double2*>(&sm.Ms[(a0 + e) * LD + k]);
#pragma unroll
                for (int f = 0; f < BS; f++) tb[f] = *reinterpret_cast<const double2*>(&Ts[(b0 + f) * LD + k]);
#pragma unroll
                for (int e = 0; e < BS; e++)
#pragma unroll
                    for (int f = 0; f < BS; f++) acc[e][f] = fma(ma[e].y, tb[f].y, fma(ma[e].x, tb[f].x, acc[e][f]));
            }
#pragma unroll
            for (int e = 0; e < BS; e++)
#pragma unroll
                for (int f = 0; f < BS; f++) {
                    p_own[e][f] = acc[e][f];
                    Ps[(a0 + e) * LD + b0 + f] = acc[e][f];
                }
        }
        __syncwarp();
        if (ic_ == TC_CHUNK - 1 || i == T) {
            // limits of stages i0 .. i: -inverse_cdf sqrt(.) for both signs of A = [-I; I], written along the stage axis
            const int i0 = i - ic_, cnt = ic_ + 1, cnt_u = min(cnt, T - i0);
            if (active) {
                for (int e = hl; e < nx * cnt; e += 16) {
                    const int row = e / cnt, j = e % cnt;
                    const double v = -a.inv_cdf * sqrt(clamp0(hx[j * NX + row]));
                    sc[(long)row * (T + 1) + i0 + j] = v;
                    sc[(long)(nx + row) * (T + 1) + i0 + j] = v;
                }
                for (int e = hl; e < nu * cnt_u; e += 16) {
                    const int row = e / cnt_u, j = e % cnt_u;
                    const double v = -a.inv_cdf * sqrt(clamp0(hu[j * TC_MAX_NU + row]));
                    ic[(long)row * T + i0 + j] = v;
                    ic[(long)(nu + row) * T + i0 + j] = v;
                }
            }
            __syncwarp();
        }
    }
}

template <int BS>
static int launch_tighten(const TightenArgs& a, cudaStream_t st) {
    const int smem = (int)sizeof(TightenSmem<BS>);
    static bool configured = false;
    if (!configured) {
        GPK_CUDA(cudaFuncSetAttribute(tighten_kernel<BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = true;
    }
    const unsigned grid = (unsigned)((a.R + TC_RPC - 1) / TC_RPC);
    tighten_kernel<BS><<<grid, TC_WARPS * 32, smem, st>>>(a);
    GPK_LAUNCHED();
    return 0;
}

}  // namespace gpk

using namespace gpk;

extern "C" int gpk_tighten_constraints(const double* var, long stride_var, int var_has_noise, const double* Z,
                                       long ldz, int idx_phi, int idx_theta, const double* hyp, const double* A,
                                       const double* B, const double* K, const double* Bd, int nx, int nu, int T,
                                       long rollouts, double dt, double inverse_cdf, double* state_constraint,
                                       double* input_constraint, void* stream) {
    GPK_ARG(nx >= 1 && nx <= TC_MAX_NX, "nx must be in 1..16");
    GPK_ARG(nu >= 1 && nu <= TC_MAX_NU && nu <= nx, "nu must be in 1..8 and <= nx");
    GPK_ARG(T >= 0 && rollouts >= 0 && rollouts < (1L << 31), "T >= 0, 0 <= rollouts < 2^31");
    GPK_ARG(idx_phi >= 0 && idx_theta >= 0 && idx_phi < ldz && idx_theta < ldz, "angle columns outside Z");
    if (rollouts == 0) return 0;
    GPK_ARG(state_constraint && (T == 0 || input_constraint), "null output");
    GPK_ARG(A && B && K && Bd && hyp, "null matrix");
    GPK_ARG(T == 0 || (var && Z), "null var / Z");
    TightenArgs a{var, stride_var, Z, ldz, idx_phi, idx_theta, hyp, A, B, K, Bd, nx, nu, T, rollouts, dt, inverse_cdf,
                  var_has_noise, state_constraint, input_constraint};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (nx <= 8) return launch_tighten<2>(a, st);
    if (nx <= 12) return launch_tighten<3>(a, st);
    return launch_tighten<4>(a, st);
}
