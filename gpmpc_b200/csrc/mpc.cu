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

extern "C" long gpk_mean_jac_work_bytes(int dtype, long q, long n, int d, int batch);
extern "C" int gpk_predict_mean_jac(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                                    const void* pack, long stride_pack, long n, void* mean, long stride_mean, void* jac,
                                    long ldj, long stride_jac, void* work, long work_bytes, int batch, void* stream);

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
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& configured = attr_dev[current_device()];
    if (!configured) {
        GPK_CUDA(cudaFuncSetAttribute(tighten_kernel<BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = true;
    }
    const unsigned grid = (unsigned)((a.R + TC_RPC - 1) / TC_RPC);
    tighten_kernel<BS><<<grid, TC_WARPS * 32, smem, st>>>(a);
    GPK_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Batched discrete stage dynamics with sensitivities: what acados evaluates per stage for the
// model built by GPMPC.setup_acados_model (gpmpc/gpmpc.py:166-224),
//   f_cont(x, u) = prior(x, u) + res_dyn(x, u),    res_dyn = (0, T c(phi) s(theta), 0, -T s(phi), 0, T c(phi) c(theta),
//                  0, 0, 0, R, P, 0)                                                          [gpmpc.py:193-199]
//   T, R, P      = GP means at z[idx_T] = (u0), z[idx_R] = (phi, dphi, u1), z[idx_P] = (theta, dtheta, u2)   [172, 187-191]
//   f_disc       = x + dt/6 (k1 + 2 k2 + 2 k3 + k4), classical RK4 of f_cont with u held                  [201-206]
// together with A = d f_disc / dx and B = d f_disc / du (CasADi AD inside acados), for q independent (x, u) pairs.
// The prior is the attitude model the reference takes from crazyflow.sim.symbolic.symbolic_attitude (absent here;
// restated from its use at gpmpc.py:128-146 and the parameter names of scripts/gp_mpc_config.yaml:9-17):
//   acc = a u0 + b along the same rotation column as res_dyn, minus g on z;  ddphi = c phi + d dphi + e u1;
//   ddtheta = f theta + h dtheta + l u2;  ddpsi = yc psi + yd dpsi + ye u3 (zero unless given).
// Each of the four RK4 sub-steps is one launch of gpk_predict_mean_jac (three GPs at the first sub-step, then only the
// roll and pitch GPs: the thrust GP's input u0 is constant along the step; the GP sums dominate) and one launch of
// stage_step_kernel: sixteen threads per pair, thread c carrying column c of d k_s / d(x, u).
// ------------------------------------------------------------------------------------------------------------
constexpr int ST_NX = 12, ST_NU = 4, ST_NC = ST_NX + ST_NU, ST_GD = 3;
enum { PR_A = 0, PR_B, PR_C, PR_D, PR_E, PR_F, PR_H, PR_L, PR_G, PR_YC, PR_YD, PR_YE, PR_COUNT };

struct StageArgs {
    const double* x; long ldx;      // [q][12]
    const double* u; long ldu;      // [q][4]
    long q;
    double prior[PR_COUNT];
    double dt;
    const double* xs_in;            // [q][12]  sub-state of the current sub-step
    double* xs_out;                 // [q][12]  sub-state of the next one (ping-pong: no intra-warp ordering needed)
    double* Zg;                     // [3][q][3] GP queries of the current sub-step
    const double* mean;             // [3][q]
    const double* jac;              // [3][q][3]
    double* D;                      // [q][12][16] d k_s / d(x, u) of the previous sub-step
    double* Acc;                    // [q][12][16] weighted sum of the d k_s / d(x, u)
    double* ksum;                   // [q][12] weighted sum of the k_s
    double* x_next; double* A; double* B;
};

// S = 0: set up sub-step 1 (xs = x, GP queries).  S = 1..4: consume the GP output at xs_S, accumulate k_S and its
// sensitivities, then prepare sub-step S + 1 or, for S = 4, write x_next, A, B.
template <int S>
__global__ void __launch_bounds__(256) stage_step_kernel(StageArgs a) {
    const long gi = (long)blockIdx.x * 16 + (threadIdx.x >> 4);
    const int c = threadIdx.x & 15;
    if (gi >= a.q) return;
    const double* x = a.x + gi * a.ldx;
    const double* u = a.u + gi * a.ldu;
    const long q = a.q;
    if constexpr (S == 0) {
        if (c < ST_NX) a.xs_out[gi * ST_NX + c] = x[c];
        if (c < 9) {
            // GP g, column j: T (u0, 0, 0); R (phi, dphi, u1); P (theta, dtheta, u2)
            const int g = c / 3, j = c % 3;
            double v = 0.0;
            if (g == 0) v = j == 0 ? u[0] : 0.0;
            else if (j == 2) v = u[g];
            else v = x[(g == 1 ? 6 : 7) + 3 * j];
            a.Zg[((long)g * q + gi) * ST_GD + j] = v;
        }
        return;
    } else {
        constexpr double cs_cur = S == 1 ? 0.0 : S == 4 ? 1.0 : 0.5;       // xs_S = x + cs_cur dt k_{S-1}
        constexpr double cs_next = S == 3 ? 1.0 : 0.5;
        constexpr double w = (S == 1 || S == 4) ? 1.0 : 2.0;
        const double* xs = a.xs_in + gi * ST_NX;
        const double* P_ = a.prior;
        const double phi = xs[6], th = xs[7];
        double sp, cp, st, ct;
        sincos(phi, &sp, &cp);
        sincos(th, &st, &ct);
        const double mT = a.mean[gi], mR = a.mean[q + gi], mP = a.mean[2 * q + gi];
        const double* jT = a.jac + gi * ST_GD;
        const double* jR = a.jac + (q + gi) * ST_GD;
        const double* jP = a.jac + (2 * q + gi) * ST_GD;
        const double Tt = fma(P_[PR_A], u[0], P_[PR_B]) + mT;             // prior thrust acceleration + GP residual
        const double dTt = P_[PR_A] + jT[0];

        // column c of G = d xs_S / d(x, u) = [I | 0] + cs_cur dt D_{S-1}
        double g[ST_NX];
        double* Dq = a.D + gi * ST_NX * ST_NC;
        double* Aq = a.Acc + gi * ST_NX * ST_NC;
#pragma unroll
        for (int r = 0; r < ST_NX; r++) {
            double v = (r == c) ? 1.0 : 0.0;
            if constexpr (S > 1) v = fma(cs_cur * a.dt, Dq[r * ST_NC + c], v);
            g[r] = v;
        }
        // column c of D_S = F_x G + F_u, with the sparse F_x, F_u of the model above
        double d[ST_NX];
        d[0] = g[1]; d[2] = g[3]; d[4] = g[5]; d[6] = g[9]; d[7] = g[10]; d[8] = g[11];
        d[1] = fma(-Tt * sp * st, g[6], (Tt * cp * ct) * g[7]);
        d[3] = (-Tt * cp) * g[6];
        d[5] = fma(-Tt * sp * ct, g[6], (-Tt * cp * st) * g[7]);
        d[9] = fma(P_[PR_C] + jR[0], g[6], (P_[PR_D] + jR[1]) * g[9]);
        d[10] = fma(P_[PR_F] + jP[0], g[7], (P_[PR_H] + jP[1]) * g[10]);
        d[11] = fma(P_[PR_YC], g[8], P_[PR_YD] * g[11]);
        if (c == ST_NX) { d[1] += dTt * cp * st; d[3] -= dTt * sp; d[5] += dTt * cp * ct; }
        if (c == ST_NX + 1) d[9] += P_[PR_E] + jR[2];
        if (c == ST_NX + 2) d[10] += P_[PR_L] + jP[2];
        if (c == ST_NX + 3) d[11] += P_[PR_YE];
#pragma unroll
        for (int r = 0; r < ST_NX; r++) {
            double acc = w * d[r];
            if constexpr (S > 1) acc += Aq[r * ST_NC + c];
            if constexpr (S < 4) {
                Dq[r * ST_NC + c] = d[r];
                Aq[r * ST_NC + c] = acc;
            } else {
                const double v = a.dt / 6.0 * acc;
                if (c < ST_NX) a.A[(gi * ST_NX + r) * ST_NX + c] = (r == c ? 1.0 : 0.0) + v;
                else a.B[(gi * ST_NX + r) * ST_NU + (c - ST_NX)] = v;
            }
        }
        if (c < ST_NX) {
            // component c of k_S = f_cont(xs_S, u)
            double k;
            switch (c) {
                case 0: k = xs[1]; break;
                case 1: k = Tt * cp * st; break;
                case 2: k = xs[3]; break;
                case 3: k = -Tt * sp; break;
                case 4: k = xs[5]; break;
                case 5: k = Tt * cp * ct - P_[PR_G]; break;
                case 6: k = xs[9]; break;
                case 7: k = xs[10]; break;
                case 8: k = xs[11]; break;
                case 9: k = fma(P_[PR_C], phi, fma(P_[PR_D], xs[9], P_[PR_E] * u[1])) + mR; break;
                case 10: k = fma(P_[PR_F], th, fma(P_[PR_H], xs[10], P_[PR_L] * u[2])) + mP; break;
                default: k = fma(P_[PR_YC], xs[8], fma(P_[PR_YD], xs[11], P_[PR_YE] * u[3])); break;
            }
            double ks = w * k;
            if constexpr (S > 1) ks += a.ksum[gi * ST_NX + c];
            if constexpr (S < 4) {
                a.ksum[gi * ST_NX + c] = ks;
                const double xn = fma(cs_next * a.dt, k, x[c]);
                a.xs_out[gi * ST_NX + c] = xn;
                if (c == 6) a.Zg[((long)1 * q + gi) * ST_GD + 0] = xn;
                if (c == 9) a.Zg[((long)1 * q + gi) * ST_GD + 1] = xn;
                if (c == 7) a.Zg[((long)2 * q + gi) * ST_GD + 0] = xn;
                if (c == 10) a.Zg[((long)2 * q + gi) * ST_GD + 1] = xn;
            } else {
                a.x_next[gi * ST_NX + c] = fma(a.dt / 6.0, ks, x[c]);
            }
        }
    }
}

template <int S>
static int launch_stage_step(const StageArgs& a, cudaStream_t st) {
    stage_step_kernel<S><<<(unsigned)((a.q + 15) / 16), 256, 0, st>>>(a);
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

extern "C" long gpk_stage_rk4_work_bytes(long q_chunk, long n, int batch_unused) {
    (void)batch_unused;
    if (q_chunk <= 0 || n <= 0) return -1;
    const long per_q = (2 * ST_NX + 3 * ST_GD + 3 + 3 * ST_GD + 2 * ST_NX * ST_NC + ST_NX) * (long)sizeof(double);
    long mj = gpk_mean_jac_work_bytes(1, q_chunk, n, ST_GD, 3);
    if (mj < 0) mj = 0;
    return per_q * q_chunk + mj + 256;
}

extern "C" int gpk_stage_rk4(const double* x, long ldx, const double* u, long ldu, long q, const double* prior,
                             int n_prior, double dt, const double* hyp, const void* pack, long stride_pack, long n,
                             double* x_next, double* A, double* B, void* work, long work_bytes, void* stream) {
    GPK_ARG(q >= 0 && ldx >= ST_NX && ldu >= ST_NU, "q / ldx / ldu");
    if (q == 0) return 0;
    GPK_ARG(x && u && prior && hyp && pack && x_next && A && B && work, "null pointer");
    GPK_ARG(n_prior >= PR_G + 1 && n_prior <= PR_COUNT, "prior holds a, b, c, d, e, f, h, l, g [, yaw c, d, e]");
    GPK_ARG(n > 0, "n");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // largest query chunk the scratch holds
    long qc = q;
    while (qc > 1 && gpk_stage_rk4_work_bytes(qc, n, 3) > work_bytes) qc = (qc + 1) / 2;
    GPK_ARG(gpk_stage_rk4_work_bytes(qc, n, 3) <= work_bytes, "scratch too small for one query");
    StageArgs a{};
    for (int i = 0; i < PR_COUNT; i++) a.prior[i] = i < n_prior ? prior[i] : 0.0;
    a.dt = dt; a.ldx = ldx; a.ldu = ldu;
    double* w = static_cast<double*>(work);
    double* xs0 = w;               w += qc * ST_NX;
    double* xs1 = w;               w += qc * ST_NX;
    double* Zg = w;                w += 3 * qc * ST_GD;
    double* mean = w;              w += 3 * qc;
    double* jac = w;               w += 3 * qc * ST_GD;
    double* D = w;                 w += qc * ST_NX * ST_NC;
    double* Acc = w;               w += qc * ST_NX * ST_NC;
    double* ksum = w;              w += qc * ST_NX;
    void* mj_work = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(w) + 255) & ~uintptr_t(255));
    const long mj_bytes = work_bytes - (static_cast<char*>(mj_work) - static_cast<char*>(work));
    for (long q0 = 0; q0 < q; q0 += qc) {
        const long qn = q - q0 < qc ? q - q0 : qc;
        a.x = x + q0 * ldx; a.u = u + q0 * ldu; a.q = qn;
        a.Zg = Zg; a.mean = mean; a.jac = jac; a.D = D; a.Acc = Acc; a.ksum = ksum;
        a.x_next = x_next + q0 * ST_NX; a.A = A + q0 * ST_NX * ST_NX; a.B = B + q0 * ST_NX * ST_NU;
        a.xs_in = xs1; a.xs_out = xs0;
        GPK_TRY(launch_stage_step<0>(a, st));
        for (int s = 1; s <= 4; s++) {
            // the thrust GP's input z[idx_T] = u0 does not move along the RK4 sub-steps: it is evaluated once, the
            // roll and pitch GPs (inputs phi / theta and their rates) at every sub-step
            const int g0 = s == 1 ? 0 : 1;
            const size_t esz = sizeof(double);
            GPK_TRY(gpk_predict_mean_jac(1, Zg + g0 * qn * ST_GD, ST_GD, qn, qn * ST_GD, ST_GD, hyp + g0 * HYP_STRIDE,
                                         static_cast<const char*>(pack) + g0 * stride_pack * esz, stride_pack, n,
                                         mean + g0 * qn, qn, jac + g0 * qn * ST_GD, ST_GD, qn * ST_GD,
                                         mj_bytes > 0 ? mj_work : nullptr, mj_bytes > 0 ? mj_bytes : 0, 3 - g0, st));
            a.xs_in = (s & 1) ? xs0 : xs1;
            a.xs_out = (s & 1) ? xs1 : xs0;
            switch (s) {
                case 1: GPK_TRY(launch_stage_step<1>(a, st)); break;
                case 2: GPK_TRY(launch_stage_step<2>(a, st)); break;
                case 3: GPK_TRY(launch_stage_step<3>(a, st)); break;
                default: GPK_TRY(launch_stage_step<4>(a, st)); break;
            }
        }
    }
    return 0;
}
