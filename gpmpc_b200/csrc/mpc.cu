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
// One CTA per rollout, one thread per entry of P; the stage loop is sequential (T steps of two small products),
// the batch of rollouts fills the machine.
#include "common.cuh"

namespace gpk {

constexpr int TC_ND = 5;        // uncertain directions: dx, dy, dz from the thrust GP, dphi, dtheta (gpmpc.py:68)
constexpr int TC_NGP = 3;       // thrust, roll, pitch GPs (gpmpc.py:59)
constexpr int TC_MAX_NX = 16, TC_MAX_NU = 8;

struct TightenArgs {
    const double* var; long stride_var;    // [3][R*T] GP variances
    const double* Z; long ldz;             // [R*T][ldz] stage inputs z = (x, u); columns idx_phi / idx_theta are read
    int idx_phi, idx_theta;
    const double* hyp;                     // 3 hyper-parameter blocks (likelihood noise)
    const double *A, *B, *K, *Bd;          // row-major nx x nx, nx x nu, nu x nx, nx x 5
    int nx, nu, T;
    double dt, inv_cdf;
    int var_has_noise;                     // 1: var is the predictive variance (latent + noise); 0: latent var_f
    double* state_c;                       // [R][2 nx][T + 1]
    double* input_c;                       // [R][2 nu][T]
};

// a variance that rounding pushed below zero is 0 (numpy's sqrt would give NaN there); NaN inputs stay NaN
__device__ __forceinline__ double clamp0(double v) { return v < 0.0 ? 0.0 : v; }

template <int NX>
__global__ void __launch_bounds__(NX* NX) tighten_kernel(TightenArgs a) {
    constexpr int LD = NX + 1;
    __shared__ double P[NX * LD], T1[NX * LD], KP[TC_MAX_NU * LD], Ks[TC_MAX_NU * LD], Bds[NX * TC_ND], c[TC_ND];
    __shared__ double Ms[NX * LD];
    const int t = threadIdx.x, ra = t / NX, cb = t % NX;
    const int nx = a.nx, nu = a.nu, T = a.T;
    const long r = blockIdx.x;

    // closed-loop matrix M = A + B K, zero-padded to NX x NX; K and Bd zero-padded
    {
        double m = 0.0;
        if (ra < nx && cb < nx) {
            m = a.A[ra * nx + cb];
            for (int j = 0; j < nu; j++) m = fma(a.B[ra * nu + j], a.K[j * nx + cb], m);
        }
        Ms[ra * LD + cb] = m;
        P[ra * LD + cb] = 0.0;
        if (ra < TC_MAX_NU) Ks[ra * LD + cb] = (ra < nu && cb < nx) ? a.K[ra * nx + cb] : 0.0;
        if (cb < TC_ND) Bds[ra * TC_ND + cb] = (ra < nx) ? a.Bd[ra * TC_ND + cb] : 0.0;
    }
    __syncthreads();
    double Ma[NX], Mb[NX], q_bd[TC_ND];
#pragma unroll
    for (int k = 0; k < NX; k++) { Ma[k] = Ms[ra * LD + k]; Mb[k] = Ms[cb * LD + k]; }
#pragma unroll
    for (int m = 0; m < TC_ND; m++) q_bd[m] = Bds[ra * TC_ND + m] * Bds[cb * TC_ND + m];
    const double noise_mult = a.var_has_noise ? 1.0 : 2.0;
    const double dt2 = a.dt * a.dt;
    double* sc = a.state_c + r * 2 * nx * (long)(T + 1);
    double* ic = a.input_c + r * 2 * nu * (long)T;

    for (int i = 0; i < T; i++) {
        // phase 1 (reads P): T1 = P M^T, KP = K P, stage disturbance c_i, state tightening from diag P
        {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < NX; k++) s = fma(P[ra * LD + k], Mb[k], s);
            T1[ra * LD + cb] = s;
        }
        if (ra < nu) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < NX; k++) s = fma(Ks[ra * LD + k], P[k * LD + cb], s);
            KP[ra * LD + cb] = s;
        }
        if (t < TC_ND) {
            const long zi = r * T + i;
            const double phi = a.Z[zi * a.ldz + a.idx_phi], th = a.Z[zi * a.ldz + a.idx_theta];
            const int g = t < 3 ? 0 : t - 2;
            const double v = a.var[g * a.stride_var + zi] + noise_mult * a.hyp[g * HYP_STRIDE + HYP_NOISE];
            const double cp = cos(phi), sp = sin(phi), ct = cos(th), st = sin(th);
            const double f = t == 0 ? cp * st * st : t == 1 ? sp * sp : t == 2 ? (cp * ct) * (cp * ct) : 1.0;
            c[t] = v * f * dt2;
        }
        if (t < nx) {
            const double v = -a.inv_cdf * sqrt(clamp0(P[t * LD + t]));
            sc[(long)t * (T + 1) + i] = v;
            sc[(long)(nx + t) * (T + 1) + i] = v;
        }
        __syncthreads();
        // phase 2 (reads T1, KP, c; writes P): P' = M T1 + Bd diag(c) Bd^T, input tightening from diag(K P K^T)
        {
            double s = 0.0;
#pragma unroll
            for (int m = 0; m < TC_ND; m++) s = fma(q_bd[m], c[m], s);
#pragma unroll
            for (int k = 0; k < NX; k++) s = fma(Ma[k], T1[k * LD + cb], s);
            P[ra * LD + cb] = s;
        }
        if (t < nu) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < NX; k++) s = fma(KP[t * LD + k], Ks[t * LD + k], s);
            const double v = -a.inv_cdf * sqrt(clamp0(s));
            ic[(long)t * T + i] = v;
            ic[(long)(nu + t) * T + i] = v;
        }
        __syncthreads();
    }
    if (t < nx) {
        const double v = -a.inv_cdf * sqrt(clamp0(P[t * LD + t]));
        sc[(long)t * (T + 1) + T] = v;
        sc[(long)(nx + t) * (T + 1) + T] = v;
    }
}

template <int NX>
static int launch_tighten(const TightenArgs& a, long R, cudaStream_t st) {
    tighten_kernel<NX><<<(unsigned)R, NX * NX, 0, st>>>(a);
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
    TightenArgs a{var, stride_var, Z, ldz, idx_phi, idx_theta, hyp, A, B, K, Bd, nx, nu, T, dt, inverse_cdf,
                  var_has_noise, state_constraint, input_constraint};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (nx <= 8) return launch_tighten<8>(a, rollouts, st);
    if (nx <= 12) return launch_tighten<12>(a, rollouts, st);
    return launch_tighten<16>(a, rollouts, st);
}
