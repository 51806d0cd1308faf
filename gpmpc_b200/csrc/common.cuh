// Shared helpers for the gp-mpc GP hot path kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "gpmpc_b200 kernels are written for sm_100a (B200) only"
#endif

namespace gpk {

extern thread_local char g_err[512];
extern std::atomic<long> g_launches;

int fail(int code, const char* fmt, ...);

// Layout of the per-GP hyper-parameter block in device memory (doubles).
constexpr int HYP_MAXD = 32;
constexpr int HYP_CENTER = 0;             // center[32]
constexpr int HYP_INV_ELL = HYP_MAXD;     // 1/ell[32]
constexpr int HYP_SF2 = 2 * HYP_MAXD;     // signal variance
constexpr int HYP_NOISE = 2 * HYP_MAXD + 1;
constexpr int HYP_STRIDE = 2 * HYP_MAXD + 8;

constexpr int TILE = 128;                 // every padded matrix dimension is a multiple of this

#define GPK_ARG(cond, msg)                                                      \
    do { if (!(cond)) return gpk::fail(-1, "%s: bad argument: %s", __func__, msg); } while (0)
#define GPK_CUDA(x)                                                             \
    do { cudaError_t e_ = (x); if (e_ != cudaSuccess)                           \
        return gpk::fail(-(1000 + (int)e_), "%s: %s -> %s", __func__, #x, cudaGetErrorString(e_)); } while (0)
#define GPK_LAUNCHED()                                                          \
    do { gpk::g_launches.fetch_add(1, std::memory_order_relaxed);               \
         cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess)            \
        return gpk::fail(-(1000 + (int)e_), "%s: kernel launch -> %s", __func__, cudaGetErrorString(e_)); } while (0)
#define GPK_TRY(x) do { int r_ = (x); if (r_ != 0) return r_; } while (0)

// index of the current device, clamped to the size of the per-device static tables
inline int current_device() {
    int d = 0;
    cudaGetDevice(&d);
    return d < 0 || d >= 64 ? 0 : d;
}

// ---- programmatic dependent launch (PDL) ----------------------------------------------------
// The fit is a chain of ~170 (N = 2048) to ~1800 (N = 16384) small dependent kernels; between two of them the GPU
// idles for ~2-3 us (grid completion -> launch of the next grid).  Kernels launched with launch_pdl may be SCHEDULED
// as soon as every CTA of the previous kernel in the stream has called pdl_launch_dependents() (or exited); they must
// call pdl_wait() before touching anything the previous kernel wrote -- it returns when that kernel has completed
// and its writes are visible.  Without the launch attribute both calls are no-ops.  GPK_PDL=0 turns the attribute off.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- small device helpers -----------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// cp.async (LDGSTS), 16 bytes, L2-only
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// ---- mbarrier + 1-D bulk TMA copy (cp.async.bulk -> SASS UBLKCP) -----------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra.uni WAIT_DONE;\n"
        "bra.uni WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity));
}
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// exp(a) for a <~ 0 to ~2^-50 relative error, branch-free, 9 FP64 operations: a log2(e) 1024 = m + f with
// m = 1024 n + j, so exp(a) = 2^n * T[j] * exp(r), T[j] = 2^(j/1024) from a shared-memory table, r = a - m ln2/1024
// (Cody-Waite, |r| <= 3.4e-4: the cubic's truncation error r^4/24 is 5.6e-16) evaluated as T + (T r)(1 + r/2 + r^2/6);
// 2^n goes straight into the exponent field.  Results below 2^-1021 are flushed to 0 by an integer compare (where
// the library exp is subnormal or 0); arguments a few ulps above 0 (rounding of -r^2/2) are fine.
constexpr int EXP_TAB = 1024;
__device__ __forceinline__ double exp_neg_tab(double a0, const double* __restrict__ tab) {
    const double a = fmax(a0, -745.0);                   // keeps m inside int32 for any input, NaN-free
    const double MAGIC = 6755399441055744.0;             // 1.5 * 2^52
    const double tm = fma(a, 1477.3197218702985, MAGIC); // 1024 / ln 2
    const double md = tm - MAGIC;
    double r = fma(md, -0.0006769015433292225, a);       // ln2/1024 hi (32 significant bits: exact products)
    r = fma(md, -1.8634911418658083e-13, r);             // ln2/1024 lo
    const int m = __double2loint(tm);                    // integer part of tm sits in the low mantissa word
    const double T = tab[m & (EXP_TAB - 1)];
    double q = fma(r, 1.66666666666666666667e-01, 0.5);
    q = fma(q, r, 1.0);
    const double p = fma(T * r, q, T);
    const int n = m >> 10;
    const double e = __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
    return n < -1021 ? 0.0 : e;
}
// the table in device memory (built by gpk_init, or lazily on first use; defined in ozaki.cu)
const double* exp_table(cudaStream_t st);
// per-device library state released by gpk_shutdown (ozaki.cu, linalg.cu, ozgemm.cu)
void exp_table_release();
int fit_side_prepare(int events);
void fit_side_release();
void ozg_items_release();
int oz_side_prepare();
void oz_side_release();

}  // namespace gpk
