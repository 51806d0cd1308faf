// fp64 "NT" GEMM of the fit path on the tcgen05 int8 tensor pipe (error-free digit slicing, see ozaki.cuh):
//     C[i,j] = beta * C[i,j] + alpha * sum_{k in range(i)} A[i,k] * B[j,k]
// Same contract as the DMMA GEMM of dgemm_nt.cuh (triangular k-ranges a_tri, lower tiles only c_lower); it takes
// over the large GEMMs of the recursive Cholesky / triangular inverse / K^-1 (gpmpc/gp.py:62 and :46 through
// gpytorch -> torch LAPACK in the reference), where the DMMA pipe (37 TFLOP/s) is the bound.
//
//   operand (fp64, rows x K) -> digit image  [row block of 64][k block of 64][digit s][64 x 64 B tile]   (int8)
//                               + power-of-two scales per (row, k-CHUNK of KC = 64 k-blocks = 4096 k)
//     x[i][k] = 2^e(i,chunk) * sum_s dig_s[i][k] 2^(-6-8s)
//   Per-chunk scales keep the fixed-point grid relative to the row's magnitude INSIDE the chunk (rows of a Cholesky
//   factor decay along k), and bound the int32 accumulation length (KC * 64 * S * 2^14 < 2^31).
//   C tile (128 x 128, one CTA pair) = sum over chunks of  2^-12 sA sB sum_g 256^-g D_g,  D_g exact in TMEM;
//   the chunks of one tile are walked by the same CTA pair, which accumulates them into C in place (fp64).
//
// Kernel = the variance kernel's pipeline (bulk-TMA producer warp, one-thread tcgen05.mma issuer, 8 epilogue warps,
// CTA pairs with the A tile multicast into both shared memories) with a storing epilogue.
#include "ozaki.cuh"
#include "gemm_params.cuh"
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

namespace gpk {

// k-blocks (of 64) per scale chunk (runtime tunable for experiments: GPK_OZG_KC); int32 accumulation stays exact while
// kc * 64 * S * 2^14 < 2^31 (kc <= 256 at S = 7)
static int g_ozg_kc = 64;

// ------------------------------------------------------------------------------------------------
// operand slicing, two HBM-bound passes:
//   ozg_scale_kernel   one warp per (row, chunk): 2^e > max |x| over the chunk's valid k-range
//   ozg_digit_kernel   one CTA per (k block, 64-row block): digits of x / 2^e into the 64 x 64 B tiles
// tri: 0 full, 1 lower (k < (tile128 + 1) * 128), 2 upper (k >= tile128 * 128) -- blocks outside are never read
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ozg_valid_kb(int row, int kb_total, int tri, int& kb0, int& kb1) {
    const int t128 = row >> 7;
    kb0 = tri == 2 ? min(kb_total, t128 * 2) : 0;
    kb1 = tri == 1 ? min(kb_total, (t128 + 1) * 2) : kb_total;
}

__global__ void __launch_bounds__(256)
ozg_scale_kernel(const double* __restrict__ src, long ld, int rows, int kb_total, int tri, int OZG_KC, double* __restrict__ scale) {
    const int chunk = blockIdx.x;
    const int row = blockIdx.y * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    int v0, v1;
    ozg_valid_kb(row, kb_total, tri, v0, v1);
    const int kb0 = max(v0, chunk * OZG_KC), kb1 = min(v1, (chunk + 1) * OZG_KC);
    double m = 0.0;
    if (kb0 < kb1) {
        const double2* p = reinterpret_cast<const double2*>(src + (long)row * ld + (long)kb0 * OZ_BK);
        const int n2 = (kb1 - kb0) * (OZ_BK / 2);
        for (int i = lane; i < n2; i += 32) {
            const double2 x = p[i];
            m = fmax(m, fmax(fabs(x.x), fabs(x.y)));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) {
        int e = 0;
        if (m > 0.0) frexp(m, &e);
        scale[(long)chunk * rows + row] = m > 0.0 ? ldexp(1.0, e) : 1.0;
    }
}

template <int S>
__global__ void __launch_bounds__(256)
ozg_digit_kernel(const double* __restrict__ src, long ld, int rows, int kb_total, int tri, int OZG_KC,
                 const double* __restrict__ scale, int8_t* __restrict__ img) {
    const int kb = blockIdx.x, ib = blockIdx.y;
    int v0, v1;
    ozg_valid_kb(ib * OZ_WB, kb_total, tri, v0, v1);
    if (kb < v0 || kb >= v1) return;
    const int kc = threadIdx.x & 3, r = threadIdx.x >> 2;
    const long grow = (long)ib * OZ_WB + r;
    const double inv = 1.0 / scale[(long)(kb / OZG_KC) * rows + grow];        // power of two: exact
    const double2* p = reinterpret_cast<const double2*>(src + grow * ld + (long)kb * OZ_BK + kc * 16);
    int8_t* out = img + ((long)ib * kb_total + kb) * S * OZ_B_SLICE;
    if constexpr (S <= 6) {
        // the fixed-point integer fits the mantissa: digits straight from one FMA per value (oz_pack_digits4)
        uint32_t w4[4][S];
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const double2 x0 = p[2 * v], x1 = p[2 * v + 1];
            const double e[4] = {x0.x * inv, x0.y * inv, x1.x * inv, x1.y * inv};
            oz_pack_digits4<S>(e, w4[v]);
        }
#pragma unroll
        for (int s = 0; s < S; s++)
            *reinterpret_cast<uint4*>(out + (long)s * OZ_B_SLICE + oz_tile_off(r, kc * 16)) =
                make_uint4(w4[0][s], w4[1][s], w4[2][s], w4[3][s]);
    } else {
        int8_t dig[16][S];
#pragma unroll
        for (int v = 0; v < 8; v++) {
            const double2 x = p[v];
            signed_digits<S>(to_fixed<S>(x.x * inv), dig[2 * v]);
            signed_digits<S>(to_fixed<S>(x.y * inv), dig[2 * v + 1]);
        }
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

template <int S>
static int ozg_slice(const double* src, long ld, int rows, int kb_total, int tri, double* scale, int8_t* img, cudaStream_t st) {
    const int OZG_KC = g_ozg_kc;
    const int nchunks = (kb_total + OZG_KC - 1) / OZG_KC;
    ozg_scale_kernel<<<dim3((unsigned)nchunks, (unsigned)(rows / 8)), 256, 0, st>>>(src, ld, rows, kb_total, tri, OZG_KC, scale);
    GPK_LAUNCHED();
    ozg_digit_kernel<S><<<dim3((unsigned)kb_total, (unsigned)(rows / OZ_WB)), 256, 0, st>>>(src, ld, rows, kb_total, tri, OZG_KC, scale, img);
    GPK_LAUNCHED();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// GEMM kernel
// ------------------------------------------------------------------------------------------------
struct OzGemmArgs {
    const int8_t* imgA; const int8_t* imgB;
    const double* scaleA; const double* scaleB;   // [chunk][rows]
    int rowsA, rowsB;
    double* C; long ldc;
    double alpha, beta;
    int kb_total, a_tri, kc;
    const int2* items; int total_items;           // (A tile of 128 rows, B row-block pair = C tile column)
};

__device__ __forceinline__ void ozg_krange(const OzGemmArgs& a, int qt, int& kb0, int& kb1) {
    kb0 = a.a_tri == 2 ? min(a.kb_total, 2 * qt) : 0;
    kb1 = a.a_tri == 1 ? min(a.kb_total, 2 * qt + 2) : a.kb_total;
}

template <int S>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 + 32 * 8, 1) ozg_gemm_kernel(const OzGemmArgs a) {
    const int OZG_KC = a.kc;
    constexpr int EPI_WARPS = 8, PRODUCER_WARP = 8, MMA_WARP = 9;
    constexpr int NST = oz_stages(S);
    constexpr uint32_t STAGE_BYTES = oz_stage_bytes(S);
    constexpr uint32_t A_BYTES = S * OZ_A_SLICE, B_BYTES = S * OZ_B_SLICE;
    // TMEM double buffering as in oz_var_kernel: two buffers at columns 0 and SPARE that share NOV digit groups
    constexpr int NOV = 2 * S - 8;
    constexpr bool DB = NOV <= 2;
    constexpr uint32_t SPARE = 512 - 64 * S;
    extern __shared__ __align__(1024) uint8_t ozg_smem[];
    __shared__ __align__(8) uint64_t full_bar[NST], empty_bar[NST], tmem_full_bar, tmem_empty_bar;
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NST; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 2); }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&tmem_empty_bar, 32 * EPI_WARPS);
        mbar_fence_init();
    }
    if (warp == MMA_WARP) tmem_alloc(&tmem_base_smem, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == PRODUCER_WARP) {
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            const int2 it = a.items[item];
            int kb0, kb1;
            ozg_krange(a, it.x, kb0, kb1);
            // this CTA loads rows [64 rank, 64 rank + 64) of the A tile (one 4 KB tile per digit) for BOTH CTAs
            const int8_t* Ag = a.imgA + (long)(2 * it.x + rank) * a.kb_total * B_BYTES;
            const int8_t* Bg = a.imgB + (long)(2 * it.y + rank) * a.kb_total * B_BYTES;
            for (int kb = kb0; kb < kb1; kb++) {
                mbar_wait(&empty_bar[st], ph ^ 1);
                mbar_expect_tx_pred(&full_bar[st], STAGE_BYTES, leader);
                const uint32_t dst = smem_u32(ozg_smem) + (uint32_t)st * STAGE_BYTES;
#pragma unroll
                for (int t = 0; t < S; t++)
                    tma_bulk_g2s_mcast_pred(dst + t * OZ_A_SLICE + rank * OZ_B_SLICE, Ag + (long)kb * B_BYTES + t * OZ_B_SLICE,
                                            OZ_B_SLICE, &full_bar[st], (uint16_t)3, leader);
                tma_bulk_g2s_pred(dst + A_BYTES, Bg + (long)kb * B_BYTES, B_BYTES, &full_bar[st], leader);
                if (++st == NST) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == MMA_WARP) {
        const uint32_t leader = elect_one() ? 1u : 0u;
        int st = 0; uint32_t ph = 0, tph = 0, par = 0;
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            const int2 it = a.items[item];
            int kb0, kb1;
            ozg_krange(a, it.x, kb0, kb1);
            for (int c0 = kb0; c0 < kb1;) {
                const int c1 = min(kb1, (c0 / OZG_KC + 1) * OZG_KC);
                mbar_wait(&tmem_empty_bar, tph ^ 1);
                tc_fence_after();
                const uint32_t dbase = tmem_base + (DB ? par * SPARE : 0u);
                par ^= 1;
                for (int kb = c0; kb < c1; kb++) {
                    mbar_wait(&full_bar[st], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(ozg_smem) + (uint32_t)st * STAGE_BYTES;
                    const uint32_t b_base = a_base + A_BYTES;
#pragma unroll
                    for (int j = 0; j < OZ_BK / 32; j++) {
#pragma unroll
                        for (int t = 0; t < S; t++) {
                            const uint64_t adesc = umma_smem_desc(a_base + t * OZ_A_SLICE + j * 256, 128, OZ_BK * 8);
                            const uint32_t acc = (kb > c0 || j > 0 || t > 0) ? 1u : 0u;
                            const int ntot = OZ_WB * (S - t);
                            const int nparts = ntot > 256 ? 2 : 1;
                            const int npart = ntot / nparts;
#pragma unroll
                            for (int h = 0; h < nparts; h++) {
                                const uint64_t bdesc = umma_smem_desc(b_base + h * npart * OZ_BK + j * 256, 128, OZ_BK * 8);
                                umma_i8_pred(dbase + (uint32_t)(t * OZ_WB + h * npart), adesc, bdesc, umma_idesc_i8(npart), acc, leader);
                            }
                        }
                    }
                    tc_commit_mcast_pred(&empty_bar[st], (uint16_t)3, leader);
                    if (++st == NST) { st = 0; ph ^= 1; }
                }
                tc_commit_pred(&tmem_full_bar, leader);
                tph ^= 1;
                c0 = c1;
            }
        }
    } else {
        // ===== epilogue: lane quarter = warp % 4 (TMEM lane restriction), column half = warp / 4 =====
        const int quarter = warp & 3;
        const int half = warp >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t tbase = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 32);
        uint32_t tph = 0, par = 0;
        // 8 recombined columns -> C (row-major): 64 contiguous bytes per thread
        // the old C values of a chunk (read-modify-write) are fetched BEFORE the wait for the accumulators, so that the
        // HBM / L2 latency of that read is not part of the epilogue the tensor pipe waits for (6 digits: one TMEM buffer)
        double cold[4][8];
        auto emit8 = [&](const double (&r8)[8], const double (&c8)[8], double* cp, const double* sb, double sa, double bfac, bool readc) {
            double o[8];
            if (readc) {
#pragma unroll
                for (int v = 0; v < 8; v++) o[v] = c8[v] * bfac;
            } else {
#pragma unroll
                for (int v = 0; v < 8; v++) o[v] = 0.0;
            }
#pragma unroll
            for (int v = 0; v < 8; v++) o[v] = fma(r8[v] * sa, __ldg(sb + v), o[v]);
#pragma unroll
            for (int v = 0; v < 4; v++) *reinterpret_cast<double2*>(cp + 2 * v) = make_double2(o[2 * v], o[2 * v + 1]);
        };
        auto tile_db = [&](auto par_tag, double* cp, const double* sb, double sa, double bfac, bool readc) {
            constexpr int PAR = decltype(par_tag)::value;
            constexpr int NREST = S - NOV;
            constexpr int OV0 = PAR == 0 ? S - NOV : 0;
            constexpr int RS0 = PAR == 0 ? 0 : NOV;
            const uint32_t tb = tbase + (PAR ? SPARE : 0u);
            int32_t vo[NOV > 0 ? NOV : 1][4][8];
            int32_t v[2][NREST > 0 ? NREST : 1][8];
#pragma unroll
            for (int g = 0; g < NOV; g++)
#pragma unroll
                for (int c = 0; c < 4; c++) tmem_ld8(tb + (uint32_t)((OV0 + g) * OZ_WB + c * 8), vo[g][c]);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&tmem_empty_bar);
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
                double r8[8];
#pragma unroll
                for (int j = 0; j < 8; j++) r8[j] = oz_recombine<S>(vv, j);
                emit8(r8, cold[c], cp + c * 8, sb + c * 8, sa, bfac, readc);
            }
        };
        for (int item = pair_id; item < a.total_items; item += n_pairs) {
            const int2 it = a.items[item];
            int kb0, kb1;
            ozg_krange(a, it.x, kb0, kb1);
            const long grow = (long)it.x * OZ_QT + row;
            const long gcol = (long)(2 * it.y + rank) * OZ_WB + half * 32;
            double* cp = a.C + grow * a.ldc + gcol;
            bool first = true;
            for (int c0 = kb0; c0 < kb1;) {
                const int chunk = c0 / OZG_KC;
                const int c1 = min(kb1, (chunk + 1) * OZG_KC);
                const double sa = a.scaleA[(long)chunk * a.rowsA + grow] * a.alpha * (1.0 / 4096.0);
                const double* sb = a.scaleB + (long)chunk * a.rowsB + gcol;
                const double bfac = first ? a.beta : 1.0;
                const bool readc = !(first && a.beta == 0.0);
                if (readc) {
#pragma unroll
                    for (int c = 0; c < 4; c++)
#pragma unroll
                        for (int v = 0; v < 4; v++) {
                            const double2 x = *reinterpret_cast<const double2*>(cp + c * 8 + 2 * v);
                            cold[c][2 * v] = x.x;
                            cold[c][2 * v + 1] = x.y;
                        }
                }
                mbar_wait(&tmem_full_bar, tph);
                tph ^= 1;
                tc_fence_after();
                if constexpr (DB) {
                    if (par) tile_db(std::integral_constant<int, 1>{}, cp, sb, sa, bfac, readc);
                    else tile_db(std::integral_constant<int, 0>{}, cp, sb, sa, bfac, readc);
                    par ^= 1;
                } else {
                    int32_t v[2][S][8];
#pragma unroll
                    for (int g = 0; g < S; g++) tmem_ld8(tbase + (uint32_t)(g * OZ_WB), v[0][g]);
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        tmem_ld_wait();
                        if (c < 3) {
#pragma unroll
                            for (int g = 0; g < S; g++) tmem_ld8(tbase + (uint32_t)(g * OZ_WB + (c + 1) * 8), v[(c + 1) & 1][g]);
                        } else {
                            tc_fence_before();
                            mbar_arrive(&tmem_empty_bar);
                        }
                        double r8[8];
#pragma unroll
                        for (int j = 0; j < 8; j++) r8[j] = oz_recombine<S>(v[c & 1], j);
                        emit8(r8, cold[c], cp + c * 8, sb + c * 8, sa, bfac, readc);
                    }
                }
                first = false;
                c0 = c1;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct OzgWorkspace { void* ptr = nullptr; long bytes = 0; int digits = 0; };
static OzgWorkspace g_ozg_ws[64];
static long g_ozg_min_dim = 1024;          // smallest M, N for which the int8 path is used
static double g_ozg_min_work = 2e9;          // useful multiply-adds.  5e8 (every 1024-sized GEMM with one triangular operand or a
                                             // lower-only result) is 4 % faster at N = 2048-4096 (GPK_OZG_MIN_WORK=5e8: 1.27 -> 1.22 ms)
                                             // but puts 6-digit GEMMs into the N = 2048 factorisation: L then differs from the fp64
                                             // oracle by 1.7e-11 instead of < 1e-11 (tests/test_gpu_kernels.py) -- not the default

static int g_ozg_gm = 8, g_ozg_gn = 9, g_ozg_pairs = 0;
static void ozg_env() {
    static bool done = false;
    if (done) return;
    done = true;
    if (getenv("GPK_OZG_KC")) g_ozg_kc = atoi(getenv("GPK_OZG_KC"));
    if (getenv("GPK_OZG_GM")) g_ozg_gm = atoi(getenv("GPK_OZG_GM"));
    if (getenv("GPK_OZG_GN")) g_ozg_gn = atoi(getenv("GPK_OZG_GN"));
    if (getenv("GPK_OZG_PAIRS")) g_ozg_pairs = atoi(getenv("GPK_OZG_PAIRS"));
}

static int ozg_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev < 0 || dev >= 64 ? 0 : dev;
}

static int ozg_sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

// Item list (A tile, C tile column) of a GEMM shape in the order the CTA pairs walk it (static round robin):
// super-tiles of GM x GN tiles, so the ~74 items in flight share GM A tiles and GN B block pairs (L2 reuse); the
// super-tile rows with the longest k-ranges come first.  Cached on the device per shape.
using OzgItemKey = std::tuple<int, int, int, int, int>;
static std::map<OzgItemKey, std::pair<int2*, int>> g_ozg_item_cache;
static std::mutex g_ozg_item_mutex;

static const int2* ozg_items(int tiles_m, int tiles_n, int a_tri, int c_lower, int* total, cudaStream_t st) {
    std::lock_guard<std::mutex> lock(g_ozg_item_mutex);
    const OzgItemKey key{ozg_device(), tiles_m, tiles_n, a_tri, c_lower};
    auto f = g_ozg_item_cache.find(key);
    if (f != g_ozg_item_cache.end()) { *total = f->second.second; return f->second.first; }
    const int GM = g_ozg_gm, GN = g_ozg_gn;
    std::vector<int2> items;
    items.reserve((size_t)tiles_m * tiles_n);
    const int srows = (tiles_m + GM - 1) / GM;
    for (int sr0 = 0; sr0 < srows; sr0++) {
        const int sr = a_tri == 1 ? srows - 1 - sr0 : sr0;            // heavy rows first
        for (int sc = 0; sc * GN < tiles_n; sc++)
            for (int dm = 0; dm < GM; dm++)
                for (int dn = 0; dn < GN; dn++) {
                    const int tm = sr * GM + dm, tn = sc * GN + dn;
                    if (tm >= tiles_m || tn >= tiles_n) continue;
                    if (c_lower && tn > tm) continue;
                    items.push_back(make_int2(tm, tn));
                }
    }
    // first sight of a shape on a device: one cudaMalloc + copy + synchronisation of `st` (`items` dies at return);
    // a fit of one size sees ~10 shapes, all during its first iteration.  Freed by gpk_shutdown.
    int2* dev = nullptr;
    if (cudaMalloc(&dev, items.size() * sizeof(int2)) != cudaSuccess) return nullptr;
    if (cudaMemcpyAsync(dev, items.data(), items.size() * sizeof(int2), cudaMemcpyHostToDevice, st) != cudaSuccess) {
        cudaFree(dev);
        return nullptr;
    }
    cudaStreamSynchronize(st);
    g_ozg_item_cache[key] = {dev, (int)items.size()};
    *total = (int)items.size();
    return dev;
}

static long ozg_img_bytes(long rows, long K, int S) { return rows * K * S; }

// A may be B itself (SYRK, K^-1 = U U^T) or a window of B's rows (row-split K^-1): one image then serves both operands.
// Returns the first row of A inside B, or -1 when the operands are unrelated.
static long ozg_a_row_in_b(const GemmParams& p) {
    if (p.lda != p.ldb) return -1;
    const uintptr_t a = (uintptr_t)p.A, b = (uintptr_t)p.B;
    if (a < b || (a - b) % ((uintptr_t)p.ldb * 8)) return -1;
    const long r = (long)((a - b) / ((uintptr_t)p.ldb * 8));
    if (r % OZ_QT || r + p.M > p.N) return -1;
    return r;
}
static long ozg_scale_bytes(long rows, long K) { return ((K / OZ_BK + g_ozg_kc - 1) / g_ozg_kc) * rows * 8; }

template <int S>
static int ozg_run(const GemmParams& p, uint8_t* work, cudaStream_t st) {
    ozg_env();
    const int kb_total = p.K / OZ_BK;
    const long arow = ozg_a_row_in_b(p);
    int8_t *imgA, *imgB;
    double *scaleA, *scaleB;
    int rowsA = p.M;
    if (arow >= 0) {
        // shared image; the triangular k-range may only be skipped while slicing when no tile reads outside it
        const int tri = (arow == 0 && p.M == p.N && p.a_tri == 2 && p.c_lower) ? 2 : 0;
        imgB = reinterpret_cast<int8_t*>(work);
        const long off = (ozg_img_bytes(p.N, p.K, S) + 255) / 256 * 256;
        scaleB = reinterpret_cast<double*>(work + off);
        GPK_TRY(ozg_slice<S>(p.B, p.ldb, p.N, kb_total, tri, scaleB, imgB, st));
        imgA = imgB + (arow / OZ_WB) * (long)kb_total * S * OZ_B_SLICE;
        scaleA = scaleB + arow;
        rowsA = p.N;
    } else {
        imgA = reinterpret_cast<int8_t*>(work);
        long off = ozg_img_bytes(p.M, p.K, S);
        imgB = reinterpret_cast<int8_t*>(work + off);
        off = (off + ozg_img_bytes(p.N, p.K, S) + 255) / 256 * 256;
        scaleA = reinterpret_cast<double*>(work + off);
        scaleB = reinterpret_cast<double*>(work + off + ozg_scale_bytes(p.M, p.K));
        GPK_TRY(ozg_slice<S>(p.A, p.lda, p.M, kb_total, p.a_tri, scaleA, imgA, st));
        GPK_TRY(ozg_slice<S>(p.B, p.ldb, p.N, kb_total, 0, scaleB, imgB, st));
    }
    OzGemmArgs a{};
    a.imgA = imgA; a.imgB = imgB; a.scaleA = scaleA; a.scaleB = scaleB;
    a.rowsA = rowsA; a.rowsB = p.N;
    a.C = p.C; a.ldc = p.ldc; a.alpha = p.alpha; a.beta = p.beta;
    a.kb_total = kb_total; a.a_tri = p.a_tri; a.kc = g_ozg_kc;
    a.items = ozg_items(p.M / OZ_QT, p.N / (2 * OZ_WB), p.a_tri, p.c_lower, &a.total_items, st);
    GPK_ARG(a.items != nullptr, "could not build the item list");
    constexpr int smem = oz_stages(S) * oz_stage_bytes(S);
    static bool attr_dev[64] = {};      // cudaFuncSetAttribute is per device
    bool& attr_done = attr_dev[current_device()];
    if (!attr_done) {
        GPK_CUDA(cudaFuncSetAttribute(ozg_gemm_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    int hw_pairs = ozg_sm_count() / 2;
    if (g_ozg_pairs > 0 && g_ozg_pairs < hw_pairs) hw_pairs = g_ozg_pairs;
    const int grid = 2 * (a.total_items < hw_pairs ? a.total_items : hw_pairs);
    ozg_gemm_kernel<S><<<grid, 64 + 32 * 8, smem, st>>>(a);
    GPK_LAUNCHED();
    return 0;
}

static long ozg_need_bytes(const GemmParams& p, int S) {
    long b = ozg_img_bytes(p.N, p.K, S) + ozg_scale_bytes(p.N, p.K) + 512;
    if (ozg_a_row_in_b(p) < 0) b += ozg_img_bytes(p.M, p.K, S) + ozg_scale_bytes(p.M, p.K);
    return b;
}

static bool ozg_shape_ok(const GemmParams& p) {
    if (p.M % OZ_QT || p.N % (2 * OZ_WB) || p.K % OZ_BK) return false;
    if (p.a_tri != 0 && p.M > p.K) return false;
    if (p.C == p.A || p.C == p.B) return false;
    if (p.lda % 2 || p.ldb % 2 || p.ldc % 2) return false;
    if (((uintptr_t)p.A | (uintptr_t)p.B | (uintptr_t)p.C) % 16) return false;
    return true;
}

static int ozg_dispatch(const GemmParams& p, int S, uint8_t* work, cudaStream_t st) {
    switch (S) {
        case 5: return ozg_run<5>(p, work, st);
        case 6: return ozg_run<6>(p, work, st);
        case 7: return ozg_run<7>(p, work, st);
    }
    return fail(-1, "digits must be 5, 6 or 7");
}

// called by launch_dgemm_nt (linalg.cu): 1 = handled on the int8 pipe, 0 = not eligible (use the DMMA kernel), < 0 error
// Two enqueuing contexts may be in flight at once (gpk_potrf_trtri: the factorisation on the caller's stream, the
// inverse factor on a side stream): each gets its own half of the registered scratch.  0 = whole scratch.
static thread_local int g_ozg_part = 0;
void ozg_set_partition(int part) { g_ozg_part = part; }

int ozg_try(const GemmParams& p, int batch, cudaStream_t st) {
    OzgWorkspace ws = g_ozg_ws[ozg_device()];
    if (ws.ptr != nullptr && g_ozg_part != 0) {
        const long half = ws.bytes / 2 / 256 * 256;
        ws.ptr = static_cast<uint8_t*>(ws.ptr) + (g_ozg_part == 2 ? half : 0);
        ws.bytes = half;
    }
    if (ws.ptr == nullptr || ws.digits == 0 || batch != 1) return 0;
    if (p.M < g_ozg_min_dim || p.N < g_ozg_min_dim || p.K < 1024) return 0;
    // below the threshold the slicing passes and the 74-pair tile quantisation eat the gain (tools/ozgemm_check.py:
    // 1024^3 triangular 0.049 vs 0.063 ms DMMA; 2048 x 2048 x 4096 full 0.33 vs 1.22 ms; GPK_OZG_MIN_WORK overrides)
    const double work = (double)p.M * p.N * p.K * (p.a_tri ? 0.5 : 1.0) * (p.c_lower ? 0.5 : 1.0);
    static const double env_min_work = getenv("GPK_OZG_MIN_WORK") ? atof(getenv("GPK_OZG_MIN_WORK")) : -1.0;   // tuning
    if (work < (env_min_work > 0.0 ? env_min_work : g_ozg_min_work)) return 0;
    if (!ozg_shape_ok(p) || ozg_need_bytes(p, ws.digits) > ws.bytes) return 0;
    const int r = ozg_dispatch(p, ws.digits, static_cast<uint8_t*>(ws.ptr), st);
    return r == 0 ? 1 : r;
}

// gpk_shutdown: drop the current device's cached item lists and forget its registered workspace
void ozg_items_release() {
    std::lock_guard<std::mutex> lock(g_ozg_item_mutex);
    const int dev = ozg_device();
    for (auto it = g_ozg_item_cache.begin(); it != g_ozg_item_cache.end();) {
        if (std::get<0>(it->first) == dev) {
            cudaFree(it->second.first);
            it = g_ozg_item_cache.erase(it);
        } else {
            ++it;
        }
    }
    g_ozg_ws[dev] = OzgWorkspace{};
}

}  // namespace gpk

using namespace gpk;

extern "C" {

// scratch for the int8 fit GEMMs at padded size npad: one shared image (A == B, K^-1 = U U^T) or two half-size ones
long gpk_oz_gemm_work_bytes(long npad, int digits) {
    return ozg_img_bytes(npad, npad, digits) + 2 * ozg_scale_bytes(npad, npad) + 1024;
}

// Registers (ptr != NULL, digits in {5,6,7}) or clears (ptr == NULL or digits == 0) the per-device scratch that switches
// the large GEMMs of gpk_potrf / gpk_trtri / gpk_kinv / gpk_dgemm_nt to the tcgen05 int8 path.  The caller owns the memory.
int gpk_oz_gemm_set_workspace(void* ptr, long bytes, int digits, long min_dim) {
    GPK_ARG(ptr == nullptr || digits == 0 || (digits >= 5 && digits <= 7), "digits must be 5, 6 or 7");
    OzgWorkspace& ws = g_ozg_ws[ozg_device()];
    ws.ptr = digits ? ptr : nullptr;
    ws.bytes = ws.ptr ? bytes : 0;
    ws.digits = ws.ptr ? digits : 0;
    if (min_dim > 0) g_ozg_min_dim = min_dim < 256 ? 256 : min_dim;
    if (min_dim < 0) g_ozg_min_work = (double)(-min_dim) * 1e9;      // tuning hook: negative = work threshold in 1e9 MACs
    return 0;
}

// explicit entry (tests, benchmarks): same contract as gpk_dgemm_nt, always on the int8 path
int gpk_oz_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc, long M, long N, long K,
                    double alpha, double beta, int a_tri, int c_lower, int digits, void* work, long work_bytes,
                    void* stream) {
    GemmParams p{};
    p.A = A; p.B = B; p.C = C; p.lda = lda; p.ldb = ldb; p.ldc = ldc;
    p.M = (int)M; p.N = (int)N; p.K = (int)K; p.alpha = alpha; p.beta = beta; p.a_tri = a_tri; p.c_lower = c_lower;
    GPK_ARG(A && B && C && work, "null pointer");
    GPK_ARG(M > 0 && N > 0 && K > 0, "empty GEMM");
    GPK_ARG(ozg_shape_ok(p), "M, N multiples of 128, K of 64; a_tri needs M <= K; C must not alias A or B; 16-byte alignment");
    GPK_ARG(digits >= 5 && digits <= 7, "digits must be 5, 6 or 7");
    GPK_ARG(work_bytes >= ozg_need_bytes(p, digits), "scratch too small");
    return ozg_dispatch(p, digits, static_cast<uint8_t*>(work), (cudaStream_t)stream);
}

long gpk_oz_dgemm_nt_work_bytes(long M, long N, long K, int same, int digits) {
    long b = ozg_img_bytes(N, K, digits) + ozg_scale_bytes(N, K) + 512;     // same: A is B or a row window of B
    if (!same) b += ozg_img_bytes(M, K, digits) + ozg_scale_bytes(M, K);
    return b;
}

}  // extern "C"
