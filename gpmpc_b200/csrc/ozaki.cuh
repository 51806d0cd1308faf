// Error-free int8 slicing ("Ozaki scheme") of the fp64 variance contraction V = W K_ZX^T so that it runs on
// the 5th-generation tensor cores (tcgen05.mma kind::i8, int32 accumulators in TMEM) instead of the FP64
// DMMA pipe.  tcgen05 has no f64 kind; B200's int8 pipe is ~120x wider than its DMMA pipe, and
// int8 x int8 -> int32 accumulation is exact, so a product of two S-digit fixed-point operands needs
// S(S+1)/2 exact int8 GEMMs and one fp64 recombination:
//
//   W[i][k]  = 2^ew_i * sum_s a_s[i][k] * 2^(-6-7s)         a_s in [-64, 64]  (signed digits, int8)
//   E[q][k]  =          sum_t b_t[q][k] * 2^(-6-7t)         E = exp(-r^2/2) in (0, 1]
//   V[i][q]  = sf2 * 2^ew_i * 2^-12 * sum_g 2^(-7g) D_g[i][q],   D_g = sum_{s+t=g} a_s b_t^T   (int32, exact)
//
// truncated at g < S (S = 5/6/7: |error| ~ 2^(-7S) relative to max|W_i| -- S = 6 reproduces the fp64
// variance to ~2e-10 relative at N = 2048, see tests).  Everything but the final Horner sum is exact.
//
// Shared-memory / HBM operand layout ("tile images"): operands are stored in HBM exactly as the tensor core
// wants them in shared memory (K-major, no swizzle: 8-row x 16-byte core matrices), so a pipeline stage is
// filled by ONE 1-D bulk TMA copy (cp.async.bulk, SASS UBLKCP) per operand and no tensor map is needed:
//   byte offset of element (row r, k) inside a R x 64 slice tile:  (r/8)*512 + (k/16)*128 + (r%8)*16 + k%16
//   W image : [gp][tile_m][kb][slice s][128 x 64 B]   (tile_m = 128 rows, kb = 64-wide k block)
//   E image : [gp][qtile][kb][slice t][ 64 x 64 B]    (qtile = 64 queries)
#pragma once
#include "common.cuh"

namespace gpk {

constexpr int OZ_BM = 128;            // rows of W per CTA tile (TMEM lanes)
constexpr int OZ_BN = 64;             // queries per CTA tile (TMEM columns per digit group)
constexpr int OZ_BK = 64;             // k bytes per pipeline stage (2 MMA k-steps of 32)
constexpr int OZ_A_SLICE = OZ_BM * OZ_BK;   // 8 KB
constexpr int OZ_B_SLICE = OZ_BN * OZ_BK;   // 4 KB
constexpr int OZ_MAX_S = 7;
constexpr int OZ_SMEM_LIMIT = 227 * 1024;

__host__ __device__ constexpr int oz_stage_bytes(int S) { return S * (OZ_A_SLICE + OZ_B_SLICE); }
__host__ __device__ constexpr int oz_stages(int S) {
    return (OZ_SMEM_LIMIT - 2048) / oz_stage_bytes(S) > 4 ? 4 : (OZ_SMEM_LIMIT - 2048) / oz_stage_bytes(S);
}
// byte offset of (row r, k byte) inside one R x 64 slice tile
__host__ __device__ constexpr int oz_tile_off(int r, int k) {
    return (r >> 3) * (OZ_BK * 8) + (k >> 4) * 128 + (r & 7) * 16 + (k & 15);
}

// ---- tcgen05 / TMEM PTX wrappers ---------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_dst)), "r"(cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(addr), "r"(cols));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
// mbarrier arrive when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, int8 x int8 -> int32, one elected thread issues
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle; lbo = bytes between the two 16-byte k chunks of one
// MMA, sbo = bytes between 8-row groups; version 1 (Blackwell)
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor: D = s32, A = B = signed int8, both K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t umma_idesc_i8(int n) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(OZ_BM >> 4) << 24);
}
// 32 lanes x 16 consecutive 32-bit columns: thread i of the warp receives lane (base lane + i)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

}  // namespace gpk
