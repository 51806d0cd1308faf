// Error-free int8 slicing ("Ozaki scheme") of the fp64 variance contraction V = W K_ZX^T so that it runs on
// the 5th-generation tensor cores (tcgen05.mma kind::i8, int32 accumulators in TMEM) instead of the FP64
// DMMA pipe.  tcgen05 has no f64 kind; B200's int8 pipe is ~120x wider than its DMMA pipe, and
// int8 x int8 -> int32 accumulation is exact, so a product of two S-digit fixed-point operands needs
// S(S+1)/2 exact int8 GEMMs and one fp64 recombination:
//
//   W[i][k]  = 2^ew_i * sum_s a_s[i][k] * 2^(-6-8s)         a_s in [-128, 127]  (balanced base-256 digits, int8;
//   E[q][k]  =          sum_t b_t[q][k] * 2^(-6-8t)          leading digit in [-64, 64]);  E = exp(-r^2/2) in (0, 1]
//   V[i][q]  = sf2 * 2^ew_i * 2^-12 * sum_g 2^(-8g) D_g[i][q],   D_g = sum_{s+t=g} a_s b_t^T   (int32, exact)
//
// truncated at g < S (S = 4/5/6 digits = 30/38/46 bits: |error| ~ 2^(-8S+2) relative to max|W_i| -- S = 5
// reproduces the fp64 variance to ~2e-9 relative at N = 2048, see tests).  Everything but the final
// recombination is exact; the int32 accumulators cannot overflow while N (16384 + (S-2) 16256) <= 2^31
// (gpk_oz_max_npad: N <= 32961 at S = 5; the leading digits are <= 64 and E > 0).
//
// Shared-memory / HBM operand layout ("tile images"): operands are stored in HBM exactly as the tensor core
// wants them in shared memory (K-major, no swizzle: 8-row x 16-byte core matrices), so a pipeline stage is
// filled by ONE 1-D bulk TMA copy (cp.async.bulk, SASS UBLKCP) per operand and no tensor map is needed:
//   byte offset of element (row r, k) inside a R x 64 slice tile:  (r/8)*512 + (k/16)*128 + (r%8)*16 + k%16
//   E image : [gp][qtile][kb][digit t][128 x 64 B]   (qtile = 128 queries = the MMA M side = TMEM lanes)
//   W image : [gp][ib]   [kb][digit s][ 64 x 64 B]   (ib = block of 64 rows of W = the MMA N side; kb <= ib)
// The queries sit on the TMEM lanes so that the sum over the rows of W of V^2 is a per-thread sum in the
// epilogue (no cross-lane reduction), accumulated in registers over all the W row blocks a CTA walks through.
#pragma once
#include "common.cuh"

namespace gpk {

constexpr int OZ_QT = 128;            // queries per CTA tile (MMA M, TMEM lanes)
constexpr int OZ_WB = 64;             // rows of W per CTA tile (MMA N per digit group, TMEM columns)
constexpr int OZ_BK = 64;             // k bytes per pipeline stage (2 MMA k-steps of 32)
constexpr int OZ_A_SLICE = OZ_QT * OZ_BK;   // 8 KB  (E digit tile)
constexpr int OZ_B_SLICE = OZ_WB * OZ_BK;   // 4 KB  (W digit tile)
constexpr int OZ_MIN_S = 4, OZ_MAX_S = 6;
constexpr int OZ_MAX_PIECES = 8;
constexpr int OZ_SMEM_LIMIT = 227 * 1024;

__host__ __device__ constexpr int oz_stage_bytes(int S) { return S * (OZ_A_SLICE + OZ_B_SLICE); }
__host__ __device__ constexpr int oz_stages(int S) {
    return (OZ_SMEM_LIMIT - 4096) / oz_stage_bytes(S) > 4 ? 4 : (OZ_SMEM_LIMIT - 4096) / oz_stage_bytes(S);
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
__device__ __forceinline__ void tc_commit_pred(uint64_t* bar, uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %1, 0;\n"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(issue) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_pred(uint64_t* bar, uint32_t bytes, uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %2, 0;\n"
        "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(bytes), "r"(issue) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s_pred(uint32_t smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar,
                                                  uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %4, 0;\n"
        "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
        "}\n" ::"r"(smem_dst), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "r"(issue) : "memory");
}
// ---- CTA-pair (cluster of 2) variants: one copy feeds both CTAs' shared memory, one commit frees both ----
__device__ __forceinline__ void tma_bulk_g2s_mcast_pred(uint32_t smem_dst, const void* gmem_src, uint32_t bytes,
                                                        uint64_t* bar, uint16_t cta_mask, uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %5, 0;\n"
        "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;\n"
        "}\n" ::"r"(smem_dst), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "h"(cta_mask), "r"(issue) : "memory");
}
__device__ __forceinline__ void tc_commit_mcast_pred(uint64_t* bar, uint16_t cta_mask, uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %2, 0;\n"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
        "}\n" ::"r"(smem_u32(bar)), "h"(cta_mask), "r"(issue) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
// warp-uniform election of one lane (SASS ELECT): the single-thread tcgen05 / bulk-TMA instructions take their
// operands from uniform registers, so the issuing code must stay warp-uniform and only be PREDICATED on the
// elected lane -- inside a divergent `if (lane == 0)` region ptxas wraps every such instruction in a
// convergence loop (R2UR + BRA.U.ANY) that costs ~150 cycles per instruction (measured, tools/peaks/umma_bench)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.b32 %0, 1, 0, p;\n"
        "}\n" : "=r"(pred));
    return pred != 0;
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
// the same, predicated inside the asm on `issue` (the elected lane): no C++ branch, so the surrounding code stays
// warp-uniform and straight-line (no BSSY/BSYNC per instruction)
__device__ __forceinline__ void umma_i8_pred(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate, uint32_t issue) {
    asm volatile(
        "{\n"
        ".reg .pred p, q;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "setp.ne.b32 q, %5, 0;\n"
        "@q tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(issue)
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
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(OZ_QT >> 4) << 24);
}
// 32 lanes x 8 consecutive 32-bit columns: thread i of the warp receives lane (base lane + i)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr));
}
// 32 lanes x 4 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, int32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr));
}
// 32 lanes x 16 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
// balanced base-256 digits of a signed fixed-point integer (most significant digit first)
//   I = sum_s dig[s] * 256^(S-1-s),  dig[s] in [-128, 127]  (leading digit in [-64, 64] by the scaling)
// ------------------------------------------------------------------------------------------------
template <int S>
__device__ __forceinline__ void signed_digits(long long I, int8_t (&dig)[S]) {
#pragma unroll
    for (int s = S - 1; s > 0; s--) {
        const int low = (int)(I & 255);
        const int dgt = low >= 128 ? low - 256 : low;    // in [-128, 127]
        dig[s] = (int8_t)dgt;
        I = (I - dgt) >> 8;
    }
    dig[0] = (int8_t)I;                                  // |I| <= 64 by construction of the scaling
}

// x in [-1, 1] -> nearest integer of x * 2^(6 + 8 (S-1))  (S <= 6: 46 bits)
template <int S>
__device__ __forceinline__ long long to_fixed(double x) {
    constexpr double SCALE = (double)(1ull << (6 + 8 * (S - 1)));
    return __double2ll_rn(x * SCALE);
}


__device__ __forceinline__ double oz_ll2d(long long v) {
    // exact int64 -> double for |v| < 2^51 without the conversion pipe: bit pattern of 1.5 * 2^52 + v, minus the bias
    return __longlong_as_double(0x4338000000000000LL + v) - 6755399441055744.0;
}
// R = sum_g 256^-g D_g: digit groups are combined three at a time in exact int64 arithmetic (integer pipe),
// so the FP64 pipe only sees one conversion add per triple and the Horner steps between triples
template <int S, int W>
__device__ __forceinline__ double oz_recombine(const int32_t (&v)[S][W], int j) {
    constexpr int NT = (S + 2) / 3;
    double acc = 0.0;
#pragma unroll
    for (int tr = NT - 1; tr >= 0; tr--) {
        const int g0 = 3 * tr;
        const int cnt = (S - g0) < 3 ? (S - g0) : 3;
        long long t = v[g0][j];
#pragma unroll
        for (int u = 1; u < cnt; u++) t = t * 256 + v[g0 + u][j];
        // t = sum_u D_{g0+u} 256^(cnt-1-u); its weight relative to the triple's leading digit is 256^-(cnt-1)
        const double dv = oz_ll2d(t) * (cnt == 3 ? 1.0 / 65536.0 : cnt == 2 ? 1.0 / 256.0 : 1.0);
        acc = (tr == NT - 1) ? dv : fma(acc, 1.0 / 16777216.0, dv);
    }
    return acc;
}


// Balanced base-256 digits of four values in (-1, 1) (S <= 6) without integer division or carries: ONE fused multiply-add
// x * 2^(8S-2) + 1.5 * 2^52 leaves the rounded fixed-point integer in the mantissa; a bias of 128 per digit turns the
// balanced digits into the plain BYTES of that integer (two's complement digit = byte ^ 0x80); byte permutes gather
// digit s of the four values into one 32-bit word (most significant digit first).  Same digits as signed_digits(to_fixed(x)).
template <int S>
__device__ __forceinline__ void oz_pack_digits4(const double (&e)[4], uint32_t (&out)[S]) {
    constexpr double SC = (double)(1ull << (6 + 8 * (S - 1)));
    long long bias = 0;
#pragma unroll
    for (int i = 0; i < S; i++) bias += 128ll << (8 * i);
    bias -= 0x4338000000000000LL;
    uint32_t lo[4], hi[4];
#pragma unroll
    for (int v = 0; v < 4; v++) {
        const long long b = __double_as_longlong(fma(e[v], SC, 6755399441055744.0)) + bias;   // sum_i (d_i + 128) 256^i
        lo[v] = (uint32_t)b;
        hi[v] = (uint32_t)(b >> 32);
    }
#pragma unroll
    for (int i = 0; i < S; i++) {                       // i = 0: least significant digit = byte i of the integer
        const uint32_t sel = (uint32_t)((i & 3) | (((i & 3) + 4) << 4));
        const uint32_t p01 = __byte_perm(i < 4 ? lo[0] : hi[0], i < 4 ? lo[1] : hi[1], sel);
        const uint32_t p23 = __byte_perm(i < 4 ? lo[2] : hi[2], i < 4 ? lo[3] : hi[3], sel);
        out[S - 1 - i] = __byte_perm(p01, p23, 0x5410) ^ 0x80808080u;
    }
}


}  // namespace gpk
