// tcgen05 / TMEM primitives for the tensor-core kernels of libmgb200.so (sm_100a only).
// Inline PTX; bit layouts of the shared-memory matrix descriptor and the instruction
// descriptor follow the PTX ISA "tcgen05 matrix descriptors" (K-major, no swizzle).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "mgb_device.cuh"

// ---------------------------------------------------------------------------
// TF32 split: x = hi + lo with hi, lo exactly representable in TF32 (10-bit mantissa),
// so  a.b ~= a_hi.b_hi + a_lo.b_hi + a_hi.b_lo  to ~2^-21 relative ("3xTF32").
// ---------------------------------------------------------------------------
__device__ __forceinline__ float mgb_tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// Device-side split WITHOUT cvt: the tensor core reads only the upper 19 bits of a tf32 operand (the low 13 mantissa bits are
// ignored), so hi = x with those bits cleared (one LOP3) and lo = x - hi (exact; the hardware drops its bits beyond tf32,
// 2^-10 |lo| <= 2^-21 |x|) give the same 3xTF32 accuracy as two cvt.rna - and cvt runs at a quarter of the FMA rate: it
// was ~100 conversions per candidate in the A fills of the step kernel and the producers' limit in the covariance kernel.
__device__ __forceinline__ void mgb_tf32_split(float x, float& hi, float& lo) {
    hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    lo = x - hi;
}

// ---------------------------------------------------------------------------
// mbarrier helpers beyond mgb_device.cuh
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mgb_mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mgb_smem_u32(bar)) : "memory");
}
// arrival without release semantics: does not order (or wait for) this thread's earlier global stores;
// TMEM reads are ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync instead
__device__ __forceinline__ void mgb_mbar_arrive_relaxed(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(mgb_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mgb_mbar_expect_tx(unsigned long long* bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mgb_smem_u32(bar)), "r"(bytes) : "memory");
}
// bulk copy global -> shared that only adds its bytes to an already armed barrier
__device__ __forceinline__ void mgb_bulk_g2s(void* dst_smem, const void* src_gmem, unsigned int bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(mgb_smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(mgb_smem_u32(bar))
                 : "memory");
}

// ---------------------------------------------------------------------------
// TMEM allocation (one warp, all 32 lanes), fences
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mgb_tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(mgb_smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void mgb_tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void mgb_tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void mgb_tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// one lane of a converged warp (elect.sync): the issuer of the warp's tcgen05.mma / tcgen05.commit
__device__ __forceinline__ bool mgb_elect_one() {
    unsigned int pred;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
}

// ---------------------------------------------------------------------------
// descriptors
// ---------------------------------------------------------------------------
// K-major operand without swizzle: 8 x 16-byte core matrices (8 rows of 4 tf32), each 128
// contiguous bytes.  lbo = byte distance between core matrices adjacent in K,
// sbo = byte distance between core matrices adjacent in M / N.
__device__ __forceinline__ uint64_t mgb_umma_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
    d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
    return d;                // base offset 0, lbo mode 0, layout type 0 = no swizzle
}
// kind::tf32, fp32 accumulate, both operands K-major, dense
__host__ __device__ constexpr uint32_t mgb_umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] . B[smem]^T, one 128 x N x 8 tf32 MMA; issued by ONE thread
__device__ __forceinline__ void mgb_umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all MMAs issued so far by this thread complete -> one arrival on bar (implies fence::before_thread_sync)
__device__ __forceinline__ void mgb_umma_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mgb_smem_u32(bar)) : "memory");
}

// ---------------------------------------------------------------------------
// TMEM -> registers: lane = 32 * (warp % 4) + laneid, 32 / 16 / 8 consecutive columns
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mgb_tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void mgb_tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void mgb_tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// canonical (K-major, no swizzle) byte offset of element (row, k) in an operand block of
// `rows` rows: core matrices K-chunk-major  ->  LBO = rows * 16, SBO = 128
__host__ __device__ __forceinline__ uint32_t mgb_canon_off(uint32_t row, uint32_t k, uint32_t rows) {
    return (k >> 2) * (rows * 16u) + (row >> 3) * 128u + (row & 7u) * 16u + (k & 3u) * 4u;
}
