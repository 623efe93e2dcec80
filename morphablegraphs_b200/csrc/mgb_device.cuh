// Device-side building blocks shared by the kernels of libmgb200.so:
// Philox4x32 + Box-Muller GMM sampling of one latent row, quaternion-form
// forward kinematics, 1-D TMA bulk copies with mbarrier completion.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define MGB_QUAT_EPS 8.881784197001252e-16f /* 4 * DBL_EPSILON: transformations.quaternion_matrix returns identity below it */

// ---------------------------------------------------------------------------
// Philox4x32 (Salmon et al. 2011); counter = (global index lo, hi, block, tag), key = seed
// ---------------------------------------------------------------------------
// Rounds: 7 - the smallest count for which Salmon et al. report Philox4x32 to pass BigCrush ("Crush-resistant", SC'11 table 2);
// 10 is the library default with a safety margin.  The draws feed a Monte-Carlo search whose candidates are checked
// distributionally (tests), and Philox is ~25 % of the S step kernel's instruction stream.
#ifndef MGB_PHILOX_ROUNDS
#define MGB_PHILOX_ROUNDS 7
#endif
__device__ __forceinline__ uint4 mgb_philox4x32(uint4 ctr, uint2 key) {
    const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < MGB_PHILOX_ROUNDS; ++r) {
        const unsigned int hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const unsigned int hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

__device__ __forceinline__ float mgb_u01(unsigned int x) {  // (0, 1]
    return fmaf((float)(x >> 8), 1.0f / 16777216.0f, 0.5f / 16777216.0f);
}

__device__ __forceinline__ float mgb_sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// four standard normals of global index gi, block b (Box-Muller on one Philox call)
__device__ __forceinline__ void mgb_normal4(unsigned long long gi, unsigned int b, uint2 key, float z[4]) {
    const uint4 r = mgb_philox4x32(make_uint4((unsigned int)gi, (unsigned int)(gi >> 32), b, 0x6d67u), key);
    // radius: sqrt.approx (one MUFU, <= 2 ulp) instead of the IEEE sqrtf sequence - the draw is a random
    // number, and both users of this function (mgb_sample, the fused kernels) share it bit for bit
    const float r0 = mgb_sqrt_approx(-2.f * __logf(mgb_u01(r.x))), r1 = mgb_sqrt_approx(-2.f * __logf(mgb_u01(r.z)));
    float s0, c0, s1, c1;
    __sincosf(6.283185307179586f * mgb_u01(r.y), &s0, &c0);
    __sincosf(6.283185307179586f * mgb_u01(r.w), &s1, &c1);
    z[0] = __fmul_rn(r0, c0);
    z[1] = __fmul_rn(r0, s0);
    z[2] = __fmul_rn(r1, c1);
    z[3] = __fmul_rn(r1, s1);
}

// s = mu_g + L_g z for one candidate, everything in registers (D <= NREG).
// L4: lower Cholesky factor, row d at L4 + d * (LS / 4) as float4 (row stride LS floats, a
// multiple of 4; entries right of the diagonal are never read); mu [D].  Both may live in
// shared or global memory.  The accumulation order is fixed (j ascending, explicit fmaf), so
// the stand-alone sampler and the fused sample+score kernel produce bit-identical rows.
template <int NREG>
__device__ __forceinline__ void mgb_sample_row_reg(const float4* __restrict__ L4, int LS, const float* __restrict__ mu, int D,
                                                   unsigned long long gi, uint2 key, float (&s)[NREG]) {
    float z[NREG];
#pragma unroll
    for (int b = 0; b < NREG / 4; ++b) {
        if (4 * b < D) {
            float q[4];
            mgb_normal4(gi, (unsigned int)b, key, q);
            z[4 * b] = q[0]; z[4 * b + 1] = q[1]; z[4 * b + 2] = q[2]; z[4 * b + 3] = q[3];
        } else {
            z[4 * b] = z[4 * b + 1] = z[4 * b + 2] = z[4 * b + 3] = 0.f;
        }
    }
    const int LS4 = LS >> 2;
#pragma unroll
    for (int d = 0; d < NREG; ++d) {
        float acc = 0.f;
        if (d < D) {
            acc = mu[d];
            const float4* Lr = L4 + d * LS4;
#pragma unroll
            for (int c = 0; c <= d / 4; ++c) {
                const float4 l = Lr[c];
                acc = fmaf(l.x, z[4 * c], acc);
                if (4 * c + 1 <= d) acc = fmaf(l.y, z[4 * c + 1], acc);
                if (4 * c + 2 <= d) acc = fmaf(l.z, z[4 * c + 2], acc);
                if (4 * c + 3 <= d) acc = fmaf(l.w, z[4 * c + 3], acc);
            }
        }
        s[d] = acc;
    }
}

// Two candidates of the same mixture component at once: every Cholesky row (and mean) is loaded ONCE for both; each candidate's
// sums keep the order of mgb_sample_row_reg, so its row has the same bits.
template <int NREG>
__device__ __forceinline__ void mgb_sample_rows_reg2(const float4* __restrict__ L4, int LS, const float* __restrict__ mu, int D,
                                                    unsigned long long gi0, unsigned long long gi1, uint2 key, float (&s0)[NREG],
                                                    float (&s1)[NREG]) {
    float z0[NREG], z1[NREG];
#pragma unroll
    for (int b = 0; b < NREG / 4; ++b) {
        if (4 * b < D) {
            float q[4];
            mgb_normal4(gi0, (unsigned int)b, key, q);
            z0[4 * b] = q[0]; z0[4 * b + 1] = q[1]; z0[4 * b + 2] = q[2]; z0[4 * b + 3] = q[3];
            mgb_normal4(gi1, (unsigned int)b, key, q);
            z1[4 * b] = q[0]; z1[4 * b + 1] = q[1]; z1[4 * b + 2] = q[2]; z1[4 * b + 3] = q[3];
        } else {
            z0[4 * b] = z0[4 * b + 1] = z0[4 * b + 2] = z0[4 * b + 3] = 0.f;
            z1[4 * b] = z1[4 * b + 1] = z1[4 * b + 2] = z1[4 * b + 3] = 0.f;
        }
    }
    const int LS4 = LS >> 2;
#pragma unroll
    for (int d = 0; d < NREG; ++d) {
        float a0 = 0.f, a1 = 0.f;
        if (d < D) {
            a0 = a1 = mu[d];
            const float4* Lr = L4 + d * LS4;
#pragma unroll
            for (int c = 0; c <= d / 4; ++c) {
                const float4 l = Lr[c];
                a0 = fmaf(l.x, z0[4 * c], a0);
                a1 = fmaf(l.x, z1[4 * c], a1);
                if (4 * c + 1 <= d) { a0 = fmaf(l.y, z0[4 * c + 1], a0); a1 = fmaf(l.y, z1[4 * c + 1], a1); }
                if (4 * c + 2 <= d) { a0 = fmaf(l.z, z0[4 * c + 2], a0); a1 = fmaf(l.z, z1[4 * c + 2], a1); }
                if (4 * c + 3 <= d) { a0 = fmaf(l.w, z0[4 * c + 3], a0); a1 = fmaf(l.w, z1[4 * c + 3], a1); }
            }
        }
        s0[d] = a0;
        s1[d] = a1;
    }
}

// same with the row in memory (any D): z is written to row[], then s in place from the last row up
__device__ __forceinline__ void mgb_sample_row_mem(const float* __restrict__ L, int LS, const float* __restrict__ mu, int D,
                                                   unsigned long long gi, uint2 key, float* row) {
    for (int d0 = 0; d0 < D; d0 += 4) {
        float q[4];
        mgb_normal4(gi, (unsigned int)(d0 >> 2), key, q);
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (d0 + c < D) row[d0 + c] = q[c];
    }
    for (int d = D - 1; d >= 0; --d) {
        float acc = mu[d];
        const float* Lr = L + (size_t)d * LS;
        for (int j = 0; j <= d; ++j) acc = fmaf(Lr[j], row[j], acc);
        row[d] = acc;
    }
}

// ---------------------------------------------------------------------------
// quaternion-form rigid transform: x -> R(q) x + t with R(q) the rotation of the
// (not necessarily unit) quaternion q = (w, x, y, z), exactly transformations.quaternion_matrix:
// scale invariant, identity when |q|^2 < eps.  s2 = 2 / |q|^2 is carried along.
// ---------------------------------------------------------------------------
struct MgbXf {
    float w, x, y, z;
    float s2;
    float tx, ty, tz;
};

__device__ __forceinline__ void mgb_xf_identity(MgbXf& G) {
    G.w = 1.f; G.x = G.y = G.z = 0.f;
    G.s2 = 2.f;
    G.tx = G.ty = G.tz = 0.f;
}

__device__ __forceinline__ void mgb_xf_set_rot(MgbXf& G, float w, float x, float y, float z) {
    const float n = fmaf(w, w, fmaf(x, x, fmaf(y, y, z * z)));
    const bool ok = n >= MGB_QUAT_EPS;
    G.w = ok ? w : 1.f;
    G.x = ok ? x : 0.f;
    G.y = ok ? y : 0.f;
    G.z = ok ? z : 0.f;
    G.s2 = ok ? __fdividef(2.f, n) : 2.f;
}

// r = R(G) v
__device__ __forceinline__ void mgb_xf_rotate(const MgbXf& G, float vx, float vy, float vz, float& rx, float& ry, float& rz) {
    const float cx = fmaf(G.y, vz, -G.z * vy), cy = fmaf(G.z, vx, -G.x * vz), cz = fmaf(G.x, vy, -G.y * vx);
    const float dx = fmaf(G.y, cz, -G.z * cy), dy = fmaf(G.z, cx, -G.x * cz), dz = fmaf(G.x, cy, -G.y * cx);
    rx = fmaf(G.s2, fmaf(G.w, cx, dx), vx);
    ry = fmaf(G.s2, fmaf(G.w, cy, dy), vy);
    rz = fmaf(G.s2, fmaf(G.w, cz, dz), vz);
}

// G = G * [R(q) | off]   (child joint with local rotation q and bone offset off)
__device__ __forceinline__ void mgb_xf_child(MgbXf& G, float qw, float qx, float qy, float qz, float ox, float oy, float oz) {
    float rx, ry, rz;
    mgb_xf_rotate(G, ox, oy, oz, rx, ry, rz);
    G.tx += rx; G.ty += ry; G.tz += rz;
    const float n = fmaf(qw, qw, fmaf(qx, qx, fmaf(qy, qy, qz * qz)));
    if (n >= MGB_QUAT_EPS) {  // local rotation is identity below eps (quaternion_matrix)
        const float w = G.w * qw - G.x * qx - G.y * qy - G.z * qz;
        const float x = G.w * qx + G.x * qw + G.y * qz - G.z * qy;
        const float y = G.w * qy - G.x * qz + G.y * qw + G.z * qx;
        const float z = G.w * qz + G.x * qy - G.y * qx + G.z * qw;
        G.w = w; G.x = x; G.y = y; G.z = z;
        const float m = fmaf(w, w, fmaf(x, x, fmaf(y, y, z * z)));
        G.s2 = __fdividef(2.f, m);
    }
}

__device__ __forceinline__ void mgb_xf_translate(MgbXf& G, float ox, float oy, float oz) {
    float rx, ry, rz;
    mgb_xf_rotate(G, ox, oy, oz, rx, ry, rz);
    G.tx += rx; G.ty += ry; G.tz += rz;
}

// ---------------------------------------------------------------------------
// 1-D TMA bulk copy global -> shared with mbarrier transaction completion (sm_90+; SASS: UBLKCP)
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned int mgb_smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mgb_mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mgb_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mgb_fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mgb_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mgb_tma_load_1d(void* dst_smem, const void* src_gmem, unsigned int bytes, unsigned long long* bar) {
    const unsigned int b = mgb_smem_u32(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(mgb_smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(b)
                 : "memory");
}

// suspend-time hint of try_wait: a waiting warp sleeps in hardware (until the phase completes or the hint expires)
// instead of re-issuing the probe - in the warp-specialised kernels the waiters share their scheduler with the
// one working warp, and 58 % of all executed instructions were probes (profiles/README.md, r01_z)
#ifndef MGB_MBAR_SUSPEND_NS
#define MGB_MBAR_SUSPEND_NS 2000u
#endif
__device__ __forceinline__ void mgb_mbar_wait(unsigned long long* bar, unsigned int parity) {
    const unsigned int b = mgb_smem_u32(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MGB_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra MGB_DONE_%=;\n"
        "bra MGB_WAIT_%=;\n"
        "MGB_DONE_%=:\n"
        "}\n" ::"r"(b),
        "r"(parity), "r"(MGB_MBAR_SUSPEND_NS)
        : "memory");
}

// 1-D TMA bulk store shared -> global (bulk async-group completion)
__device__ __forceinline__ void mgb_tma_store_1d(void* dst_gmem, const void* src_smem, unsigned int bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(mgb_smem_u32(src_smem)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void mgb_tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
