// K6 / K5a: fused keyframe score kernels (sm_100a).
//
// One thread scores one candidate.  The constraint list was compiled by
// mgb_constraints_pack into (a) a table of "quads": for every constrained time
// slot and every needed group of 4 pose channels (root translation, one joint
// quaternion) the spatial basis pre-contracted with that slot's 4 B-spline
// weights, laid out [quad][k][4] so one 16-byte broadcast load feeds 4 FMAs, and
// (b) a small op program: a depth-first walk over the needed part of the
// skeleton (forward kinematics with a 3x4 transform held in registers) with the
// constraint terms attached to the nodes they read.
//
// What it replaces in the reference, per candidate:
//   MotionPrimitive.back_project_spatial_coeffs   motion_primitive.py:236-256
//   MotionSpline.evaluate (P splev calls)         motion_spline.py:89-92
//   anim_utils SkeletonNode.get_global_matrix     (call sites global_transform_constraint.py:103 ...)
//   constraint.evaluate_motion_spline, weighted sum   motion_primitive_constraints.py:118-122
//   naturalness -log p(s)                         objective_functions.py:256
//   + 2|s|                                        mg_state_planner.py:110-129
//   first-minimum arg-min                         motion_primitive_generator.py:251-257
//
// Memory: s tile staged through shared memory with coalesced 128-bit loads
// (rows are D floats = 52 B at the small shape, so a thread-per-row global read
// would touch 13 lines per request); quad table, op program and constants are
// copied to shared memory once per CTA (persistent, grid-stride over tiles);
// block arg-min by warp shuffles, last-arriving CTA folds the per-CTA partials.
#include <float.h>
#include <math.h>

#include "mgb_internal.cuh"

#define MGB_BLOCK 256
#define MGB_QUAT_EPS 8.881784197001252e-16f /* 4 * DBL_EPSILON, transformations.quaternion_matrix */

namespace {

struct Xf {
    float R[9];
    float t[3];
};

__device__ __forceinline__ void quat_to_R(float w, float x, float y, float z, float* R) {
    float n = w * w + x * x + y * y + z * z;
    if (n < MGB_QUAT_EPS) {
        R[0] = 1.f; R[1] = 0.f; R[2] = 0.f;
        R[3] = 0.f; R[4] = 1.f; R[5] = 0.f;
        R[6] = 0.f; R[7] = 0.f; R[8] = 1.f;
        return;
    }
    float s = __fdividef(2.0f, n);
    float xs = x * s, ys = y * s, zs = z * s;
    float wx = w * xs, wy = w * ys, wz = w * zs;
    float xx = x * xs, xy = x * ys, xz = x * zs;
    float yy = y * ys, yz = y * zs, zz = z * zs;
    R[0] = 1.f - (yy + zz); R[1] = xy - wz;         R[2] = xz + wy;
    R[3] = xy + wz;         R[4] = 1.f - (xx + zz); R[5] = yz - wx;
    R[6] = xz - wy;         R[7] = yz + wx;         R[8] = 1.f - (xx + yy);
}

// G = G * [Rl | tl]
__device__ __forceinline__ void xf_mul(Xf& G, const float* Rl, float tx, float ty, float tz) {
    G.t[0] = fmaf(G.R[0], tx, fmaf(G.R[1], ty, fmaf(G.R[2], tz, G.t[0])));
    G.t[1] = fmaf(G.R[3], tx, fmaf(G.R[4], ty, fmaf(G.R[5], tz, G.t[1])));
    G.t[2] = fmaf(G.R[6], tx, fmaf(G.R[7], ty, fmaf(G.R[8], tz, G.t[2])));
    float N[9];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            N[i * 3 + j] = fmaf(G.R[i * 3 + 0], Rl[0 + j], fmaf(G.R[i * 3 + 1], Rl[3 + j], G.R[i * 3 + 2] * Rl[6 + j]));
        }
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) G.R[i] = N[i];
}

__device__ __forceinline__ void xf_translate(Xf& G, float tx, float ty, float tz) {
    G.t[0] = fmaf(G.R[0], tx, fmaf(G.R[1], ty, fmaf(G.R[2], tz, G.t[0])));
    G.t[1] = fmaf(G.R[3], tx, fmaf(G.R[4], ty, fmaf(G.R[5], tz, G.t[1])));
    G.t[2] = fmaf(G.R[6], tx, fmaf(G.R[7], ty, fmaf(G.R[8], tz, G.t[2])));
}

#define MGB_FMA4(k)                                  \
    {                                                \
        float4 e = row[k];                           \
        acc.x = fmaf(s[k], e.x, acc.x);              \
        acc.y = fmaf(s[k], e.y, acc.y);              \
        acc.z = fmaf(s[k], e.z, acc.z);              \
        acc.w = fmaf(s[k], e.w, acc.w);              \
    }
#define MGB_CASE(k) \
    case (k) + 1:   \
        if ((k) < NREG) MGB_FMA4((k) < NREG ? (k) : 0)

// acc = row[bias] + sum_{k < cnt} s[k] * row[k]; s lives in registers, the
// runtime count is dispatched once through a fall-through switch
template <int NREG>
__device__ __forceinline__ float4 dotquad_reg(const float4* __restrict__ row, int bias, int cnt, const float (&s)[NREG > 0 ? NREG : 1]) {
    float4 acc = row[bias];
    switch (cnt) {
        MGB_CASE(31) MGB_CASE(30) MGB_CASE(29) MGB_CASE(28) MGB_CASE(27) MGB_CASE(26) MGB_CASE(25) MGB_CASE(24)
        MGB_CASE(23) MGB_CASE(22) MGB_CASE(21) MGB_CASE(20) MGB_CASE(19) MGB_CASE(18) MGB_CASE(17) MGB_CASE(16)
        MGB_CASE(15) MGB_CASE(14) MGB_CASE(13) MGB_CASE(12) MGB_CASE(11) MGB_CASE(10) MGB_CASE(9) MGB_CASE(8)
        MGB_CASE(7) MGB_CASE(6) MGB_CASE(5) MGB_CASE(4) MGB_CASE(3) MGB_CASE(2) MGB_CASE(1) MGB_CASE(0)
        default: break;
    }
    return acc;
}

// same with s read from the thread's shared-memory row (large D)
__device__ __forceinline__ float4 dotquad_smem(const float4* __restrict__ row, int bias, int cnt, const float* __restrict__ s) {
    float4 acc = row[bias];
    for (int k = 0; k < cnt; ++k) {
        float4 e = row[k];
        float sk = s[k];
        acc.x = fmaf(sk, e.x, acc.x);
        acc.y = fmaf(sk, e.y, acc.y);
        acc.z = fmaf(sk, e.z, acc.z);
        acc.w = fmaf(sk, e.w, acc.w);
    }
    return acc;
}

__device__ __forceinline__ float masked_dist(unsigned mask, const float* tgt, const float* p) {
    float d = 0.f;
    if (mask & 1u) { float e = tgt[0] - p[0]; d = fmaf(e, e, d); }
    if (mask & 2u) { float e = tgt[1] - p[1]; d = fmaf(e, e, d); }
    if (mask & 4u) { float e = tgt[2] - p[2]; d = fmaf(e, e, d); }
    return sqrtf(d);
}

__device__ __forceinline__ float dist3(const float* a, const float* b) {
    float x = a[0] - b[0], y = a[1] - b[1], z = a[2] - b[2];
    return sqrtf(fmaf(x, x, fmaf(y, y, z * z)));
}

__device__ __forceinline__ bool better(float c1, long long i1, float c2, long long i2) {
    return (c1 < c2) || (c1 == c2 && i1 < i2);
}

// ---------------------------------------------------------------------------
// NREG > 0 : the whole latent vector s lives in registers (D <= NREG)
// NREG == 0: s stays in the shared-memory tile (any D)
// FRAMES   : pose channels come from materialised frames instead of the
//            contracted basis (staged pipeline, K5a)
// ---------------------------------------------------------------------------
template <int NREG, bool FRAMES, bool TERMS, typename SIn>
__global__ void __launch_bounds__(MGB_BLOCK) mgb_score_kernel(ScoreParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int D = p.D, K = p.K;
    const int SROW = D | 1;  // odd row stride: conflict-free per-thread rows

    // ---- carve shared memory ------------------------------------------------
    size_t off = 0;
    float4* sm_tab = reinterpret_cast<float4*>(smem_raw + off);
    if (p.tab_in_smem) off += (size_t)p.n_quads * (K + 1) * sizeof(float4);
    float4* sm_gmm = reinterpret_cast<float4*>(smem_raw + off);
    if (p.gmm_in_smem) off += (size_t)p.G * p.NQ * (D + 1) * sizeof(float4);
    MgbOp* sm_ops = reinterpret_cast<MgbOp*>(smem_raw + off);
    off += (size_t)p.n_ops * sizeof(MgbOp);
    float* sm_cst = reinterpret_cast<float*>(smem_raw + off);
    off += (size_t)((p.n_cst + 3) & ~3) * sizeof(float);
    float* sm_s = reinterpret_cast<float*>(smem_raw + off);
    off += (size_t)MGB_BLOCK * SROW * sizeof(float);
    off = (off + 15) & ~(size_t)15;
    float* sm_redc = reinterpret_cast<float*>(smem_raw + off);
    off += 32 * sizeof(float);
    long long* sm_redi = reinterpret_cast<long long*>(smem_raw + ((off + 7) & ~(size_t)7));

    // ---- one-time staging of program + tables -------------------------------
    if (p.tab_in_smem) {
        const int n4 = p.n_quads * (K + 1);
        for (int i = tid; i < n4; i += MGB_BLOCK) sm_tab[i] = p.tab[i];
    }
    if (p.gmm_in_smem) {
        const int n4 = p.G * p.NQ * (D + 1);
        for (int i = tid; i < n4; i += MGB_BLOCK) sm_gmm[i] = p.gmm_tab[i];
    }
    for (int i = tid; i < p.n_ops; i += MGB_BLOCK) sm_ops[i] = p.ops[i];
    for (int i = tid; i < p.n_cst; i += MGB_BLOCK) sm_cst[i] = p.cst[i];
    const float4* tab = p.tab_in_smem ? sm_tab : p.tab;
    const float4* gtab = p.gmm_in_smem ? sm_gmm : p.gmm_tab;
    const bool use_s = (!FRAMES) || (p.s != nullptr);

    float best_c = INFINITY;
    long long best_i = LLONG_MAX;

    const long long n_tiles = (p.n + MGB_BLOCK - 1) / MGB_BLOCK;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long base = tile * MGB_BLOCK;
        __syncthreads();
        if (use_s) {
            // coalesced tile load: element e of the tile -> row e / D, column e % D,
            // tracked incrementally (no per-element division)
            const SIn* __restrict__ src = reinterpret_cast<const SIn*>(p.s) + base * D;
            long long remaining = (p.n - base) * (long long)D;
            const int total = MGB_BLOCK * D;
            const int step_r = MGB_BLOCK / D, step_c = MGB_BLOCK % D;
            int r = tid / D, c = tid % D;
            for (int e = tid; e < total; e += MGB_BLOCK) {
                float v = 0.f;
                if (e < remaining) v = (float)__ldg(src + e);
                sm_s[r * SROW + c] = v;
                r += step_r;
                c += step_c;
                if (c >= D) { c -= D; r += 1; }
            }
        }
        __syncthreads();

        const long long i = base + tid;
        const bool valid = i < p.n;
        const float* srow = sm_s + tid * SROW;
        float s[NREG > 0 ? NREG : 1];
        if (NREG > 0) {
#pragma unroll
            for (int k = 0; k < NREG; ++k) s[k] = (k < D && use_s) ? srow[k] : 0.f;
        }
        const float* frow = FRAMES ? (p.frames + (size_t)(valid ? i : 0) * p.n_times * p.P) : nullptr;

        // ---- interpreter state ------------------------------------------------
        Xf G;
#pragma unroll
        for (int q = 0; q < 9; ++q) G.R[q] = (q % 4 == 0) ? 1.f : 0.f;
        G.t[0] = G.t[1] = G.t[2] = 0.f;
        float al_c = 1.f, al_s = 0.f, al_dx = 0.f, al_dz = 0.f, al_qw = 1.f, al_qy = 0.f;
        float total = 0.f;
        float scratch[MGB_SCRATCH_FLOATS];
        float stack[MGB_STACK_DEPTH * 12];
        int sp = 0;

        auto quad = [&](int q) -> float4 {
            if (FRAMES) {
                const float* f = frow + q;
                return make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), __ldg(f + 3));
            } else if (NREG > 0) {
                return dotquad_reg<NREG>(tab + (size_t)q * (K + 1), K, K, s);
            } else {
                return dotquad_smem(tab + (size_t)q * (K + 1), K, K, srow);
            }
        };
        auto emit = [&](int term, float w, float e) {
            total = fmaf(w, e, total);
            if (TERMS) {
                if (valid) p.terms[(size_t)term * p.n + i] += e;
            }
        };

        for (int pc = 0; pc < p.n_ops; ++pc) {
            const MgbOp op = sm_ops[pc];
            const float* cs = sm_cst + op.c;
            switch (op.code) {
                case OP_ROOT: {
                    float4 tr = quad(op.a);
                    float4 q = quad(op.b);
                    // optional alignment to the previous motion (identity when local)
                    float tx = al_c * tr.x + al_s * tr.z + al_dx;
                    float tz = -al_s * tr.x + al_c * tr.z + al_dz;
                    float qw = al_qw * q.x - al_qy * q.z;
                    float qx = al_qw * q.y + al_qy * q.w;
                    float qy = al_qw * q.z + al_qy * q.x;
                    float qz = al_qw * q.w - al_qy * q.y;
                    quat_to_R(qw, qx, qy, qz, G.R);
                    G.t[0] = tx + cs[0];
                    G.t[1] = tr.y + cs[1];
                    G.t[2] = tz + cs[2];
                    break;
                }
                case OP_JOINT: {
                    float4 q = quad(op.a);
                    float Rl[9];
                    quat_to_R(q.x, q.y, q.z, q.w, Rl);
                    xf_mul(G, Rl, cs[0], cs[1], cs[2]);
                    break;
                }
                case OP_CONST: {
                    if (op.a) {
                        xf_translate(G, cs[9], cs[10], cs[11]);
                    } else {
                        float Rl[9];
#pragma unroll
                        for (int q = 0; q < 9; ++q) Rl[q] = cs[q];
                        xf_mul(G, Rl, cs[9], cs[10], cs[11]);
                    }
                    break;
                }
                case OP_PUSH: {
#pragma unroll
                    for (int q = 0; q < 9; ++q) stack[sp * 12 + q] = G.R[q];
#pragma unroll
                    for (int q = 0; q < 3; ++q) stack[sp * 12 + 9 + q] = G.t[q];
                    ++sp;
                    break;
                }
                case OP_POP: {
                    --sp;
#pragma unroll
                    for (int q = 0; q < 9; ++q) G.R[q] = stack[sp * 12 + q];
#pragma unroll
                    for (int q = 0; q < 3; ++q) G.t[q] = stack[sp * 12 + 9 + q];
                    break;
                }
                case OP_ALIGN: {
                    // heading of the aligning node at control point 0 (G.R), raw root
                    // translation of control point 0 (quad op.a), previous heading / root
                    float px = G.R[0] * cs[0] + G.R[1] * cs[1] + G.R[2] * cs[2];
                    float pz = G.R[6] * cs[0] + G.R[7] * cs[1] + G.R[8] * cs[2];
                    float ang = cs[3] - atan2f(px, pz);
                    const float PI = 3.14159265358979323846f;
                    if (ang > PI) ang -= 2.f * PI;
                    if (ang <= -PI) ang += 2.f * PI;
                    sincosf(ang, &al_s, &al_c);
                    sincosf(0.5f * ang, &al_qy, &al_qw);
                    float4 tr = quad(op.a);
                    al_dx = cs[4] - (al_c * tr.x + al_s * tr.z);
                    al_dz = cs[5] - (-al_s * tr.x + al_c * tr.z);
                    break;
                }
                case OP_POS: {
                    emit(op.a, cs[0], masked_dist((unsigned)op.b, cs + 1, G.t));
                    break;
                }
                case OP_ORIENT: {
                    // angle_between_vectors(v1, v2) = acos(v1.v2 / |v1||v2|), evaluated as
                    // atan2(|v1 x v2|, v1.v2): same value, but well conditioned near 0 and pi in fp32
                    float vx = G.R[2], vy = G.R[5], vz = G.R[8];
                    float cx = vy * cs[3] - vz * cs[2], cy = vz * cs[1] - vx * cs[3], cz = vx * cs[2] - vy * cs[1];
                    float d = vx * cs[1] + vy * cs[2] + vz * cs[3];
                    emit(op.a, cs[0], atan2f(sqrtf(cx * cx + cy * cy + cz * cz), d));
                    break;
                }
                case OP_REL: {
                    float ox = cs[1], oy = cs[2], oz = cs[3], ow = cs[4];
                    float pp[3];
                    pp[0] = G.R[0] * ox + G.R[1] * oy + G.R[2] * oz + G.t[0] * ow;
                    pp[1] = G.R[3] * ox + G.R[4] * oy + G.R[5] * oz + G.t[1] * ow;
                    pp[2] = G.R[6] * ox + G.R[7] * oy + G.R[8] * oz + G.t[2] * ow;
                    emit(op.a, cs[0], dist3(cs + 5, pp));
                    break;
                }
                case OP_DIR2D: {
                    // degrees(acos(clamp(t.d / (|t||d|)))) evaluated as atan2(|t x d|, t.d) (same value,
                    // well conditioned for small angles in fp32); d = (x, z) of R_global * ref_dir
                    float px = G.R[0] * cs[1] + G.R[1] * cs[2] + G.R[2] * cs[3];
                    float pz = G.R[6] * cs[1] + G.R[7] * cs[2] + G.R[8] * cs[3];
                    float cr = cs[4] * pz - cs[5] * px;
                    float dt = cs[4] * px + cs[5] * pz;
                    emit(op.a, cs[0], atan2f(fabsf(cr), dt) * 57.29577951308232f);
                    break;
                }
                case OP_LOOKAT: {
                    float tx = cs[1] - G.t[0], ty = cs[2] - G.t[1], tz = cs[3] - G.t[2];
                    float ti = rsqrtf(tx * tx + ty * ty + tz * tz);
                    tx *= ti; ty *= ti; tz *= ti;
                    float hx = G.R[2], hy = G.R[5], hz = G.R[8];
                    float hi = rsqrtf(hx * hx + hy * hy + hz * hz);
                    hx *= hi; hy *= hi; hz *= hi;
                    float cx = hy * tz - hz * ty, cy = hz * tx - hx * tz, cz = hx * ty - hy * tx;
                    float sn = sqrtf(cx * cx + cy * cy + cz * cz);
                    float cn = hx * tx + hy * ty + hz * tz;
                    emit(op.a, cs[0], fabsf(atan2f(sn, cn)));
                    break;
                }
                case OP_STORE: {
                    scratch[op.a] = G.t[0];
                    scratch[op.a + 1] = G.t[1];
                    scratch[op.a + 2] = G.t[2];
                    break;
                }
                case OP_TWOHANDS: {
                    const float* lp = scratch + op.b;
                    const float* rp = scratch + op.b + 3;
                    float mid[3] = {lp[0] + 0.5f * (rp[0] - lp[0]), lp[1] + 0.5f * (rp[1] - lp[1]), lp[2] + 0.5f * (rp[2] - lp[2])};
                    float e = dist3(cs + 1, mid) + dist3(cs + 4, lp) + dist3(cs + 7, rp);
                    emit(op.a, cs[0], e);
                    break;
                }
                case OP_FEET: {
                    const float* lp = scratch + op.b;
                    const float* rp = scratch + op.b + 3;
                    float e = (dist3(cs + 1, lp) + dist3(cs + 4, rp)) * cs[7];
                    emit(op.a, cs[0], e);
                    break;
                }
                case OP_POSE: {
                    // weighted 2-D rigid alignment of the candidate's point cloud (b) onto the
                    // target cloud (a), then mean point distance + root velocity error
                    const int np = (int)cs[1];
                    const float* pts = scratch + op.b;
                    const float W = cs[6], sax = cs[7], saz = cs[8];
                    const float* tg = cs + 9;  // (ax, ay, az, w) * np
                    float num_l = 0.f, den_l = 0.f, sbx = 0.f, sbz = 0.f;
                    for (int j = 0; j < np; ++j) {
                        float ax = tg[4 * j], az = tg[4 * j + 2], w = tg[4 * j + 3];
                        float bx = pts[3 * j], bz = pts[3 * j + 2];
                        num_l = fmaf(w, ax * bz - bx * az, num_l);
                        den_l = fmaf(w, ax * bx + az * bz, den_l);
                        sbx = fmaf(w, bx, sbx);
                        sbz = fmaf(w, bz, sbz);
                    }
                    float invW = 1.f / W;
                    float num = num_l - (sax * sbz - sbx * saz) * invW;
                    float den = den_l - (sax * sbx + saz * sbz) * invW;
                    float theta = atan2f(num, den);
                    float st, ct;
                    sincosf(theta, &st, &ct);
                    float ox = (sax - sbx * ct - sbz * st) * invW;
                    float oz = (saz + sbx * st - sbz * ct) * invW;
                    float dsum = 0.f;
                    for (int j = 0; j < np; ++j) {
                        float bx = pts[3 * j], by = pts[3 * j + 1], bz = pts[3 * j + 2];
                        float tx = ct * bx + st * bz + ox;
                        float tz = -st * bx + ct * bz + oz;
                        float ex = tg[4 * j] - tx, ey = tg[4 * j + 1] - by, ez = tg[4 * j + 2] - tz;
                        dsum += sqrtf(fmaf(ex, ex, fmaf(ey, ey, ez * ez)));
                    }
                    float e = dsum / (float)np;
                    if (cs[2] != 0.f) {
                        const float* r2 = pts + 3 * np;
                        float vx = cs[3] - (r2[0] - pts[0]), vy = cs[4] - (r2[1] - pts[1]), vz = cs[5] - (r2[2] - pts[2]);
                        e += sqrtf(fmaf(vx, vx, fmaf(vy, vy, vz * vz)));
                    }
                    emit(op.a, cs[0], e);
                    break;
                }
                case OP_CA_BEGIN: {
                    scratch[op.a] = INFINITY;
                    break;
                }
                case OP_CA_ACC: {
                    scratch[op.a] = fminf(scratch[op.a], masked_dist((unsigned)op.b, cs, G.t));
                    break;
                }
                case OP_CA_END: {
                    emit(op.a, cs[0], scratch[op.b]);
                    break;
                }
                default:
                    break;
            }
        }

        float cost = p.err_scale * total;
        if (p.quality_scale != 0.f) {
            // GMM log-density from the precision Cholesky factors; column quads of the
            // upper-triangular factor only need rows k <= 4q+3
            float mx = -INFINITY, ssum = 0.f;
            for (int g = 0; g < p.G; ++g) {
                float qf = 0.f;
                for (int jq = 0; jq < p.NQ; ++jq) {
                    const float4* row = gtab + (size_t)(g * p.NQ + jq) * (D + 1);
                    int cnt = min(D, 4 * jq + 4);
                    float4 y = (NREG > 0) ? dotquad_reg<NREG>(row, D, cnt, s) : dotquad_smem(row, D, cnt, srow);
                    qf = fmaf(y.x, y.x, fmaf(y.y, y.y, fmaf(y.z, y.z, fmaf(y.w, y.w, qf))));
                }
                float lp = p.gmm_c[g] - 0.5f * qf;
                if (lp > mx) {
                    ssum = ssum * __expf(mx - lp) + 1.f;
                    mx = lp;
                } else {
                    ssum += __expf(lp - mx);
                }
            }
            cost = fmaf(p.quality_scale, -(mx + __logf(ssum)), cost);
        }
        if (p.magnitude_scale != 0.f) {
            float m2 = 0.f;
            if (NREG > 0) {
#pragma unroll
                for (int k = 0; k < NREG; ++k) m2 = fmaf(s[k], s[k], m2);
            } else {
                for (int k = 0; k < D; ++k) m2 = fmaf(srow[k], srow[k], m2);
            }
            cost = fmaf(p.magnitude_scale, sqrtf(m2), cost);
        }
        if (valid) {
            if (p.cost) p.cost[i] = cost;
            const long long gi = p.index_base + i;
            if (better(cost, gi, best_c, best_i)) { best_c = cost; best_i = gi; }
        }
    }

    if (!p.want_best) return;
    // ---- block arg-min: warp shuffles, then one warp over the warp results ---
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
        long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
    }
    __syncthreads();
    if ((tid & 31) == 0) { sm_redc[tid >> 5] = best_c; sm_redi[tid >> 5] = best_i; }
    __syncthreads();
    if (tid < 32) {
        best_c = (tid < MGB_BLOCK / 32) ? sm_redc[tid] : INFINITY;
        best_i = (tid < MGB_BLOCK / 32) ? sm_redi[tid] : LLONG_MAX;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
            long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
        }
    }
    __shared__ bool is_last;
    if (tid == 0) {
        p.part_cost[blockIdx.x] = best_c;
        p.part_idx[blockIdx.x] = best_i;
        __threadfence();
        unsigned int t = atomicAdd(p.done_count, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    best_c = INFINITY;
    best_i = LLONG_MAX;
    for (int b = tid; b < (int)gridDim.x; b += MGB_BLOCK) {
        float oc = __ldcg(p.part_cost + b);
        long long oi = __ldcg(p.part_idx + b);
        if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
        long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
    }
    __syncthreads();
    if ((tid & 31) == 0) { sm_redc[tid >> 5] = best_c; sm_redi[tid >> 5] = best_i; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < MGB_BLOCK / 32; ++w)
            if (better(sm_redc[w], sm_redi[w], best_c, best_i)) { best_c = sm_redc[w]; best_i = sm_redi[w]; }
        if (best_i == LLONG_MAX) best_i = p.index_base;  // nothing finite: reference keeps index 0
        p.best_cost[0] = best_c;
        p.best_idx[0] = best_i;
        *p.done_count = 0u;  // ready for the next launch
    }
}

// ---------------------------------------------------------------------------
// K5c: k smallest of a cost array (k passes; k is 1..64 in every caller)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(MGB_BLOCK) mgb_topk_partial(const float* __restrict__ cost, long long n, long long index_base,
                                                             const float* __restrict__ prev_cost, const long long* __restrict__ prev_idx,
                                                             int have_prev, float* part_cost, long long* part_idx) {
    // candidates strictly "after" the previous pick in (cost, index) order
    float pc = have_prev ? prev_cost[0] : -INFINITY;
    long long pi = have_prev ? prev_idx[0] : -1;
    float best_c = INFINITY;
    long long best_i = LLONG_MAX;
    for (long long i = (long long)blockIdx.x * MGB_BLOCK + threadIdx.x; i < n; i += (long long)gridDim.x * MGB_BLOCK) {
        float c = cost[i];
        long long gi = index_base + i;
        bool after = !have_prev || (c > pc) || (c == pc && gi > pi);
        if (after && better(c, gi, best_c, best_i)) { best_c = c; best_i = gi; }
    }
    __shared__ float rc[MGB_BLOCK / 32];
    __shared__ long long ri[MGB_BLOCK / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
        long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
    }
    if ((threadIdx.x & 31) == 0) { rc[threadIdx.x >> 5] = best_c; ri[threadIdx.x >> 5] = best_i; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < MGB_BLOCK / 32; ++w)
            if (better(rc[w], ri[w], best_c, best_i)) { best_c = rc[w]; best_i = ri[w]; }
        part_cost[blockIdx.x] = best_c;
        part_idx[blockIdx.x] = best_i;
    }
}

__global__ void mgb_topk_final(const float* part_cost, const long long* part_idx, int nparts, long long index_base,
                               float* out_cost, long long* out_idx) {
    float best_c = INFINITY;
    long long best_i = LLONG_MAX;
    for (int b = threadIdx.x; b < nparts; b += 32) {
        if (better(part_cost[b], part_idx[b], best_c, best_i)) { best_c = part_cost[b]; best_i = part_idx[b]; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
        long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
    }
    if (threadIdx.x == 0) {
        if (best_i == LLONG_MAX) best_i = index_base;
        out_cost[0] = best_c;
        out_idx[0] = best_i;
    }
}

int g_num_sms = 0;
int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

template <int NREG, bool FRAMES, bool TERMS, typename SIn>
int launch_variant(const ScoreParams& p, size_t smem, cudaStream_t stream) {
    auto kern = mgb_score_kernel<NREG, FRAMES, TERMS, SIn>;
    static bool configured = false;
    static size_t configured_smem = 0;
    if (!configured || smem > configured_smem) {
        MGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
        configured = true;
        configured_smem = 200 * 1024;
    }
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, MGB_BLOCK, smem));
    if (occ < 1) occ = 1;
    long long n_tiles = (p.n + MGB_BLOCK - 1) / MGB_BLOCK;
    long long grid = (long long)num_sms() * occ;
    if (grid > n_tiles) grid = n_tiles;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, MGB_BLOCK, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

template <bool FRAMES, bool TERMS, typename SIn>
int launch_nreg(int D, const ScoreParams& p, size_t smem, cudaStream_t stream) {
    if (D <= 16) return launch_variant<16, FRAMES, TERMS, SIn>(p, smem, stream);
    if (D <= 32) return launch_variant<32, FRAMES, TERMS, SIn>(p, smem, stream);
    return launch_variant<0, FRAMES, TERMS, SIn>(p, smem, stream);
}

}  // namespace

size_t mgb_score_grid_upper_bound() { return (size_t)num_sms() * 8; }

extern "C" size_t mgb_score_workspace_bytes(int64_t n, int32_t k) {
    (void)n;
    (void)k;
    // per-CTA partials (cost + index) + completion counter, generously sized
    size_t parts = 148 * 16;
    return parts * (sizeof(float) + sizeof(long long)) + 256;
}

int mgb_launch_score(const mgb_model* m, const mgb_constraints* c, const void* d_s, int s_is_f64,
                     const float* d_frames, int64_t n, int64_t index_base, float err_scale,
                     float quality_scale, float magnitude_scale, float* d_cost, float* d_terms, int k,
                     float* d_best_cost, int64_t* d_best_idx, void* d_ws, size_t ws_bytes,
                     cudaStream_t stream) {
    MGB_REQUIRE(m && c, "null handle");
    MGB_REQUIRE(c->model == m, "constraints were packed for a different model");
    MGB_REQUIRE(n >= 0, "negative n");
    MGB_REQUIRE(k >= 0 && k <= 64, "k must be in [0, 64]");
    const bool frames = d_frames != nullptr;
    MGB_REQUIRE(frames || d_s || n == 0, "d_s is NULL");
    MGB_REQUIRE(!(frames && s_is_f64), "frames mode takes float32 s");
    if ((quality_scale != 0.f || magnitude_scale != 0.f) && frames) MGB_REQUIRE(d_s, "logpdf / magnitude need d_s");
    if (k > 0) {
        MGB_REQUIRE(d_best_cost && d_best_idx, "k > 0 needs d_best_cost / d_best_idx");
        MGB_REQUIRE(d_ws && ws_bytes >= mgb_score_workspace_bytes(n, k), "workspace too small");
        MGB_REQUIRE(k == 1 || d_cost, "k > 1 needs d_cost");
    }
    if (n == 0) {
        if (k > 0) {
            // empty batch: reference returns index 0 / inf
            float inf = INFINITY;
            for (int j = 0; j < k; ++j) {
                MGB_CUDA(cudaMemcpyAsync(d_best_cost + j, &inf, sizeof(float), cudaMemcpyHostToDevice, stream));
                MGB_CUDA(cudaMemcpyAsync(d_best_idx + j, &index_base, sizeof(int64_t), cudaMemcpyHostToDevice, stream));
            }
            MGB_CUDA(cudaStreamSynchronize(stream));
        }
        return MGB_OK;
    }
    ScoreParams p;
    p.tab = c->d_tab;
    p.ops = frames ? c->d_ops_frames : c->d_ops_fused;
    p.cst = c->d_cst;
    p.n_ops = c->n_ops;
    p.n_cst = c->n_cst;
    p.n_quads = c->n_quads;
    p.K = m->K;
    p.D = m->D;
    p.P = m->P;
    p.n_times = c->n_times;
    p.n_terms = c->n_terms;
    p.gmm_tab = m->d_gmm_tab;
    p.gmm_c = m->d_gmm_c;
    p.G = m->G;
    p.NQ = m->NQ;
    p.s = d_s;
    p.frames = d_frames;
    p.n = n;
    p.index_base = index_base;
    p.err_scale = err_scale;
    p.quality_scale = quality_scale;
    p.magnitude_scale = magnitude_scale;
    p.cost = d_cost;
    p.terms = d_terms;
    p.want_best = (k == 1);
    unsigned char* ws = reinterpret_cast<unsigned char*>(d_ws);
    const size_t parts = 148 * 16;
    p.part_idx = reinterpret_cast<long long*>(ws);
    p.part_cost = reinterpret_cast<float*>(ws + parts * sizeof(long long));
    p.done_count = reinterpret_cast<unsigned int*>(ws + parts * (sizeof(long long) + sizeof(float)));
    p.best_cost = d_best_cost;
    p.best_idx = reinterpret_cast<long long*>(d_best_idx);

    // shared-memory budget: tables go to shared memory when they fit
    size_t tab_bytes = frames ? 0 : (size_t)c->n_quads * (m->K + 1) * sizeof(float4);
    size_t gmm_bytes = (quality_scale != 0.f) ? (size_t)m->G * m->NQ * (m->D + 1) * sizeof(float4) : 0;
    p.tab_in_smem = (!frames && tab_bytes <= 96 * 1024) ? 1 : 0;
    size_t smem = p.tab_in_smem ? tab_bytes : 0;
    p.gmm_in_smem = (gmm_bytes > 0 && smem + gmm_bytes <= 128 * 1024) ? 1 : 0;
    if (p.gmm_in_smem) smem += gmm_bytes;
    smem += (size_t)c->n_ops * sizeof(MgbOp);
    smem += (size_t)((c->n_cst + 3) & ~3) * sizeof(float);
    smem += (size_t)MGB_BLOCK * (m->D | 1) * sizeof(float);
    smem += 16 + 32 * sizeof(float) + 8 + 32 * sizeof(long long);
    MGB_REQUIRE(smem <= 200 * 1024, "constraint program too large for shared memory (%zu bytes)", smem);
    // the completion counter in the workspace must be zero on entry: the caller zero-fills a
    // fresh workspace once (include/mgb200.h) and the kernel re-arms it at the end of every launch
    if (d_terms) MGB_CUDA(cudaMemsetAsync(d_terms, 0, (size_t)c->n_terms * n * sizeof(float), stream));

    int rc;
    const int D = m->D;
    if (frames) {
        rc = d_terms ? launch_nreg<true, true, float>(D, p, smem, stream) : launch_nreg<true, false, float>(D, p, smem, stream);
    } else if (s_is_f64) {
        rc = d_terms ? launch_nreg<false, true, double>(D, p, smem, stream) : launch_nreg<false, false, double>(D, p, smem, stream);
    } else {
        rc = d_terms ? launch_nreg<false, true, float>(D, p, smem, stream) : launch_nreg<false, false, float>(D, p, smem, stream);
    }
    if (rc != MGB_OK) return rc;
    if (k > 1) {
        return mgb_topk(d_cost, n, index_base, k, d_best_cost, d_best_idx, d_ws, ws_bytes, stream);
    }
    return MGB_OK;
}

extern "C" int mgb_topk(const float* d_cost, int64_t n, int64_t index_base, int32_t k, float* d_best_cost,
                        int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(d_cost && d_best_cost && d_best_idx, "null pointer");
    MGB_REQUIRE(k >= 1 && k <= 64, "k must be in [1, 64]");
    MGB_REQUIRE(d_workspace && workspace_bytes >= mgb_score_workspace_bytes(n, k), "workspace too small");
    unsigned char* ws = reinterpret_cast<unsigned char*>(d_workspace);
    const size_t parts = 148 * 16;
    long long* part_idx = reinterpret_cast<long long*>(ws);
    float* part_cost = reinterpret_cast<float*>(ws + parts * sizeof(long long));
    long long n_tiles = (n + MGB_BLOCK - 1) / MGB_BLOCK;
    int grid = (int)((n_tiles < (long long)num_sms() * 8) ? n_tiles : (long long)num_sms() * 8);
    if (grid < 1) grid = 1;
    for (int j = 0; j < k; ++j) {
        mgb_topk_partial<<<grid, MGB_BLOCK, 0, stream>>>(d_cost, n, index_base, j ? d_best_cost + j - 1 : nullptr,
                                                        j ? reinterpret_cast<long long*>(d_best_idx) + j - 1 : nullptr, j ? 1 : 0,
                                                        part_cost, part_idx);
        mgb_topk_final<<<1, 32, 0, stream>>>(part_cost, part_idx, grid, index_base, d_best_cost + j,
                                             reinterpret_cast<long long*>(d_best_idx) + j);
        mgb_count_launch(2);
    }
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}
