// Residual-vector objective and batched Levenberg-Marquardt on the device, sm_100a.
//
// What it replaces in the reference, per seed:
//   MotionPrimitiveConstraints.get_residual_vector              constraints/motion_primitive_constraints.py:124-146
//   obj_spatial_error_residual_vector_and_naturalness           motion_generator/optimization/objective_functions.py:239-267
//   LeastSquares.run -> scipy.optimize.leastsq (MINPACK lmdif)  motion_generator/optimization/least_squares.py:35-64
//   called from _optimize_parameters_numerically                motion_generator/motion_primitive_generator.py:179-190
//
// residual r_i(s) = (error_scale * w_c * e_i(s) + quality_scale * (-log p(s))) / init_error_sum, one e_i per entry of the
// constraint classes' get_residual_vector_spline (see mgb_interp.cuh, RESID); the reference pads with zeros up to len(s),
// which changes neither J^T J nor J^T r and is left out here.
//
// K9  mgb_residuals_kernel : one thread per candidate, residual vectors of a batch -> [n, R]
// K10 mgb_lm_kernel        : one CTA per seed, the whole minimisation in one launch: forward-difference Jacobian with one
//      thread per latent dimension - a perturbation of dimension j < K moves every contracted quad by h * (its basis row j),
//      so a column costs one run of the constraint program and NO contraction; the GMM term of a column comes from the
//      y_g = (s - mu_g) P_g of the base point (|y_g + h P_g[j,:]|^2) -, J^T J and J^T r by the block, damped normal equations
//      (Marquardt scaling) solved by an in-block Cholesky factorisation, accept / reject with the damping adapted.
// MINPACK steps its forward differences by sqrt(eps_f64) |x| ~ 1.5e-8 |x|, below the resolution of a float32 residual; the
// step here is 2^-7 max(|x_j|, 1), and convergence is judged against the float32 noise floor.  The minimiser therefore
// follows a different path than lmdif and is compared with it by the error it reaches (tests), not step by step.
#include <float.h>
#include <math.h>
#include <string.h>

#include <algorithm>

#include "mgb_device.cuh"
#include "mgb_internal.cuh"
#include "mgb_interp.cuh"

namespace {

constexpr int LM_THREADS = 128;

struct LmParams {
    // constraint program (residual form) and contracted quads
    const float4* tab;      // [n_quads][1 + KP]
    const MgbOp* ops;
    const float* cst;
    int n_ops, n_cst, n_quads, K, KP, D, DP;
    int R;                  // residuals
    int RS;                 // row stride of J (>= D, odd)
    // GMM
    const float4* gmm_tab;  // [(g * NQ + q)][1 + DP]
    const float* gmm_c;
    int G, NQ;
    float err_scale, quality_scale;
    float init_error_sum;   // > 0: given; <= 0: max(|sum r(x0)|, 1) per seed (motion_primitive_generator.py:186-187)
    int max_iter;           // Jacobian evaluations
    const float* seeds;     // [n_seeds, D]
    float* out_s;           // [n_seeds, D]
    float* out_err;         // [n_seeds, 4]: sum of weighted residuals at the result (min_error), cost 0.5 |r|^2 at start, at the result, iterations
    int ops_in_smem;
};

// contracted pose channels of quad q at latent vector x (shared memory, D floats)
__device__ __forceinline__ float4 lm_quad_at(const LmParams& p, const float* x, int q) {
    const float4* row = p.tab + (size_t)q * (p.KP + 1);
    float4 acc = __ldg(row);
    for (int k = 0; k < p.K; ++k) {
        const float4 e = __ldg(row + 1 + k);
        const float sk = x[k];
        acc.x = fmaf(sk, e.x, acc.x);
        acc.y = fmaf(sk, e.y, acc.y);
        acc.z = fmaf(sk, e.z, acc.z);
        acc.w = fmaf(sk, e.w, acc.w);
    }
    return acc;
}

// y[g][d] = ((x - mu_g) P_g)[d] for the block (P_g upper triangular: column quad jq needs rows k < 4 jq + 4)
__device__ __forceinline__ void lm_gmm_y(const LmParams& p, const float* x, float* y, int tid, int nthreads) {
    const int items = p.G * p.NQ;
    for (int it = tid; it < items; it += nthreads) {
        const int g = it / p.NQ, jq = it - g * p.NQ;
        const float4* row = p.gmm_tab + (size_t)it * (p.DP + 1);
        float4 acc = __ldg(row);
        const int cnt = min(p.D, 4 * jq + 4);
        for (int k = 0; k < cnt; ++k) {
            const float4 e = __ldg(row + 1 + k);
            const float sk = x[k];
            acc.x = fmaf(sk, e.x, acc.x);
            acc.y = fmaf(sk, e.y, acc.y);
            acc.z = fmaf(sk, e.z, acc.z);
            acc.w = fmaf(sk, e.w, acc.w);
        }
        float* yo = y + (size_t)g * p.DP + 4 * jq;
        yo[0] = acc.x; yo[1] = acc.y; yo[2] = acc.z; yo[3] = acc.w;
    }
}

// -log p(x) from y (one thread); dj >= 0: of x + h e_dj instead, |y_g + h P_g[dj, :]|^2 expanded around y
__device__ __forceinline__ float lm_nll(const LmParams& p, const float* y, int dj, float h) {
    float mx = -INFINITY, ssum = 0.f;
    for (int g = 0; g < p.G; ++g) {
        const float* yg = y + (size_t)g * p.DP;
        float qf = 0.f;
        if (dj < 0) {
            for (int d = 0; d < p.DP; ++d) qf = fmaf(yg[d], yg[d], qf);
        } else {
            for (int jq = 0; jq < p.NQ; ++jq) {
                float4 pr = make_float4(0.f, 0.f, 0.f, 0.f);
                if (4 * jq + 3 >= dj) pr = __ldg(p.gmm_tab + (size_t)(g * p.NQ + jq) * (p.DP + 1) + 1 + dj);   // P_g[dj][4 jq ..], zero left of the diagonal
                const float a0 = fmaf(h, pr.x, yg[4 * jq]), a1 = fmaf(h, pr.y, yg[4 * jq + 1]);
                const float a2 = fmaf(h, pr.z, yg[4 * jq + 2]), a3 = fmaf(h, pr.w, yg[4 * jq + 3]);
                qf = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, fmaf(a3, a3, qf))));
            }
        }
        const float lp = __ldg(p.gmm_c + g) - 0.5f * qf;
        if (lp > mx) {
            ssum = ssum * __expf(mx - lp) + 1.f;
            mx = lp;
        } else {
            ssum += __expf(lp - mx);
        }
    }
    return -(mx + __logf(ssum));
}

}  // namespace

namespace {

// ---------------------------------------------------------------------------
// K9: weighted residual vectors of a batch (MotionPrimitiveConstraints.get_residual_vector, batched)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(LM_THREADS) mgb_residuals_kernel(const LmParams p, const float* __restrict__ s, long long n, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char lm_smem[];
    MgbOp* sm_ops = reinterpret_cast<MgbOp*>(lm_smem);
    float* sm_cst = reinterpret_cast<float*>(lm_smem + (size_t)p.n_ops * sizeof(MgbOp));
    if (p.ops_in_smem) {
        for (int i = threadIdx.x; i < p.n_ops; i += LM_THREADS) sm_ops[i] = p.ops[i];
        for (int i = threadIdx.x; i < p.n_cst; i += LM_THREADS) sm_cst[i] = p.cst[i];
    }
    __syncthreads();
    const MgbOp* ops = p.ops_in_smem ? sm_ops : p.ops;
    const float* cst = p.ops_in_smem ? sm_cst : p.cst;
    for (long long idx = (long long)blockIdx.x * LM_THREADS + threadIdx.x; idx < n; idx += (long long)gridDim.x * LM_THREADS) {
        const float* x = s + (size_t)idx * p.D;
        float* r = out + (size_t)idx * p.R;
        for (int i = 0; i < p.R; ++i) r[i] = 0.f;
        auto quad = [&](int q, float4 (&o)[1]) {
            const float4* row = p.tab + (size_t)q * (p.KP + 1);
            float4 acc = __ldg(row);
            for (int k = 0; k < p.K; ++k) {
                const float4 e = __ldg(row + 1 + k);
                const float sk = __ldg(x + k);
                acc.x = fmaf(sk, e.x, acc.x); acc.y = fmaf(sk, e.y, acc.y); acc.z = fmaf(sk, e.z, acc.z); acc.w = fmaf(sk, e.w, acc.w);
            }
            o[0] = acc;
        };
        auto emit = [&](int, int ridx, float w, float e) { r[ridx] += w * e; };
        MgbInterp<1, true> it;
        it.init();
        it.run(ops, 0, p.n_ops, cst, quad, emit);
    }
}

// ---------------------------------------------------------------------------
// K10: Levenberg-Marquardt, one CTA per seed
// ---------------------------------------------------------------------------
struct LmSmem {
    size_t x, xt, delta, grad, diag, bq, y, r, J, A, L, red, ops, cst, total;
};
__host__ __device__ inline LmSmem lm_smem_plan(int D, int DP, int n_quads, int G, int R, int RS, int n_ops, int n_cst, bool gmm) {
    LmSmem s;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 15) & ~(size_t)15; return o; };
    s.x = take(DP * 4); s.xt = take(DP * 4); s.delta = take(DP * 4); s.grad = take(DP * 4); s.diag = take(DP * 4);
    s.bq = take((size_t)2 * n_quads * 16);
    s.y = take(gmm ? (size_t)2 * G * DP * 4 : 0);
    s.r = take((size_t)2 * R * 4);
    s.J = take((size_t)R * RS * 4);
    s.A = take((size_t)D * D * 4);
    s.L = take((size_t)D * D * 4);
    s.red = take(LM_THREADS * 4 + 64);
    s.ops = take((size_t)n_ops * sizeof(MgbOp));
    s.cst = take((size_t)n_cst * 4);
    s.total = off;
    return s;
}

__device__ __forceinline__ float lm_block_sum(float v, float* red) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    float t = 0.f;
    for (int w = 0; w < LM_THREADS / 32; ++w) t += red[w];
    __syncthreads();
    return t;
}

__global__ void __launch_bounds__(LM_THREADS) mgb_lm_kernel(const LmParams p) {
    extern __shared__ __align__(16) unsigned char lm_smem[];
    const bool gmm = p.quality_scale != 0.f;
    const LmSmem S = lm_smem_plan(p.D, p.DP, p.n_quads, p.G, p.R, p.RS, p.ops_in_smem ? p.n_ops : 0, p.ops_in_smem ? p.n_cst : 0, gmm);
    float* x = reinterpret_cast<float*>(lm_smem + S.x);
    float* xt = reinterpret_cast<float*>(lm_smem + S.xt);
    float* delta = reinterpret_cast<float*>(lm_smem + S.delta);
    float* grad = reinterpret_cast<float*>(lm_smem + S.grad);
    float* diag = reinterpret_cast<float*>(lm_smem + S.diag);   // Marquardt scaling: running maximum of diag(J^T J), as MINPACK keeps it
    float4* bq2 = reinterpret_cast<float4*>(lm_smem + S.bq);
    float* y2 = reinterpret_cast<float*>(lm_smem + S.y);
    float* r2 = reinterpret_cast<float*>(lm_smem + S.r);
    float* J = reinterpret_cast<float*>(lm_smem + S.J);
    float* A = reinterpret_cast<float*>(lm_smem + S.A);
    float* L = reinterpret_cast<float*>(lm_smem + S.L);
    float* red = reinterpret_cast<float*>(lm_smem + S.red);
    float* scal = red + LM_THREADS;   // [0] nll, [1] cost, [2] sum of weighted residuals, [3] flag
    MgbOp* sm_ops = reinterpret_cast<MgbOp*>(lm_smem + S.ops);
    float* sm_cst = reinterpret_cast<float*>(lm_smem + S.cst);
    const int tid = threadIdx.x, D = p.D, R = p.R, RS = p.RS, NQd = p.n_quads;
    if (p.ops_in_smem) {
        for (int i = tid; i < p.n_ops; i += LM_THREADS) sm_ops[i] = p.ops[i];
        for (int i = tid; i < p.n_cst; i += LM_THREADS) sm_cst[i] = p.cst[i];
    }
    const MgbOp* ops = p.ops_in_smem ? sm_ops : p.ops;
    const float* cst = p.ops_in_smem ? sm_cst : p.cst;
    for (int d = tid; d < p.DP; d += LM_THREADS) {
        x[d] = d < D ? p.seeds[(size_t)blockIdx.x * D + d] : 0.f;
        diag[d] = 0.f;
    }
    __syncthreads();
    float norm = 1.f;

    // residuals of the point `pt` (x or xt) into buffer `b`: base quads, y, nll by the block, the program by thread 0.
    // Returns (to every thread) the cost sum r^2; scal[2] = sum of the weighted constraint residuals (min_error).
    auto eval_point = [&](const float* pt, int b, bool set_norm) -> float {
        float4* bq = bq2 + (size_t)b * NQd;
        float* y = y2 + (size_t)b * p.G * p.DP;
        float* r = r2 + (size_t)b * R;
        for (int q = tid; q < NQd; q += LM_THREADS) bq[q] = lm_quad_at(p, pt, q);
        if (gmm) lm_gmm_y(p, pt, y, tid, LM_THREADS);
        for (int i = tid; i < R; i += LM_THREADS) r[i] = 0.f;
        __syncthreads();
        if (tid == 0) {
            auto quad = [&](int q, float4 (&o)[1]) { o[0] = bq[q]; };
            auto emit = [&](int, int ridx, float w, float e) { r[ridx] += w * e; };
            MgbInterp<1, true> it;
            it.init();
            it.run(ops, 0, p.n_ops, cst, quad, emit);
            scal[0] = gmm ? lm_nll(p, y, -1, 0.f) : 0.f;
        }
        __syncthreads();
        const float nll = scal[0];
        float part = 0.f, psum = 0.f;
        for (int i = tid; i < R; i += LM_THREADS) {
            psum += r[i];
            part += fmaf(p.err_scale, r[i], p.quality_scale * nll);
        }
        const float wsum = lm_block_sum(psum, red);
        if (set_norm) {
            const float tot = lm_block_sum(part, red);
            norm = p.init_error_sum > 0.f ? p.init_error_sum : fmaxf(fabsf(tot), 1.f);
        }
        float c = 0.f;
        for (int i = tid; i < R; i += LM_THREADS) {
            const float v = fmaf(p.err_scale, r[i], p.quality_scale * nll) / norm;
            r[i] = v;
            c = fmaf(v, v, c);
        }
        c = lm_block_sum(c, red);
        if (tid == 0) scal[2] = wsum;
        __syncthreads();
        return c;
    };

    int cur = 0;
    float cost = eval_point(x, cur, true);
    const float cost_start = cost;
    float wres = scal[2];
    float lam = 1e-3f;
    int iters = 0, small = 0;
    bool done = !(cost > 0.f) || !isfinite(cost);
    while (!done && iters < p.max_iter) {
        ++iters;
        const float4* bq = bq2 + (size_t)cur * NQd;
        const float* y = y2 + (size_t)cur * p.G * p.DP;
        const float* r0 = r2 + (size_t)cur * R;
        // ---- Jacobian: thread j owns column j ------------------------------------------------------------------------------
        for (int j = tid; j < D; j += LM_THREADS) {
            const float h = 0.0078125f * fmaxf(fabsf(x[j]), 1.f);
            for (int i = 0; i < R; ++i) J[i * RS + j] = 0.f;
            if (j < p.K) {
                auto quad = [&](int q, float4 (&o)[1]) {
                    const float4 e = __ldg(p.tab + (size_t)q * (p.KP + 1) + 1 + j);
                    const float4 b = bq[q];
                    o[0] = make_float4(fmaf(h, e.x, b.x), fmaf(h, e.y, b.y), fmaf(h, e.z, b.z), fmaf(h, e.w, b.w));
                };
                auto emit = [&](int, int ridx, float w, float e) { J[ridx * RS + j] += w * e; };
                MgbInterp<1, true> it;
                it.init();
                it.run(ops, 0, p.n_ops, cst, quad, emit);
            }
            const float nll_j = gmm ? lm_nll(p, y, j, h) : 0.f;
            const float nll0 = scal[0];
            const float hinv = 1.f / h;
            for (int i = 0; i < R; ++i) {
                // columns j >= K (time parameters) move no pose channel: their constraint part is that of the base point
                const float a = (j < p.K) ? J[i * RS + j] : (r0[i] * norm - p.quality_scale * nll0) / p.err_scale;   // err_scale != 0 (checked by the launcher)
                const float rj = fmaf(p.err_scale, a, p.quality_scale * nll_j) / norm;
                J[i * RS + j] = (rj - r0[i]) * hinv;
            }
        }
        __syncthreads();
        // ---- A = J^T J, grad = J^T r ------------------------------------------------------------------------------------------
        for (int e = tid; e < D * D; e += LM_THREADS) {
            const int a = e / D, b = e - a * D;
            if (b > a) continue;
            float acc = 0.f;
            for (int i = 0; i < R; ++i) acc = fmaf(J[i * RS + a], J[i * RS + b], acc);
            A[a * D + b] = acc;
            A[b * D + a] = acc;
        }
        for (int a = tid; a < D; a += LM_THREADS) {
            float acc = 0.f;
            for (int i = 0; i < R; ++i) acc = fmaf(J[i * RS + a], r0[i], acc);
            grad[a] = acc;
        }
        __syncthreads();
        for (int a = tid; a < D; a += LM_THREADS) diag[a] = fmaxf(diag[a], A[a * D + a]);
        __syncthreads();
        float dmax = 0.f;
        for (int a = 0; a < D; ++a) dmax = fmaxf(dmax, diag[a]);
        bool accepted = false;
        float rel = 0.f;
        float nu = 2.f;
        for (int tries = 0; tries < 24 && !accepted; ++tries) {
            // ---- (A + lam diag) delta = -grad by Cholesky ------------------------------------------------------------------
            for (int e = tid; e < D * D; e += LM_THREADS) {
                const int a = e / D, b = e - a * D;
                L[e] = A[e] + ((a == b) ? lam * fmaxf(diag[a], 1e-6f * dmax) + 1e-20f : 0.f);
            }
            __syncthreads();
            for (int k = 0; k < D; ++k) {
                if (tid == 0) L[k * D + k] = sqrtf(fmaxf(L[k * D + k], 1e-30f));
                __syncthreads();
                const float dk = 1.f / L[k * D + k];
                for (int i = k + 1 + tid; i < D; i += LM_THREADS) L[i * D + k] *= dk;
                __syncthreads();
                for (int i = k + 1 + tid; i < D; i += LM_THREADS) {
                    const float lik = L[i * D + k];
                    for (int j = k + 1; j <= i; ++j) L[i * D + j] = fmaf(-lik, L[j * D + k], L[i * D + j]);
                }
                __syncthreads();
            }
            if (tid == 0) {
                for (int i = 0; i < D; ++i) {          // L z = -grad
                    float v = -grad[i];
                    for (int j = 0; j < i; ++j) v = fmaf(-L[i * D + j], delta[j], v);
                    delta[i] = v / L[i * D + i];
                }
                for (int i = D - 1; i >= 0; --i) {     // L^T delta = z
                    float v = delta[i];
                    for (int j = i + 1; j < D; ++j) v = fmaf(-L[j * D + i], delta[j], v);
                    delta[i] = v / L[i * D + i];
                }
            }
            __syncthreads();
            float pr = 0.f;   // predicted reduction of |r|^2: delta . (lam diag delta - grad)
            for (int d = tid; d < p.DP; d += LM_THREADS) {
                xt[d] = d < D ? x[d] + delta[d] : 0.f;
                if (d < D) pr = fmaf(delta[d], lam * fmaxf(diag[d], 1e-6f * dmax) * delta[d] - grad[d], pr);
            }
            pr = lm_block_sum(pr, red);
            const float ct = eval_point(xt, cur ^ 1, false);
            if (isfinite(ct) && ct < cost) {
                accepted = true;
                rel = (cost - ct) / cost;
                const float rho = (cost - ct) / fmaxf(pr, 1e-30f), t = 2.f * rho - 1.f;
                lam = fmaxf(lam * fmaxf(1.f / 3.f, 1.f - t * t * t), 1e-12f);   // gain-ratio update (Nielsen)
                cost = ct;
                wres = scal[2];
                cur ^= 1;
                for (int d = tid; d < p.DP; d += LM_THREADS) x[d] = xt[d];
                __syncthreads();
            } else {
                lam *= nu;
                nu *= 2.f;
                if (tid == 0 && gmm) scal[0] = lm_nll(p, y2 + (size_t)cur * p.G * p.DP, -1, 0.f);   // nll of the kept point (eval_point overwrote it)
                __syncthreads();
                if (lam > 1e12f) break;
            }
        }
        small = (accepted && rel < 3e-7f) ? small + 1 : 0;   // float32 residuals: a relative change of |r|^2 below ~1e-7 is rounding
        if (!accepted || small >= 3 || cost < 1e-12f) done = true;
    }
    for (int d = tid; d < D; d += LM_THREADS) p.out_s[(size_t)blockIdx.x * D + d] = x[d];
    if (tid == 0) {
        float* o = p.out_err + (size_t)blockIdx.x * 4;
        o[0] = wres;
        o[1] = 0.5f * cost_start * norm * norm;
        o[2] = 0.5f * cost * norm * norm;
        o[3] = (float)iters;
    }
}

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
static int lm_fill(const mgb_model* m, const mgb_constraints* c, float err_scale, float quality_scale, LmParams& p) {
    MGB_REQUIRE(m && c && c->model == m, "constraints were packed for a different model");
    MGB_REQUIRE(c->n_resid > 0 && c->d_ops_resid, "the constraint set has no residuals");
    memset(&p, 0, sizeof(p));
    p.tab = c->d_tab;
    p.ops = c->d_ops_resid;
    p.cst = c->d_cst;
    p.n_ops = c->n_ops;
    p.n_cst = c->n_cst;
    p.n_quads = c->n_quads;
    p.K = m->K;
    p.KP = c->KP;
    p.D = m->D;
    p.DP = m->DP;
    p.R = c->n_resid;
    p.RS = m->D | 1;
    p.gmm_tab = m->d_gmm_tab;
    p.gmm_c = m->d_gmm_c;
    p.G = m->G;
    p.NQ = m->NQ;
    p.err_scale = err_scale;
    p.quality_scale = quality_scale;
    return MGB_OK;
}

extern "C" int32_t mgb_constraints_n_residuals(const mgb_constraints* c) { return c ? c->n_resid : 0; }

extern "C" int mgb_residuals(const mgb_model* m, const mgb_constraints* c, const float* d_s, int64_t n, float* d_out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    LmParams p;
    if (int rc = lm_fill(m, c, 1.f, 0.f, p)) return rc;
    MGB_REQUIRE(n >= 0 && (n == 0 || (d_s && d_out)), "null argument");
    if (n == 0) return MGB_OK;
    size_t smem = (size_t)p.n_ops * sizeof(MgbOp) + (size_t)p.n_cst * 4;
    p.ops_in_smem = smem <= 40 * 1024 ? 1 : 0;
    if (!p.ops_in_smem) smem = 0;
    if (int rc = mgb_constraints_before_launch(c, stream)) return rc;
    const int grid = (int)std::min<int64_t>((n + LM_THREADS - 1) / LM_THREADS, (int64_t)mgb_num_sms() * 8);
    mgb_residuals_kernel<<<grid, LM_THREADS, smem, stream>>>(p, d_s, n, d_out);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return mgb_constraints_after_launch(c, stream);
}

extern "C" int mgb_lm_refine(const mgb_model* m, const mgb_constraints* c, const float* d_seeds, int32_t n_seeds, float err_scale,
                             float quality_scale, float init_error_sum, int32_t max_iter, float* d_out_s, float* d_out_err, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    LmParams p;
    if (int rc = lm_fill(m, c, err_scale, quality_scale, p)) return rc;
    MGB_REQUIRE(n_seeds >= 0 && (n_seeds == 0 || (d_seeds && d_out_s && d_out_err)), "null argument");
    MGB_REQUIRE(err_scale != 0.f, "error_scale must not be zero");
    MGB_REQUIRE(max_iter > 0, "max_iter must be positive");
    if (n_seeds == 0) return MGB_OK;
    p.init_error_sum = init_error_sum;
    p.max_iter = max_iter;
    p.seeds = d_seeds;
    p.out_s = d_out_s;
    p.out_err = d_out_err;
    const bool gmm = quality_scale != 0.f;
    const size_t limit = 200 * 1024;
    p.ops_in_smem = 1;
    size_t smem = lm_smem_plan(p.D, p.DP, p.n_quads, p.G, p.R, p.RS, p.n_ops, p.n_cst, gmm).total;
    if (smem > limit) {
        p.ops_in_smem = 0;
        smem = lm_smem_plan(p.D, p.DP, p.n_quads, p.G, p.R, p.RS, 0, 0, gmm).total;
    }
    if (smem > limit) {
        mgb_set_error("Levenberg-Marquardt working set (%zu bytes: %d residuals x %d dimensions) exceeds shared memory", smem, p.R, p.D);
        return MGB_E_UNSUPPORTED;
    }
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(mgb_lm_kernel, configured, (int)limit)) return rc;
    if (int rc = mgb_constraints_before_launch(c, stream)) return rc;
    mgb_lm_kernel<<<n_seeds, LM_THREADS, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return mgb_constraints_after_launch(c, stream);
}
