// K6 / K5a / fused K1+K6: keyframe score kernels (sm_100a).
//
// One thread scores CPT candidates.  The constraint list was compiled by
// mgb_constraints_pack into (a) a table of "quads": for every constrained time
// slot and every needed group of 4 pose channels (root translation, one joint
// quaternion) the spatial basis pre-contracted with that slot's 4 B-spline
// weights, laid out [quad][1 + KP] float4 (bias row first, K rows zero-padded to
// a multiple of 4) so one 16-byte broadcast load feeds 4 FMAs per candidate, and
// (b) a small op program: a depth-first walk over the needed part of the
// skeleton (forward kinematics in quaternion form, 8 floats of state) with the
// constraint terms attached to the nodes they read.
//
// What it replaces in the reference, per candidate:
//   GaussianMixture.sample (MODE_SAMPLE)          motion_primitive.py:182-189
//   MotionPrimitive.back_project_spatial_coeffs   motion_primitive.py:236-256
//   MotionSpline.evaluate (P splev calls)         motion_spline.py:89-92
//   anim_utils SkeletonNode.get_global_matrix     (call sites global_transform_constraint.py:103 ...)
//   constraint.evaluate_motion_spline, weighted sum   motion_primitive_constraints.py:118-122
//   naturalness -log p(s)                         objective_functions.py:256
//   + 2|s|                                        mg_state_planner.py:110-129
//   first-minimum arg-min                         motion_primitive_generator.py:251-257
//
// Memory: MODE_S reads the candidate tile (a contiguous run of TILE*D elements)
// with ONE 1-D TMA bulk copy per tile into shared memory, double-buffered on
// mbarriers so the next tile lands while the current one is scored; MODE_SAMPLE
// draws the latent vector in registers (Philox keyed by the global index) and
// touches no HBM at all; quad table, GMM table, op program and constants are
// staged in shared memory once per persistent CTA; block arg-min by warp
// shuffles, last-arriving CTA folds the per-CTA partials.
#include <float.h>
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <type_traits>

#include "mgb_device.cuh"
#include "mgb_internal.cuh"
#include "mgb_interp.cuh"

#define MGB_BLOCK 256

enum { MODE_S = 0, MODE_FRAMES = 1, MODE_SAMPLE = 2 };

namespace {

__device__ __forceinline__ bool better(float c1, long long i1, float c2, long long i2) {
    return (c1 < c2) || (c1 == c2 && i1 < i2);
}

// acc[c] = row[0] + sum_{k < 4 * chunks} s[c][k] * row[1 + k]; s in registers
template <int NREG, int CPT>
__device__ __forceinline__ void dot_rows_reg(const float4* __restrict__ row, int chunks, const float (&s)[CPT][NREG > 0 ? NREG : 1],
                                             float4 (&acc)[CPT]) {
    const float4 b = row[0];
#pragma unroll
    for (int c = 0; c < CPT; ++c) acc[c] = b;
#pragma unroll
    for (int j = 0; j < NREG / 4; ++j) {
        if (j >= chunks) break;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const float4 e = row[1 + 4 * j + kk];
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const float sk = s[c][4 * j + kk];
                acc[c].x = fmaf(sk, e.x, acc[c].x);
                acc[c].y = fmaf(sk, e.y, acc[c].y);
                acc[c].z = fmaf(sk, e.z, acc[c].z);
                acc[c].w = fmaf(sk, e.w, acc[c].w);
            }
        }
    }
}

// same with s read from memory (candidate row pointers), any D; rows beyond D are zero in the table
template <int CPT, typename ST>
__device__ __forceinline__ void dot_rows_mem(const float4* __restrict__ row, int cnt, const ST* (&srow)[CPT], float4 (&acc)[CPT]) {
    const float4 b = row[0];
#pragma unroll
    for (int c = 0; c < CPT; ++c) acc[c] = b;
    for (int k = 0; k < cnt; ++k) {
        const float4 e = row[1 + k];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            const float sk = (float)srow[c][k];
            acc[c].x = fmaf(sk, e.x, acc[c].x);
            acc[c].y = fmaf(sk, e.y, acc[c].y);
            acc[c].z = fmaf(sk, e.z, acc[c].z);
            acc[c].w = fmaf(sk, e.w, acc[c].w);
        }
    }
}

// ---------------------------------------------------------------------------
// NREG > 0 : the whole latent vector lives in registers (D <= NREG)
// NREG == 0: s is read from the shared-memory tile (any D)
// CPT      : candidates per thread (rows tid, tid + MGB_BLOCK, ... of the tile)
// MODE     : MODE_S (candidates from global memory), MODE_FRAMES (pose channels from
//            materialised frames, staged pipeline K5a), MODE_SAMPLE (fused K1: candidates drawn in-kernel)
// ---------------------------------------------------------------------------
// cta / n_cta: this CTA's position among the CTAs that share the job `p` (the whole grid for a single job; a slice of the
// grid when several jobs - the outgoing options of a planner step - run in one launch)
template <int NREG, int CPT, int MODE, typename SIn>
__device__ __forceinline__ void score_body(const ScoreParams& p, const int cta, const int n_cta, unsigned char* smem_raw) {
    constexpr int TILE = MGB_BLOCK * CPT;
    constexpr int NR = NREG > 0 ? NREG : 1;
    // element type of the shared-memory tile: large latent vectors (NREG == 0) are kept as float
    // (float64 input is converted while it is staged), small ones stay raw and are converted on the way to registers
    using TIn = typename std::conditional<NREG == 0, float, SIn>::type;
    constexpr bool RAW = sizeof(TIn) == sizeof(SIn);
    const int tid = threadIdx.x;
    const int D = p.D, KP = p.KP, DP = p.DP;
    const int SROW = D | 1;  // MODE_SAMPLE with NREG == 0: per-thread rows, odd stride

    // ---- carve shared memory ------------------------------------------------
    size_t off = 0;
    float4* sm_tab = reinterpret_cast<float4*>(smem_raw + off);
    if (p.tab_in_smem) off += (size_t)p.n_quads * (KP + 1) * sizeof(float4);
    float4* sm_gmm = reinterpret_cast<float4*>(smem_raw + off);
    if (p.gmm_in_smem) off += (size_t)p.G * p.NQ * (DP + 1) * sizeof(float4);
    float* sm_chol = reinterpret_cast<float*>(smem_raw + off);
    if (MODE == MODE_SAMPLE && p.chol_in_smem) off += ((size_t)p.G * D * p.LS + (((size_t)p.G * D + 3) & ~(size_t)3)) * sizeof(float);
    float* sm_mu = sm_chol + (size_t)p.G * D * p.LS;
    MgbOp* sm_ops = reinterpret_cast<MgbOp*>(smem_raw + off);
    off += (size_t)p.n_ops * sizeof(MgbOp);
    float* sm_cst = reinterpret_cast<float*>(smem_raw + off);
    off += (size_t)((p.n_cst + 3) & ~3) * sizeof(float);
    float* sm_redc = reinterpret_cast<float*>(smem_raw + off);
    off += 32 * sizeof(float);
    long long* sm_redi = reinterpret_cast<long long*>(smem_raw + off);
    off += 32 * sizeof(long long);
    unsigned long long* sm_bar = reinterpret_cast<unsigned long long*>(smem_raw + off);
    off += 2 * sizeof(unsigned long long);
    off = (off + 127) & ~(size_t)127;
    unsigned char* sm_tile = smem_raw + off;  // nbuf raw tiles [TILE][D] of SIn (MODE_S / MODE_FRAMES), or [TILE][SROW] floats (MODE_SAMPLE, NREG == 0)
    const size_t tile_bytes = (MODE == MODE_SAMPLE) ? (size_t)TILE * SROW * sizeof(float) : (size_t)TILE * D * sizeof(TIn);

    // ---- one-time staging of program + tables -------------------------------
    if (p.tab_in_smem) {
        const int n4 = p.n_quads * (KP + 1);
        for (int i = tid; i < n4; i += MGB_BLOCK) sm_tab[i] = p.tab[i];
    }
    if (p.gmm_in_smem) {
        const int n4 = p.G * p.NQ * (DP + 1);
        for (int i = tid; i < n4; i += MGB_BLOCK) sm_gmm[i] = p.gmm_tab[i];
    }
    if (MODE == MODE_SAMPLE && p.chol_in_smem) {
        const int nl = p.G * D * p.LS;
        for (int i = tid; i < nl; i += MGB_BLOCK) sm_chol[i] = p.chol[i];
        for (int i = tid; i < p.G * D; i += MGB_BLOCK) sm_mu[i] = p.means[i];
    }
    for (int i = tid; i < p.n_ops; i += MGB_BLOCK) sm_ops[i] = p.ops[i];
    for (int i = tid; i < p.n_cst; i += MGB_BLOCK) sm_cst[i] = p.cst[i];
    const float4* tab = p.tab_in_smem ? sm_tab : p.tab;
    const float4* gtab = p.gmm_in_smem ? sm_gmm : p.gmm_tab;
    const float* chol = (MODE == MODE_SAMPLE && p.chol_in_smem) ? sm_chol : p.chol;
    const float* mus = (MODE == MODE_SAMPLE && p.chol_in_smem) ? sm_mu : p.means;
    const bool load_s = (MODE == MODE_S) || (MODE == MODE_FRAMES && p.s != nullptr);
    const bool use_s = load_s || MODE == MODE_SAMPLE;
    const bool use_tma = RAW && load_s && p.s_aligned;
    if (use_tma && tid == 0) {
        mgb_mbar_init(&sm_bar[0], 1);
        mgb_mbar_init(&sm_bar[1], 1);
        mgb_fence_barrier_init();
    }
    __syncthreads();

    float best_c = INFINITY;
    long long best_i = LLONG_MAX;
    const long long n_tiles = (p.n + TILE - 1) / TILE;
    const SIn* __restrict__ s_glob = reinterpret_cast<const SIn*>(p.s);
    auto tile_by_tma = [&](long long tile) -> bool { return use_tma && (tile + 1) * TILE <= p.n; };
    unsigned int phase[2] = {0u, 0u};
    if (use_tma && tid == 0 && cta < n_tiles && tile_by_tma(cta))
        mgb_tma_load_1d(sm_tile, s_glob + (size_t)cta * TILE * D, (unsigned int)tile_bytes, &sm_bar[0]);

    int it = 0;
    int g_hint = 0;
    for (long long tile = cta; tile < n_tiles; tile += n_cta, ++it) {
        const long long base = tile * TILE;
        const int buf = (p.nbuf == 2) ? (it & 1) : 0;
        const TIn* sm_s = reinterpret_cast<const TIn*>(sm_tile + (size_t)buf * tile_bytes);
        if (load_s) {
            if (p.nbuf == 2) {
                const long long next = tile + n_cta;
                if (tid == 0 && next < n_tiles && tile_by_tma(next))
                    mgb_tma_load_1d(sm_tile + (size_t)(buf ^ 1) * tile_bytes, s_glob + (size_t)next * TILE * D, (unsigned int)tile_bytes,
                                    &sm_bar[buf ^ 1]);
            } else if (it > 0 && tid == 0 && tile_by_tma(tile)) {
                mgb_tma_load_1d(sm_tile, s_glob + (size_t)tile * TILE * D, (unsigned int)tile_bytes, &sm_bar[0]);
            }
            if (tile_by_tma(tile)) {
                mgb_mbar_wait(&sm_bar[buf], phase[buf]);
                phase[buf] ^= 1u;
            } else {
                // ragged / unaligned tile: cooperative coalesced copy, zero fill past the end
                TIn* dst = const_cast<TIn*>(sm_s);
                const SIn* src = s_glob + (size_t)base * D;
                const long long remaining = (p.n - base) * (long long)D;
                for (int e = tid; e < TILE * D; e += MGB_BLOCK) dst[e] = (e < remaining) ? (TIn)__ldg(src + e) : (TIn)0;
                __syncthreads();
            }
        }

        // ---- candidates of this thread ------------------------------------------
        long long idx[CPT];
        bool valid[CPT];
        float s[CPT][NR];
        int g_first = 0;                       // CPT == 2, sampling: component / global index of the thread's first candidate
        unsigned long long gi_first = 0;
        const TIn* srow[CPT];
        const float* frow[CPT];
        float* myrow[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            const int r = c * MGB_BLOCK + tid;
            idx[c] = base + r;
            valid[c] = idx[c] < p.n;
            srow[c] = sm_s + (size_t)r * D;
            frow[c] = (MODE == MODE_FRAMES) ? (p.frames + (size_t)(valid[c] ? idx[c] : 0) * p.n_times * p.P) : nullptr;
            myrow[c] = reinterpret_cast<float*>(sm_tile) + (size_t)r * SROW;
            if (MODE == MODE_SAMPLE) {
                const unsigned long long gi = (unsigned long long)(p.index_base + idx[c]);
                // mixture component of the global index (samples are stacked in component order): the search
                // resumes from the component of the thread's previous candidate - indices only grow
                while (g_hint + 1 < p.G && (long long)gi >= p.prefix.v[g_hint + 1]) ++g_hint;
                const int g = g_hint;
                const uint2 key = make_uint2((unsigned int)p.seed, (unsigned int)(p.seed >> 32));
                if (NREG > 0) {
                    // without a density or magnitude term only the K spatial entries of s are ever read: rows >= K of the
                    // Cholesky product (and the normals only they need) are dead - the entries that ARE computed keep their bits
                    const int d_need = (p.quality_scale == 0.f && p.magnitude_scale == 0.f) ? p.K : D;
                    if constexpr (CPT == 2) {
                        // both candidates of the thread are drawn together when they share a component (all but the tiles that
                        // straddle a component boundary): one load of every Cholesky row for the two
                        if (c == 0) {
                            g_first = g;
                        } else if (g == g_first) {
                            mgb_sample_rows_reg2<NR>(reinterpret_cast<const float4*>(chol + (size_t)g * D * p.LS), p.LS, mus + (size_t)g * D, d_need,
                                                     gi_first, gi, key, s[0], s[1]);
                        } else {
                            mgb_sample_row_reg<NR>(reinterpret_cast<const float4*>(chol + (size_t)g_first * D * p.LS), p.LS, mus + (size_t)g_first * D,
                                                   d_need, gi_first, key, s[0]);
                            mgb_sample_row_reg<NR>(reinterpret_cast<const float4*>(chol + (size_t)g * D * p.LS), p.LS, mus + (size_t)g * D, d_need, gi, key, s[1]);
                        }
                        gi_first = gi;
                    } else {
                        mgb_sample_row_reg<NR>(reinterpret_cast<const float4*>(chol + (size_t)g * D * p.LS), p.LS, mus + (size_t)g * D, d_need, gi, key, s[c]);
                    }
                } else {
                    mgb_sample_row_mem(chol + (size_t)g * D * p.LS, p.LS, mus + (size_t)g * D, D, gi, key, myrow[c]);
                }
            } else if (NREG > 0) {
#pragma unroll
                for (int k = 0; k < NR; ++k) s[c][k] = (k < D && use_s) ? (float)srow[c][k] : 0.f;
            }
        }

        // ---- constraint program ------------------------------------------------
        float total[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) total[c] = 0.f;
        auto quad = [&](int q, float4 (&out)[CPT]) {
            if constexpr (MODE == MODE_FRAMES) {
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    const float* f = frow[c] + q;
                    out[c] = make_float4(__ldg(f), __ldg(f + 1), __ldg(f + 2), __ldg(f + 3));
                }
            } else if constexpr (NREG > 0) {
                dot_rows_reg<NREG, CPT>(tab + (size_t)q * (KP + 1), KP >> 2, s, out);
            } else if constexpr (MODE == MODE_SAMPLE) {
                const float* rows[CPT];
#pragma unroll
                for (int c = 0; c < CPT; ++c) rows[c] = myrow[c];
                dot_rows_mem<CPT, float>(tab + (size_t)q * (KP + 1), p.K, rows, out);
            } else {
                dot_rows_mem<CPT, TIn>(tab + (size_t)q * (KP + 1), p.K, srow, out);
            }
        };
        auto emit = [&](int c, int term, float w, float e) {
            total[c] = fmaf(w, e, total[c]);
            if (p.terms != nullptr && valid[c]) p.terms[(size_t)term * p.n + idx[c]] += e;
        };

        mgb_run_program<CPT>(sm_ops, p.n_ops, sm_cst, quad, emit);

        float cost[CPT];
#pragma unroll
        for (int c = 0; c < CPT; ++c) cost[c] = p.err_scale * total[c];
        if (p.cost_base != nullptr) {
#pragma unroll
            for (int c = 0; c < CPT; ++c)
                if (valid[c]) cost[c] += p.cost_base[idx[c]];
        }
        if (p.quality_scale != 0.f) {
            // GMM log-density from the precision Cholesky factors; column quad jq of the
            // upper-triangular factor only needs rows k < 4 jq + 4
            float mx[CPT], ssum[CPT];
#pragma unroll
            for (int c = 0; c < CPT; ++c) { mx[c] = -INFINITY; ssum[c] = 0.f; }
            for (int g = 0; g < p.G; ++g) {
                float qf[CPT];
#pragma unroll
                for (int c = 0; c < CPT; ++c) qf[c] = 0.f;
                for (int jq = 0; jq < p.NQ; ++jq) {
                    const float4* row = gtab + (size_t)(g * p.NQ + jq) * (DP + 1);
                    float4 y[CPT];
                    if constexpr (NREG > 0) {
                        dot_rows_reg<NREG, CPT>(row, jq + 1, s, y);
                    } else if constexpr (MODE == MODE_SAMPLE) {
                        const float* rows[CPT];
#pragma unroll
                        for (int c = 0; c < CPT; ++c) rows[c] = myrow[c];
                        dot_rows_mem<CPT, float>(row, min(D, 4 * jq + 4), rows, y);
                    } else {
                        dot_rows_mem<CPT, TIn>(row, min(D, 4 * jq + 4), srow, y);
                    }
#pragma unroll
                    for (int c = 0; c < CPT; ++c)
                        qf[c] = fmaf(y[c].x, y[c].x, fmaf(y[c].y, y[c].y, fmaf(y[c].z, y[c].z, fmaf(y[c].w, y[c].w, qf[c]))));
                }
                const float gc = p.gmm_c[g];
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    const float lp = gc - 0.5f * qf[c];
                    if (lp > mx[c]) {
                        ssum[c] = ssum[c] * __expf(mx[c] - lp) + 1.f;
                        mx[c] = lp;
                    } else {
                        ssum[c] += __expf(lp - mx[c]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < CPT; ++c) cost[c] = fmaf(p.quality_scale, -(mx[c] + __logf(ssum[c])), cost[c]);
        }
        if (p.magnitude_scale != 0.f) {
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                float m2 = 0.f;
                if constexpr (NREG > 0) {
#pragma unroll
                    for (int k = 0; k < NR; ++k) m2 = fmaf(s[c][k], s[c][k], m2);
                } else if constexpr (MODE == MODE_SAMPLE) {
                    for (int k = 0; k < D; ++k) m2 = fmaf(myrow[c][k], myrow[c][k], m2);
                } else {
                    for (int k = 0; k < D; ++k) { const float v = (float)srow[c][k]; m2 = fmaf(v, v, m2); }
                }
                cost[c] = fmaf(p.magnitude_scale, sqrtf(m2), cost[c]);
            }
        }
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            if (valid[c]) {
                if (p.cost) p.cost[idx[c]] = cost[c];
                const long long gi = p.index_base + idx[c];
                if (better(cost[c], gi, best_c, best_i)) { best_c = cost[c]; best_i = gi; }
            }
        }
        if (load_s) {
            __syncthreads();                       // every thread is done reading this tile buffer ...
            if (tid == 0) mgb_fence_proxy_async();  // ... before the async proxy (TMA) may overwrite it
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
        p.part_cost[cta] = best_c;
        p.part_idx[cta] = best_i;
        __threadfence();
        unsigned int t = atomicAdd(p.done_count, 1u);
        is_last = (t == (unsigned int)n_cta - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    best_c = INFINITY;
    best_i = LLONG_MAX;
    for (int b = tid; b < n_cta; b += MGB_BLOCK) {
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
    if (tid < 32) {
        // warp 0: fold the warp results, then (world > 1) exchange this rank's best with the peers over NVLink - in this CTA's
        // epilogue, no second launch
        if (tid == 0) {
            for (int w = 1; w < MGB_BLOCK / 32; ++w)
                if (better(sm_redc[w], sm_redi[w], best_c, best_i)) { best_c = sm_redc[w]; best_i = sm_redi[w]; }
            if (best_i == LLONG_MAX) best_i = p.index_base;  // nothing finite: reference keeps index 0
        }
        best_c = __shfl_sync(0xffffffffu, best_c, 0);
        best_i = __shfl_sync(0xffffffffu, best_i, 0);
        if (p.xch.peer_slots != nullptr) {
            float gc;
            long long gi;
            mgb_exchange_warp(p.xch, best_c, best_i, gc, gi);
            best_c = gc;
            best_i = gi;
        }
    }
    if (tid == 0) {
        p.best_cost[0] = best_c;
        p.best_idx[0] = best_i;
        *p.done_count = 0u;  // ready for the next launch
        if (MODE == MODE_SAMPLE && p.best_row != nullptr) {
            // the winning latent vector, re-drawn from its global index (what mgb_sample writes for that index, bit for bit)
            int g = 0;
            while (g + 1 < p.G && best_i >= p.prefix.v[g + 1]) ++g;
            const uint2 key = make_uint2((unsigned int)p.seed, (unsigned int)(p.seed >> 32));
            if constexpr (NREG > 0) {
                float row[NR];
                mgb_sample_row_reg<NR>(reinterpret_cast<const float4*>(p.chol + (size_t)g * D * p.LS), p.LS, p.means + (size_t)g * D, D,
                                       (unsigned long long)best_i, key, row);
#pragma unroll
                for (int k = 0; k < NR; ++k)
                    if (k < D) p.best_row[k] = row[k];
            } else {
                mgb_sample_row_mem(p.chol + (size_t)g * D * p.LS, p.LS, p.means + (size_t)g * D, D, (unsigned long long)best_i, key, p.best_row);
            }
        }
    }
}

// 3 resident CTAs per SM (<= 85 registers): measured on B200 at S1M - uncapped the compiler takes more registers and the step
// is 93 us, capped to 3 CTAs 65 us, to 4 CTAs (<= 64 registers) 72 us (profiles/r02_o)
#ifndef MGB_SCORE_MIN_BLOCKS
#define MGB_SCORE_MIN_BLOCKS 3
#endif
template <int NREG, int CPT, int MODE, typename SIn>
__global__ void __launch_bounds__(MGB_BLOCK, MGB_SCORE_MIN_BLOCKS) mgb_score_kernel(const __grid_constant__ ScoreParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    score_body<NREG, CPT, MODE, SIn>(p, (int)blockIdx.x, (int)gridDim.x, smem_raw);
}

// Several independent sample-and-score jobs in ONE launch: the outgoing options of a planner step (reference
// graph_walk_planner.py:216-226, mg_state_planner.py:946-968 score them one after another).  jobs[j] is a complete
// ScoreParams (own model tables, constraint program, workspace slice, results); CTAs [first[j], first[j + 1]) work on job j.
struct MultiParams {
    const ScoreParams* jobs;
    const int* first;   // [n_jobs + 1]
    int n_jobs;
};

template <int NREG, int CPT>
__global__ void __launch_bounds__(MGB_BLOCK) mgb_score_multi_kernel(const MultiParams mp) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ ScoreParams sp;
    __shared__ int s_job, s_first, s_cnt;
    if (threadIdx.x == 0) {
        int j = 0;
        while (j + 1 < mp.n_jobs && (int)blockIdx.x >= __ldg(mp.first + j + 1)) ++j;
        s_job = j;
        s_first = __ldg(mp.first + j);
        s_cnt = __ldg(mp.first + j + 1) - s_first;
    }
    __syncthreads();
    {
        const unsigned int* src = reinterpret_cast<const unsigned int*>(mp.jobs + s_job);
        unsigned int* dst = reinterpret_cast<unsigned int*>(&sp);
        for (int i = threadIdx.x; i < (int)(sizeof(ScoreParams) / 4); i += MGB_BLOCK) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    score_body<NREG, CPT, MODE_SAMPLE, float>(sp, (int)blockIdx.x - s_first, s_cnt, smem_raw);
}

// ---------------------------------------------------------------------------
// K5c: k smallest of a cost array (k passes; k is 1..64 in every caller)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(MGB_BLOCK) mgb_topk_partial(const float* __restrict__ cost, long long n, long long index_base,
                                                             const float* __restrict__ prev_cost, const long long* __restrict__ prev_idx,
                                                             int have_prev, float* part_cost, long long* part_idx,
                                                             const int* __restrict__ exhausted) {
    // candidates strictly "after" the previous pick in (cost, index) order; nothing once a pass has found no entry
    float pc = have_prev ? prev_cost[0] : -INFINITY;
    long long pi = have_prev ? prev_idx[0] : -1;
    float best_c = INFINITY;
    long long best_i = LLONG_MAX;
    const long long n_eff = (have_prev && *exhausted) ? 0 : n;
    for (long long i = (long long)blockIdx.x * MGB_BLOCK + threadIdx.x; i < n_eff; i += (long long)gridDim.x * MGB_BLOCK) {
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
                               float* out_cost, long long* out_idx, int* exhausted) {
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
        if (best_i == LLONG_MAX) {
            best_i = index_base;
            *exhausted = 1;      // fewer selectable entries than picks: the remaining slots are (inf, index_base)
        }
        out_cost[0] = best_c;
        out_idx[0] = best_i;
    }
}

int num_sms() { return mgb_num_sms(); }

constexpr size_t MGB_SMEM_MAX = 220 * 1024;

struct SmemPlan {
    size_t fixed;  // tables + program + reduction scratch (everything before the tile buffers)
};

template <int NREG, int CPT, int MODE, typename SIn>
int launch_variant(ScoreParams& p, const SmemPlan& plan, cudaStream_t stream) {
    auto kern = mgb_score_kernel<NREG, CPT, MODE, SIn>;
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(kern, configured, (int)MGB_SMEM_MAX)) return rc;
    size_t tile = 0;  // bytes of one tile buffer
    if (MODE == MODE_S || (MODE == MODE_FRAMES && p.s != nullptr)) tile = (size_t)MGB_BLOCK * CPT * p.D * (NREG == 0 ? sizeof(float) : sizeof(SIn));
    if (MODE == MODE_SAMPLE && NREG == 0) tile = (size_t)MGB_BLOCK * CPT * (p.D | 1) * sizeof(float);
    tile = (tile + 127) & ~(size_t)127;
    // two tile buffers (TMA prefetch of the next tile) when they fit comfortably
    p.nbuf = (MODE != MODE_SAMPLE && tile > 0 && plan.fixed + 2 * tile <= 96 * 1024) ? 2 : 1;
    if (MODE == MODE_SAMPLE) p.nbuf = 1;
    const size_t smem = plan.fixed + (size_t)p.nbuf * tile;
    MGB_REQUIRE(smem <= MGB_SMEM_MAX, "constraint program / candidate tile too large for shared memory (%zu bytes)", smem);
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, MGB_BLOCK, smem));
    if (occ < 1) occ = 1;
    const long long n_tiles = (p.n + MGB_BLOCK * CPT - 1) / (MGB_BLOCK * CPT);
    long long grid = (long long)num_sms() * occ;
    if (grid > n_tiles) grid = n_tiles;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, MGB_BLOCK, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

constexpr int cpt_preference() { return 2; }

template <int MODE, typename SIn>
int launch_shape(int D, ScoreParams& p, const SmemPlan& plan, cudaStream_t stream) {
    if (D <= 16) {
        if (cpt_preference() == 2 && p.n >= 2LL * MGB_BLOCK * num_sms()) return launch_variant<16, 2, MODE, SIn>(p, plan, stream);
        return launch_variant<16, 1, MODE, SIn>(p, plan, stream);
    }
    if (D <= 32) return launch_variant<32, 1, MODE, SIn>(p, plan, stream);
    return launch_variant<0, 1, MODE, SIn>(p, plan, stream);
}

}  // namespace

extern "C" size_t mgb_score_workspace_bytes(int64_t n, int32_t k) {
    (void)n;
    // per-CTA partials (cost + index) + completion counter, generously sized; for k > 1 also the stage-1 partials of mgb_topk
    size_t parts = 148 * 16;
    const size_t argmin = parts * (sizeof(float) + sizeof(long long)) + 256;
    return k > 1 ? argmin + (size_t)592 * 64 * (sizeof(float) + sizeof(long long)) : argmin;
}

// Fills the kernel parameters of one job and its shared-memory plan (everything but the launch).
static int build_score_params(const mgb_model* m, const mgb_constraints* c, int mode, const void* d_s, const float* d_frames,
                              const int64_t* h_counts, uint64_t seed, int64_t n, int64_t index_base, float err_scale, float quality_scale,
                              float magnitude_scale, float* d_cost, float* d_terms, int k, float* d_best_cost, int64_t* d_best_idx,
                              float* d_best_row, void* d_ws, ScoreParams& p, SmemPlan& plan) {
    const bool frames = mode == MODE_FRAMES;
    memset(&p, 0, sizeof(p));
    if (c) {
        p.tab = c->d_tab;
        p.ops = frames ? c->d_ops_frames : c->d_ops_fused;
        p.cst = c->d_cst;
        p.n_ops = c->n_ops;
        p.n_cst = c->n_cst;
        p.n_quads = c->n_quads;
        p.n_times = c->n_times;
        p.n_terms = c->n_terms;
        p.KP = c->KP;
    }
    p.K = m->K;
    p.D = m->D;
    p.DP = m->DP;
    p.P = m->P;
    p.gmm_tab = m->d_gmm_tab;
    p.gmm_c = m->d_gmm_c;
    p.G = m->G;
    p.NQ = m->NQ;
    p.chol = m->d_chol;
    p.means = m->d_means;
    p.LS = m->LS;
    p.seed = seed;
    p.s = d_s;
    p.frames = d_frames;
    p.n = n;
    p.index_base = index_base;
    p.s_aligned = (d_s != nullptr && (reinterpret_cast<uintptr_t>(d_s) & 15u) == 0) ? 1 : 0;
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
    p.best_row = d_best_row;
    if (mode == MODE_SAMPLE) {
        MGB_REQUIRE(h_counts, "sampling needs the per-component counts");
        MGB_REQUIRE(m->G <= MGB_MAX_GMM, "at most %d mixture components are supported by the sampler", MGB_MAX_GMM);
        p.prefix.v[0] = 0;
        for (int g = 0; g < m->G; ++g) {
            MGB_REQUIRE(h_counts[g] >= 0, "negative component count");
            p.prefix.v[g + 1] = p.prefix.v[g] + h_counts[g];
        }
        for (int g = m->G + 1; g <= MGB_MAX_GMM; ++g) p.prefix.v[g] = p.prefix.v[m->G];
        MGB_REQUIRE(index_base >= 0 && index_base + n <= p.prefix.v[m->G], "component counts (%lld) do not cover [%lld, %lld)",
                    p.prefix.v[m->G], (long long)index_base, (long long)(index_base + n));
    }
    // shared-memory budget: tables go to shared memory when they fit
    const size_t tab_bytes = (frames || !c) ? 0 : (size_t)c->n_quads * (c->KP + 1) * sizeof(float4);
    const size_t gmm_bytes = (quality_scale != 0.f) ? (size_t)m->G * m->NQ * (m->DP + 1) * sizeof(float4) : 0;
    const size_t chol_bytes = (mode == MODE_SAMPLE) ? ((size_t)m->G * m->D * m->LS + (((size_t)m->G * m->D + 3) & ~(size_t)3)) * sizeof(float) : 0;
    size_t fixed = 0;
    p.tab_in_smem = (tab_bytes > 0 && tab_bytes <= 96 * 1024) ? 1 : 0;
    if (p.tab_in_smem) fixed += tab_bytes;
    p.gmm_in_smem = (gmm_bytes > 0 && fixed + gmm_bytes <= 112 * 1024) ? 1 : 0;
    if (p.gmm_in_smem) fixed += gmm_bytes;
    p.chol_in_smem = (chol_bytes > 0 && fixed + chol_bytes <= 112 * 1024) ? 1 : 0;
    if (p.chol_in_smem) fixed += chol_bytes;
    fixed += (size_t)p.n_ops * sizeof(MgbOp);
    fixed += (size_t)((p.n_cst + 3) & ~3) * sizeof(float);
    fixed += 32 * sizeof(float) + 32 * sizeof(long long) + 2 * sizeof(unsigned long long);
    fixed = (fixed + 127) & ~(size_t)127;
    plan.fixed = fixed + 128;  // slack for the base alignment of the dynamic segment
    return MGB_OK;
}

int mgb_launch_score(const mgb_model* m, const mgb_constraints* c, int mode, const void* d_s, int s_is_f64,
                     const float* d_frames, const int64_t* h_counts, uint64_t seed, int64_t n, int64_t index_base, float err_scale,
                     float quality_scale, float magnitude_scale, float* d_cost, float* d_terms, int k,
                     float* d_best_cost, int64_t* d_best_idx, void* d_ws, size_t ws_bytes,
                     cudaStream_t stream, float* d_best_row, const MgbExchange* xch) {
    MGB_REQUIRE(m, "null model handle");
    MGB_REQUIRE(!xch || (k == 1 && n > 0), "the in-kernel exchange needs k = 1 and a non-empty shard");
    MGB_REQUIRE(!c || c->model == m, "constraints were packed for a different model");
    MGB_REQUIRE(n >= 0, "negative n");
    MGB_REQUIRE(k >= 0 && k <= 64, "k must be in [0, 64]");
    MGB_REQUIRE(mode >= MODE_S && mode <= MODE_SAMPLE, "bad mode");
    const bool frames = mode == MODE_FRAMES;
    MGB_REQUIRE(!frames || d_frames, "d_frames is NULL");
    MGB_REQUIRE(mode != MODE_S || d_s || n == 0, "d_s is NULL");
    MGB_REQUIRE(!(frames && s_is_f64), "frames mode takes float32 s");
    MGB_REQUIRE(!d_terms || c, "d_terms needs a constraint set");
    MGB_REQUIRE(!d_best_row || (mode == MODE_SAMPLE && k == 1), "the winning row is re-drawn only by the fused sample + score step with k = 1");
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
            if (d_best_row) MGB_CUDA(cudaMemsetAsync(d_best_row, 0, sizeof(float) * m->D, stream));
            MGB_CUDA(cudaStreamSynchronize(stream));
        }
        return MGB_OK;
    }
    ScoreParams p;
    SmemPlan plan;
    if (int rc = build_score_params(m, c, mode, d_s, d_frames, h_counts, seed, n, index_base, err_scale, quality_scale, magnitude_scale, d_cost,
                                    d_terms, k, d_best_cost, d_best_idx, d_best_row, d_ws, p, plan))
        return rc;
    if (xch) p.xch = *xch;
    // the completion counter in the workspace must be zero on entry: the caller zero-fills a
    // fresh workspace once (include/mgb200.h) and the kernel re-arms it at the end of every launch
    if (d_terms) MGB_CUDA(cudaMemsetAsync(d_terms, 0, (size_t)c->n_terms * n * sizeof(float), stream));
    if (int rcb = mgb_constraints_before_launch(c, stream)) return rcb;   // the program's tables were uploaded on the model's pack stream

    int rc;
    const int D = m->D;
    if ((mode == MODE_S || mode == MODE_SAMPLE) && mgb_score_tc_usable(m, c, quality_scale, mode == MODE_SAMPLE ? 1 : 0)) {
        // large latent spaces: sampling, contraction and GMM log-density on the tensor cores (mgb_tc.cu)
        rc = mgb_launch_score_tc(m, c, p, s_is_f64, quality_scale, mode == MODE_SAMPLE ? 1 : 0, stream);
    } else if (mode == MODE_FRAMES) {
        if ((quality_scale != 0.f || magnitude_scale != 0.f) && d_cost && d_s && mgb_score_tc_usable(m, nullptr, quality_scale, 0)) {
            // large latent space: the density (2 G D^2 flop per candidate) on the tensor cores by a launch of its own, into d_cost;
            // the frames kernel then adds the constraint error to it (same index read and written by the same thread)
            if (int r1 = mgb_launch_score(m, nullptr, MODE_S, d_s, 0, nullptr, nullptr, 0, n, index_base, 0.f, quality_scale, magnitude_scale, d_cost,
                                          nullptr, 0, nullptr, nullptr, nullptr, 0, stream))
                return r1;
            p.quality_scale = 0.f;
            p.magnitude_scale = 0.f;
            p.gmm_in_smem = 0;
            p.s = nullptr;             // the frames kernel no longer reads the latent vectors
            p.cost_base = d_cost;
        }
        rc = launch_variant<0, 1, MODE_FRAMES, float>(p, plan, stream);
    } else if (mode == MODE_SAMPLE) {
        rc = launch_shape<MODE_SAMPLE, float>(D, p, plan, stream);
    } else if (s_is_f64) {
        rc = launch_shape<MODE_S, double>(D, p, plan, stream);
    } else {
        rc = launch_shape<MODE_S, float>(D, p, plan, stream);
    }
    if (rc != MGB_OK) return rc;
    if (int rca = mgb_constraints_after_launch(c, stream)) return rca;
    if (k > 1) {
        return mgb_topk(d_cost, n, index_base, k, d_best_cost, d_best_idx, d_ws, ws_bytes, stream);
    }
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// several jobs, one launch (mgb_sample_score_multi)
// ---------------------------------------------------------------------------
template <int NREG>
static int launch_multi(const ScoreParams* d_jobs, const int* d_first, int n_jobs, int total_ctas, size_t smem, cudaStream_t stream) {
    auto kern = mgb_score_multi_kernel<NREG, 1>;
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(kern, configured, (int)MGB_SMEM_MAX)) return rc;
    MultiParams mp;
    mp.jobs = d_jobs;
    mp.first = d_first;
    mp.n_jobs = n_jobs;
    kern<<<(unsigned)total_ctas, MGB_BLOCK, smem, stream>>>(mp);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

int mgb_launch_score_multi(const MgbJob* jobs, int n_jobs, ScoreParams* h_params /*pinned, [n_jobs]*/, int* h_first /*pinned, [n_jobs + 1]*/,
                           ScoreParams* d_params, int* d_first, unsigned char* d_ws, size_t ws_stride, cudaStream_t stream) {
    MGB_REQUIRE(n_jobs > 0, "no jobs");
    const int D0 = jobs[0].m->D;
    const int cls = D0 <= 16 ? 16 : (D0 <= 32 ? 32 : 0);
    size_t smem_max = 0;
    long long n_sum = 0;
    for (int j = 0; j < n_jobs; ++j) {
        const MgbJob& J = jobs[j];
        MGB_REQUIRE(J.m && J.n > 0, "job %d: empty", j);
        MGB_REQUIRE(!J.c || J.c->model == J.m, "job %d: constraints were packed for a different model", j);
        MGB_REQUIRE(J.m->device == jobs[0].m->device, "job %d: models live on different devices", j);
        const int D = J.m->D;
        if ((D <= 16 ? 16 : (D <= 32 ? 32 : 0)) != cls || mgb_score_tc_usable(J.m, J.c, J.quality_scale, 1)) {
            mgb_set_error("jobs of one launch must share a kernel class (latent dimension <= 16, <= 32, or larger without the tensor-core path)");
            return MGB_E_UNSUPPORTED;
        }
        SmemPlan plan;
        if (int rc = build_score_params(J.m, J.c, MODE_SAMPLE, nullptr, nullptr, J.h_counts, J.seed, J.n, J.index_base, J.err_scale, J.quality_scale,
                                        J.magnitude_scale, nullptr, nullptr, 1, J.d_best_cost, J.d_best_idx, J.d_best_row, d_ws + (size_t)j * ws_stride,
                                        h_params[j], plan))
            return rc;
        size_t tile = (cls == 0) ? (size_t)MGB_BLOCK * (D | 1) * sizeof(float) : 0;
        tile = (tile + 127) & ~(size_t)127;
        h_params[j].nbuf = 1;
        smem_max = std::max(smem_max, plan.fixed + tile);
        n_sum += J.n;
    }
    MGB_REQUIRE(smem_max <= MGB_SMEM_MAX, "constraint program / candidate tile too large for shared memory (%zu bytes)", smem_max);
    // CTAs per job: proportional to its candidates, at least one, at most its tiles and the per-job partial slots
    const int budget = num_sms() * 4;
    int total = 0;
    h_first[0] = 0;
    for (int j = 0; j < n_jobs; ++j) {
        const long long tiles = (jobs[j].n + MGB_BLOCK - 1) / MGB_BLOCK;
        long long ctas = std::max<long long>(1, (long long)budget * jobs[j].n / std::max<long long>(1, n_sum));
        ctas = std::min<long long>(std::min<long long>(ctas, tiles), 148 * 16);
        total += (int)ctas;
        h_first[j + 1] = total;
    }
    MGB_CUDA(cudaMemcpyAsync(d_params, h_params, sizeof(ScoreParams) * n_jobs, cudaMemcpyHostToDevice, stream));
    MGB_CUDA(cudaMemcpyAsync(d_first, h_first, sizeof(int) * (n_jobs + 1), cudaMemcpyHostToDevice, stream));
    for (int j = 0; j < n_jobs; ++j)
        if (int rc = mgb_constraints_before_launch(jobs[j].c, stream)) return rc;
    int rc;
    if (cls == 16) rc = launch_multi<16>(d_params, d_first, n_jobs, total, smem_max, stream);
    else if (cls == 32) rc = launch_multi<32>(d_params, d_first, n_jobs, total, smem_max, stream);
    else rc = launch_multi<0>(d_params, d_first, n_jobs, total, smem_max, stream);
    if (rc != MGB_OK) return rc;
    for (int j = 0; j < n_jobs; ++j)
        if (int rca = mgb_constraints_after_launch(jobs[j].c, stream)) return rca;
    return MGB_OK;
}
// ---------------------------------------------------------------------------
// K5c, one pass: every CTA keeps its chunk of the cost array in shared memory and extracts the chunk's k smallest
// (k block arg-mins over shared memory, a taken entry becomes NaN = "never wins"); one CTA then extracts the k smallest of
// the grid * k partials the same way.  The cost array is read ONCE (the k-pass form read it k times with 2 k launches).
// Order: (cost, global index) ascending; NaN never selected; slots beyond the finite + infinite entries get (inf, index_base).
// ---------------------------------------------------------------------------
namespace {

constexpr int TOPK_CHUNK_MAX = 49152;   // floats of shared memory per CTA (192 KB)
constexpr int TOPK_MAX_CTAS = 592;

__device__ __forceinline__ void topk_block_argmin(float c, long long i, float* red_c, long long* red_i, float& out_c, long long& out_i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(0xffffffffu, c, o);
        const long long oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (better(oc, oi, c, i)) { c = oc; i = oi; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { red_c[threadIdx.x >> 5] = c; red_i[threadIdx.x >> 5] = i; }
    __syncthreads();
    c = red_c[0];
    i = red_i[0];
    for (int w = 1; w < MGB_BLOCK / 32; ++w)
        if (better(red_c[w], red_i[w], c, i)) { c = red_c[w]; i = red_i[w]; }
    out_c = c;
    out_i = i;
}

// src_idx == nullptr: entry e of the chunk has global index index_base + first + e (stage 1); else the index is read (stage 2)
__global__ void __launch_bounds__(MGB_BLOCK) mgb_topk_select_kernel(const float* __restrict__ cost, const long long* __restrict__ src_idx, long long n,
                                                                   long long chunk, long long index_base, int k, float* __restrict__ out_cost,
                                                                   long long* __restrict__ out_idx, int final_stage) {
    extern __shared__ __align__(16) float sm_c[];
    __shared__ float red_c[MGB_BLOCK / 32];
    __shared__ long long red_i[MGB_BLOCK / 32];
    const long long first = (long long)blockIdx.x * chunk;
    const int cnt = (int)max(0LL, min(chunk, n - first));
    for (int e = threadIdx.x; e < cnt; e += MGB_BLOCK) sm_c[e] = cost[first + e];
    __syncthreads();
    for (int j = 0; j < k; ++j) {
        float bc = INFINITY;
        long long bi = LLONG_MAX;
        int be = -1;
        for (int e = threadIdx.x; e < cnt; e += MGB_BLOCK) {
            const float c = sm_c[e];
            const long long gi = src_idx ? __ldg(src_idx + first + e) : index_base + first + e;
            if (better(c, gi, bc, bi)) { bc = c; bi = gi; be = e; }
        }
        float wc;
        long long wi;
        topk_block_argmin(bc, bi, red_c, red_i, wc, wi);
        if (be >= 0 && bi == wi && bc == wc) sm_c[be] = __int_as_float(0x7fc00000);   // taken: NaN never wins again (indices are unique)
        if (threadIdx.x == 0) {
            const bool none = (wi == LLONG_MAX);
            out_cost[(size_t)blockIdx.x * k + j] = (none && !final_stage) ? __int_as_float(0x7fc00000) : wc;
            out_idx[(size_t)blockIdx.x * k + j] = none ? (final_stage ? index_base : LLONG_MAX) : wi;
        }
        __syncthreads();
    }
}

}  // namespace

extern "C" int mgb_topk(const float* d_cost, int64_t n, int64_t index_base, int32_t k, float* d_best_cost,
                        int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(d_cost && d_best_cost && d_best_idx, "null pointer");
    MGB_REQUIRE(k >= 1 && k <= 64, "k must be in [1, 64]");
    MGB_REQUIRE(d_workspace && workspace_bytes >= mgb_score_workspace_bytes(n, k), "workspace too small");
    MGB_REQUIRE(n >= 0, "negative n");
    // partials of stage 1 live in the workspace after the arg-min partials of the score kernels
    unsigned char* ws = reinterpret_cast<unsigned char*>(d_workspace);
    const size_t parts = 148 * 16;
    unsigned char* tk = ws + parts * (sizeof(long long) + sizeof(float)) + 256;
    long long ctas = std::max<long long>(1, std::min<long long>((n + 4095) / 4096, TOPK_MAX_CTAS));
    long long chunk = std::max<long long>(1, (n + ctas - 1) / ctas);
    if (k > 1 && chunk <= TOPK_CHUNK_MAX && ctas * k <= TOPK_CHUNK_MAX) {
        long long* part_idx = reinterpret_cast<long long*>(tk);
        float* part_cost = reinterpret_cast<float*>(tk + (size_t)TOPK_MAX_CTAS * 64 * sizeof(long long));
        static bool configured[MGB_MAX_DEVICES] = {false};
        if (int rc = mgb_opt_in_smem(mgb_topk_select_kernel, configured, TOPK_CHUNK_MAX * (int)sizeof(float))) return rc;
        mgb_topk_select_kernel<<<(unsigned)ctas, MGB_BLOCK, (size_t)chunk * sizeof(float), stream>>>(d_cost, nullptr, n, chunk, index_base, k, part_cost,
                                                                                                  part_idx, 0);
        const long long m = ctas * k;
        mgb_topk_select_kernel<<<1, MGB_BLOCK, (size_t)m * sizeof(float), stream>>>(part_cost, part_idx, m, m, index_base, k, d_best_cost,
                                                                                  reinterpret_cast<long long*>(d_best_idx), 1);
        mgb_count_launch(2);
        MGB_CUDA(cudaGetLastError());
        return MGB_OK;
    }
    // k = 1, or arrays too large for one shared-memory chunk per CTA (n > 592 x 49 152 = 29 M): k passes, each a grid arg-min over
    // the entries "after" the previous pick; a pass that finds nothing raises a flag (in the pad behind the completion counter)
    // and the later passes fill their slots with (inf, index_base)
    int* d_exhausted = reinterpret_cast<int*>(ws + parts * (sizeof(long long) + sizeof(float)) + 128);
    MGB_CUDA(cudaMemsetAsync(d_exhausted, 0, sizeof(int), stream));
    long long* part_idx = reinterpret_cast<long long*>(ws);
    float* part_cost = reinterpret_cast<float*>(ws + parts * sizeof(long long));
    long long n_tiles = (n + MGB_BLOCK - 1) / MGB_BLOCK;
    int grid = (int)((n_tiles < (long long)num_sms() * 8) ? n_tiles : (long long)num_sms() * 8);
    if (grid < 1) grid = 1;
    for (int j = 0; j < k; ++j) {
        mgb_topk_partial<<<grid, MGB_BLOCK, 0, stream>>>(d_cost, n, index_base, j ? d_best_cost + j - 1 : nullptr,
                                                        j ? reinterpret_cast<long long*>(d_best_idx) + j - 1 : nullptr, j ? 1 : 0,
                                                        part_cost, part_idx, d_exhausted);
        mgb_topk_final<<<1, 32, 0, stream>>>(part_cost, part_idx, grid, index_base, d_best_cost + j,
                                             reinterpret_cast<long long*>(d_best_idx) + j, d_exhausted);
        mgb_count_launch(2);
    }
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// mgb_topk_noalloc: K5c for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
// Same two launches as mgb_topk's one-pass form, but the stage-1 partials live in a file-scope device array instead of
// a caller workspace: MGB_TOPK_NOALLOC_SCRATCH_BYTES (444 KB) held by every GPU that loads the library, until exit.
// Nothing in the array has to start at zero (stage 1 writes all ctas * k slots stage 2 reads), so nothing is cleared;
// every call shares it, so two calls that overlap on the GPU (different streams, or different host threads) race.
// ---------------------------------------------------------------------------
namespace {
constexpr size_t TOPK_NOALLOC_IDX_BYTES = (size_t)TOPK_MAX_CTAS * 64 * sizeof(long long);
constexpr size_t TOPK_NOALLOC_BYTES = TOPK_NOALLOC_IDX_BYTES + (size_t)TOPK_MAX_CTAS * 64 * sizeof(float);
static_assert(TOPK_NOALLOC_BYTES == MGB_TOPK_NOALLOC_SCRATCH_BYTES, "header and kernel disagree on the scratch size");
static_assert((long long)MGB_TOPK_NOALLOC_MAX_N <= (long long)TOPK_MAX_CTAS * TOPK_CHUNK_MAX, "a chunk must fit shared memory");
}  // namespace

static __device__ __align__(256) unsigned char mgb_topk_noalloc_scratch[TOPK_NOALLOC_BYTES];

// Device address of the scratch array on the CURRENT device.  The array's name is only a host-side stand-in in host code;
// the address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
static int topk_noalloc_scratch_ptr(unsigned char** out) {
    static unsigned char* addr[MGB_MAX_DEVICES] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    MGB_CUDA(cudaGetDevice(&dev));
    MGB_REQUIRE(dev >= 0 && dev < MGB_MAX_DEVICES, "device ordinal %d beyond the %d the library keeps per-device state for", dev, MGB_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        MGB_CUDA(cudaGetSymbolAddress(&p, mgb_topk_noalloc_scratch));
        MGB_REQUIRE(p, "cudaGetSymbolAddress returned a null address for the top-k scratch");
        addr[dev] = reinterpret_cast<unsigned char*>(p);
    }
    *out = addr[dev];
    return MGB_OK;
}

extern "C" int mgb_topk_noalloc(const float* d_cost, int64_t n, int64_t index_base, int32_t k, float* d_best_cost,
                                int64_t* d_best_idx, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(d_cost && d_best_cost && d_best_idx, "null pointer");
    MGB_REQUIRE(k >= 1 && k <= 64, "k must be in [1, 64]");
    MGB_REQUIRE(n >= 0, "negative n");
    MGB_REQUIRE(n <= (int64_t)MGB_TOPK_NOALLOC_MAX_N, "mgb_topk_noalloc handles at most %lld costs per call (got %lld): use mgb_topk with a workspace",
                (long long)MGB_TOPK_NOALLOC_MAX_N, (long long)n);
    unsigned char* tk = nullptr;
    if (int rc = topk_noalloc_scratch_ptr(&tk)) return rc;
    const long long ctas = std::max<long long>(1, std::min<long long>((n + 4095) / 4096, TOPK_MAX_CTAS));
    const long long chunk = std::max<long long>(1, (n + ctas - 1) / ctas);
    MGB_REQUIRE(chunk <= TOPK_CHUNK_MAX && ctas * k <= TOPK_CHUNK_MAX, "internal: chunk %lld does not fit shared memory", chunk);
    long long* part_idx = reinterpret_cast<long long*>(tk);
    float* part_cost = reinterpret_cast<float*>(tk + TOPK_NOALLOC_IDX_BYTES);
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(mgb_topk_select_kernel, configured, TOPK_CHUNK_MAX * (int)sizeof(float))) return rc;
    mgb_topk_select_kernel<<<(unsigned)ctas, MGB_BLOCK, (size_t)chunk * sizeof(float), stream>>>(d_cost, nullptr, n, chunk, index_base, k, part_cost,
                                                                                              part_idx, 0);
    const long long m = ctas * k;
    mgb_topk_select_kernel<<<1, MGB_BLOCK, (size_t)m * sizeof(float), stream>>>(part_cost, part_idx, m, m, index_base, k, d_best_cost,
                                                                              reinterpret_cast<long long*>(d_best_idx), 1);
    mgb_count_launch(2);
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}
