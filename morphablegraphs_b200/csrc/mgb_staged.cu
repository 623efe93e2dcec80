// K1 (GMM sampler), K2 (fp32 SIMT back-projection), K3 (spline evaluation),
// K4 (forward kinematics), K5b (GMM log-density), K7 (k-means level step).
// These are the HBM-bound staged kernels: each stages its tile through shared
// memory so that every global access is a coalesced 128-bit (or 32-bit
// unit-stride) transaction, and sizes its grid from the SM count.
#include <float.h>
#include <math.h>

#include <algorithm>

#include "mgb_device.cuh"
#include "mgb_internal.cuh"

namespace {

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

#define SAMPLE_BLOCK 256

// K1.  One thread draws one candidate (Philox4x32 keyed by the global index, Box-Muller,
// s = mu_g + L_g z with the Cholesky rows broadcast from shared memory as float4), writes its
// row into a densely packed shared-memory tile [SAMPLE_BLOCK][D]; the tile (one contiguous
// run of global memory) leaves with ONE TMA bulk store, double-buffered so the store of
// tile i overlaps the arithmetic of tile i + 1.
template <int NREG>
__global__ void __launch_bounds__(SAMPLE_BLOCK) mgb_sample_kernel(const float* __restrict__ means, const float* __restrict__ chol, int LS,
                                                                 const __grid_constant__ MgbGmmPrefix pre, int G, int D, long long n,
                                                                 long long index_base, unsigned long long seed, int in_smem, int aligned,
                                                                 float* __restrict__ out) {
    extern __shared__ __align__(128) unsigned char sm_raw1[];
    constexpr int NR = NREG > 0 ? NREG : 1;
    const int SB = blockDim.x;  // rows per tile (<= SB; fewer for large D so that two tiles fit)
    const size_t tile_bytes = ((size_t)SB * D * sizeof(float) + 127) & ~(size_t)127;
    float* sm_tile = reinterpret_cast<float*>(sm_raw1);
    float* sm_chol = reinterpret_cast<float*>(sm_raw1 + 2 * tile_bytes);
    float* sm_mu = sm_chol + (size_t)G * D * LS;
    if (in_smem) {
        for (int i = threadIdx.x; i < G * D * LS; i += SB) sm_chol[i] = chol[i];
        for (int i = threadIdx.x; i < G * D; i += SB) sm_mu[i] = means[i];
    }
    const float* L = in_smem ? sm_chol : chol;
    const float* mus = in_smem ? sm_mu : means;
    const uint2 key = make_uint2((unsigned int)seed, (unsigned int)(seed >> 32));
    const long long n_tiles = (n + SB - 1) / SB;
    int it = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        float* buf = reinterpret_cast<float*>(sm_raw1 + (size_t)(it & 1) * tile_bytes);
        if (threadIdx.x == 0) mgb_tma_store_wait_read<1>();  // the store that last read this buffer (two tiles ago) is done
        __syncthreads();
        const long long i = tile * SB + threadIdx.x;
        const unsigned long long gi = (unsigned long long)(index_base + i);
        float* row = buf + (size_t)threadIdx.x * D;
        if (i < n) {
            int g = 0;  // component of global index gi in the component-ordered stacking
            while (g + 1 < G && (long long)gi >= pre.v[g + 1]) ++g;
            if (NREG > 0) {
                float s[NR];
                mgb_sample_row_reg<NR>(reinterpret_cast<const float4*>(L + (size_t)g * D * LS), LS, mus + (size_t)g * D, D, gi, key, s);
#pragma unroll
                for (int d = 0; d < NR; ++d)
                    if (d < D) row[d] = s[d];
            } else {
                mgb_sample_row_mem(L + (size_t)g * D * LS, LS, mus + (size_t)g * D, D, gi, key, row);
            }
        }
        const long long base = tile * SB;
        const bool full = aligned && base + SB <= n;
        if (full) mgb_fence_proxy_async();  // generic-proxy writes of the tile visible to the async proxy
        __syncthreads();
        if (full) {
            if (threadIdx.x == 0) mgb_tma_store_1d(out + base * D, buf, (unsigned int)(SB * D * sizeof(float)));
        } else {
            const long long remaining = (n - base) * (long long)D;
            float* dst = out + base * D;
            for (int e = threadIdx.x; e < SB * D; e += SB)
                if (e < remaining) dst[e] = buf[e];
        }
    }
    if (threadIdx.x == 0) mgb_tma_store_wait_read<0>();
}

// ---------------------------------------------------------------------------
// K1 for large latent spaces (32 < D <= 128): one CTA = a tile of SG_ROWS candidates of ONE mixture component, drawn as a
// register-tiled product  s[r][d] = mu[d] + sum_j L[d][j] z[r][j]  out of shared memory (z transposed, L transposed and
// padded to a conflict-free stride).  The per-row loop above spends two loads and a dependent FMA per term (15 ms per 4M
// candidates at D = 103); here a thread owns 8 rows x D/16 columns and every output is still accumulated by ONE thread in
// ascending j starting from mu[d] - the terms above the diagonal multiply a stored zero - so the rows are bit for bit those
// of the per-row loop (and of the reference-ordered oracle).
// ---------------------------------------------------------------------------
constexpr int SG_ROWS = 128;
constexpr int SG_THREADS = 256;
constexpr int SG_CMAX = 8;          // columns per thread: d = cg + 16 c, D <= 128

struct SgTiles {
    int first[MGB_MAX_GMM + 1];     // tiles before component g (tiles never straddle components)
};

template <int CN>
__global__ void __launch_bounds__(SG_THREADS) mgb_sample_gemm_kernel(const float* __restrict__ means, const float* __restrict__ chol, int LS,
                                                                    const __grid_constant__ MgbGmmPrefix pre, const __grid_constant__ SgTiles tl,
                                                                    int G, int D, long long n, long long index_base, unsigned long long seed,
                                                                    float* __restrict__ out) {
    extern __shared__ __align__(128) float sg_sm[];
    const int D4 = (D + 3) & ~3;                 // z columns drawn (4 per Philox call)
    constexpr int ZS = SG_ROWS + 4;              // zt[j][row]
    constexpr int LT = 16 * CN + 1;              // lt[j][d], odd stride: the transposing stores do not conflict
    float* zt = sg_sm;                           // [D4][ZS]; later the finished rows [SG_ROWS][D]
    float* lt = sg_sm + (size_t)D4 * ZS;         // [D4][LT]
    const int tid = threadIdx.x;
    int g = 0;
    while (g + 1 < G && (int)blockIdx.x >= tl.first[g + 1]) ++g;
    const long long lo = max(pre.v[g], index_base), hi = min(pre.v[g + 1], index_base + n);
    const long long row0 = lo + (long long)((int)blockIdx.x - tl.first[g]) * SG_ROWS;      // global index of the tile's first row
    const int rows = (int)min((long long)SG_ROWS, hi - row0);
    const uint2 key = make_uint2((unsigned int)seed, (unsigned int)(seed >> 32));
    // L_g transposed (zero above the diagonal and beyond D)
    const float* Lg = chol + (size_t)g * D * LS;
    for (int e = tid; e < D4 * 16 * CN; e += SG_THREADS) {
        const int d = e / D4, j = e - d * D4;
        lt[j * LT + d] = (d < D && j <= d) ? __ldg(Lg + (size_t)d * LS + j) : 0.f;
    }
    // normals of the tile, transposed
    const int nb = D4 >> 2;
    for (int e = tid; e < SG_ROWS * nb; e += SG_THREADS) {
        const int r = e % SG_ROWS, b = e / SG_ROWS;
        float q[4] = {0.f, 0.f, 0.f, 0.f};
        if (r < rows) mgb_normal4((unsigned long long)(row0 + r), (unsigned int)b, key, q);
#pragma unroll
        for (int c = 0; c < 4; ++c) zt[(4 * b + c) * ZS + r] = q[c];
    }
    __syncthreads();
    const int rg = tid >> 4, cg = tid & 15;
    float acc[8][CN];
#pragma unroll
    for (int c = 0; c < CN; ++c) {
        const int d = cg + 16 * c;
        const float mu = d < D ? __ldg(means + (size_t)g * D + d) : 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r][c] = mu;
    }
    // j in blocks of 16: L is lower triangular, so the columns d < 16 jb (c < jb) have nothing but zeros in block jb and are left
    // out (57 % of the square at D = 103); a skipped term would have added 0 * z
#pragma unroll
    for (int jb = 0; jb < CN; ++jb) {
        const int j1 = min(16 * jb + 16, D);
        for (int j = 16 * jb; j < j1; ++j) {
            const float4 za = *reinterpret_cast<const float4*>(zt + j * ZS + rg * 8);
            const float4 zb = *reinterpret_cast<const float4*>(zt + j * ZS + rg * 8 + 4);
            const float z[8] = {za.x, za.y, za.z, za.w, zb.x, zb.y, zb.z, zb.w};
#pragma unroll
            for (int c = jb; c < CN; ++c) {
                const float l = lt[j * LT + cg + 16 * c];
#pragma unroll
                for (int r = 0; r < 8; ++r) acc[r][c] = fmaf(l, z[r], acc[r][c]);
            }
        }
    }
    __syncthreads();                             // every thread is done with zt: it becomes the output tile [rows][D]
    float* tile = zt;
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            const int d = cg + 16 * c;
            if (d < D) tile[(size_t)(rg * 8 + r) * D + d] = acc[r][c];
        }
    float* dst = out + (size_t)(row0 - index_base) * D;
    const unsigned int bytes = (unsigned int)((size_t)rows * D * sizeof(float));
    const bool bulk = ((reinterpret_cast<uintptr_t>(dst) | bytes) & 15u) == 0;
    if (bulk) mgb_fence_proxy_async();
    __syncthreads();
    if (bulk) {
        if (tid == 0) {
            mgb_tma_store_1d(dst, tile, bytes);
            mgb_tma_store_wait_read<0>();
        }
    } else {
        for (int e = tid; e < rows * D; e += SG_THREADS) dst[e] = tile[e];
    }
}

template <int CN>
static int launch_sample_gemm(const mgb_model* m, const MgbGmmPrefix& prefix, const SgTiles& tl, int total, int64_t n, int64_t index_base,
                              uint64_t seed, float* d_s, cudaStream_t stream) {
    auto kern = mgb_sample_gemm_kernel<CN>;
    const int D4 = (m->D + 3) & ~3;
    const size_t smem = ((size_t)D4 * (SG_ROWS + 4) + (size_t)D4 * (16 * CN + 1)) * sizeof(float);
    MGB_REQUIRE((size_t)SG_ROWS * m->D <= (size_t)D4 * (SG_ROWS + 4), "internal: output tile larger than the z tile");
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(kern, configured, 200 * 1024)) return rc;
    MGB_REQUIRE(smem <= 200 * 1024, "latent dimension too large for the sampler");
    kern<<<(unsigned)total, SG_THREADS, smem, stream>>>(m->d_means, m->d_chol, m->LS, prefix, tl, m->G, m->D, n, index_base,
                                                        (unsigned long long)seed, d_s);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// K2 (SIMT fp32): coeffs[n, M] = alpha[n, K] . E'[K, M] + mean'[M]
// CTA = 256 threads = 256 consecutive columns x BP_ROWS candidates; alpha tile
// transposed in shared memory so one 128-bit broadcast load feeds 4 FMAs;
// every store instruction of a warp writes 128 contiguous bytes.
// ---------------------------------------------------------------------------
#define BP_ROWS 32
#define BP_COLS 256
__global__ void __launch_bounds__(BP_COLS) mgb_backproject_simt(const float* __restrict__ s, int D, int K, const float* __restrict__ E,
                                                                const float* __restrict__ mean, int M, long long n, float* __restrict__ out) {
    extern __shared__ __align__(16) float sm_a[];  // [K][BP_ROWS]
    const int col = blockIdx.x * BP_COLS + threadIdx.x;
    const bool col_ok = col < M;
    const float mu = col_ok ? __ldg(mean + col) : 0.f;
    const long long n_row_tiles = (n + BP_ROWS - 1) / BP_ROWS;
    for (long long rt = blockIdx.y; rt < n_row_tiles; rt += gridDim.y) {
        const long long r0 = rt * BP_ROWS;
        __syncthreads();
        for (int e = threadIdx.x; e < BP_ROWS * K; e += BP_COLS) {
            int r = e / K, k = e - r * K;
            long long gr = r0 + r;
            sm_a[k * BP_ROWS + r] = (gr < n) ? __ldg(s + gr * D + k) : 0.f;
        }
        __syncthreads();
        float acc[BP_ROWS];
#pragma unroll
        for (int r = 0; r < BP_ROWS; ++r) acc[r] = mu;
        for (int k = 0; k < K; ++k) {
            const float e = col_ok ? __ldg(E + (size_t)k * M + col) : 0.f;
            const float4* a4 = reinterpret_cast<const float4*>(sm_a + k * BP_ROWS);
#pragma unroll
            for (int r4 = 0; r4 < BP_ROWS / 4; ++r4) {
                float4 a = a4[r4];
                acc[4 * r4 + 0] = fmaf(a.x, e, acc[4 * r4 + 0]);
                acc[4 * r4 + 1] = fmaf(a.y, e, acc[4 * r4 + 1]);
                acc[4 * r4 + 2] = fmaf(a.z, e, acc[4 * r4 + 2]);
                acc[4 * r4 + 3] = fmaf(a.w, e, acc[4 * r4 + 3]);
            }
        }
        if (col_ok) {
#pragma unroll
            for (int r = 0; r < BP_ROWS; ++r) {
                long long gr = r0 + r;
                if (gr < n) __stcs(out + gr * M + col, acc[r]);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// K3: frames[n, T, P] = B[T, Nb] . coeffs[n, Nb, P]  (4 non-zeros per row of B)
// One thread owns one (candidate, channel) pair and walks the T evaluation
// times with a sliding window of 4 coefficients in registers: a coefficient is
// loaded once per window move, not once per output (Nb loads and T stores per
// thread for a full motion).  Lanes of a warp hold consecutive flat indices
// i * P + p, so every load and store instruction of a warp covers one
// contiguous 128-byte run.  Spans / weights are broadcast from shared memory.
// ---------------------------------------------------------------------------
#define K3_PARAM_T 96
struct SplineTimes {  // small time lists travel as a kernel argument: no allocation, copy or sync per call
    float4 w[K3_PARAM_T];
    int span[K3_PARAM_T];
};

template <bool IN_PARAM>
__global__ void __launch_bounds__(256) mgb_eval_frames_kernel(const float* __restrict__ coeffs, int Nb, int P, long long n,
                                                              const int* __restrict__ spans, const float4* __restrict__ w4, int T,
                                                              float* __restrict__ frames, const __grid_constant__ SplineTimes st) {
    extern __shared__ __align__(16) unsigned char sm_raw3[];
    float4* sm_w = reinterpret_cast<float4*>(sm_raw3);
    int* sm_span = reinterpret_cast<int*>(sm_w + T);
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        sm_w[t] = IN_PARAM ? st.w[t] : w4[t];
        sm_span[t] = IN_PARAM ? st.span[t] : spans[t];
    }
    __syncthreads();
    const long long total = n * (long long)P;
    const size_t M = (size_t)Nb * P;
    const size_t TP = (size_t)T * P;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long i = e / P;
        const int pch = (int)(e - i * P);
        const float* __restrict__ src = coeffs + (size_t)i * M + pch;
        float* __restrict__ dst = frames + (size_t)i * TP + pch;
        int cur = -8;
        float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
        for (int t = 0; t < T; ++t) {
            const int sp = sm_span[t];
            if (sp != cur) {
                if (sp == cur + 1) {
                    c0 = c1; c1 = c2; c2 = c3;
                    c3 = __ldcs(src + (size_t)(sp + 3) * P);
                } else {
                    c0 = __ldcs(src + (size_t)sp * P);
                    c1 = __ldcs(src + (size_t)(sp + 1) * P);
                    c2 = __ldcs(src + (size_t)(sp + 2) * P);
                    c3 = __ldcs(src + (size_t)(sp + 3) * P);
                }
                cur = sp;
            }
            const float4 w = sm_w[t];
            float v = w.x * c0;
            v = fmaf(w.y, c1, v);
            v = fmaf(w.z, c2, v);
            v = fmaf(w.w, c3, v);
            __stcs(dst + (size_t)t * P, v);
        }
    }
}

// ---------------------------------------------------------------------------
// K4: forward kinematics of listed nodes.  A tile of frames is staged in shared memory (rows of P floats, odd stride).
// ---------------------------------------------------------------------------
#define FK_BLOCK 128

// K4: the chain root -> joint of every requested joint is walked in quaternion form with the transform in registers (the helpers of
// the fused kernels: same arithmetic as their forward kinematics).  Shared ancestors are recomputed per requested joint (about 1.5x
// the arithmetic of a walk over the needed part of the tree with one 3 x 4 matrix per node), but nothing lives in local memory:
// the matrix walk kept 7.7 KB per thread there and took 1.40 ms where this takes 0.98 ms (32768 x 50 frames, 10 joints).
struct FkChainStep {
    int kind;        // 0 root, 1 animated joint, 2 fixed rotation + offset, 3 offset only
    int channel;
    float off[3];
    float q[4];
};

__global__ void __launch_bounds__(FK_BLOCK) mgb_fk_chain_kernel(const float* __restrict__ frames, int P, long long n_frames,
                                                            const FkChainStep* __restrict__ steps, const int* __restrict__ chain_first,
                                                            int n_steps, int n_out, float* __restrict__ pos) {
    extern __shared__ __align__(16) float smf[];
    const int FKB = blockDim.x;  // frames per tile = threads (chosen by the launcher to fit shared memory)
    const int SROW = P | 1;
    const int OROW = (n_out * 3) | 1;
    float* sm_frames = smf;                                  // [FKB][SROW]
    float* sm_out = smf + (size_t)FKB * SROW;                // [FKB][OROW]
    FkChainStep* sm_steps = reinterpret_cast<FkChainStep*>(sm_out + (size_t)FKB * OROW + 1);
    int* sm_first = reinterpret_cast<int*>(sm_steps + n_steps);
    for (int e = threadIdx.x; e < n_steps; e += FKB) sm_steps[e] = steps[e];
    for (int e = threadIdx.x; e <= n_out; e += FKB) sm_first[e] = chain_first[e];
    const long long n_tiles = (n_frames + FKB - 1) / FKB;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long base = tile * FKB;
        __syncthreads();
        {
            const float* src = frames + base * P;
            const long long remaining = (n_frames - base) * (long long)P;
            const int total = FKB * P;
            const int step_r = FKB / P, step_c = FKB % P;
            int r = threadIdx.x / P, c = threadIdx.x % P;
            for (int e = threadIdx.x; e < total; e += FKB) {
                sm_frames[r * SROW + c] = (e < remaining) ? __ldcs(src + e) : 0.f;
                r += step_r;
                c += step_c;
                if (c >= P) { c -= P; r += 1; }
            }
        }
        __syncthreads();
        for (int w = threadIdx.x; w < FKB * n_out; w += FKB) {
            const int r = w / n_out, j = w - r * n_out;
            const float* f = sm_frames + r * SROW;
            MgbXf G;
            mgb_xf_identity(G);
            for (int si = sm_first[j]; si < sm_first[j + 1]; ++si) {
                const FkChainStep& st = sm_steps[si];
                if (st.kind == 0) {
                    mgb_xf_set_rot(G, f[3], f[4], f[5], f[6]);
                    G.tx = f[0] + st.off[0];
                    G.ty = f[1] + st.off[1];
                    G.tz = f[2] + st.off[2];
                } else if (st.kind == 1) {
                    const float* q = f + 3 + 4 * st.channel;
                    mgb_xf_child(G, q[0], q[1], q[2], q[3], st.off[0], st.off[1], st.off[2]);
                } else if (st.kind == 2) {
                    mgb_xf_child(G, st.q[0], st.q[1], st.q[2], st.q[3], st.off[0], st.off[1], st.off[2]);
                } else {
                    mgb_xf_translate(G, st.off[0], st.off[1], st.off[2]);
                }
            }
            float* o = sm_out + r * OROW + 3 * j;
            o[0] = G.tx; o[1] = G.ty; o[2] = G.tz;
        }
        __syncthreads();
        {
            const int W = n_out * 3;
            float* dst = pos + base * W;
            const long long remaining = (n_frames - base) * (long long)W;
            const int total = FKB * W;
            const int step_r = FKB / W, step_c = FKB % W;
            int r = threadIdx.x / W, c = threadIdx.x % W;
            for (int e = threadIdx.x; e < total; e += FKB) {
                if (e < remaining) __stcs(dst + e, sm_out[r * OROW + c]);
                r += step_r;
                c += step_c;
                if (c >= W) { c -= W; r += 1; }
            }
        }
    }
}

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
template <int NREG>
static int launch_sample(const mgb_model* m, const MgbGmmPrefix& prefix, int64_t n, int64_t index_base, uint64_t seed, float* d_s,
                         cudaStream_t stream) {
    auto kern = mgb_sample_kernel<NREG>;
    int sb = SAMPLE_BLOCK;
    while (sb > 32 && 2 * (size_t)sb * m->D * sizeof(float) > 64 * 1024) sb >>= 1;
    const size_t tile_bytes = ((size_t)sb * m->D * sizeof(float) + 127) & ~(size_t)127;
    const size_t chol_bytes = ((size_t)m->G * m->D * m->LS + (size_t)m->G * m->D) * sizeof(float);
    const int in_smem = (2 * tile_bytes + chol_bytes <= 96 * 1024) ? 1 : 0;
    const size_t smem = 2 * tile_bytes + (in_smem ? chol_bytes : 0) + 128;
    MGB_REQUIRE(smem <= 200 * 1024, "latent dimension too large for the sampler");
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(kern, configured, 200 * 1024)) return rc;
    long long n_tiles = (n + sb - 1) / sb;
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, sb, smem));
    long long grid = (long long)sm_count() * (occ > 0 ? occ : 1);
    if (grid > n_tiles) grid = n_tiles;
    const int aligned = ((reinterpret_cast<uintptr_t>(d_s) & 15u) == 0 && ((size_t)sb * m->D * sizeof(float)) % 16 == 0) ? 1 : 0;
    kern<<<(unsigned)grid, sb, smem, stream>>>(m->d_means, m->d_chol, m->LS, prefix, m->G, m->D, n, index_base,
                                                        (unsigned long long)seed, in_smem, aligned, d_s);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

extern "C" int mgb_sample(const mgb_model* m, int64_t n, int64_t index_base, const int64_t* h_counts, uint64_t seed,
                          float* d_s, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(m && h_counts && (d_s || n == 0), "null argument");
    MGB_REQUIRE(n >= 0 && index_base >= 0, "negative size");
    if (n == 0) return MGB_OK;
    MGB_REQUIRE(m->G <= MGB_MAX_GMM, "at most %d mixture components are supported by the sampler", MGB_MAX_GMM);
    MgbGmmPrefix prefix;  // passed by value as a kernel argument: no allocation, copy or sync per call
    prefix.v[0] = 0;
    for (int g = 0; g < m->G; ++g) {
        MGB_REQUIRE(h_counts[g] >= 0, "negative component count");
        prefix.v[g + 1] = prefix.v[g] + h_counts[g];
    }
    for (int g = m->G + 1; g <= MGB_MAX_GMM; ++g) prefix.v[g] = prefix.v[m->G];
    MGB_REQUIRE(index_base + n <= prefix.v[m->G], "component counts (%lld) do not cover [%lld, %lld)", prefix.v[m->G],
                (long long)index_base, (long long)(index_base + n));
    if (m->D <= 16) return launch_sample<16>(m, prefix, n, index_base, seed, d_s, stream);
    if (m->D <= 32) return launch_sample<32>(m, prefix, n, index_base, seed, d_s, stream);
    if (m->D <= 16 * SG_CMAX && n >= 4 * SG_ROWS) {
        // tiles of one component each
        SgTiles tl;
        int total = 0;
        for (int g = 0; g < m->G; ++g) {
            tl.first[g] = total;
            const long long lo = std::max<long long>(prefix.v[g], index_base), hi = std::min<long long>(prefix.v[g + 1], index_base + n);
            if (hi > lo) total += (int)((hi - lo + SG_ROWS - 1) / SG_ROWS);
        }
        for (int g = m->G; g <= MGB_MAX_GMM; ++g) tl.first[g] = total;
        switch ((m->D + 15) / 16) {
            case 3: return launch_sample_gemm<3>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
            case 4: return launch_sample_gemm<4>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
            case 5: return launch_sample_gemm<5>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
            case 6: return launch_sample_gemm<6>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
            case 7: return launch_sample_gemm<7>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
            default: return launch_sample_gemm<8>(m, prefix, tl, total, n, index_base, seed, d_s, stream);
        }
    }
    return launch_sample<0>(m, prefix, n, index_base, seed, d_s, stream);
}

extern "C" int mgb_backproject(const mgb_model* m, const float* d_s, int64_t n, float* d_coeffs, int32_t impl, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(m && (n == 0 || (d_s && d_coeffs)), "null argument");
    MGB_REQUIRE(impl >= 0 && impl <= 2, "impl must be 0 (auto), 1 (simt) or 2 (tensor core)");
    if (n == 0) return MGB_OK;
    if (impl == 2 || (impl == 0 && m->d_Bsplit != nullptr)) return mgb_backproject_tc(m, d_s, n, d_coeffs, stream);
    size_t smem = (size_t)m->K * BP_ROWS * sizeof(float);
    MGB_REQUIRE(smem <= 48 * 1024, "K too large for the SIMT back-projection (use impl=2)");
    dim3 grid((m->M + BP_COLS - 1) / BP_COLS, 1, 1);
    long long row_tiles = (n + BP_ROWS - 1) / BP_ROWS;
    long long gy = ((long long)sm_count() * 8 + grid.x - 1) / grid.x;
    if (gy > row_tiles) gy = row_tiles;
    if (gy > 65535) gy = 65535;
    grid.y = (unsigned)gy;
    mgb_backproject_simt<<<grid, BP_COLS, smem, stream>>>(d_s, m->D, m->K, m->d_E, m->d_mean, m->M, n, d_coeffs);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

extern "C" int mgb_eval_frames(const mgb_model* m, const float* d_coeffs, int64_t n, const double* times, int32_t n_times,
                               const int32_t* spans, const double* weights, float* d_frames, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(m && times && n_times > 0, "null argument");
    MGB_REQUIRE(n == 0 || (d_coeffs && d_frames), "null device pointer");
    if (n == 0) return MGB_OK;
    SplineTimes st;
    std::vector<int> sp(n_times);
    std::vector<float4> w4(n_times);
    for (int t = 0; t < n_times; ++t) {
        double w[4];
        int span;
        if (spans && weights) {
            span = spans[t];
            for (int q = 0; q < 4; ++q) w[q] = weights[4 * t + q];
        } else {
            int rc = mgb_spline_basis(m->knots.data(), (int)m->knots.size(), times[t], &span, w);
            if (rc != MGB_OK) return rc;
        }
        MGB_REQUIRE(span >= 0 && span + 4 <= m->Nb, "span out of range");
        sp[t] = span;
        w4[t] = make_float4((float)w[0], (float)w[1], (float)w[2], (float)w[3]);
        if (t < K3_PARAM_T) {
            st.span[t] = span;
            st.w[t] = w4[t];
        }
    }
    const bool in_param = n_times <= K3_PARAM_T;
    int* d_sp = nullptr;
    float4* d_w4 = nullptr;
    if (!in_param) {
        MGB_CUDA(cudaMallocAsync((void**)&d_sp, sizeof(int) * n_times, stream));
        MGB_CUDA(cudaMallocAsync((void**)&d_w4, sizeof(float4) * n_times, stream));
        MGB_CUDA(cudaMemcpyAsync(d_sp, sp.data(), sizeof(int) * n_times, cudaMemcpyHostToDevice, stream));
        MGB_CUDA(cudaMemcpyAsync(d_w4, w4.data(), sizeof(float4) * n_times, cudaMemcpyHostToDevice, stream));
        MGB_CUDA(cudaStreamSynchronize(stream));  // sp / w4 are stack-owned host buffers
    }
    size_t smem = (size_t)n_times * (sizeof(float4) + sizeof(int)) + 16;
    MGB_REQUIRE(smem <= 200 * 1024, "too many evaluation times for shared memory");
    auto kern = in_param ? mgb_eval_frames_kernel<true> : mgb_eval_frames_kernel<false>;
    if (smem > 48 * 1024) MGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem));
    long long grid = (long long)sm_count() * (occ > 0 ? occ : 1);
    const long long blocks = (n * (long long)m->P + 255) / 256;
    if (grid > blocks) grid = blocks;
    kern<<<(unsigned)grid, 256, smem, stream>>>(d_coeffs, m->Nb, m->P, n, d_sp, d_w4, n_times, d_frames, st);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    if (!in_param) {
        MGB_CUDA(cudaFreeAsync(d_sp, stream));
        MGB_CUDA(cudaFreeAsync(d_w4, stream));
    }
    return MGB_OK;
}

extern "C" int mgb_fk(const mgb_model* m, const float* d_frames, int64_t n_frames, const int32_t* h_nodes, int32_t n_out,
                      float* d_pos, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(m && h_nodes && n_out > 0, "null argument");
    MGB_REQUIRE(n_frames == 0 || (d_frames && d_pos), "null device pointer");
    if (n_frames == 0) return MGB_OK;
    // one chain root -> node per requested output
    std::vector<FkChainStep> steps;
    std::vector<int> first(n_out + 1, 0);
    for (int j = 0; j < n_out; ++j) {
        int nd = h_nodes[j];
        MGB_REQUIRE(nd >= 0 && nd < m->n_nodes, "node index %d out of range", nd);
        std::vector<int> chain;
        for (; nd >= 0; nd = m->parent[nd]) chain.push_back(nd);
        first[j] = (int)steps.size();
        for (int i = (int)chain.size() - 1; i >= 0; --i) {
            const int c = chain[i];
            FkChainStep st;
            st.channel = m->channel[c];
            for (int q = 0; q < 3; ++q) st.off[q] = (float)m->offset[c * 3 + q];
            bool ident = true;
            for (int q = 0; q < 4; ++q) {
                const double r = (m->type[c] == MGB_NODE_END_SITE) ? (q == 0 ? 1.0 : 0.0) : m->fixed_quat[(size_t)4 * c + q];
                st.q[q] = (float)r;
                if (fabs(fabs(r) - (q == 0 ? 1.0 : 0.0)) > 0.0) ident = false;
            }
            if (m->type[c] == MGB_NODE_ROOT) st.kind = 0;
            else if (m->type[c] == MGB_NODE_JOINT && m->channel[c] >= 0) st.kind = 1;
            else st.kind = ident ? 3 : 2;
            steps.push_back(st);
        }
    }
    first[n_out] = (int)steps.size();
    const size_t plan_bytes = sizeof(FkChainStep) * steps.size() + sizeof(int) * (n_out + 1);
    MGB_REQUIRE(plan_bytes <= 48 * 1024, "FK plan of %d steps too large", (int)steps.size());
    unsigned char* d_plan = nullptr;
    MGB_CUDA(cudaMallocAsync((void**)&d_plan, plan_bytes, stream));
    // pageable sources: the runtime stages them before the calls return, so the vectors may go out of scope without a synchronisation
    MGB_CUDA(cudaMemcpyAsync(d_plan, steps.data(), sizeof(FkChainStep) * steps.size(), cudaMemcpyHostToDevice, stream));
    MGB_CUDA(cudaMemcpyAsync(d_plan + sizeof(FkChainStep) * steps.size(), first.data(), sizeof(int) * (n_out + 1), cudaMemcpyHostToDevice, stream));
    // tile of fkb frames and as many threads, each walking the n_out chains of a frame one after the other (measured: 0.98 ms per
    // 32768 x 50 frames x 10 joints; one thread per (frame, joint) in 512-thread CTAs of 64 frames: 1.11 ms)
    int fkb = FK_BLOCK;
    size_t smem = 0;
    for (;; fkb >>= 1) {
        smem = ((size_t)fkb * (m->P | 1) + (size_t)fkb * ((n_out * 3) | 1) + 2) * sizeof(float) + plan_bytes;
        if (smem <= 100 * 1024 || fkb == 32) break;
    }
    MGB_REQUIRE(smem <= 200 * 1024, "frame tile too large for shared memory");
    const int threads = fkb;
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(mgb_fk_chain_kernel, configured, 200 * 1024)) return rc;
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, mgb_fk_chain_kernel, threads, smem));
    long long n_tiles = (n_frames + fkb - 1) / fkb;
    long long grid = (long long)sm_count() * (occ > 0 ? occ : 1);
    if (grid > n_tiles) grid = n_tiles;
    mgb_fk_chain_kernel<<<(unsigned)grid, threads, smem, stream>>>(d_frames, m->P, n_frames, reinterpret_cast<const FkChainStep*>(d_plan),
                                                                    reinterpret_cast<const int*>(d_plan + sizeof(FkChainStep) * steps.size()),
                                                                    (int)steps.size(), n_out, d_pos);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    MGB_CUDA(cudaFreeAsync(d_plan, stream));
    return MGB_OK;
}

// ---------------------------------------------------------------------------
// diagnostic: achievable fp32 FMA rate of this device (denominator of the
// compute-bound fused score kernel's roofline)
// ---------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) mgb_fma_peak_kernel(float* out, int iters, float a, float b) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;  // never true in practice; keeps the chain alive
}
}  // namespace

extern "C" int mgb_fp32_fma_peak(int32_t device, double* tflops) {
    MGB_REQUIRE(tflops, "null argument");
    MGB_CUDA(cudaSetDevice(device));
    float* d = nullptr;
    MGB_CUDA(cudaMalloc((void**)&d, sizeof(float)));
    cudaEvent_t e0, e1;
    MGB_CUDA(cudaEventCreate(&e0));
    MGB_CUDA(cudaEventCreate(&e1));
    const int iters = 4096, blocks = sm_count() * 8;
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        MGB_CUDA(cudaEventRecord(e0, 0));
        mgb_fma_peak_kernel<<<blocks, 256>>>(d, iters, 0.999f, 1e-3f);
        MGB_CUDA(cudaEventRecord(e1, 0));
        MGB_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        MGB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double fl = 2.0 * 16.0 * iters * 256.0 * blocks;
        if (rep > 0 && ms > 0.f) best = std::max(best, fl / (ms * 1e-3) * 1e-12);
    }
    mgb_count_launch(6);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return MGB_OK;
}
