// Tensor-core (tcgen05 / TMEM) kernels of libmgb200.so, sm_100a.
//
// K2-TC  mgb_backproject_tc: coeffs[n, M] = s[n, :K] . E'[K, M] + mean'[M]
//        (MotionPrimitive.back_project_spatial_coeffs, reference motion_primitive.py:236-256)
//        as a 3xTF32 GEMM: every fp32 operand is split into two TF32 numbers (hi + lo) and
//        the product is accumulated as a_hi.b_hi + a_lo.b_hi + a_hi.b_lo in the fp32 TMEM
//        accumulator, which restores ~fp32 accuracy (2^-21 relative per product) - the
//        north star's 1e-4 parity bound cannot be met by single-pass TF32 (2^-11).
//
// Structure of every kernel here (one persistent CTA per SM, 192 threads):
//   warp 0      B producer: the constant operand (basis / GMM tables) was packed on the host
//               into per-stage images that already have the UMMA canonical K-major layout,
//               so a stage is ONE 1-D TMA bulk copy (UBLKCP) from L2 into shared memory,
//               ring of NST stages on full/empty mbarriers.
//   warp 1      MMA issuer: one lane issues tcgen05.mma (SASS UTCMMA), 128 x NC x 8 per
//               instruction, accumulators double-buffered in TMEM (2 x 256 columns) so the
//               epilogue of N-chunk c overlaps the MMAs of chunk c + 1; tcgen05.commit
//               releases the B stage / publishes the accumulator.
//   warps 2-5   A producers + epilogue: thread r owns candidate row r of the 128-row tile
//               (= TMEM lane r): writes the hi/lo split of its latent vector into the
//               canonical A layout, later reads its accumulator row with tcgen05.ld (LDTM).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "mgb_internal.cuh"
#include "mgb_interp.cuh"
#include "mgb_tc.cuh"

#define MGB_HAVE_TC 1
#ifndef MGB_ST
#define MGB_ST(p, v) (*(p) = (v))
#endif

namespace {

constexpr int TC_THREADS = 320;    // warp 0 B producer, warp 1 MMA issuer, warps 2-5 / 6-9 epilogue groups of TMEM buffer 0 / 1
constexpr int TC_M = 128;        // rows (candidates) per tile = TMEM lanes
constexpr int TC_MAX_STAGES = 4;
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_BUF_COLS = 256;  // one accumulator buffer

// packed constant operand (device)
struct TcB {
    const float* data;
    int Kp;  // K padded to a multiple of 8
    int Np;  // N padded to a multiple of 16
    int NC;  // N-chunk (multiple of 16, <= 256)
    int KC;  // K-chunk (multiple of 8)
    __host__ __device__ int n_nc() const { return (Np + NC - 1) / NC; }
    __host__ __device__ int n_kc() const { return (Kp + KC - 1) / KC; }
    __host__ __device__ int wc(int c) const { return min(NC, Np - c * NC); }
    __host__ __device__ int wk(int kc) const { return min(KC, Kp - kc * KC); }
    // float offset of stage image (c, kc): [hi block wc x wk][lo block wc x wk]
    __host__ __device__ size_t off(int c, int kc) const { return 2 * ((size_t)Kp * c * NC + (size_t)wc(c) * kc * KC); }
    __host__ __device__ size_t total() const { return 2 * (size_t)Kp * Np; }
};

struct TcBackprojectParams {
    TcB B;
    const float* s;     // [n, D]
    const float* mean;  // [M]
    float* out;         // [n, M]
    long long n;
    int D, K, M;
    int nst;            // B stages
    int W;              // epilogue staging width: a warp writes W consecutive floats of a row back to back
    int nbuf;           // staging buffers (1 or 2)
};

// shared-memory carve-up, identical on host and device
struct TcSmem {
    size_t a_hi, a_lo, b_stage, out_stage, bars, tmem_ptr, total;
    size_t stage_bytes;
};
// W: columns of the epilogue's row staging (0 = none); a staged row has W + 4 floats (16-byte aligned, conflict-free 128-bit accesses)
__host__ __device__ inline TcSmem tc_smem_plan(int Kp, int NC, int KC, int nst, int W = 0, int nbuf = 1) {
    TcSmem p;
    size_t off = 0;
    p.a_hi = off; off += (size_t)TC_M * Kp * 4;
    p.a_lo = off; off += (size_t)TC_M * Kp * 4;
    p.stage_bytes = (size_t)2 * NC * KC * 4;
    p.b_stage = off; off += p.stage_bytes * nst;
    p.out_stage = off; off += (W > 0) ? (size_t)nbuf * TC_M * (W + 4) * 4 : 0;
    p.bars = off; off += 8 * (2 * TC_MAX_STAGES + 2 + 6);
    p.tmem_ptr = off; off += 16;
    p.total = off;
    return p;
}

struct TcBars {
    unsigned long long* b_full;   // [TC_MAX_STAGES]
    unsigned long long* b_empty;  // [TC_MAX_STAGES]
    unsigned long long* a_full;
    unsigned long long* a_empty;
    unsigned long long* t_full;   // [2 owners][2 buffers]: accumulator buffer b holds a finished chunk for consumer group `owner`
    unsigned long long* t_empty;  // [2 buffers]
};

__device__ __forceinline__ TcBars tc_bars(unsigned char* base) {
    TcBars b;
    unsigned long long* p = reinterpret_cast<unsigned long long*>(base);
    b.b_full = p;
    b.b_empty = p + TC_MAX_STAGES;
    b.a_full = p + 2 * TC_MAX_STAGES;
    b.a_empty = p + 2 * TC_MAX_STAGES + 1;
    b.t_full = p + 2 * TC_MAX_STAGES + 2;
    b.t_empty = p + 2 * TC_MAX_STAGES + 6;
    return b;
}

// ---- role bodies shared by the tensor-core kernels ---------------------------------------

// The constant operands of one tile pass: up to two packed matrices that share the A tile
// (e.g. the contracted constraint quads, then the GMM precision factors), streamed in order.
struct TcSegs {
    TcB seg[3];
    int n;
    int samp_seg;                       // >= 0: this operand (one N-chunk per GMM component) is multiplied first, by the tile's own
                                        // component only, against the first of TWO fills of the A tile (z, then s = mu + L z)
    int tile_first[MGB_MAX_GMM + 1];    // sampling mode: first tile of every component (tiles never straddle components)
    int G;
    int il_q, il_g;                     // both >= 0: the chunks of these two operands are issued INTERLEAVED (q0 q1 g.. q2 g.. ...):
                                        // il_q's chunks are consumed by epilogue group 0, il_g's by group 1
};

// Accumulators: every chunk of a tile, whatever its operand, goes to the two 256-column TMEM buffers in turn (chunk number
// & 1).  A chunk has an OWNER - the epilogue group that consumes it: 0 = group 0 (quads), 1 = group 1 (GMM), 2 = both (the
// sampling chunk: both groups read their half of s and refill the A tile; group 0 releases the buffer).  Every owner has its
// own "full" barrier per buffer, so a group only ever waits on barriers all of whose phases it observes (a parity wait is
// ambiguous for a waiter that can fall two phases behind).
enum { TC_OWN_G0 = 0, TC_OWN_G1 = 1, TC_OWN_BOTH = 2 };

// chunk schedule of one tile, shared by the B producer, the MMA issuer and the consumer groups: f(segment, chunk, owner,
// is_sampling).  Interleaved mode: the quad stream is issued one chunk ahead and the GMM chunks are spread evenly between
// the quad chunks, so that the quad consumer (the constraint program: the longer of the two) always has its next chunk ready.
template <typename F>
__device__ __forceinline__ void tc_for_each_chunk(const TcSegs& S, int gcomp, F f) {
    const bool il = S.il_q >= 0 && S.il_g >= 0;
    for (int sg = 0; sg < S.n; ++sg) {
        if (sg == S.samp_seg) {
            f(sg, gcomp, (int)TC_OWN_BOTH, true);
        } else if (!il) {
            for (int c = 0; c < S.seg[sg].n_nc(); ++c) f(sg, c, sg == S.il_g ? (int)TC_OWN_G1 : (int)TC_OWN_G0, false);
        }
    }
    if (il) {
        const int nq = S.seg[S.il_q].n_nc(), ng = S.seg[S.il_g].n_nc();
        int qi = 0, gi = 0;
        if (nq > 0) f(S.il_q, qi++, (int)TC_OWN_G0, false);
        while (qi < nq || gi < ng) {
            if (qi < nq) f(S.il_q, qi++, (int)TC_OWN_G0, false);
            // the GMM chunks are spread evenly over the quad chunks (ng / nq per quad chunk), the rest after the last one
            while (gi < ng && (qi >= nq || (long long)gi * nq < (long long)(qi - 1) * ng)) f(S.il_g, gi++, (int)TC_OWN_G1, false);
        }
    }
}

// phase bit of "full" barrier (owner, buf), kept as one bit each of a 4-bit mask
__device__ __forceinline__ unsigned int tc_full_phase(unsigned int& mask, int owner, unsigned int buf) {
    const unsigned int bit = 1u << (2 * owner + buf);
    const unsigned int ph = (mask & bit) ? 1u : 0u;
    mask ^= bit;
    return ph;
}

// component of tile T (sampling mode)
__device__ __forceinline__ int tc_tile_comp(const int* tile_first, int G, long long T) {
    int g = 0;
    while (g + 1 < G && T >= tile_first[g + 1]) ++g;
    return g;
}

// warp 0, one lane: stream every stage image of every operand once per tile.  The stage ring position (st, ph) and the
// source pointer advance by additions only: this lane and the MMA lane below are single threads on the critical path.
__device__ __forceinline__ void tc_b_producer(const TcSegs& S, unsigned char* sm_stage, size_t stage_bytes, int nst, const TcBars& bars,
                                              long long my_tiles, const int* tile_first = nullptr) {
    int st = 0;
    unsigned int ph = 0;
    unsigned char* dst = sm_stage;
    for (long long t = 0; t < my_tiles; ++t) {
        const int gcomp = (S.samp_seg >= 0) ? tc_tile_comp(tile_first, S.G, blockIdx.x + t * gridDim.x) : 0;
        tc_for_each_chunk(S, gcomp, [&](int sg, int c, int, bool) {
            const TcB& B = S.seg[sg];
            const int KC = B.KC, Kp = B.Kp;
            const int wc = B.wc(c);
            const float* src = B.data + B.off(c, 0);
            for (int k0 = 0; k0 < Kp; k0 += KC) {
                const int wk = min(KC, Kp - k0);
                mgb_mbar_wait(&bars.b_empty[st], ph ^ 1u);
                const unsigned int bytes = (unsigned int)(2 * wc * wk * 4);
                mgb_mbar_expect_tx(&bars.b_full[st], bytes);
                mgb_bulk_g2s(dst, src, bytes, &bars.b_full[st]);
                src += 2 * wc * wk;      // stage images of one N-chunk are consecutive: [hi wc x wk][lo wc x wk]
                dst += stage_bytes;
                if (++st == nst) { st = 0; ph ^= 1u; dst = sm_stage; }
            }
        });
    }
}

// Cycle counters of CTA 0 (tools/tc_cycles.py) exist only in a -DMGB_TC_PROFILE build; the product build carries
// neither the counters nor the clock reads.
#ifdef MGB_TC_PROFILE
__device__ long long g_tc_cycles[16];
#endif

// warp 1, ALL lanes converged, the MMAs and commits issued by one elected lane (a loop run by a single lane inside a
// divergent branch costs ~60 % more issue time per K-step: tools/micro/mma_issue2.cu): all MMAs of this CTA.  Chunk number cc accumulates into TMEM buffer cc & 1; the finished chunk is
// published on the "full" barrier(s) of its owner(s).  all_own0: every chunk is published on owner 0's barriers (the
// back-projection kernel, whose epilogue warps all drain every chunk).
__device__ __forceinline__ void tc_mma_issuer(const TcSegs& S, uint32_t a_hi_addr, uint32_t a_lo_addr, uint32_t stage_addr, size_t stage_bytes,
                                              int nst, const TcBars& bars, uint32_t tmem_base, long long my_tiles,
                                              const int* tile_first = nullptr, bool all_own0 = false) {
    unsigned int cc = 0, fill = 0;
    const bool leader = mgb_elect_one();
    const uint32_t A_LBO = TC_M * 16, SBO = 128;
    // descriptors differ only in their 14-bit start-address field (bytes >> 4): build them once, then add
    const uint64_t da_hi0 = mgb_umma_desc(a_hi_addr, A_LBO, SBO), da_lo0 = mgb_umma_desc(a_lo_addr, A_LBO, SBO);
    const uint32_t stage16 = (uint32_t)(stage_bytes >> 4), ring16 = stage_addr >> 4;
    int st = 0;
    unsigned int ph = 0;
    uint32_t b16 = ring16;               // (address of stage st) >> 4
#ifdef MGB_TC_PROFILE
    const bool dbgi = S.samp_seg >= 0 && blockIdx.x == 0;
    long long w_te = 0, w_bf = 0, w_af = 0, tq = 0;
#define MGB_TC_W(acc, stmt) { if (dbgi) tq = clock64(); stmt; if (dbgi) acc += clock64() - tq; }
#else
#define MGB_TC_W(acc, stmt) { stmt; }
#endif
    for (long long t = 0; t < my_tiles; ++t) {
        const int gcomp = (S.samp_seg >= 0) ? tc_tile_comp(tile_first, S.G, blockIdx.x + t * gridDim.x) : 0;
        MGB_TC_W(w_af, mgb_mbar_wait(bars.a_full, fill & 1u))
        ++fill;
        mgb_tc_fence_after();
        tc_for_each_chunk(S, gcomp, [&](int sg, int c, int owner, bool samp) {
            const TcB& B = S.seg[sg];
            const int KC = B.KC, Kp = B.Kp;
            const int wc = B.wc(c);
            const unsigned int buf = cc & 1u;
            MGB_TC_W(w_te, mgb_mbar_wait(&bars.t_empty[buf], ((cc >> 1) & 1u) ^ 1u))
            ++cc;
            mgb_tc_fence_after();
            const uint32_t d_addr = tmem_base + buf * TC_BUF_COLS;
            const uint32_t idesc = mgb_umma_idesc_tf32(TC_M, wc);
            const uint64_t db0 = mgb_umma_desc(0u, (uint32_t)wc * 16, SBO);
            uint32_t acc = 0;
            uint32_t a16 = 0;            // (byte offset of the K-slice in the A tile) >> 4: 8 columns of K = 2 * A_LBO bytes
            for (int k0 = 0; k0 < Kp; k0 += KC) {
                const int wk = min(KC, Kp - k0);
                MGB_TC_W(w_bf, mgb_mbar_wait(&bars.b_full[st], ph))
                mgb_tc_fence_after();
                const uint32_t lo16 = (uint32_t)(wc * wk) >> 2;      // the lo block follows the hi block: wc * wk * 4 bytes
                uint32_t bj16 = b16;
                for (int j = 0; j < wk; j += 8, a16 += 2 * A_LBO / 16, bj16 += 2 * wc) {
                    const uint64_t da_hi = da_hi0 + a16, da_lo = da_lo0 + a16;
                    const uint64_t db_hi = db0 + bj16, db_lo = db_hi + lo16;
                    if (leader) {
                        mgb_umma_tf32(d_addr, da_lo, db_hi, idesc, acc);  // small terms first
                        mgb_umma_tf32(d_addr, da_hi, db_lo, idesc, 1u);
                        mgb_umma_tf32(d_addr, da_hi, db_hi, idesc, 1u);
                    }
                    acc = 1u;
                }
                if (leader) mgb_umma_commit(&bars.b_empty[st]);
                __syncwarp();
                b16 += stage16;
                if (++st == nst) { st = 0; ph ^= 1u; b16 = ring16; }
            }
            if (leader) {
                if (all_own0 || owner != TC_OWN_G1) mgb_umma_commit(&bars.t_full[buf]);
                if (!all_own0 && owner != TC_OWN_G0) mgb_umma_commit(&bars.t_full[2 + buf]);
            }
            __syncwarp();
            if (samp) {  // the A tile is refilled with the drawn latent vectors before the other operands
                MGB_TC_W(w_af, mgb_mbar_wait(bars.a_full, fill & 1u))
                ++fill;
                mgb_tc_fence_after();
            }
        });
        if (leader) mgb_umma_commit(bars.a_empty);
        __syncwarp();
    }
#ifdef MGB_TC_PROFILE
    if (dbgi && leader) { g_tc_cycles[13] += w_te; g_tc_cycles[14] += w_bf; g_tc_cycles[15] += w_af; }
#endif
#undef MGB_TC_W
}

// one row of A: hi / lo split of `cnt` values (zero padded to Kp), written as float4 chunks
// into the canonical layout (threads of a warp write 512 contiguous bytes per chunk)
template <typename Load>
__device__ __forceinline__ void tc_write_a_row(unsigned char* a_hi, unsigned char* a_lo, int row, int Kp, int cnt, Load load) {
    for (int k0 = 0; k0 < Kp; k0 += 4) {
        float h[4], l[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float v = (k0 + q < cnt) ? load(k0 + q) : 0.f;
            mgb_tf32_split(v, h[q], l[q]);
        }
        const uint32_t off = mgb_canon_off((uint32_t)row, (uint32_t)k0, TC_M);
        *reinterpret_cast<float4*>(a_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<float4*>(a_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
    }
}


// staging [32][SS] -> global rows: row after row, consecutive instructions cover consecutive bytes of a row;
// 8 independent shared loads are in flight before their 8 stores issue
template <typename V>
__device__ __forceinline__ void tc_store_rows(const float* stg, int SS, float* obase, int M, int n_rows, int ncol, int W, int lane) {
    constexpr int VW = sizeof(V) / 4;
    const int steps = max(1, W / (32 * VW));          // store instructions per row (power of two)
    const int sh = 31 - __clz(steps);
    const int total = 16 << sh;  // 16 rows per warp
    for (int i0 = 0; i0 < total; i0 += 8) {
        V tmp[8];
        bool ok[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int i = i0 + u, r = i >> sh, ci = VW * (lane + 32 * (i & (steps - 1)));
            ok[u] = r < n_rows && ci < ncol;
            if (ok[u]) tmp[u] = *reinterpret_cast<const V*>(stg + r * SS + ci);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int i = i0 + u, r = i >> sh, ci = VW * (lane + 32 * (i & (steps - 1)));
            if (ok[u]) __stcs(reinterpret_cast<V*>(obase + (size_t)r * M + ci), tmp[u]);
        }
    }
}

// ---------------------------------------------------------------------------
// K2-TC
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TC_THREADS, 1) mgb_backproject_tc_kernel(const __grid_constant__ TcBackprojectParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const TcSmem plan = tc_smem_plan(p.B.Kp, p.B.NC, p.B.KC, p.nst, p.W, p.nbuf);
    const TcBars bars = tc_bars(smem + plan.bars);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + plan.tmem_ptr);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_tiles = (p.n + TC_M - 1) / TC_M;
    const long long my_tiles = (n_tiles > blockIdx.x) ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (threadIdx.x == 0) {
        for (int i = 0; i < TC_MAX_STAGES; ++i) {
            mgb_mbar_init(&bars.b_full[i], 1);
            mgb_mbar_init(&bars.b_empty[i], 1);
        }
        mgb_mbar_init(bars.a_full, 128);
        mgb_mbar_init(bars.a_empty, 1);
        for (int i = 0; i < 2; ++i) {
            mgb_mbar_init(&bars.t_full[i], 1);
            mgb_mbar_init(&bars.t_empty[i], 256);
        }
        mgb_fence_barrier_init();
    }
    if (warp == 1) mgb_tmem_alloc(tmem_ptr, TC_TMEM_COLS);
    mgb_tc_fence_before();
    __syncthreads();
    mgb_tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {
            TcSegs S;
            S.seg[0] = p.B;
            S.n = 1;
            S.samp_seg = -1;
            S.il_q = S.il_g = -1;
            tc_b_producer(S, smem + plan.b_stage, plan.stage_bytes, p.nst, bars, my_tiles);
        }
    } else if (warp == 1) {
        {
            TcSegs S;
            S.seg[0] = p.B;
            S.n = 1;
            S.samp_seg = -1;
            S.il_q = S.il_g = -1;
            tc_mma_issuer(S, mgb_smem_u32(smem + plan.a_hi), mgb_smem_u32(smem + plan.a_lo), mgb_smem_u32(smem + plan.b_stage),
                          plan.stage_bytes, p.nst, bars, tmem_base, my_tiles, nullptr, true);
        }
    } else {
        const int wq = warp & 3;            // TMEM lane quarter of this warp
        const int ew = warp - 2;            // epilogue warp 0..7
        const int h = ew >> 2;              // which half of a column group this warp moves out of TMEM
        const int row = 32 * wq + lane;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * wq) << 16);
        const int W = p.W, SS = W + 4;
        float* stg0 = reinterpret_cast<float*>(smem + plan.out_stage);  // nbuf x [128][W + 4]
        const int n_nc = p.B.n_nc();
        const int M = p.M;
        const int vec = (M % 4 == 0) ? 4 : (M % 2 == 0) ? 2 : 1;
        unsigned int cc = 0, gi = 0;
        // A row of tile t: [s[:K] | 1] so that row K of the packed operand (the mean) is added by the GEMM itself
        auto write_a = [&](long long t) {
            const long long g = (blockIdx.x + t * gridDim.x) * TC_M + row;
            const bool valid = g < p.n;
            mgb_mbar_wait(bars.a_empty, (unsigned int)(t & 1) ^ 1u);
            const float* srow = p.s + (size_t)(valid ? g : 0) * p.D;
            const int K = p.K;
            tc_write_a_row(smem + plan.a_hi, smem + plan.a_lo, row, p.B.Kp, valid ? K + 1 : 0, [&](int k) { return k < K ? __ldg(srow + k) : 1.f; });
            mgb_fence_proxy_async();
            mgb_mbar_arrive(bars.a_full);
        };
        if (h == 0 && my_tiles > 0) write_a(0);
        for (long long t = 0; t < my_tiles; ++t) {
            const long long tile0 = (blockIdx.x + t * gridDim.x) * TC_M;
            // store phase: this warp writes rows 16 ew .. 16 ew + 15 of the tile, so the CTA's 8 warps cover one
            // compact run of 128 rows x W columns at a time (what the memory system writes back best)
            const int n_rows = (int)max((long long)0, min((long long)16, p.n - tile0 - 16 * ew));
            for (int c = 0; c < n_nc; ++c, ++cc) {
                // refill A for the next tile at this tile's last chunk: its MMAs are all issued by then
                if (h == 0 && c == n_nc - 1 && t + 1 < my_tiles) write_a(t + 1);
                const unsigned int buf = cc & 1u;
                mgb_mbar_wait(&bars.t_full[buf], (cc >> 1) & 1u);
                mgb_tc_fence_after();
                const int wc = p.B.wc(c);
                for (int w0 = 0; w0 < wc; w0 += W, ++gi) {
                    const int col0 = c * p.B.NC + w0;   // first column of this group
                    const int ncol = min(W, M - col0);  // valid columns of the group (CTA-uniform)
                    const bool last = (w0 + W >= wc) || (col0 + W >= M);
                    float* stg = stg0 + (size_t)(p.nbuf == 2 ? (gi & 1u) : 0u) * TC_M * SS;
                    if (ncol > 0) {
                        if (p.nbuf == 1) asm volatile("bar.sync 1, 256;" ::: "memory");  // stores of the previous group have read the staging
                        // TMEM -> registers -> staging, thread-per-row: this warp moves columns [h W/2, (h+1) W/2) of its 32 rows
                        for (int b0 = h * (W >> 1); b0 < (h + 1) * (W >> 1) && b0 < ncol; b0 += 32) {
                            float v[32];
                            mgb_tmem_ld32(lane_addr + buf * TC_BUF_COLS + w0 + b0, v);
#pragma unroll
                            for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(stg + row * SS + b0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                        }
                    }
                    if (last) {  // accumulator buffer drained: hand it back before the stores
                        mgb_tc_fence_before();
                        mgb_mbar_arrive_relaxed(&bars.t_empty[buf]);
                    }
                    if (ncol <= 0) break;
                    asm volatile("bar.sync 1, 256;" ::: "memory");  // staging of this group complete
                    float* obase = p.out + (size_t)(tile0 + 16 * ew) * M + col0;
                    const float* srows = stg + (size_t)16 * ew * SS;
                    if (vec == 4) {
                        tc_store_rows<float4>(srows, SS, obase, M, n_rows, ncol, W, lane);
                    } else if (vec == 2) {
                        tc_store_rows<float2>(srows, SS, obase, M, n_rows, ncol, W, lane);
                    } else {
                        for (int r = 0; r < n_rows; ++r)
                            for (int ci = lane; ci < ncol; ci += 32) __stcs(obase + (size_t)r * M + ci, srows[r * SS + ci]);
                    }
                    if (last) {
                        ++gi;
                        break;
                    }
                }
            }
        }
    }
    mgb_tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        mgb_tc_fence_after();
        mgb_tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
}

}  // namespace

// ===========================================================================
// host side
// ===========================================================================
namespace {

// cvt.rna.tf32.f32 on the host: round to nearest, ties away from zero, 13 bits dropped
float tf32_rna_host(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7F800000u) == 0x7F800000u) return x;
    u = (u + 0x1000u) & 0xFFFFE000u;
    float r;
    memcpy(&r, &u, 4);
    return r;
}

// Pack logical B[k][n] (fp64, k < K, n < N; element(k, n) = get(k, n)) into per-stage images
// in the UMMA canonical K-major layout, hi block then lo block (TcB::off).
template <typename Get>
std::vector<float> tc_pack_b(TcB& B, int K, int N, Get get) {
    std::vector<float> out(B.total(), 0.f);
    for (int c = 0; c < B.n_nc(); ++c) {
        const int wc = B.wc(c);
        for (int kc = 0; kc < B.n_kc(); ++kc) {
            const int wk = B.wk(kc);
            float* hi = out.data() + B.off(c, kc);
            float* lo = hi + (size_t)wc * wk;
            for (int nn = 0; nn < wc; ++nn) {
                const int n = c * B.NC + nn;
                for (int kk = 0; kk < wk; ++kk) {
                    const int k = kc * B.KC + kk;
                    const double v = (k < K && n < N) ? get(k, n) : 0.0;
                    const float h = tf32_rna_host((float)v);
                    const float l = tf32_rna_host((float)(v - (double)h));
                    const size_t o = mgb_canon_off((uint32_t)nn, (uint32_t)kk, (uint32_t)wc) / 4;
                    hi[o] = h;
                    lo[o] = l;
                }
            }
        }
    }
    return out;
}

constexpr size_t TC_SMEM_BUDGET = 220 * 1024;
constexpr size_t TC_SCORE_SMEM_MAX = 227 * 1024;   // cudaDevAttrMaxSharedMemoryPerBlockOptin on sm_100

// choose the chunking for an operand with K rows / N columns next to a resident A tile
bool tc_plan_b(TcB& B, int K, int N, int* nst_out, int* W_out = nullptr, int* nbuf_out = nullptr) {
    B.Kp = (K + 7) & ~7;
    B.Np = (N + 15) & ~15;
    B.NC = std::min(256, B.Np);
    // epilogue staging (width, buffers) in order of preference, then B stage depth / size
    const int cand[5][2] = {{128, 2}, {256, 1}, {128, 1}, {64, 2}, {64, 1}};
    int forced_w = 0, forced_nb = 0;
#ifdef MGB_TC_PROFILE
    if (const char* e = getenv("MGB_TC_W")) forced_w = atoi(e);
    if (const char* e = getenv("MGB_TC_NBUF")) forced_nb = atoi(e);
#endif
    for (int wi = 0; wi < (W_out ? 5 : 1); ++wi) {
        const int W = W_out ? cand[wi][0] : 0, nbuf = W_out ? cand[wi][1] : 1;
        if (W_out && ((forced_w && W != forced_w) || (forced_nb && nbuf != forced_nb))) continue;
        const int KCs[2] = {16, 8};
        for (int ki = 0; ki < 2; ++ki) {
            B.KC = std::min(KCs[ki], B.Kp);
            for (int nst = TC_MAX_STAGES; nst >= 2; --nst) {
                if (tc_smem_plan(B.Kp, B.NC, B.KC, nst, W, nbuf).total + 1024 <= TC_SMEM_BUDGET) {
                    *nst_out = nst;
                    if (W_out) *W_out = W;
                    if (nbuf_out) *nbuf_out = nbuf;
                    return true;
                }
            }
        }
    }
    return false;
}

int tc_num_sms() { return mgb_num_sms(); }

}  // namespace

// Packs E' once per model (hi/lo TF32 stage images); thread-compatible with the rest of the
// handle: called from mgb_model_create, so the handle stays immutable afterwards.
int mgb_backproject_tc_prepare(mgb_model* m) {
    TcB B;
    int nst = 0;
    int W = 0, nbuf = 1;
    if (!tc_plan_b(B, m->K + 1, m->M, &nst, &W, &nbuf)) {
        mgb_set_error("tensor-core back-projection: K = %d does not fit shared memory", m->K);
        return MGB_E_UNSUPPORTED;
    }
    const int M = m->M;
    const int K = m->K;  // row K of the operand is the mean: the A tile carries a column of ones
    std::vector<float> packed = tc_pack_b(B, K + 1, M, [&](int k, int n) { return k < K ? m->E[(size_t)k * M + n] : m->mean[n]; });
    MGB_CUDA(cudaMalloc((void**)&m->d_Bsplit, packed.size() * sizeof(float)));
    MGB_CUDA(cudaMemcpy(m->d_Bsplit, packed.data(), packed.size() * sizeof(float), cudaMemcpyHostToDevice));
    m->tc_Kp = B.Kp; m->tc_Np = B.Np; m->tc_NC = B.NC; m->tc_KC = B.KC; m->tc_nst = nst; m->tc_W = W; m->tc_nbuf = nbuf;
    return MGB_OK;
}

int mgb_backproject_tc(const mgb_model* m, const float* d_s, int64_t n, float* d_coeffs, cudaStream_t stream) {
    MGB_REQUIRE(m->d_Bsplit != nullptr, "tensor-core back-projection operand was not prepared for this model");
    MGB_REQUIRE((reinterpret_cast<uintptr_t>(d_coeffs) & 15u) == 0, "d_coeffs must be 16-byte aligned");
    TcBackprojectParams p;
    p.B.data = m->d_Bsplit;
    p.B.Kp = m->tc_Kp; p.B.Np = m->tc_Np; p.B.NC = m->tc_NC; p.B.KC = m->tc_KC;
    p.nst = m->tc_nst;
    p.W = m->tc_W;
    p.nbuf = m->tc_nbuf;
    p.s = d_s; p.mean = m->d_mean; p.out = d_coeffs;
    p.n = n; p.D = m->D; p.K = m->K; p.M = m->M;
    const size_t smem = tc_smem_plan(p.B.Kp, p.B.NC, p.B.KC, p.nst, p.W, p.nbuf).total;
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(mgb_backproject_tc_kernel, configured, (int)TC_SMEM_BUDGET)) return rc;
    const long long n_tiles = (n + TC_M - 1) / TC_M;
    const int grid = (int)std::min<long long>(n_tiles, tc_num_sms());
    mgb_backproject_tc_kernel<<<grid, TC_THREADS, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

// ===========================================================================
// K6-TC: fused keyframe score for large latent spaces (D > 32) on the tensor cores.
//
// Per 128-candidate tile one A operand [s | 1] (hi / lo TF32 split) feeds two packed
// constant operands in turn:
//   quads : the spatial basis pre-contracted with the B-spline weights of every constrained
//           time slot (mgb_constraints_pack), bias (mean) in the row that meets the 1
//           -> pose channels at the constrained keyframes        (back-projection + spline)
//   GMM   : the precision Cholesky factors P_g of all components side by side, bias -mu_g.P_g
//           -> y_g = (s - mu_g) . P_g, log-density by logsumexp   (naturalness term)
// Epilogue group 0 (warps 2-5) prepares A, drains the quad chunk into shared memory and runs
// the constraint program (forward kinematics + terms) on it; group 1 (warps 6-9) drains the
// GMM chunks (sum of squares per component straight from TMEM, thread = candidate) while the
// next chunk is being multiplied.  Replaces the same reference functions as mgb_score.cu.
// ===========================================================================
namespace {

struct TcScoreParams {
    ScoreParams sp;
    TcSegs segs;
    int seg_q, seg_g;  // index of the quad / GMM operand in segs, -1 if absent
    int nst;
    int Kp;
    int QS;            // floats per candidate row of the quad-value area (also the raw-sample staging)
    const int* phase;  // [quad chunks + 1] first op of every chunk's part of the program
    int ops_in_smem;   // program + constants staged in shared memory (else read through L1)
    int draw_ahead;    // sampling mode, short constraint programs: group 0 draws the next tile's normals into the quad-value area
                       // while it waits for this tile's GMM chunks (the draw is ~9 k of a ~99 k-cycle tile otherwise serial with the pipe)
    int CW;            // GMM columns per component (a multiple of 8; components are packed densely and may straddle N-chunks)
    int seg_s;         // sampling operand (fused K1), -1 if the candidates come from memory
    int CWs;           // its columns per component
    long long n_tiles; // tiles of this launch (sampling mode: per-component tiling)
};

struct TcScoreSmem {
    size_t a_hi, a_lo, b_stage, qv, ops, cst, lp, redc, redi, bars, sbar, tmem_ptr, flag, tfirst, total;
    size_t stage_bytes;
};
__host__ __device__ inline TcScoreSmem tc_score_smem_plan(int Kp, size_t stage_bytes, int nst, int QS, int n_ops, int n_cst) {
    TcScoreSmem p;
    size_t off = 0;
    p.a_hi = off; off += (size_t)TC_M * Kp * 4;
    p.a_lo = off; off += (size_t)TC_M * Kp * 4;
    p.stage_bytes = stage_bytes;
    p.b_stage = off; off += stage_bytes * nst;
    p.qv = off; off += ((size_t)TC_M * QS * 4 + 15) & ~(size_t)15;
    p.ops = off; off += (size_t)n_ops * sizeof(MgbOp);
    p.cst = off; off += (size_t)((n_cst + 3) & ~3) * 4;
    p.lp = off; off += 2 * TC_M * 4;   // -log p(s) and group 1's share of |s|^2
    p.redc = off; off += 16;
    p.redi = off; off += 32;
    p.bars = off; off += 8 * (2 * TC_MAX_STAGES + 2 + 6);
    p.sbar = off; off += 8;
    p.tmem_ptr = off; off += 8;
    p.flag = off; off += 8;
    p.tfirst = off; off += (MGB_MAX_GMM + 1) * 4 + 12;
    p.total = off;
    return p;
}

// cycle counters of CTA 0 (diagnostic: mgb_debug_tc_cycles, tools/tc_cycles.py).  First group-0 thread: 0 wait a_empty,
// 1 draw + fill z, 2 wait sampling MMA, 3 refill [s|1], 4 wait quad chunks, 5 drain quad chunks, 6 constraint program,
// 7 wait GMM group, 8 tiles, 9 total; first group-1 thread: 10 wait chunk, 11 consume, 12 total; MMA lane: 13 wait
// t_empty, 14 wait b_full, 15 wait a_full
#ifdef MGB_TC_PROFILE
#define MGB_TC_T(i)                                     \
    if (dbg) {                                          \
        const long long now_ = clock64();               \
        g_tc_cycles[i] += now_ - tmark;                 \
        tmark = now_;                                   \
    }
#else
#define MGB_TC_T(i)
#endif

__device__ __forceinline__ bool tc_better(float c1, long long i1, float c2, long long i2) { return (c1 < c2) || (c1 == c2 && i1 < i2); }

template <typename SIn, bool SAMPLE>
__global__ void __launch_bounds__(TC_THREADS, 1) mgb_score_tc_kernel(const __grid_constant__ TcScoreParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const ScoreParams& sp = p.sp;
    size_t stage_bytes = 0;
    for (int i = 0; i < p.segs.n; ++i) stage_bytes = max(stage_bytes, (size_t)2 * p.segs.seg[i].NC * p.segs.seg[i].KC * 4);
    const TcScoreSmem L = tc_score_smem_plan(p.Kp, stage_bytes, p.nst, p.QS, p.ops_in_smem ? sp.n_ops : 0, p.ops_in_smem ? sp.n_cst : 0);
    const TcBars bars = tc_bars(smem + L.bars);
    unsigned long long* s_full = reinterpret_cast<unsigned long long*>(smem + L.sbar);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L.tmem_ptr);
    MgbOp* sm_ops = reinterpret_cast<MgbOp*>(smem + L.ops);
    float* sm_cst = reinterpret_cast<float*>(smem + L.cst);
    float* sm_lp = reinterpret_cast<float*>(smem + L.lp);
    float* qv = reinterpret_cast<float*>(smem + L.qv);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_tiles = p.n_tiles;
    const long long my_tiles = (n_tiles > blockIdx.x) ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    int* sm_tile_first = reinterpret_cast<int*>(smem + L.tfirst);
    if (SAMPLE)
        for (int i = threadIdx.x; i <= MGB_MAX_GMM; i += TC_THREADS) sm_tile_first[i] = p.segs.tile_first[i];

    if (p.ops_in_smem) {
        for (int i = threadIdx.x; i < sp.n_ops; i += TC_THREADS) sm_ops[i] = sp.ops[i];
        for (int i = threadIdx.x; i < sp.n_cst; i += TC_THREADS) sm_cst[i] = sp.cst[i];
    }
    const MgbOp* ops = p.ops_in_smem ? sm_ops : sp.ops;
    const float* cst = p.ops_in_smem ? sm_cst : sp.cst;
    const bool g1_active = SAMPLE || p.seg_g >= 0;   // group 1 shares the A fills of the sampling mode and owns the GMM chunks
    if (threadIdx.x == 0) {
        for (int i = 0; i < TC_MAX_STAGES; ++i) {
            mgb_mbar_init(&bars.b_full[i], 1);
            mgb_mbar_init(&bars.b_empty[i], 1);
        }
        mgb_mbar_init(bars.a_full, SAMPLE ? 256 : 128);
        mgb_mbar_init(bars.a_empty, 1);
        for (int i = 0; i < 4; ++i) mgb_mbar_init(&bars.t_full[i], 1);
        for (int i = 0; i < 2; ++i) mgb_mbar_init(&bars.t_empty[i], 128);
        mgb_mbar_init(s_full, 1);
        mgb_fence_barrier_init();
    }
    if (warp == 1) mgb_tmem_alloc(tmem_ptr, TC_TMEM_COLS);
    mgb_tc_fence_before();
    __syncthreads();
    mgb_tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    float* sm_m2 = sm_lp + TC_M;   // group 1's share of |s|^2 (sampling mode)

    // ---- sampling mode, both epilogue groups: half h of row `row` of the two A fills --------------------------------
    // K columns [k_lo, k_hi) of the A tile belong to group h: group 0 takes the first Kh columns (a multiple of 8), group 1 the rest
    const int Kh = ((p.Kp >> 1) + 7) & ~7;
    // fill 1: [z | 1], z = the standard normals mgb_sample draws for global index gi
    // pre: the row's normals were drawn one tile ahead into the quad-value area (draw_ahead below): only split + store is left
    auto fill_z = [&](int row, int k_lo, int k_hi, bool valid, long long gi, const float* pre) {
        const int D = sp.D;
        const uint2 key = make_uint2((unsigned int)sp.seed, (unsigned int)(sp.seed >> 32));
        // four independent Philox / Box-Muller chains per iteration: a warp has its scheduler almost to itself here, and one
        // chain is a sequence of dependent integer multiplies
        for (int k0 = k_lo; k0 < k_hi; k0 += 16) {
            float z[4][4];
            if (pre != nullptr) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int q = 0; q < 4; ++q) z[u][q] = (k0 + 4 * u + q < D) ? pre[k0 + 4 * u + q] : 0.f;
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    mgb_normal4((unsigned long long)gi, (unsigned int)((k0 >> 2) + u), key, z[u]);   // unconditional: the chains interleave
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (k0 + 4 * u < k_hi) {
                    float h[4], l[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int k = k0 + 4 * u + q;
                        const float v = !valid ? 0.f : (k < D ? z[u][q] : (k == D ? 1.f : 0.f));
                        mgb_tf32_split(v, h[q], l[q]);
                    }
                    const uint32_t off = mgb_canon_off((uint32_t)row, (uint32_t)(k0 + 4 * u), TC_M);
                    *reinterpret_cast<float4*>(smem + L.a_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
                    *reinterpret_cast<float4*>(smem + L.a_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
                }
            }
        }
    };
    // all D normals of global index gi into dst (shared memory): run by group 0 for the NEXT tile while it would otherwise wait
    // for the GMM chunks of this one
    auto draw_ahead = [&](long long gi, float* dst) {
        const int D = sp.D;
        const uint2 key = make_uint2((unsigned int)sp.seed, (unsigned int)(sp.seed >> 32));
        for (int k0 = 0; k0 < D; k0 += 16) {
            float z[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) mgb_normal4((unsigned long long)gi, (unsigned int)((k0 >> 2) + u), key, z[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (k0 + 4 * u + q < D) dst[k0 + 4 * u + q] = z[u][q];
        }
    };
    // fill 2: s = mu_g + L_g z comes back in TMEM (columns [k_lo, k_hi) of this thread's lane, a multiple of 8); A becomes
    // [s | 1]; returns the thread's share of |s|^2
    auto fill_s = [&](int row, uint32_t tbase, int k_lo, int k_hi, bool valid) -> float {
        const int D = sp.D;
        float m2 = 0.f;
        for (int b0 = k_lo; b0 < k_hi; b0 += 16) {
            float v[16];
            if (b0 + 16 <= k_hi) {
                mgb_tmem_ld16(tbase + b0, v);
            } else {
                float w8[8];
                mgb_tmem_ld8(tbase + b0, w8);
#pragma unroll
                for (int i = 0; i < 8; ++i) { v[i] = w8[i]; v[8 + i] = 0.f; }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (b0 + 4 * u < k_hi) {
                    float h[4], l[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int k = b0 + 4 * u + q;
                        const float x = !valid ? 0.f : (k < D ? v[4 * u + q] : (k == D ? 1.f : 0.f));
                        if (k < D) m2 = fmaf(x, x, m2);
                        mgb_tf32_split(x, h[q], l[q]);
                    }
                    const uint32_t off = mgb_canon_off((uint32_t)row, (uint32_t)(b0 + 4 * u), TC_M);
                    *reinterpret_cast<float4*>(smem + L.a_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
                    *reinterpret_cast<float4*>(smem + L.a_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
                }
            }
        }
        return m2;
    };
    // candidate of (tile T, row) in sampling mode: a tile holds (up to) 128 consecutive candidates of ONE mixture component
    auto sample_row = [&](long long T, int row, int& g, long long& gi, bool& valid) {
        g = tc_tile_comp(sm_tile_first, sp.G, T);
        const long long lo = max(sp.prefix.v[g], sp.index_base), hi = min(sp.prefix.v[g + 1], sp.index_base + sp.n);
        gi = lo + (T - sm_tile_first[g]) * TC_M + row;
        valid = gi < hi;
    };

    if (warp == 0) {
        if (lane == 0) tc_b_producer(p.segs, smem + L.b_stage, L.stage_bytes, p.nst, bars, my_tiles, sm_tile_first);
    } else if (warp == 1) {
        tc_mma_issuer(p.segs, mgb_smem_u32(smem + L.a_hi), mgb_smem_u32(smem + L.a_lo), mgb_smem_u32(smem + L.b_stage), L.stage_bytes,
                          p.nst, bars, tmem_base, my_tiles, sm_tile_first);
    } else if (warp < 6) {
        // ================= group 0: A producer, quad drain, constraint program, selection =================
        const int wq = warp & 3;
        const int row = 32 * wq + lane;  // candidate row of the tile = TMEM lane
        const int tid0 = threadIdx.x - 64;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * wq) << 16);
        const int D = sp.D, QS = p.QS;
        constexpr int NH = sizeof(SIn) / 4;   // the raw tile is staged in NH pieces of RP rows (it must fit the quad-value area)
        constexpr int RP = TC_M / NH;
        const SIn* __restrict__ s_glob = reinterpret_cast<const SIn*>(sp.s);
        SIn* raw = reinterpret_cast<SIn*>(qv);
        float best_c = INFINITY;
        long long best_i = LLONG_MAX;
        unsigned int cc = 0, sload = 0, fmask = 0;   // cc: chunks of this CTA so far (all owners); fmask: phase bits of the full barriers
#ifdef MGB_TC_PROFILE
        const bool dbg = SAMPLE && blockIdx.x == 0 && tid0 == 0;
        long long tmark = clock64();
        const long long tstart = tmark;
#endif
        for (long long t = 0; t < my_tiles; ++t) {
            const long long T = blockIdx.x + t * gridDim.x;
            long long g0 = T * TC_M, idx, gi = 0;
            bool valid;
            float m2 = 0.f;
            int gcomp = 0;
            if constexpr (SAMPLE) {
                // ---- fused K1, fill 1 of A (this group's half of the row) -----------------------------------------
                sample_row(T, row, gcomp, gi, valid);
                idx = gi - sp.index_base;
                mgb_mbar_wait(bars.a_empty, (unsigned int)(t & 1) ^ 1u);
                MGB_TC_T(0)
                fill_z(row, 0, Kh, valid, gi, (p.draw_ahead && t > 0) ? qv + (size_t)row * QS : nullptr);
            } else {
                idx = g0 + row;
                valid = idx < sp.n;
                // ---- A tile: [s | 1] split into TF32 hi / lo -------------------------------------------
                for (int piece = 0; piece < NH; ++piece) {
                    const long long r0 = g0 + (long long)piece * RP;
                    const long long rows_here = max(0LL, min((long long)RP, sp.n - r0));
                    asm volatile("bar.sync 3, 128;" ::: "memory");  // the area is free: program of the previous tile / previous piece done
                    if (sp.s_aligned && rows_here == RP) {
                        if (tid0 == 0) {
                            const unsigned int bytes = (unsigned int)(RP * D * sizeof(SIn));
                            mgb_fence_proxy_async();  // generic-proxy accesses of the area are ordered before the async-proxy write
                            mgb_mbar_expect_tx(s_full, bytes);
                            mgb_bulk_g2s(raw, s_glob + (size_t)r0 * D, bytes, s_full);
                        }
                        mgb_mbar_wait(s_full, sload & 1u);
                        ++sload;
                    } else {
                        const long long remaining = rows_here * D;
                        for (int e = tid0; e < RP * D; e += 128) raw[e] = (e < remaining) ? s_glob[(size_t)r0 * D + e] : (SIn)0;
                        asm volatile("bar.sync 3, 128;" ::: "memory");
                    }
                    if (piece == 0) mgb_mbar_wait(bars.a_empty, (unsigned int)(t & 1) ^ 1u);
                    if (row / RP == piece) {
                        const SIn* mine = raw + (size_t)(row - piece * RP) * D;
                        tc_write_a_row(smem + L.a_hi, smem + L.a_lo, row, p.Kp, valid ? D + 1 : 0, [&](int k) {
                            const float v = k < D ? (float)mine[k] : 1.f;
                            if (k < D) m2 = fmaf(v, v, m2);
                            return v;
                        });
                    }
                }
            }
            mgb_fence_proxy_async();
            mgb_mbar_arrive(bars.a_full);
            MGB_TC_T(1)
            // ---- this group's chunks: the sampling chunk (second fill of A), then the quad chunks: TMEM -> this thread's
            //      row of the quad-value area, then the part of the constraint program that reads them (every chunk holds
            //      whole time slots; the kinematic state carries over) --------------------------------------------------
            float total = 0.f;
            MgbInterp<1> interp;
            interp.init();
            const float* myq = qv + (size_t)row * QS;
            auto quad = [&](int q, float4 (&out)[1]) { out[0] = make_float4(myq[4 * q], myq[4 * q + 1], myq[4 * q + 2], myq[4 * q + 3]); };
            auto emit = [&](int, int term, float w, float e) {
                total = fmaf(w, e, total);
                if (sp.terms != nullptr && valid) sp.terms[(size_t)term * sp.n + idx] += e;
            };
            tc_for_each_chunk(p.segs, gcomp, [&](int sg, int c, int owner, bool samp) {
                const unsigned int buf = cc & 1u;
                ++cc;
                if (owner == TC_OWN_G1) return;
                mgb_mbar_wait(&bars.t_full[buf], tc_full_phase(fmask, 0, buf));
                mgb_tc_fence_after();
                const uint32_t tbase = lane_addr + buf * TC_BUF_COLS;
                if (samp) {
                    MGB_TC_T(2)
                    m2 = fill_s(row, tbase, 0, Kh, valid);
                    mgb_tc_fence_before();
                    mgb_mbar_arrive_relaxed(&bars.t_empty[buf]);   // group 1's reads of this buffer are ordered by its a_full arrival
                    mgb_fence_proxy_async();
                    mgb_mbar_arrive(bars.a_full);
                    MGB_TC_T(3)
                    return;
                }
                MGB_TC_T(4)
                const int wc = p.segs.seg[sg].wc(c);
                for (int b0 = 0; b0 < wc; b0 += 32) {
                    float v[32];
                    mgb_tmem_ld32(tbase + b0, v);
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (b0 + i < wc) qv[(size_t)row * QS + b0 + i] = v[i];
                }
                mgb_tc_fence_before();
                mgb_mbar_arrive_relaxed(&bars.t_empty[buf]);
                MGB_TC_T(5)
                interp.run(ops, __ldg(p.phase + c), __ldg(p.phase + c + 1), cst, quad, emit);
                MGB_TC_T(6)
            });
            float cost = sp.err_scale * total;
            if constexpr (SAMPLE) {
                if (p.draw_ahead && t + 1 < my_tiles) {
                    // the quad values of this tile are consumed: the area takes the next tile's normals (both groups read them
                    // after the barrier below, before the next quad chunk is drained into it)
                    int g_;
                    long long gi_;
                    bool v_;
                    sample_row(T + gridDim.x, row, g_, gi_, v_);
                    if (v_) draw_ahead(gi_, qv + (size_t)row * QS);
                }
            }
            if (g1_active) {
                asm volatile("bar.sync 2, 256;" ::: "memory");  // group 1 has published -log p(s) and its share of |s|^2 of this tile
                if (p.seg_g >= 0) cost = fmaf(sp.quality_scale, sm_lp[row], cost);
                if (SAMPLE) m2 += sm_m2[row];
            }
            MGB_TC_T(7)
#ifdef MGB_TC_PROFILE
            if (dbg) g_tc_cycles[8] += 1;
#endif
            if (sp.magnitude_scale != 0.f) cost = fmaf(sp.magnitude_scale, sqrtf(m2), cost);
            if (valid) {
                if (sp.cost) sp.cost[idx] = cost;
                const long long gidx = sp.index_base + idx;
                if (tc_better(cost, gidx, best_c, best_i)) { best_c = cost; best_i = gidx; }
            }
        }
#ifdef MGB_TC_PROFILE
        if (dbg) g_tc_cycles[9] += clock64() - tstart;
#endif
        // ---- selection: group-0 arg-min, per-CTA partial, last CTA folds (first minimum wins) ----------
        if (sp.want_best) {
            float* sm_redc = reinterpret_cast<float*>(smem + L.redc);
            long long* sm_redi = reinterpret_cast<long long*>(smem + L.redi);
            int* sm_flag = reinterpret_cast<int*>(smem + L.flag);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
                const long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
                if (tc_better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
            }
            if (lane == 0) { sm_redc[wq] = best_c; sm_redi[wq] = best_i; }
            asm volatile("bar.sync 3, 128;" ::: "memory");
            if (tid0 == 0) {
                for (int w = 0; w < 4; ++w)
                    if (tc_better(sm_redc[w], sm_redi[w], best_c, best_i)) { best_c = sm_redc[w]; best_i = sm_redi[w]; }
                sp.part_cost[blockIdx.x] = best_c;
                sp.part_idx[blockIdx.x] = best_i;
                __threadfence();
                const unsigned int done = atomicAdd(sp.done_count, 1u);
                *sm_flag = (done == gridDim.x - 1) ? 1 : 0;
            }
            asm volatile("bar.sync 3, 128;" ::: "memory");
            if (*sm_flag) {
                __threadfence();
                best_c = INFINITY;
                best_i = LLONG_MAX;
                for (int b = tid0; b < (int)gridDim.x; b += 128) {
                    const float oc = __ldcg(sp.part_cost + b);
                    const long long oi = __ldcg(sp.part_idx + b);
                    if (tc_better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float oc = __shfl_xor_sync(0xffffffffu, best_c, o);
                    const long long oi = __shfl_xor_sync(0xffffffffu, best_i, o);
                    if (tc_better(oc, oi, best_c, best_i)) { best_c = oc; best_i = oi; }
                }
                asm volatile("bar.sync 3, 128;" ::: "memory");
                if (lane == 0) { sm_redc[wq] = best_c; sm_redi[wq] = best_i; }
                asm volatile("bar.sync 3, 128;" ::: "memory");
                if (tid0 < 32) {   // the group's first warp: final fold, then (world > 1) the exchange with the peers in this epilogue
                    if (tid0 == 0) {
                        for (int w = 0; w < 4; ++w)
                            if (tc_better(sm_redc[w], sm_redi[w], best_c, best_i)) { best_c = sm_redc[w]; best_i = sm_redi[w]; }
                        if (best_i == LLONG_MAX) best_i = sp.index_base;  // nothing finite: reference keeps index 0
                    }
                    best_c = __shfl_sync(0xffffffffu, best_c, 0);
                    best_i = __shfl_sync(0xffffffffu, best_i, 0);
                    if (sp.xch.peer_slots != nullptr) {
                        float gc;
                        long long gi2;
                        mgb_exchange_warp(sp.xch, best_c, best_i, gc, gi2);
                        best_c = gc;
                        best_i = gi2;
                    }
                }
                if (tid0 == 0) {
                    sp.best_cost[0] = best_c;
                    sp.best_idx[0] = best_i;
                    *sp.done_count = 0u;  // ready for the next launch
                    if (SAMPLE && sp.best_row != nullptr) {
                        // the winning latent vector as mgb_sample writes it for that global index (fp32 Cholesky product; the
                        // tile scored its tensor-core product, equal to ~1e-6)
                        int g = 0;
                        while (g + 1 < sp.G && best_i >= sp.prefix.v[g + 1]) ++g;
                        mgb_sample_row_mem(sp.chol + (size_t)g * sp.D * sp.LS, sp.LS, sp.means + (size_t)g * sp.D, sp.D, (unsigned long long)best_i,
                                           make_uint2((unsigned int)sp.seed, (unsigned int)(sp.seed >> 32)), sp.best_row);
                    }
                }
            }
        }
    } else if (g1_active) {
        // ================= group 1: its half of the A fills (sampling mode); GMM log-density straight from TMEM ==========
        const int wq = warp & 3;
        const int row = 32 * wq + lane;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * wq) << 16);
        const int CW = p.CW, G = sp.G;
        unsigned int cc = 0, fmask = 0;
#ifdef MGB_TC_PROFILE
        const bool dbg1 = SAMPLE && blockIdx.x == 0 && threadIdx.x == 64 + 128;
        long long tc1 = 0;
        const long long t1start = clock64();
#endif
        for (long long t = 0; t < my_tiles; ++t) {
            const long long T = blockIdx.x + t * gridDim.x;
            int gcomp = 0;
            long long gi = 0;
            bool valid = false;
            float m2 = 0.f;
            if constexpr (SAMPLE) {
                sample_row(T, row, gcomp, gi, valid);
                mgb_mbar_wait(bars.a_empty, (unsigned int)(t & 1) ^ 1u);
                fill_z(row, Kh, p.Kp, valid, gi, (p.draw_ahead && t > 0) ? qv + (size_t)row * p.QS : nullptr);
                mgb_fence_proxy_async();
                mgb_mbar_arrive(bars.a_full);
            }
            // GMM columns are packed densely: component g owns columns [g CW, (g + 1) CW) of the operand, CW a multiple of 8, and a
            // component may straddle two chunks - the running sum of squares carries over
            float mx = -INFINITY, ssum = 0.f, qf0 = 0.f, qf1 = 0.f;
            int g = 0, left = CW;
            tc_for_each_chunk(p.segs, gcomp, [&](int sg, int c, int owner, bool samp) {
                const unsigned int buf = cc & 1u;
                ++cc;
                if (owner == TC_OWN_G0) return;
#ifdef MGB_TC_PROFILE
                long long tw0 = dbg1 ? clock64() : 0;
#endif
                mgb_mbar_wait(&bars.t_full[2 + buf], tc_full_phase(fmask, 1, buf));
#ifdef MGB_TC_PROFILE
                if (dbg1) { const long long now_ = clock64(); g_tc_cycles[10] += now_ - tw0; tc1 = now_; }
#endif
                mgb_tc_fence_after();
                const uint32_t tbase = lane_addr + buf * TC_BUF_COLS;
                if (samp) {
                    m2 = fill_s(row, tbase, Kh, p.Kp, valid);
                    mgb_tc_fence_before();      // this thread's TMEM reads are complete before the arrival below
                    mgb_fence_proxy_async();
                    mgb_mbar_arrive(bars.a_full);
                    return;
                }
                const int wc = p.segs.seg[sg].wc(c);   // a multiple of 16
                for (int b0 = 0; b0 < wc; b0 += 32) {
                    float v[32];
                    if (b0 + 32 <= wc) {
                        mgb_tmem_ld32(tbase + b0, v);
                    } else {
                        float w16[16];
                        mgb_tmem_ld16(tbase + b0, w16);
#pragma unroll
                        for (int i = 0; i < 16; ++i) { v[i] = w16[i]; v[16 + i] = 0.f; }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (b0 + 8 * u < wc) {
#pragma unroll
                            for (int i = 0; i < 8; i += 2) {
                                qf0 = fmaf(v[8 * u + i], v[8 * u + i], qf0);
                                qf1 = fmaf(v[8 * u + i + 1], v[8 * u + i + 1], qf1);
                            }
                            left -= 8;
                            if (left == 0) {   // y = (s - mu_g) . P_g complete (padding columns are exact zeros)
                                if (g < G) {
                                    const float lp = __ldg(sp.gmm_c + g) - 0.5f * (qf0 + qf1);
                                    if (lp > mx) {
                                        ssum = ssum * __expf(mx - lp) + 1.f;
                                        mx = lp;
                                    } else {
                                        ssum += __expf(lp - mx);
                                    }
                                }
                                ++g;
                                left = CW;
                                qf0 = qf1 = 0.f;
                            }
                        }
                    }
                }
                mgb_tc_fence_before();
                mgb_mbar_arrive_relaxed(&bars.t_empty[buf]);
#ifdef MGB_TC_PROFILE
                if (dbg1) g_tc_cycles[11] += clock64() - tc1;
#endif
            });
            sm_lp[row] = -(mx + __logf(ssum));
            sm_m2[row] = m2;
            asm volatile("bar.sync 2, 256;" ::: "memory");
        }
#ifdef MGB_TC_PROFILE
        if (dbg1) g_tc_cycles[12] += clock64() - t1start;
#endif
    }
    mgb_tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        mgb_tc_fence_after();
        mgb_tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
}

}  // namespace

// ---- host side of K6-TC -------------------------------------------------------------------
namespace {
constexpr int TC_SCORE_MIN_D = 33;  // smaller latent spaces stay on the register-resident SIMT kernel
// K rows per stage: a stage slot holds 2 (hi, lo) x 256 x 8 floats = 16 KB; operands of <= 128 columns put 16 rows into it, which
// halves the number of barrier round trips of their (latency-bound) chunks
constexpr int tc_score_kc(int NC) { return NC <= 128 ? 16 : 8; }
constexpr int TC_SCORE_QCHUNK = 128;  // columns of one chunk of the quad operand = 32 quads: what one pass of the program sees

bool tc_disabled() {
#ifdef MGB_TC_PROFILE
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("MGB_NO_TC");
        v = (e && atoi(e) != 0) ? 1 : 0;
    }
    return v == 1;
#else
    return false;
#endif
}
}  // namespace

int mgb_score_tc_prepare_model(mgb_model* m) {
    if (m->D < TC_SCORE_MIN_D) return MGB_OK;
    const int D = m->D, G = m->G;
    const int CW = (D + 7) & ~7;
    TcB B;
    B.data = nullptr;
    B.Kp = (D + 1 + 7) & ~7;
    B.Np = (G * CW + 15) & ~15;
    B.NC = std::min(TC_BUF_COLS, B.Np);   // N = 256: the lane that issues the MMAs needs ~65 cycles per MMA, the pipe N / 2
    B.KC = tc_score_kc(B.NC);
    std::vector<float> packed = tc_pack_b(B, D + 1, G * CW, [&](int k, int n) -> double {
        const int g = n / CW, j = n % CW;
        if (j >= D) return 0.0;
        return k < D ? m->prec_chol[((size_t)g * D + k) * D + j] : m->gmm_bias[(size_t)g * D + j];
    });
    MGB_CUDA(cudaMalloc((void**)&m->d_Bgmm, packed.size() * sizeof(float)));
    MGB_CUDA(cudaMemcpy(m->d_Bgmm, packed.data(), packed.size() * sizeof(float), cudaMemcpyHostToDevice));
    m->tcg_Kp = B.Kp; m->tcg_Np = B.Np; m->tcg_NC = B.NC; m->tcg_KC = B.KC; m->tcg_CW = CW;
    // sampling operand: s = z . L_g^T + mu_g, one N-chunk per component (row D meets the 1 of the A tile)
    const int CWs = (D + 15) & ~15;
    if (CWs <= TC_BUF_COLS && G <= MGB_MAX_GMM) {
        TcB S;
        S.data = nullptr;
        S.Kp = B.Kp;
        S.Np = G * CWs;
        S.NC = CWs;
        S.KC = tc_score_kc(S.NC);
        std::vector<float> ps = tc_pack_b(S, D + 1, G * CWs, [&](int k, int n) -> double {
            const int g = n / CWs, j = n % CWs;
            if (j >= D) return 0.0;
            if (k == D) return m->means64[(size_t)g * D + j];
            return k <= j ? m->chol64[((size_t)g * D + j) * D + k] : 0.0;
        });
        MGB_CUDA(cudaMalloc((void**)&m->d_Bsamp, ps.size() * sizeof(float)));
        MGB_CUDA(cudaMemcpy(m->d_Bsamp, ps.data(), ps.size() * sizeof(float), cudaMemcpyHostToDevice));
        m->tcs_Kp = S.Kp; m->tcs_Np = S.Np; m->tcs_NC = S.NC; m->tcs_KC = S.KC;
    }
    return MGB_OK;
}

int mgb_score_tc_prepare_constraints(const mgb_model* m, mgb_constraints* c) {
    if (m->D < TC_SCORE_MIN_D || c->n_quads == 0) return MGB_OK;
    const int D = m->D, K = m->K, KP = c->KP;
    const int NQC = TC_SCORE_QCHUNK / 4;  // quads per N-chunk
    // column layout: quads in program order, but a slot never straddles a chunk (the interpreter runs chunk by chunk)
    std::vector<int> pos(c->n_quads);  // position (in quads) of every quad in the padded layout
    int cur = 0;
    for (int q = 0; q < c->n_quads;) {
        int e = q;
        while (e < c->n_quads && c->quad_slot[e] == c->quad_slot[q]) ++e;
        const int cnt = e - q;
        if (cnt > NQC) {
            mgb_set_error("tensor-core score: a time slot needs %d quads (max %d per chunk)", cnt, NQC);
            return MGB_E_UNSUPPORTED;
        }
        if (cur / NQC != (cur + cnt - 1) / NQC) cur = (cur / NQC + 1) * NQC;
        for (int i = q; i < e; ++i) pos[i] = cur++;
        q = e;
    }
    const int n_pos = cur;
    const int n_chunks = (n_pos + NQC - 1) / NQC;
    std::vector<int> quad_at(n_pos, -1);
    for (int q = 0; q < c->n_quads; ++q) quad_at[pos[q]] = q;
    TcB B;
    B.data = nullptr;
    B.Kp = (D + 1 + 7) & ~7;
    B.Np = (4 * n_pos + 15) & ~15;
    B.NC = std::min(TC_SCORE_QCHUNK, B.Np);
    B.KC = tc_score_kc(B.NC);
    // row k < K: basis row k of the quad; row D (the one that meets the 1 of the A tile): the mean
    std::vector<float> packed = tc_pack_b(B, D + 1, 4 * n_pos, [&](int k, int n) -> double {
        const int q = quad_at[n >> 2], ch = n & 3;
        if (q < 0) return 0.0;
        if (k < K) return c->tab64[((size_t)q * (KP + 1) + 1 + k) * 4 + ch];
        if (k == D) return c->tab64[((size_t)q * (KP + 1)) * 4 + ch];
        return 0.0;
    });
    // program: quad fields -> position inside the chunk; phase table
    c->ops_tc = c->ops_fused;
    c->phase_tc.assign(n_chunks + 1, 0);
    int chunk = 0;
    for (size_t i = 0; i < c->ops_tc.size(); ++i) {
        MgbOp& o = c->ops_tc[i];
        // quad operands: a of ROOT / JOINT / ALIGN, b of ROOT / JROT
        const bool qa = o.code == OP_ROOT || o.code == OP_JOINT || o.code == OP_ALIGN, qb = o.code == OP_ROOT || o.code == OP_JROT;
        if (!qa && !qb) continue;
        const int ca = qa ? pos[o.a] / NQC : pos[o.b] / NQC, cb = qb ? pos[o.b] / NQC : ca;
        if (ca != cb || ca < chunk) {
            mgb_set_error("tensor-core score: program is not monotone in its quads");
            return MGB_E_UNSUPPORTED;
        }
        while (chunk < ca) c->phase_tc[++chunk] = (int)i;
        if (qa) o.a = pos[o.a] % NQC;
        if (qb) o.b = pos[o.b] % NQC;
    }
    while (chunk < n_chunks) c->phase_tc[++chunk] = (int)c->ops_tc.size();
    c->phase_tc[n_chunks] = (int)c->ops_tc.size();  // ops after the last quad (finals) run with the last chunk
    MGB_CUDA(cudaMalloc((void**)&c->d_Bq, packed.size() * sizeof(float)));
    MGB_CUDA(cudaMemcpy(c->d_Bq, packed.data(), packed.size() * sizeof(float), cudaMemcpyHostToDevice));
    MGB_CUDA(cudaMalloc((void**)&c->d_ops_tc, c->ops_tc.size() * sizeof(MgbOp)));
    MGB_CUDA(cudaMemcpy(c->d_ops_tc, c->ops_tc.data(), c->ops_tc.size() * sizeof(MgbOp), cudaMemcpyHostToDevice));
    MGB_CUDA(cudaMalloc((void**)&c->d_phase_tc, c->phase_tc.size() * sizeof(int)));
    MGB_CUDA(cudaMemcpy(c->d_phase_tc, c->phase_tc.data(), c->phase_tc.size() * sizeof(int), cudaMemcpyHostToDevice));
    c->tcq_Kp = B.Kp; c->tcq_Np = B.Np; c->tcq_NC = B.NC; c->tcq_KC = B.KC;
    return MGB_OK;
}

static bool tc_score_plan(const mgb_model* m, const mgb_constraints* c, float quality_scale, int sample, TcScoreParams* out, size_t* smem_out) {
    TcScoreParams p;
    memset(&p, 0, sizeof(p));
    p.seg_q = p.seg_g = p.seg_s = -1;
    p.segs.n = 0;
    p.segs.samp_seg = -1;
    p.segs.G = m->G;
    p.Kp = (m->D + 1 + 7) & ~7;
    if (sample) {  // must be the first operand: it is multiplied against the first fill of the A tile
        if (!m->d_Bsamp) return false;
        TcB& B = p.segs.seg[p.segs.n];
        B.data = m->d_Bsamp; B.Kp = m->tcs_Kp; B.Np = m->tcs_Np; B.NC = m->tcs_NC; B.KC = m->tcs_KC;
        p.seg_s = p.segs.samp_seg = p.segs.n++;
        p.CWs = m->tcs_NC;
    }
    const bool use_q = c && c->n_quads > 0;
    const bool use_g = quality_scale != 0.f;
    if (use_q) {
        if (!c->d_Bq) return false;
        TcB& B = p.segs.seg[p.segs.n];
        B.data = c->d_Bq; B.Kp = c->tcq_Kp; B.Np = c->tcq_Np; B.NC = c->tcq_NC; B.KC = c->tcq_KC;
        p.seg_q = p.segs.n++;
        p.phase = c->d_phase_tc;
    }
    if (use_g) {
        if (!m->d_Bgmm) return false;
        TcB& B = p.segs.seg[p.segs.n];
        B.data = m->d_Bgmm; B.Kp = m->tcg_Kp; B.Np = m->tcg_Np; B.NC = m->tcg_NC; B.KC = m->tcg_KC;
        p.seg_g = p.segs.n++;
        p.CW = m->tcg_CW;
    }
    if (p.segs.n == (sample ? 1 : 0)) return false;
    // quads and GMM both present: interleave their chunks, one TMEM buffer and one consumer group each
    p.segs.il_q = (use_q && use_g) ? p.seg_q : -1;
    p.segs.il_g = use_g ? p.seg_g : -1;   // also names the operand whose chunks belong to epilogue group 1
    p.QS = std::max(use_q ? std::min(c->tcq_Np, TC_SCORE_QCHUNK) : 0, m->D) | 1;
    // worth it only when group 0 has time to spare: GMM on, and at most two chunks of constraint program per tile
    p.draw_ahead = (sample && use_g && (!use_q || (c->tcq_Np + TC_SCORE_QCHUNK - 1) / TC_SCORE_QCHUNK <= 2)) ? 1 : 0;
    size_t stage_bytes = 0;
    for (int i = 0; i < p.segs.n; ++i) stage_bytes = std::max(stage_bytes, (size_t)2 * p.segs.seg[i].NC * p.segs.seg[i].KC * 4);
    const int n_ops = c ? c->n_ops : 0, n_cst = c ? c->n_cst : 0;
    // The kernel has no static shared memory: the whole 227 KB opt-in maximum is available to the carve-up.  Short constraint
    // programs (<= 2 quad chunks per tile: the operand stream is the critical path, it runs at ~70 % of the L2 throughput and a
    // deeper ring hides its latency) take as many operand stages as fit, program in shared memory only if room is left
    // (L4M: 4 stages 11.3 ms, 3 stages 11.6 ms); long programs keep program + constants in shared memory first (the stress
    // set loses 12 % when they are read through L1).
    const bool stages_first = p.draw_ahead != 0 || !use_q || (c->tcq_Np + TC_SCORE_QCHUNK - 1) / TC_SCORE_QCHUNK <= 2;
    for (int pass = 0; pass < 2 * (TC_MAX_STAGES - 1); ++pass) {
        int nst, in_smem;
        if (stages_first) { nst = TC_MAX_STAGES - pass / 2; in_smem = 1 - (pass & 1); }
        else { in_smem = pass < (TC_MAX_STAGES - 1) ? 1 : 0; nst = TC_MAX_STAGES - pass % (TC_MAX_STAGES - 1); }
        const size_t total = tc_score_smem_plan(p.Kp, stage_bytes, nst, p.QS, in_smem ? n_ops : 0, in_smem ? n_cst : 0).total;
        if (total <= TC_SCORE_SMEM_MAX) {
            p.nst = nst;
            p.ops_in_smem = in_smem;
            *out = p;
            *smem_out = total;
            return true;
        }
    }
    return false;
}

bool mgb_score_tc_usable(const mgb_model* m, const mgb_constraints* c, float quality_scale, int sample) {
    if (tc_disabled() || m->D < TC_SCORE_MIN_D) return false;
    TcScoreParams p;
    size_t smem;
    return tc_score_plan(m, c, quality_scale, sample, &p, &smem);
}

int mgb_launch_score_tc(const mgb_model* m, const mgb_constraints* c, const ScoreParams& sp, int s_is_f64, float quality_scale,
                        int sample, cudaStream_t stream) {
    TcScoreParams p;
    size_t smem = 0;
    MGB_REQUIRE(tc_score_plan(m, c, quality_scale, sample, &p, &smem), "tensor-core score path is not available for this model / constraint set");
    p.sp = sp;
    p.sp.ops = c ? c->d_ops_tc : nullptr;
    p.n_tiles = (sp.n + TC_M - 1) / TC_M;
    if (sample) {
        // tiles never straddle mixture components: component g owns ceil(len_g / 128) tiles of this shard
        int acc = 0;
        for (int g = 0; g < m->G; ++g) {
            p.segs.tile_first[g] = acc;
            const long long lo = std::max<long long>(sp.prefix.v[g], sp.index_base), hi = std::min<long long>(sp.prefix.v[g + 1], sp.index_base + sp.n);
            if (hi > lo) acc += (int)((hi - lo + TC_M - 1) / TC_M);
        }
        for (int g = m->G; g <= MGB_MAX_GMM; ++g) p.segs.tile_first[g] = acc;
        p.n_tiles = acc;
    }
    static bool cfg_a[MGB_MAX_DEVICES] = {false}, cfg_b[MGB_MAX_DEVICES] = {false}, cfg_c[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(mgb_score_tc_kernel<float, false>, cfg_a, (int)TC_SCORE_SMEM_MAX)) return rc;
    if (int rc = mgb_opt_in_smem(mgb_score_tc_kernel<double, false>, cfg_b, (int)TC_SCORE_SMEM_MAX)) return rc;
    if (int rc = mgb_opt_in_smem(mgb_score_tc_kernel<float, true>, cfg_c, (int)TC_SCORE_SMEM_MAX)) return rc;
    const int grid = (int)std::max<long long>(1, std::min<long long>(p.n_tiles, tc_num_sms()));
    if (sample)
        mgb_score_tc_kernel<float, true><<<grid, TC_THREADS, smem, stream>>>(p);
    else if (s_is_f64)
        mgb_score_tc_kernel<double, false><<<grid, TC_THREADS, smem, stream>>>(p);
    else
        mgb_score_tc_kernel<float, false><<<grid, TC_THREADS, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}


// diagnostic of a -DMGB_TC_PROFILE build (not part of include/mgb200.h): read and clear the cycle counters
#ifdef MGB_TC_PROFILE
extern "C" int mgb_debug_tc_cycles(long long* out16) {
    long long zero[16] = {0};
    MGB_CUDA(cudaMemcpyFromSymbol(out16, g_tc_cycles, sizeof(zero)));
    MGB_CUDA(cudaMemcpyToSymbol(g_tc_cycles, zero, sizeof(zero)));
    return MGB_OK;
}
#endif
