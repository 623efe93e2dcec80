// K7: level-synchronous k-means of the cluster-tree build.
//
// The reference subdivides one node at a time: sklearn KMeans(n_clusters=4).fit_predict on
// data[indices, :dim] (space_partitioning/cluster_tree_node_builder.py:83-103,
// space_partitioning/clustering.py:74-80), recursively.  Here every node of one tree level is
// subdivided by the same launches: the samples are kept sorted by node (the host re-sorts them
// after every level), one kernel does the E-step for every sample and accumulates the M-step
// sums of every (node, cluster), a second small kernel turns the sums into centroids and tests
// every node's convergence (sklearn's rule: squared centre shift <= tol * mean variance).
//
// HBM-bound: per sample and Lloyd iteration 4*ld + 4 bytes read (row + node id), 4 bytes written
// (label).  The accumulation never goes through per-sample atomics: a warp walks a contiguous range
// of samples 32 at a time, lane = sample for the distance part, then lane = (group, column) for the
// accumulation part (the tile lives in shared memory; the 32 / W lane groups split the samples of a
// run of equal node ids), keeping the running sums of its current node in registers and flushing
// them with one vector RED per (node, cluster) when the node changes.  Member counts come from
// ballots, the inertia from a warp reduction at flush time.
#include <cuda_pipeline.h>
#include <math.h>

#include <algorithm>

#include "mgb_device.cuh"
#include "mgb_internal.cuh"

namespace {

int sm_count_tree() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

constexpr int KM_WARPS = 4;
constexpr int KM_SMALL_STAGES = 3;

// CH = ceil(dim_acc / 32): columns handled per lane.  W (power of two, >= dim_acc when CH == 1, else 32)
// is the width of a lane group; the 32 / W groups of a warp split the samples of a run between them.
template <int NS, int CH>
__global__ void __launch_bounds__(KM_WARPS * 32) mgb_kmeans_level_kernel(
    const float* __restrict__ x, long long n, int dim, int dim_acc, int ld, const int* __restrict__ node,
    const float* __restrict__ cent, int n_sub, int* __restrict__ label, float* __restrict__ sums,
    int* __restrict__ counts, float* __restrict__ node_inertia, float* __restrict__ inertia, long long tiles_per_warp, int W,
    const int* __restrict__ skip) {
    // skip: per-node flags (converged); a tile ALL of whose samples belong to flagged nodes is neither fetched nor processed -
    // labels, sums and counts of such nodes are those of an earlier iteration and are not read again before the final pass
    extern __shared__ float sm_tiles[];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float* tile = sm_tiles + (size_t)warp * 32 * ld;
    const long long gwarp = (long long)blockIdx.x * KM_WARPS + warp;
    const long long first = gwarp * tiles_per_warp * 32;
    long long last = first + tiles_per_warp * 32;
    if (last > n) last = n;
    const int G = 32 / W;          // lane groups
    const int g = lane / W;        // my group
    const int col0 = lane % W;     // my column (CH == 1) / first column (CH > 1: W == 32)

    float acc[NS][CH];
    int cnt[NS];
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        cnt[k] = 0;
#pragma unroll
        for (int c = 0; c < CH; ++c) acc[k][c] = 0.f;
    }
    float inertia_part = 0.f;      // per lane, of the current node's run(s)
    float total_inertia = 0.f;     // lane 0
    int cur = -1;

    auto flush = [&]() {           // warp-uniform call
        if (cur < 0) return;
#pragma unroll
        for (int k = 0; k < NS; ++k) {
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                float v = acc[k][c];
                for (int o = W; o < 32; o <<= 1) v += __shfl_xor_sync(FULL, v, o);
                const int col = col0 + 32 * c;
                if (g == 0 && k < n_sub && col < dim_acc && v != 0.f) atomicAdd(sums + ((size_t)cur * n_sub + k) * dim_acc + col, v);
                acc[k][c] = 0.f;
            }
            if (lane == 0 && k < n_sub && cnt[k] != 0) atomicAdd(counts + (size_t)cur * n_sub + k, cnt[k]);
            cnt[k] = 0;
        }
        float t = inertia_part;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
        if (lane == 0) {
            if (node_inertia && t != 0.f) atomicAdd(node_inertia + cur, t);
            total_inertia += t;
        }
        inertia_part = 0.f;
    };

    // double-buffered tiles: the next tile's rows are in flight (cp.async) while this one is processed
    float* const tile0 = tile;
    const size_t tstride = (size_t)KM_WARPS * 32 * ld;
    auto fetch = [&](long long t0, float* dst) {
        if (t0 < last) {
            const int rows = (int)((last - t0 < 32) ? (last - t0) : 32);
            const float* src = x + (size_t)t0 * ld;
            const int total = rows * ld;
            for (int e = lane; e < total; e += 32) __pipeline_memcpy_async(dst + e, src + e, sizeof(float));
        }
        __pipeline_commit();
    };
    auto ids_of = [&](long long t0) -> int { return (t0 + lane < last) ? (node ? __ldg(node + t0 + lane) : 0) : -1; };
    auto tile_skippable = [&](int nd_lane) -> bool {
        return skip != nullptr && __all_sync(FULL, nd_lane < 0 || __ldg(skip + nd_lane) != 0);
    };
    // node ids run two tiles ahead of the tile being processed: whether tile t0 + 32 is fetched at all is decided from them
    int nd_a = ids_of(first), nd_b = ids_of(first + 32);
    bool skip_a = tile_skippable(nd_a);
    if (!skip_a) fetch(first, tile0);
    else __pipeline_commit();
    int pb = 0;
    for (long long t0 = first; t0 < last; t0 += 32, pb ^= 1) {
        const int rows = (int)((last - t0 < 32) ? (last - t0) : 32);
        const bool skip_b = (t0 + 32 < last) ? tile_skippable(nd_b) : true;
        if (!skip_b) fetch(t0 + 32, tile0 + (pb ? 0 : tstride));
        else __pipeline_commit();
        const int nd_c = ids_of(t0 + 64);
        __pipeline_wait_prior(1);
        __syncwarp();
        tile = tile0 + (pb ? tstride : 0);
        const int nd_cur = nd_a;
        const bool skip_cur = skip_a;
        nd_a = nd_b; nd_b = nd_c; skip_a = skip_b;
        if (skip_cur) continue;
        // E-step: lane = sample; the NS distances advance together so that NS independent centroid loads are
        // in flight per column.
        int nd = -1, lb = 0;
        float best = 0.f;
        if (lane < rows) {
            nd = nd_cur;
            const float* xi = tile + lane * ld;
            const float* cb = cent + (size_t)nd * n_sub * dim;
            float d[NS];
#pragma unroll
            for (int k = 0; k < NS; ++k) d[k] = 0.f;
#pragma unroll 2
            for (int j = 0; j < dim; ++j) {
                const float xv = xi[j];
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    if (k < n_sub) {
                        const float e = xv - __ldg(cb + k * dim + j);
                        d[k] = fmaf(e, e, d[k]);
                    }
                }
            }
            best = INFINITY;
#pragma unroll
            for (int k = 0; k < NS; ++k)
                if (k < n_sub && d[k] < best) { best = d[k]; lb = k; }
            label[t0 + lane] = lb;
        }
        // M-step sums, one run of equal node ids at a time (the samples are sorted by node)
        const int prev = __shfl_up_sync(FULL, nd, 1);
        unsigned heads = __ballot_sync(FULL, lane < rows && (lane == 0 || nd != prev));
        while (heads) {
            const int a = __ffs(heads) - 1;
            heads &= heads - 1;
            const int b = heads ? (__ffs(heads) - 1) : rows;
            const int run_nd = __shfl_sync(FULL, nd, a);
            if (run_nd != cur) {
                flush();
                cur = run_nd;
            }
            const bool in_run = lane >= a && lane < b;
#pragma unroll
            for (int k = 0; k < NS; ++k) cnt[k] += __popc(__ballot_sync(FULL, in_run && lb == k));
            if (in_run) inertia_part += best;
            const int iters = (b - a + G - 1) / G;
            for (int it = 0; it < iters; ++it) {
                const int s = a + it * G + g;
                const bool on = s < b;
                const int slb = __shfl_sync(FULL, lb, on ? s : a);
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const int col = col0 + 32 * c;
                    const float v = (on && col < dim_acc) ? tile[s * ld + col] : 0.f;
#pragma unroll
                    for (int k = 0; k < NS; ++k) acc[k][c] += (k == slb) ? v : 0.f;
                }
            }
        }
        __syncwarp();
    }
    flush();
    if (inertia && lane == 0 && total_inertia != 0.f) atomicAdd(inertia, total_inertia);
}

// The same step for narrow rows (dim_acc <= DA <= 16, n_sub <= NS = 4: the latent space of the S primitive): lane = sample
// for the E-step AND the accumulation.  The row, the NS centroids of the current node and the NS x DA running sums live in
// registers; a sample adds its row to the sums of its label with predicated adds (no shuffles, no shared-memory traffic per
// sample), and the 32 lanes' sums are folded through a padded shared-memory transpose only when the node changes.  About a
// third of the instructions of the general kernel per sample.
template <int NS, int DA, int DD>
__global__ void __launch_bounds__(KM_WARPS * 32, 4) mgb_kmeans_level_small_kernel(
    const float* __restrict__ x, long long n, int dim, int dim_acc, int ld, const int* __restrict__ node,
    const float* __restrict__ cent, int n_sub, int* __restrict__ label, float* __restrict__ sums,
    int* __restrict__ counts, float* __restrict__ node_inertia, float* __restrict__ inertia, long long tiles_per_warp,
    const int* __restrict__ skip) {
    extern __shared__ __align__(16) float sm_tiles[];
    constexpr int ST = KM_SMALL_STAGES;          // tile buffers per warp: the rows of ST - 1 tiles are in flight while one is processed
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int tile_floats = (32 * ld + 1 + 3) & ~3;  // 16-byte aligned tile buffers; + 1: an odd dim_acc reads one float past its row
    float* const tile0 = sm_tiles + (size_t)warp * ST * tile_floats;
    // DA = dim_acc rounded up to even (the spare column of an odd dim_acc carries the next row's first float and is dropped at the
    // flush), DD = dim exactly (even): no column guards in the loop.  Per-warp scratch: flush transpose, then the NS
    // centroids of the current node, NEGATED, zero for clusters beyond n_sub (read as float2 broadcasts: every offset below
    // is an even number of floats, so they are 8-byte aligned), then the ring of node ids
    static_assert(DA % 2 == 0 && DD % 2 == 0 && DD <= DA && NS % 2 == 0, "even sizes");
    float* const scr = sm_tiles + (size_t)KM_WARPS * ST * tile_floats + (size_t)warp * (NS * DA * 33 + NS * DD + 5 * 32 + 2 * KM_SMALL_STAGES);
    float* const cen = scr + NS * DA * 33;
    int* const ids_s = reinterpret_cast<int*>(cen + NS * DD);      // [5][32] node ids of the tiles t .. t + 4
    unsigned long long* const bars = reinterpret_cast<unsigned long long*>(ids_s + 5 * 32);   // one mbarrier per row buffer
    if (lane == 0) {
        for (int b = 0; b < ST; ++b) mgb_mbar_init(bars + b, 1);
        mgb_fence_barrier_init();
    }
    __syncwarp();
    const long long gwarp = (long long)blockIdx.x * KM_WARPS + warp;
    const long long first = gwarp * tiles_per_warp * 32;
    long long last = first + tiles_per_warp * 32;
    if (last > n) last = n;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) && ((32 * ld) % 4 == 0);   // every full tile starts 16-byte aligned

    static_assert(DA % 2 == 0, "columns are handled in pairs (packed fp32x2 arithmetic)");
    constexpr int DP = DA / 2;
    float2 acc[NS][DP];
    int cnt[NS];
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        cnt[k] = 0;
#pragma unroll
        for (int j = 0; j < DP; ++j) acc[k][j] = make_float2(0.f, 0.f);
    }
    float inertia_part = 0.f, total_inertia = 0.f;
    int cur = -1;

    auto flush = [&]() {           // warp-uniform call
        if (cur < 0) return;
#pragma unroll
        for (int k = 0; k < NS; ++k)
#pragma unroll
            for (int j = 0; j < DP; ++j) {
                scr[(k * DA + 2 * j) * 33 + lane] = acc[k][j].x;
                scr[(k * DA + 2 * j + 1) * 33 + lane] = acc[k][j].y;
                acc[k][j] = make_float2(0.f, 0.f);
            }
        __syncwarp();
#pragma unroll
        for (int v0 = 0; v0 < NS * DA; v0 += 32) {
            const int v = v0 + lane;
            if (v < NS * DA) {
                const float* row = scr + v * 33;
                float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
#pragma unroll
                for (int i = 0; i < 32; i += 4) { t0 += row[i]; t1 += row[i + 1]; t2 += row[i + 2]; t3 += row[i + 3]; }
                const float t = (t0 + t1) + (t2 + t3);
                const int k = v / DA, j = v - k * DA;
                if (k < n_sub && j < dim_acc && t != 0.f) atomicAdd(sums + ((size_t)cur * n_sub + k) * dim_acc + j, t);
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < NS; ++k) {
            if (lane == 0 && k < n_sub && cnt[k] != 0) atomicAdd(counts + (size_t)cur * n_sub + k, cnt[k]);
            cnt[k] = 0;
        }
        float t = inertia_part;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
        if (lane == 0) {
            if (node_inertia && t != 0.f) atomicAdd(node_inertia + cur, t);
            total_inertia += t;
        }
        inertia_part = 0.f;
    };
    // rows of a tile -> buffer b.  A full, 16-byte aligned tile is ONE bulk copy issued by lane 0 and completes on the buffer's
    // mbarrier (returns true); the last, partial tile goes through 4-byte asynchronous copies that complete with the caller's
    // copy group (no commit here)
    auto fetch = [&](long long t0, int b) -> bool {
        float* dst = tile0 + b * tile_floats;
        if (t0 >= last) return false;
        const int rows = (int)((last - t0 < 32) ? (last - t0) : 32);
        const float* src = x + (size_t)t0 * ld;
        if (rows == 32 && vec_ok) {
            if (lane == 0) {
                mgb_fence_proxy_async();        // the buffer's previous rows were read through the generic proxy
                mgb_tma_load_1d(dst, src, (unsigned int)(32 * ld * sizeof(float)), bars + b);
            }
            return true;
        }
        const int total = rows * ld;
        for (int e = lane; e < total; e += 32) __pipeline_memcpy_async(dst + e, src + e, sizeof(float));
        return false;
    };
    auto tile_skippable = [&](long long t0, int nd_lane) -> bool {
        if (t0 >= last) return true;
        return skip != nullptr && __all_sync(FULL, nd_lane < 0 || __ldg(skip + nd_lane) != 0);
    };
    // Node ids travel through the same asynchronous-copy groups as the rows, four tiles ahead of the tile being processed, into
    // a ring of five slots: whether a tile's rows are fetched at all is decided from ids that have long arrived (a register
    // prefetch of the ids stalled the warp on every tile: the copy of an in-flight load waits for it).
    //   group of iteration i = rows of tile i + 2, ids of tile i + 4;  before it is issued, all groups but one are awaited,
    //   which completes the rows of tile i and the ids of tile i + 2.
    static_assert(ST == 3, "the id ring below assumes three row buffers");
    for (int q = 0; q < 4; ++q) {
        const long long tq = first + 32 * q;
        ids_s[q * 32 + lane] = (tq + lane < last) ? (node ? __ldg(node + tq + lane) : 0) : -1;
    }
    __syncwarp();
    bool skq[2], blq[2];                           // tiles t, t + 1: skipped / fetched by a bulk copy
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        skq[q] = tile_skippable(first + 32 * q, ids_s[q * 32 + lane]);
        blq[q] = skq[q] ? false : fetch(first + 32 * q, q);
        __pipeline_commit();
    }
    int bi = 0, s5 = 0;                            // row buffer / id slot of the tile being processed
    unsigned int ph = 0;                           // bit b: parity of the next completion of buffer b's mbarrier
    for (long long t0 = first; t0 < last; t0 += 32) {
        const int rows = (int)((last - t0 < 32) ? (last - t0) : 32);
        __pipeline_wait_prior(1);
        __syncwarp();
        bool sk2, bl2;
        {
            const long long tq = t0 + 64;
            int bq = bi + 2, sq = s5 + 2, sa = s5 + 4;
            if (bq >= 3) bq -= 3;
            if (sq >= 5) sq -= 5;
            if (sa >= 5) sa -= 5;
            sk2 = tile_skippable(tq, (tq + lane < last) ? ids_s[sq * 32 + lane] : -1);
            bl2 = sk2 ? false : fetch(tq, bq);
            const long long ta = t0 + 128;
            if (ta + lane < last) {
                if (node) __pipeline_memcpy_async(ids_s + sa * 32 + lane, node + ta + lane, sizeof(int));
                else ids_s[sa * 32 + lane] = 0;
            } else {
                ids_s[sa * 32 + lane] = -1;
            }
            __pipeline_commit();
        }
        const float* tile = tile0 + bi * tile_floats;
        const int nd = (lane < rows) ? ids_s[s5 * 32 + lane] : -1;
        const bool skip_cur = skq[0], bulk_cur = blq[0];
        skq[0] = skq[1]; blq[0] = blq[1];
        skq[1] = sk2; blq[1] = bl2;
        if (bulk_cur) {
            mgb_mbar_wait(bars + bi, (ph >> bi) & 1u);
            ph ^= 1u << bi;
        }
        if (++bi == 3) bi = 0;
        if (++s5 == 5) s5 = 0;
        if (skip_cur) continue;
        // lanes beyond the last row of a partial tile read its last row: they belong to no run and add it with factor 0
        float2 xr[DP];
        const bool have = lane < rows;
        const float* xrow = tile + min(lane, rows - 1) * ld;
#pragma unroll
        for (int j = 0; j < DP; ++j) {
            xr[j].x = xrow[2 * j];
            xr[j].y = xrow[2 * j + 1];
        }
        const int prev = __shfl_up_sync(FULL, nd, 1);
        unsigned heads = __ballot_sync(FULL, have && (lane == 0 || nd != prev));
        while (heads) {
            const int a = __ffs(heads) - 1;
            heads &= heads - 1;
            const int b = heads ? (__ffs(heads) - 1) : rows;
            const int run_nd = __shfl_sync(FULL, nd, a);
            if (run_nd != cur) {
                flush();
                cur = run_nd;
                const float* cb = cent + (size_t)cur * n_sub * dim;
                __syncwarp();
                for (int v = lane; v < NS * DD; v += 32) cen[v] = (v < n_sub * DD) ? -__ldg(cb + v) : 0.f;
                __syncwarp();
            }
            const bool in_run = lane >= a && lane < b;
            // squared distances over the first DD columns, two columns per instruction (FADD2 / FFMA2: each half rounded like
            // the scalar instruction)
            float2 d2[NS];
#pragma unroll
            for (int k = 0; k < NS; ++k) d2[k] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < DD / 2; ++j) {
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    const float2 e = __fadd2_rn(xr[j], *reinterpret_cast<const float2*>(cen + k * DD + 2 * j));
                    d2[k] = __ffma2_rn(e, e, d2[k]);
                }
            }
            float d[NS];
#pragma unroll
            for (int k = 0; k < NS; ++k) d[k] = d2[k].x + d2[k].y;
            float best = INFINITY;
            int lb = 0;
#pragma unroll
            for (int k = 0; k < NS; ++k)
                if (k < n_sub && d[k] < best) { best = d[k]; lb = k; }
            if (in_run) {
                label[t0 + lane] = lb;
                inertia_part += best;
            }
            // the row joins the sums of its label: one FMA per (label, column) with a 0 / 1 factor (rows beyond the tile are zero)
#pragma unroll
            for (int k = 0; k < NS; ++k) {
                const bool mine = in_run && lb == k;
                cnt[k] += __popc(__ballot_sync(FULL, mine));
                const float m = mine ? 1.f : 0.f;
                const float2 mm = make_float2(m, m);
#pragma unroll
                for (int j = 0; j < DP; ++j) acc[k][j] = __ffma2_rn(mm, xr[j], acc[k][j]);
            }
        }
        __syncwarp();
    }
    flush();
    if (inertia && lane == 0 && total_inertia != 0.f) atomicAdd(inertia, total_inertia);
}

template <int NS, int DA, int DD>
int launch_level_small(const float* d_x, int64_t n, int dim, int dim_acc, int ld, const int* d_node, const float* d_cent, int n_sub,
                       int* d_label, float* d_sums, int* d_counts, float* d_node_inertia, float* d_inertia, cudaStream_t stream,
                       const int* d_skip) {
    auto kern = mgb_kmeans_level_small_kernel<NS, DA, DD>;
    const size_t smem = ((size_t)KM_SMALL_STAGES * KM_WARPS * ((32 * ld + 1 + 3) & ~3) + (size_t)KM_WARPS * (NS * DA * 33 + NS * DD + 5 * 32 + 2 * KM_SMALL_STAGES)) * sizeof(float);
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int rc = mgb_opt_in_smem(kern, configured, 160 * 1024)) return rc;
    const long long tiles = (n + 31) / 32;
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, KM_WARPS * 32, smem));
    if (occ < 1) occ = 1;
    // one resident wave: every warp folds its register sums once per node it meets, so fewer, longer warps are cheaper
    long long warps = (long long)sm_count_tree() * occ * KM_WARPS;
    if (warps > tiles) warps = tiles;
    const long long tpw = (tiles + warps - 1) / warps;
    const long long used_warps = (tiles + tpw - 1) / tpw;
    const long long grid = (used_warps + KM_WARPS - 1) / KM_WARPS;
    kern<<<(unsigned)grid, KM_WARPS * 32, smem, stream>>>(d_x, n, dim, dim_acc, ld, d_node, d_cent, n_sub, d_label, d_sums, d_counts,
                                                           d_node_inertia, d_inertia, tpw, d_skip);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

// M-step finish + convergence test, one thread per node.  Empty clusters keep their centroid.
__global__ void __launch_bounds__(128) mgb_kmeans_centroid_kernel(int n_nodes, int n_sub, int dim, int dim_acc,
                                                                 const float* __restrict__ sums, const int* __restrict__ counts,
                                                                 const float* __restrict__ tol, float* __restrict__ cent,
                                                                 int* __restrict__ converged, int* __restrict__ n_active,
                                                                 int* __restrict__ n_active_next, float* __restrict__ sums_rw,
                                                                 int* __restrict__ counts_rw) {
    // n_active / n_active_next: two counters used alternately, this launch clears the one the NEXT iteration adds to (no memset
    // between iterations); sums_rw / counts_rw: the sums this thread consumed are zeroed for the next E-step likewise
    const int nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd == 0) *n_active_next = 0;
    if (nd >= n_nodes) return;
    if (converged[nd]) return;
    float shift = 0.f;
    for (int k = 0; k < n_sub; ++k) {
        const int cnt = counts[(size_t)nd * n_sub + k];
        if (cnt <= 0) continue;
        const float inv = 1.f / (float)cnt;
        float* c = cent + ((size_t)nd * n_sub + k) * dim;
        const float* s = sums + ((size_t)nd * n_sub + k) * dim_acc;
        for (int j = 0; j < dim; ++j) {
            const float v = s[j] * inv;
            const float e = v - c[j];
            shift = fmaf(e, e, shift);
            c[j] = v;
        }
    }
    for (int k = 0; k < n_sub; ++k) {
        counts_rw[(size_t)nd * n_sub + k] = 0;
        float* sz = sums_rw + ((size_t)nd * n_sub + k) * dim_acc;
        for (int j = 0; j < dim_acc; ++j) sz[j] = 0.f;
    }
    if (shift <= (tol ? tol[nd] : 0.f)) converged[nd] = 1;
    else atomicAdd(n_active, 1);
}

template <int NS, int CH>
int launch_level(const float* d_x, int64_t n, int dim, int dim_acc, int ld, const int* d_node, const float* d_cent, int n_sub,
                 int* d_label, float* d_sums, int* d_counts, float* d_node_inertia, float* d_inertia, cudaStream_t stream,
                 const int* d_skip) {
    auto kern = mgb_kmeans_level_kernel<NS, CH>;
    const size_t smem = (size_t)2 * KM_WARPS * 32 * ld * sizeof(float);
    MGB_REQUIRE(smem <= 200 * 1024, "row stride too large for the k-means kernel (ld <= 200)");
    if (smem > 48 * 1024) MGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = (n + 31) / 32;
    int occ = 1;
    MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, KM_WARPS * 32, smem));
    if (occ < 1) occ = 1;
    long long warps = (long long)sm_count_tree() * occ * KM_WARPS;
    if (warps > tiles) warps = tiles;
    const long long tpw = (tiles + warps - 1) / warps;
    const long long used_warps = (tiles + tpw - 1) / tpw;
    const long long grid = (used_warps + KM_WARPS - 1) / KM_WARPS;
    int W = 32;
    if (CH == 1) {
        W = 4;
        while (W < dim_acc) W <<= 1;
    }
    kern<<<(unsigned)grid, KM_WARPS * 32, smem, stream>>>(d_x, n, dim, dim_acc, ld, d_node, d_cent, n_sub, d_label, d_sums, d_counts,
                                                           d_node_inertia, d_inertia, tpw, W, d_skip);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

int dispatch_level(const float* d_x, int64_t n, int dim, int dim_acc, int ld, const int* d_node, const float* d_cent, int n_sub,
                   int* d_label, float* d_sums, int* d_counts, float* d_node_inertia, float* d_inertia, cudaStream_t stream,
                   const int* d_skip = nullptr) {
    const int ch = (dim_acc + 31) / 32;
    MGB_REQUIRE(n_sub >= 1 && n_sub <= 8, "n_sub must be in 1..8");
    if (n_sub <= 4 && ld <= 64) {
        // narrow rows (the latent space of the S primitive: 14 columns, clustered on the first 10): registers + packed fp32x2
#define MGB_KM_SMALL(DA, DD)                                                                                                            \
    if (((dim_acc + 1) & ~1) == DA && dim == DD)                                                                                        \
        return launch_level_small<4, DA, DD>(d_x, n, dim, dim_acc, ld, d_node, d_cent, n_sub, d_label, d_sums, d_counts, d_node_inertia, \
                                             d_inertia, stream, d_skip)
        MGB_KM_SMALL(14, 10);
        MGB_KM_SMALL(14, 14);
        MGB_KM_SMALL(16, 16);
        MGB_KM_SMALL(16, 10);
        MGB_KM_SMALL(12, 12);
        MGB_KM_SMALL(12, 10);
        MGB_KM_SMALL(10, 10);
        MGB_KM_SMALL(8, 8);
#undef MGB_KM_SMALL
    }
#define MGB_KM_CASE(NS, CH)                                                                                              \
    return launch_level<NS, CH>(d_x, n, dim, dim_acc, ld, d_node, d_cent, n_sub, d_label, d_sums, d_counts, d_node_inertia, \
                                d_inertia, stream, d_skip)
    if (n_sub <= 4) {
        switch (ch) {
            case 1: MGB_KM_CASE(4, 1);
            case 2: MGB_KM_CASE(4, 2);
            case 3: MGB_KM_CASE(4, 3);
            default: MGB_KM_CASE(4, 4);
        }
    }
    switch (ch) {
        case 1: MGB_KM_CASE(8, 1);
        case 2: MGB_KM_CASE(8, 2);
        case 3: MGB_KM_CASE(8, 3);
        default: MGB_KM_CASE(8, 4);
    }
#undef MGB_KM_CASE
}

}  // namespace

// k-means++ seeding helper: d2[i] = min(d2[i], |x[i, :dim] - centers[node[i]]|^2) (or just the distance when first != 0), in
// float64 like the seeding of sklearn's KMeans (cluster/_kmeans.py _kmeans_plusplus: closest_dist_sq).  One pass over the rows
// of a level in place of six element-wise library kernels per chosen centre.
namespace {
__global__ void __launch_bounds__(256) mgb_min_sqdist_kernel(const float* __restrict__ x, long long n, int dim, int ld, const int* __restrict__ node,
                                                            const float* __restrict__ centers, double* __restrict__ d2, int first) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float* xi = x + (size_t)i * ld;
        const float* c = centers + (size_t)(node ? __ldg(node + i) : 0) * dim;
        float acc = 0.f;
        for (int j = 0; j < dim; ++j) {
            const float e = __ldg(xi + j) - __ldg(c + j);
            acc += e * e;                         // float32 terms summed in order, as the element-wise form did
        }
        const double v = (double)acc;
        d2[i] = first ? v : fmin(d2[i], v);
    }
}
}  // namespace

extern "C" int mgb_min_sqdist(const float* d_x, int64_t n, int32_t dim, int32_t ld, const int32_t* d_node, const float* d_centers, double* d_d2,
                              int32_t first, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(n >= 0 && dim > 0 && ld >= dim, "bad shape");
    if (n == 0) return MGB_OK;
    MGB_REQUIRE(d_x && d_centers && d_d2, "null argument");
    const long long blocks = std::min<long long>((n + 255) / 256, (long long)sm_count_tree() * 16);
    mgb_min_sqdist_kernel<<<(unsigned)blocks, 256, 0, stream>>>(d_x, n, dim, ld, d_node, d_centers, d_d2, first);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

// k-means++ seeding, the pick: for every node (rows [start, start + count) of x) choose a row - uniformly when cs == NULL
// (first centre: start + min(floor(u count), count - 1)), else by D^2 sampling over the GLOBAL inclusive prefix sums cs of the
// closest squared distances: the first row of the node whose prefix sum exceeds base + u * (mass of the node) - and copy its
// first dim columns to centers[node].  One thread per node (a binary search), all nodes of a level per launch.
namespace {
__global__ void __launch_bounds__(128) mgb_kmeanspp_pick_kernel(const float* __restrict__ x, int dim, int ld, const long long* __restrict__ starts,
                                                               const long long* __restrict__ counts, int n_nodes, const double* __restrict__ u,
                                                               int u_stride, const double* __restrict__ cs, float* __restrict__ centers,
                                                               int c_stride) {
    const int nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= n_nodes) return;
    const long long start = starts[nd], count = counts[nd];
    const double un = u[(size_t)nd * u_stride];
    long long pick;
    if (cs == nullptr) {
        pick = start + min((long long)(un * (double)count), count - 1);
    } else {
        const double base = start > 0 ? cs[start - 1] : 0.0;
        const double target = base + un * (cs[start + count - 1] - base);
        long long lo = start, hi = start + count;          // first index in [start, start + count) with cs[i] > target
        while (lo < hi) {
            const long long mid = (lo + hi) >> 1;
            if (cs[mid] > target) hi = mid;
            else lo = mid + 1;
        }
        pick = min(lo, start + count - 1);
    }
    const float* row = x + (size_t)pick * ld;
    float* c = centers + (size_t)nd * c_stride;
    for (int j = 0; j < dim; ++j) c[j] = row[j];
}
}  // namespace

extern "C" int mgb_kmeanspp_pick(const float* d_x, int32_t dim, int32_t ld, const int64_t* d_starts, const int64_t* d_counts, int32_t n_nodes,
                                 const double* d_u, int32_t u_stride, const double* d_prefix, float* d_centers, int32_t center_stride,
                                 void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(n_nodes >= 0 && dim > 0 && ld >= dim && center_stride >= dim && u_stride >= 1, "bad shape");
    if (n_nodes == 0) return MGB_OK;
    MGB_REQUIRE(d_x && d_starts && d_counts && d_u && d_centers, "null argument");
    mgb_kmeanspp_pick_kernel<<<(n_nodes + 127) / 128, 128, 0, stream>>>(d_x, dim, ld, reinterpret_cast<const long long*>(d_starts),
                                                                        reinterpret_cast<const long long*>(d_counts), n_nodes, d_u, u_stride,
                                                                        d_prefix, d_centers, center_stride);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}

extern "C" int mgb_kmeans_assign_update(const float* d_x, int64_t n, int32_t dim, int32_t ld, const int32_t* d_node,
                                        const float* d_centroids, int32_t n_sub, int32_t* d_label, float* d_sums,
                                        int32_t* d_counts, float* d_inertia, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(n == 0 || (d_x && d_centroids && d_label && d_sums && d_counts), "null argument");
    MGB_REQUIRE(dim > 0 && ld >= dim && n_sub > 0, "bad shape");
    if (n == 0) return MGB_OK;
    return dispatch_level(d_x, n, dim, dim, ld, d_node, d_centroids, n_sub, d_label, d_sums, d_counts, nullptr, d_inertia, stream);
}

extern "C" int mgb_kmeans_level(const float* d_x, int64_t n, int32_t dim, int32_t dim_acc, int32_t ld, const int32_t* d_node,
                                int32_t n_nodes, int32_t n_sub, float* d_centroids, const float* d_tol, int32_t max_iter,
                                int32_t check_every, int32_t* d_label, float* d_sums, int32_t* d_counts, float* d_node_inertia,
                                int32_t* d_converged, int32_t* iters_done, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(n >= 0 && n_nodes >= 0, "negative size");
    if (iters_done) *iters_done = 0;
    if (n == 0 || n_nodes == 0) return MGB_OK;
    MGB_REQUIRE(d_x && d_centroids && d_label && d_sums && d_counts && d_converged, "null argument");
    MGB_REQUIRE(dim > 0 && dim_acc >= dim && ld >= dim_acc && n_sub > 0, "bad shape");
    MGB_REQUIRE(max_iter >= 0, "max_iter < 0");
    if (check_every < 1) check_every = 4;
    static thread_local int* h_active = nullptr;
    static thread_local int* d_active = nullptr;
    if (!h_active) {
        MGB_CUDA(cudaHostAlloc((void**)&h_active, sizeof(int), cudaHostAllocDefault));
    }
    // the counter lives with the call's device (one handle per GPU, one device per process in practice)
    static thread_local int d_active_dev = -1;
    int dev = 0;
    MGB_CUDA(cudaGetDevice(&dev));
    if (!d_active || d_active_dev != dev) {
        MGB_CUDA(cudaMalloc((void**)&d_active, 2 * sizeof(int)));
        d_active_dev = dev;
    }
    const size_t sums_bytes = (size_t)n_nodes * n_sub * dim_acc * sizeof(float);
    const size_t counts_bytes = (size_t)n_nodes * n_sub * sizeof(int);
    // fresh = true: all accumulators cleared by memsets and every tile processed (first and final pass); otherwise the centroid
    // kernel of the previous iteration has cleared what it consumed and tiles of converged nodes are skipped
    auto e_step = [&](bool fresh) -> int {
        if (fresh) {
            MGB_CUDA(cudaMemsetAsync(d_sums, 0, sums_bytes, stream));
            MGB_CUDA(cudaMemsetAsync(d_counts, 0, counts_bytes, stream));
            if (d_node_inertia) MGB_CUDA(cudaMemsetAsync(d_node_inertia, 0, (size_t)n_nodes * sizeof(float), stream));
        }
        return dispatch_level(d_x, n, dim, dim_acc, ld, d_node, d_centroids, n_sub, d_label, d_sums, d_counts, fresh ? d_node_inertia : nullptr,
                              nullptr, stream, fresh ? nullptr : d_converged);
    };
    int it = 0;
    bool done = (max_iter == 0);
    MGB_CUDA(cudaMemsetAsync(d_active, 0, 2 * sizeof(int), stream));
    while (!done) {
        int rc = e_step(it == 0);
        if (rc != MGB_OK) return rc;
        mgb_kmeans_centroid_kernel<<<(n_nodes + 127) / 128, 128, 0, stream>>>(n_nodes, n_sub, dim, dim_acc, d_sums, d_counts, d_tol,
                                                                              d_centroids, d_converged, d_active + (it & 1),
                                                                              d_active + ((it + 1) & 1), d_sums, d_counts);
        mgb_count_launch();
        MGB_CUDA(cudaGetLastError());
        ++it;
        if (it >= max_iter) break;
        if (it % check_every == 0) {
            MGB_CUDA(cudaMemcpyAsync(h_active, d_active + ((it - 1) & 1), sizeof(int), cudaMemcpyDeviceToHost, stream));
            MGB_CUDA(cudaStreamSynchronize(stream));
            if (*h_active == 0) done = true;
        }
    }
    // labels, sums, counts and inertia that belong to the final centroids (sklearn re-runs the
    // E-step after a tolerance stop; _kmeans_single_lloyd)
    int rc = e_step(true);
    if (rc != MGB_OK) return rc;
    if (iters_done) *iters_done = it;
    return MGB_OK;
}
