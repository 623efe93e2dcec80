// K11: covariance of the euclidean-feature PCA on the tensor cores (sm_100a, tcgen05 / TMEM).
//
//   C += (X - mean)^T (X - mean),   X [n, d] float32 row-major (d = frames x joints x 3 = 1500 for the S model, n = 1M)
//
// What it replaces: the covariance inside sklearn.decomposition.PCA(n_components=0.95).fit_transform as the reference calls it
// from map_motions_to_euclidean_pca (space_partitioning/features.py:125-130) - a 2 n d^2 = 4.5 TFLOP contraction at n = 1M,
// the dominant cost of the euclidean-feature cluster-tree build.
//
// 3xTF32 (a_lo.b_hi + a_hi.b_lo + a_hi.b_hi, fp32 accumulate in TMEM), flushed to a float64 matrix per work item.
// Work item = (128 x 256 tile of the upper triangle, slice of the samples); persistent CTAs walk the item list.
//   warps 0-7   producers: coalesced loads of X (a warp reads 32 consecutive features of one sample), mean subtracted,
//               TF32 hi / lo split, written TRANSPOSED into the K-major canonical UMMA layout (one thread = one feature,
//               four consecutive samples = one 16-byte store: conflict-free)
//   warp 8      MMA issuer (converged warp, elected lane): 128 x 256 x 8 tf32, 6 per stage of 16 samples, blocks of 512
//               samples alternating between two TMEM accumulators
//   warps 9-12  epilogue: every finished block (tcgen05.ld) added to a float32 matrix in shared memory, the matrix to the
//               float64 result with atomic adds at the end of the item
#include <math.h>
#include <string.h>

#include <algorithm>

#include "mgb_internal.cuh"
#include "mgb_tc.cuh"

namespace {

constexpr int SY_M = 128, SY_N = 256, SY_KS = 16;      // tile rows, tile columns, samples per stage
constexpr int SY_STAGES = 2;
constexpr int SY_BLK = 512;                            // samples per TMEM accumulation block (see below)
constexpr int SY_PRODUCERS = 256;
constexpr int SY_EPI = 128;
constexpr int SY_THREADS = SY_PRODUCERS + 32 + SY_EPI;   // warps 0-7 producers, 8 MMA issuer, 9-12 epilogue
constexpr size_t SY_A_BYTES = (size_t)SY_M * SY_KS * 4;     // one of hi / lo
constexpr size_t SY_B_BYTES = (size_t)SY_N * SY_KS * 4;
constexpr size_t SY_STAGE_BYTES = 2 * (SY_A_BYTES + SY_B_BYTES);   // [A hi][A lo][B hi][B lo] = 48 KB
constexpr size_t SY_ACC_BYTES = (size_t)SY_M * SY_N * 4;           // float32 block sums [column][row] = 128 KB

struct SyrkParams {
    const float* X;
    const float* mean;   // [d] or NULL
    double* C;           // [d, d] row-major, upper triangle (i <= j) accumulated
    long long n;
    int d, ld;
    int n_tiles, n_slices;
    long long slice_len;             // samples per slice (a multiple of SY_BLK)
    const int* tile_rc;              // [n_tiles][2] first row / first column of every tile
};

// The tensor core adds into its fp32 accumulator with truncation: a sum of n same-sign products (a variance) comes out low by
// ~n / 8 x 3 x 2^-25 relative - 5e-5 after 5 000 samples, 7e-4 after 66 000.  The accumulation is therefore cut into blocks of
// SY_BLK samples that alternate between two TMEM accumulators; the epilogue warps add every finished block into a float32
// matrix in shared memory with round-to-nearest (block bias ~6e-6), and that matrix goes to the float64 result once per item.
__global__ void __launch_bounds__(SY_THREADS, 1) mgb_syrk_kernel(const __grid_constant__ SyrkParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ unsigned long long full[SY_STAGES], empty[SY_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    float* acc_sm = reinterpret_cast<float*>(smem + SY_STAGES * SY_STAGE_BYTES);   // [SY_N][SY_M]
    if (tid == 0) {
        for (int i = 0; i < SY_STAGES; ++i) {
            mgb_mbar_init(&full[i], SY_PRODUCERS);
            mgb_mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mgb_mbar_init(&acc_full[i], 1);
            mgb_mbar_init(&acc_empty[i], SY_EPI);
        }
        mgb_fence_barrier_init();
    }
    if (warp == 8) mgb_tmem_alloc(&tmem_ptr, 2 * SY_N);
    mgb_tc_fence_before();
    __syncthreads();
    mgb_tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    const int n_items = p.n_tiles * p.n_slices;

    if (warp < 8) {
        // ================= producers =================
        const int ia = tid & 127, kqa = tid >> 7;       // A: feature ia of the tile's rows, sample quads kqa and kqa + 2
        const int jb = tid;                             // B: feature jb of the tile's columns, all four sample quads
        unsigned int st = 0, ph = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int tile = item % p.n_tiles, slice = item / p.n_tiles;
            const int i0 = __ldg(p.tile_rc + 2 * tile), j0 = __ldg(p.tile_rc + 2 * tile + 1);
            const long long s0 = (long long)slice * p.slice_len, s1 = min(p.n, s0 + p.slice_len);
            const bool va = i0 + ia < p.d, vb = j0 + jb < p.d;
            const float ma = (va && p.mean) ? __ldg(p.mean + i0 + ia) : 0.f, mb = (vb && p.mean) ? __ldg(p.mean + j0 + jb) : 0.f;
            const float* xa = p.X + i0 + ia;
            const float* xb = p.X + j0 + jb;
            // software pipeline: the 24 loads of stage k + 1 are in flight while stage k is split and stored
            float a[2][4], b[4][4], an[2][4], bn[4][4];
            auto load_stage = [&](long long k0, float (&ra)[2][4], float (&rb)[4][4]) {
#pragma unroll
                for (int q = 0; q < 2; ++q)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const long long k = k0 + 4 * (kqa + 2 * q) + kk;
                        ra[q][kk] = (va && k < s1) ? __ldg(xa + (size_t)k * p.ld) - ma : 0.f;
                    }
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const long long k = k0 + 4 * q + kk;
                        rb[q][kk] = (vb && k < s1) ? __ldg(xb + (size_t)k * p.ld) - mb : 0.f;
                    }
            };
            load_stage(s0, an, bn);
            for (long long k0 = s0; k0 < s1; k0 += SY_KS) {
#pragma unroll
                for (int q = 0; q < 2; ++q)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) a[q][kk] = an[q][kk];
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) b[q][kk] = bn[q][kk];
                if (k0 + SY_KS < s1) load_stage(k0 + SY_KS, an, bn);
                mgb_mbar_wait(&empty[st], ph ^ 1u);
                unsigned char* sa_hi = smem + (size_t)st * SY_STAGE_BYTES;
                unsigned char* sa_lo = sa_hi + SY_A_BYTES;
                unsigned char* sb_hi = sa_lo + SY_A_BYTES;
                unsigned char* sb_lo = sb_hi + SY_B_BYTES;
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    float h[4], l[4];
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) mgb_tf32_split(a[q][kk], h[kk], l[kk]);
                    const uint32_t off = mgb_canon_off((uint32_t)ia, (uint32_t)(4 * (kqa + 2 * q)), SY_M);
                    *reinterpret_cast<float4*>(sa_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
                    *reinterpret_cast<float4*>(sa_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float h[4], l[4];
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) mgb_tf32_split(b[q][kk], h[kk], l[kk]);
                    const uint32_t off = mgb_canon_off((uint32_t)jb, (uint32_t)(4 * q), SY_N);
                    *reinterpret_cast<float4*>(sb_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
                    *reinterpret_cast<float4*>(sb_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
                }
                mgb_fence_proxy_async();
                mgb_mbar_arrive(&full[st]);
                if (++st == SY_STAGES) { st = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 8) {
        // ================= MMA issuer =================
        const bool leader = mgb_elect_one();
        const uint32_t idesc = mgb_umma_idesc_tf32(SY_M, SY_N);
        const uint32_t base = mgb_smem_u32(smem);
        unsigned int st = 0, ph = 0, blocks = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int slice = item / p.n_tiles;
            const long long s0 = (long long)slice * p.slice_len, s1 = min(p.n, s0 + p.slice_len);
            for (long long b0 = s0; b0 < s1; b0 += SY_BLK, ++blocks) {
                const unsigned int ab = blocks & 1u;
                mgb_mbar_wait(&acc_empty[ab], ((blocks >> 1) & 1u) ^ 1u);      // this accumulator's previous block has been drained
                mgb_tc_fence_after();
                const uint32_t d_addr = tmem_base + ab * SY_N;
                uint32_t acc = 0;
                const long long b1 = min(s1, b0 + SY_BLK);
                for (long long k0 = b0; k0 < b1; k0 += SY_KS) {
                    mgb_mbar_wait(&full[st], ph);
                    mgb_tc_fence_after();
                    if (leader) {
                        const uint32_t a_hi = base + (uint32_t)(st * SY_STAGE_BYTES), a_lo = a_hi + (uint32_t)SY_A_BYTES;
                        const uint32_t b_hi = a_lo + (uint32_t)SY_A_BYTES, b_lo = b_hi + (uint32_t)SY_B_BYTES;
#pragma unroll
                        for (int s = 0; s < SY_KS / 8; ++s) {       // one K slab = 8 samples = two 16-byte sample quads
                            const uint64_t da_hi = mgb_umma_desc(a_hi + s * 2 * SY_M * 16, SY_M * 16, 128);
                            const uint64_t da_lo = mgb_umma_desc(a_lo + s * 2 * SY_M * 16, SY_M * 16, 128);
                            const uint64_t db_hi = mgb_umma_desc(b_hi + s * 2 * SY_N * 16, SY_N * 16, 128);
                            const uint64_t db_lo = mgb_umma_desc(b_lo + s * 2 * SY_N * 16, SY_N * 16, 128);
                            mgb_umma_tf32(d_addr, da_lo, db_hi, idesc, s == 0 ? acc : 1u);
                            mgb_umma_tf32(d_addr, da_hi, db_lo, idesc, 1u);
                            mgb_umma_tf32(d_addr, da_hi, db_hi, idesc, 1u);
                        }
                        mgb_umma_commit(&empty[st]);
                    }
                    acc = 1u;
                    __syncwarp();
                    if (++st == SY_STAGES) { st = 0; ph ^= 1u; }
                }
                if (leader) mgb_umma_commit(&acc_full[ab]);
                __syncwarp();
            }
        }
    } else {
        // ================= epilogue: block sums -> shared float32 matrix -> float64 result =================
        const int wq = warp & 3;                        // TMEM lane quarter of this warp (warps 9..12 -> 1, 2, 3, 0)
        const int row = 32 * wq + lane;
        unsigned int blocks = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int tile = item % p.n_tiles, slice = item / p.n_tiles;
            const int i0 = __ldg(p.tile_rc + 2 * tile), j0 = __ldg(p.tile_rc + 2 * tile + 1);
            const long long s0 = (long long)slice * p.slice_len, s1 = min(p.n, s0 + p.slice_len);
            for (int c = 0; c < SY_N; ++c) acc_sm[c * SY_M + row] = 0.f;
            for (long long b0 = s0; b0 < s1; b0 += SY_BLK, ++blocks) {
                const unsigned int ab = blocks & 1u;
                mgb_mbar_wait(&acc_full[ab], (blocks >> 1) & 1u);
                mgb_tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(32 * wq) << 16) + ab * SY_N;
                for (int c0 = 0; c0 < SY_N; c0 += 32) {
                    float v[32];
                    mgb_tmem_ld32(taddr + c0, v);
#pragma unroll
                    for (int c = 0; c < 32; ++c) acc_sm[(c0 + c) * SY_M + row] += v[c];
                }
                mgb_tc_fence_before();
                mgb_mbar_arrive(&acc_empty[ab]);
            }
            const int gi = i0 + row;
            if (gi < p.d)
                for (int c = 0; c < SY_N; ++c) {
                    const int gj = j0 + c;
                    if (gj < p.d && gj >= gi) atomicAdd(p.C + (size_t)gi * p.d + gj, (double)acc_sm[c * SY_M + row]);
                }
        }
    }
    mgb_tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        mgb_tc_fence_after();
        mgb_tmem_dealloc(tmem_base, 2 * SY_N);
    }
}

}  // namespace

// C[d, d] (float64, device, row-major) += (X - mean)^T (X - mean) for the upper triangle i <= j; the caller mirrors it.
extern "C" int mgb_syrk_upper(const float* d_X, int64_t n, int32_t d, int32_t ld, const float* d_mean, double* d_C, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    MGB_REQUIRE(d_X && d_C && n >= 0 && d > 0 && ld >= d, "bad argument");
    if (n == 0) return MGB_OK;
    std::vector<int> rc;
    for (int j0 = 0; j0 < d; j0 += SY_N)
        for (int i0 = 0; i0 < d && i0 < j0 + SY_N; i0 += SY_M) {
            rc.push_back(i0);
            rc.push_back(j0);
        }
    SyrkParams p;
    memset(&p, 0, sizeof(p));
    p.X = d_X; p.mean = d_mean; p.C = d_C; p.n = n; p.d = d; p.ld = ld;
    p.n_tiles = (int)rc.size() / 2;
    const int sms = mgb_num_sms();
    // enough slices that the persistent CTAs balance (>= 4 items per CTA), slices of at least 4096 samples
    long long want = std::max<long long>(1, (4LL * sms + p.n_tiles - 1) / p.n_tiles);
    want = std::min<long long>(want, std::max<long long>(1, n / 4096));
    long long slice = (n + want - 1) / want;
    slice = (slice + SY_BLK - 1) / SY_BLK * SY_BLK;
    p.slice_len = slice;
    p.n_slices = (int)((n + slice - 1) / slice);
    int* d_rc = nullptr;
    MGB_CUDA(cudaMallocAsync((void**)&d_rc, rc.size() * sizeof(int), stream));
    MGB_CUDA(cudaMemcpyAsync(d_rc, rc.data(), rc.size() * sizeof(int), cudaMemcpyHostToDevice, stream));
    MGB_CUDA(cudaStreamSynchronize(stream));     // rc is a host temporary
    p.tile_rc = d_rc;
    const size_t smem = SY_STAGES * SY_STAGE_BYTES + SY_ACC_BYTES;
    static bool configured[MGB_MAX_DEVICES] = {false};
    if (int r = mgb_opt_in_smem(mgb_syrk_kernel, configured, (int)smem)) return r;
    const int grid = std::min(sms, p.n_tiles * p.n_slices);
    mgb_syrk_kernel<<<grid, SY_THREADS, smem, stream>>>(p);
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    MGB_CUDA(cudaFreeAsync(d_rc, stream));
    return MGB_OK;
}
