// Micro-benchmark 2: the K-step loop of the tensor-core score kernel (wait on a full barrier, fence, 3 MMAs of 128 x N x 8
// tf32, commit) issued (a) by one lane inside `if (lane == 0)` and (b) by a converged warp with an elected lane.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I morphablegraphs_b200/csrc -o tools/micro/mma_issue2 tools/micro/mma_issue2.cu
#include <cstdio>
#include <cuda_runtime.h>

#include "mgb_tc.cuh"

__device__ __forceinline__ bool elect_one() {
    unsigned int pred;
    asm volatile(
        "{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
}

template <int MODE>
__device__ __forceinline__ void k_loop(uint32_t tmem_base, uint32_t a_addr, uint32_t b_addr, int N, int n_steps, unsigned long long* bar_ready,
                                       unsigned long long* bar_misc, bool leader) {
    const uint32_t idesc = mgb_umma_idesc_tf32(128, N);
    const uint64_t da_hi0 = mgb_umma_desc(a_addr, 128 * 16, 128), da_lo0 = mgb_umma_desc(a_addr + 53248, 128 * 16, 128);
    const uint64_t db0 = mgb_umma_desc(0u, (uint32_t)N * 16, 128);
    const uint32_t ring16 = b_addr >> 4, stage16 = (uint32_t)(N * 8 * 8) >> 4, lo16 = (uint32_t)(N * 8) >> 2;
    int st = 0;
    uint32_t b16 = ring16, a16 = 0, acc = 0;
    for (int i = 0; i < n_steps; ++i) {
        mgb_mbar_wait(&bar_ready[st], 0);
        mgb_tc_fence_after();
        if (MODE == 0 || leader) {
            const uint64_t da_hi = da_hi0 + a16, da_lo = da_lo0 + a16;
            const uint64_t db_hi = db0 + b16, db_lo = db_hi + lo16;
            mgb_umma_tf32(tmem_base, da_lo, db_hi, idesc, acc);
            mgb_umma_tf32(tmem_base, da_hi, db_lo, idesc, 1u);
            mgb_umma_tf32(tmem_base, da_hi, db_hi, idesc, 1u);
            mgb_umma_commit(&bar_misc[st]);
        }
        if (MODE == 1) __syncwarp();
        acc = 1u;
        a16 += 256; if (a16 >= 13 * 256) { a16 = 0; }
        b16 += stage16;
        if (++st == 3) { st = 0; b16 = ring16; }
    }
}

template <int MODE>
__global__ void __launch_bounds__(128, 1) mma_loop(int N, int n_steps, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ unsigned long long bar_done, bar_misc[4], bar_ready[4];
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.f;
    if (threadIdx.x == 0) {
        mgb_mbar_init(&bar_done, 1);
        for (int i = 0; i < 4; ++i) { mgb_mbar_init(&bar_misc[i], 1); mgb_mbar_init(&bar_ready[i], 1); }
        mgb_fence_barrier_init();
    }
    if (warp == 0) mgb_tmem_alloc(&tmem_ptr, 512);
    mgb_tc_fence_before();
    __syncthreads();
    mgb_tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    if (threadIdx.x == 0) for (int i = 0; i < 4; ++i) mgb_mbar_arrive(&bar_ready[i]);
    __syncthreads();
    if (warp == 1) {
        const bool leader = elect_one();
        if (MODE == 1 || lane == 0) {
            mgb_fence_proxy_async();
            const long long t0 = clock64();
            k_loop<MODE>(tmem_base, mgb_smem_u32(smem), mgb_smem_u32(smem + 112 * 1024), N, n_steps, bar_ready, bar_misc, leader);
            const long long t1 = clock64();
            if (MODE == 0 || leader) mgb_umma_commit(&bar_done);
            mgb_mbar_wait(&bar_done, 0);
            const long long t2 = clock64();
            if (blockIdx.x == 0 && lane == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
        }
    }
    mgb_tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        mgb_tc_fence_after();
        mgb_tmem_dealloc(tmem_base, 512);
    }
}

int main() {
    long long* d_out;
    cudaMalloc(&d_out, 16);
    cudaFuncSetAttribute(mma_loop<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(mma_loop<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int n_steps = 169;
    for (int N : {128, 256})
        for (int mode = 0; mode < 2; ++mode) {
            long long h[2];
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) mma_loop<0><<<148, 128, 200 * 1024>>>(N, n_steps, d_out);
                else mma_loop<1><<<148, 128, 200 * 1024>>>(N, n_steps, d_out);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost);
            printf("mode %d (%s) N %d: issue %.1f cycles per K-step (3 MMAs + commit), total %.1f; pipe floor %.1f\n", mode,
                   mode ? "converged warp, elected lane" : "one lane in a divergent branch", N, (double)h[0] / n_steps, (double)h[1] / n_steps,
                   3.0 * 128.0 * N / 256.0);
        }
    return 0;
}
