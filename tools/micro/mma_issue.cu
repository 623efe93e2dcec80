// Micro-benchmark: how fast can ONE lane issue tcgen05.mma kind::tf32 (128 x N x 8, no-swizzle K-major operands in
// shared memory), and what do commit / fence / barrier round trips add?  Measured per SM with clock64.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I morphablegraphs_b200/csrc -o /tmp/mma_issue tools/micro/mma_issue.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "mgb_tc.cuh"

// variant 0: back-to-back MMAs, one commit at the end
// variant 1: commit to a barrier after every `per` MMAs (nobody waits on it)
// variant 2: as 1, plus tcgen05.fence::after_thread_sync before every group
// variant 3: as 1, plus a try_wait on an already completed barrier before every group
__global__ void __launch_bounds__(128, 1) mma_issue(int N, int n_mma, int per, int variant, int a_stride16, int b_stride16, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ unsigned long long bar_done, bar_misc[8], bar_ready;
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.f;
    if (threadIdx.x == 0) {
        mgb_mbar_init(&bar_done, 1);
        for (int i = 0; i < 8; ++i) mgb_mbar_init(&bar_misc[i], 1);
        mgb_mbar_init(&bar_ready, 1);
        mgb_fence_barrier_init();
    }
    if (warp == 0) mgb_tmem_alloc(&tmem_ptr, 512);
    mgb_tc_fence_before();
    __syncthreads();
    mgb_tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    if (threadIdx.x == 0) mgb_mbar_arrive(&bar_ready);  // phase 0 of bar_ready completes: a wait on parity 0 succeeds at once
    __syncthreads();
    if (warp == 1 && lane == 0) {
        mgb_fence_proxy_async();
        const uint32_t a_addr = mgb_smem_u32(smem), b_addr = mgb_smem_u32(smem + 64 * 1024);
        const uint64_t da0 = mgb_umma_desc(a_addr, 128 * 16, 128), db0 = mgb_umma_desc(b_addr, (uint32_t)N * 16, 128);
        const uint32_t idesc = mgb_umma_idesc_tf32(128, N);
        const long long t0 = clock64();
        uint32_t a16 = 0, b16 = 0;
        int cnt = 0, grp = 0;
        for (int i = 0; i < n_mma; ++i) {
            if (variant >= 1 && cnt == 0) {
                if (variant == 3) mgb_mbar_wait(&bar_ready, 0);
                if (variant >= 2) mgb_tc_fence_after();
            }
            mgb_umma_tf32(tmem_base, da0 + a16, db0 + b16, idesc, i ? 1u : 0u);
            a16 += a_stride16; if (a16 >= 2048) a16 = 0;
            b16 += b_stride16; if (b16 >= 4096) b16 = 0;
            if (++cnt == per) {
                cnt = 0;
                if (variant >= 1) mgb_umma_commit(&bar_misc[grp]);
                grp = (grp + 1) & 7;
            }
        }
        const long long t1 = clock64();
        mgb_umma_commit(&bar_done);
        mgb_mbar_wait(&bar_done, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    mgb_tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        mgb_tc_fence_after();
        mgb_tmem_dealloc(tmem_base, 512);
    }
}

int main(int argc, char** argv) {
    long long* d_out;
    cudaMalloc(&d_out, 16);
    cudaFuncSetAttribute(mma_issue, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int n_mma = 390;
    printf("%-8s %-8s %-4s %-8s %-10s %-12s %-12s %-10s\n", "grid", "variant", "N", "per", "strides", "issue cyc/mma", "total cyc/mma", "floor");
    for (int N : {128, 256})
        for (int per : {1, 2, 3, 6, 13, 39})
            for (int variant = 0; variant < 4; ++variant) {
                const int a_s = 256, b_s = 2 * N;
                long long h[2] = {0, 0};
                for (int rep = 0; rep < 2; ++rep) {
                    mma_issue<<<148, 128, 200 * 1024>>>(N, n_mma, per, variant, a_s, b_s, d_out);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                }
                cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost);
                printf("%-8d %-8d %-4d %-8d %-10s %-12.1f %-12.1f %-10.1f\n", 148, variant, N, per, "advance", (double)h[0] / n_mma,
                       (double)h[1] / n_mma, 128.0 * N / 256.0);
            }
    return 0;
}
