// Micro-benchmark: HBM write bandwidth of the store patterns a [n, M] row-major tile writer can use.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o store_patterns store_patterns.cu
#include <cstdio>
#include <cuda_runtime.h>
// each CTA owns tiles of 128 rows x M floats (contiguous); 8 warps
// mode 0: linear float2 stores over the tile (address order)
// mode 1: strips: warp w handles 32-col blocks (b % 2 == w / 4), rows 32 * (w % 4) ..+31, 2 rows x 128 B per instruction
// mode 2: like 1 but a warp owns a 256-col chunk quarter: rows 32*(w%4).., all blocks of a chunk in sequence (1 KB run per row over time)
// mode 3: warp per row: each warp writes full rows (256 B per instruction, consecutive)
__global__ void __launch_bounds__(256) k(float* out, long long n, int M, int mode) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_tiles = n / 128;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        float* base = out + t * 128 * (long long)M;
        const float2 v = make_float2((float)t, (float)lane);
        if (mode == 0) {
            const long long tot2 = 128LL * M / 2;
            for (long long e = threadIdx.x; e < tot2; e += 256) __stcs(reinterpret_cast<float2*>(base) + e, v);
        } else if (mode == 1 || mode == 2) {
            const int wq = warp & 3, half = warp >> 2;
            const int nblk = M / 32;
            for (int c = 0; c * 8 < nblk; ++c) {
                for (int bb = half; bb < 8; bb += 2) {
                    const int b = (mode == 1) ? c * 8 + bb : c * 8 + half * 4 + (bb >> 1);
                    if (b >= nblk) continue;
                    float* ob = base + (size_t)(32 * wq) * M + b * 32;
                    const int ci = 2 * (lane & 15);
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int r = 2 * q + (lane >> 4);
                        __stcs(reinterpret_cast<float2*>(ob + (size_t)r * M + ci), v);
                    }
                }
                __syncthreads();
            }
        } else if (mode == 7 || mode == 8 || mode == 9) {
            // replica of the K2-TC epilogue: group = warp / 4 takes chunks (256 cols) of its parity, warp % 4 = 32-row quarter,
            // groups of W = 128 cols: rows one after the other, 2 x 256 B per row (mode 7); mode 8: same, 16 rows per
            // warp-step interleaved between the two groups' warps (all 8 warps on the same chunk); mode 9: W = 256
            const int wq = warp & 3, grp = warp >> 2;
            const int W = (mode == 9) ? 256 : 128;
            const int nchunk = (M + 255) / 256;
            for (int c = (mode == 8 ? 0 : grp); c < nchunk; c += (mode == 8 ? 1 : 2)) {
                for (int w0 = 0; w0 < 256; w0 += W) {
                    const int col0 = c * 256 + w0;
                    const int ncol = min(W, M - col0);
                    if (ncol <= 0) break;
                    const int rbase = (mode == 8) ? 32 * wq + 16 * grp : 32 * wq;
                    const int nr = (mode == 8) ? 16 : 32;
                    float* ob = base + (size_t)rbase * M + col0;
                    for (int r = 0; r < nr; ++r)
                        for (int ci = 2 * lane; ci < ncol; ci += 64) __stcs(reinterpret_cast<float2*>(ob + (size_t)r * M + ci), v);
                    __syncwarp();
                }
            }
        } else if (mode == 5) {
            // warp w: rows 32 (w % 4) .. +31, chunks (256 cols) of parity w / 4, blocks of a chunk in sequence, 2 rows x 128 B per instruction
            const int wq = warp & 3, par = warp >> 2;
            const int nchunk = (M + 255) / 256;
            for (int c = par; c < nchunk; c += 2) {
                for (int bb = 0; bb < 8; ++bb) {
                    const int col0 = c * 256 + bb * 32;
                    if (col0 + 32 > M) break;
                    float* ob = base + (size_t)(32 * wq) * M + col0;
                    const int ci = 2 * (lane & 15);
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        const int r = 2 * q + (lane >> 4);
                        __stcs(reinterpret_cast<float2*>(ob + (size_t)r * M + ci), v);
                    }
                    __syncwarp();
                }
            }
        } else if (mode >= 64) {
            // warp w owns rows 16 w .. 16 w + 15; per column chunk of `mode` floats it writes each row's run at once
            const int W = mode;
            for (int c0 = 0; c0 < M; c0 += W) {
                const int w2 = min(W, M - c0) / 2;
                for (int r = 16 * warp; r < 16 * warp + 16; ++r) {
                    float* ob = base + (size_t)r * M + c0;
                    for (int e = lane; e < w2; e += 32) __stcs(reinterpret_cast<float2*>(ob) + e, v);
                }
                __syncthreads();
            }
        } else {
            for (int r = warp; r < 128; r += 8) {
                float* ob = base + (size_t)r * M;
                for (int e = lane; e < M / 2; e += 32) __stcs(reinterpret_cast<float2*>(ob) + e, v);
            }
        }
    }
}
int main() {
    const long long n = 1000000 / 128 * 128;
    for (int M : {790, 5660}) {
        const long long nn = (M < 1000) ? n : n / 4;
        float* d;
        cudaMalloc(&d, nn * M * sizeof(float));
        for (int mode : {0, 3, 7, 8, 9, 128, 256}) {
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            float best = 1e9;
            for (int rep = 0; rep < 4; ++rep) {
                cudaEventRecord(e0);
                k<<<148, 256>>>(d, nn, M, mode);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (ms < best) best = ms;
            }
            printf("M=%d mode=%d grid=148: %.3f ms  %.0f GB/s\n", M, mode, best, nn * M * 4.0 / best / 1e6);
        }
        cudaFree(d);
    }
    return 0;
}
