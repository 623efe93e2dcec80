// The per-GPU best (cost, global index) exchange over NVLink peer memory as a DEVICE function, so that the CTA of a score
// kernel that folds the per-CTA minima can run it in its own epilogue: one launch per step instead of two (the separate
// mgb_exchange_best launch cost 5 - 9 us of an ~75 us step at N = 2 .. 8).  Protocol: mgb_exchange.cu.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

struct __align__(16) MgbBestSlot {
    float cost;
    unsigned int seq;
    long long idx;
};

// device-side description of one exchange (all-zero = no exchange)
struct MgbExchange {
    void* const* peer_slots;   // [world] slot arrays of every rank (peer memory), this rank's own included
    int rank, world;
    unsigned int seq;          // step sequence number (> 0); slots are double-buffered by its parity
    long long spin_limit;
    int* err;                  // [1] set to 1 + peer when a record never arrived
};

__device__ __forceinline__ void mgb_st_release_sys_u32(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int mgb_ld_acquire_sys_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Called by ONE full warp (32 converged lanes).  (mc, mi): this rank's best.  Returns the job-wide winner in (c, i) of every
// lane: lowest cost, lowest global index on ties, NaN never wins, (inf, 0) when nothing is finite.
__device__ __forceinline__ void mgb_exchange_warp(const MgbExchange& x, float mc, long long mi, float& c, long long& i) {
    const int lane = threadIdx.x & 31;
    const int buf = (int)(x.seq & 1u) * x.world;
    c = INFINITY;
    i = 0x7fffffffffffffffLL;
    bool have = false;
    if (lane < x.world) {
        MgbBestSlot* dst = reinterpret_cast<MgbBestSlot*>(x.peer_slots[lane]) + buf + x.rank;   // my record in peer `lane`'s array
        *reinterpret_cast<volatile float*>(&dst->cost) = mc;
        *reinterpret_cast<volatile long long*>(&dst->idx) = mi;
        mgb_st_release_sys_u32(&dst->seq, x.seq);
        const MgbBestSlot* src = reinterpret_cast<const MgbBestSlot*>(x.peer_slots[x.rank]) + buf + lane;   // peer `lane`'s record in MY array
        long long spins = 0;
        while (mgb_ld_acquire_sys_u32(&src->seq) != x.seq) {
            if (++spins > x.spin_limit) {
                atomicExch(x.err, 1 + lane);
                break;
            }
            __nanosleep(40);
        }
        c = *reinterpret_cast<const volatile float*>(&src->cost);
        i = *reinterpret_cast<const volatile long long*>(&src->idx);
        have = (c == c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(0xffffffffu, c, o);
        const long long oi = __shfl_xor_sync(0xffffffffu, i, o);
        const bool oh = __shfl_xor_sync(0xffffffffu, have ? 1 : 0, o) != 0;
        const bool take = oh && (!have || oc < c || (oc == c && oi < i));
        if (take) { c = oc; i = oi; have = true; }
    }
    if (!have) { c = INFINITY; i = 0; }
}
