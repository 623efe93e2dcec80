// Exchange of the per-GPU best (cost, global index) over NVLink peer memory.
//
// The reference is single-process; its arg-min is the serial loop of
// motion_primitive_generator.py:251-257 (strict '<', first minimum wins).  With the candidates of a step
// sharded over the GPUs of one box the only exchange of the path is every rank's best (cost, index).
// A step at S takes ~76 us on one B200, so a library all-gather (tens of us of launch + protocol
// latency) would halve the throughput; here every rank stores its 16-byte record straight into every
// peer's slot array (P2P stores through NVLink / NVSwitch), then waits for the peers' records in its
// OWN memory and reduces them - one 32-thread launch, no host round trip.
//
// Slot layout per rank: 2 * world records (double-buffered by the parity of the step sequence number: a
// rank can run at most one step ahead of the slowest peer, because it needs that peer's record to finish
// its own step).  A record is {cost, seq, idx}; payload first, fence, then seq (the flag).
#include <math.h>

#include "mgb_exchange.cuh"
#include "mgb_internal.cuh"

namespace {

__global__ void __launch_bounds__(32) mgb_exchange_best_kernel(const float* __restrict__ my_cost, const long long* __restrict__ my_idx,
                                                              void* const* __restrict__ peer_slots, int rank, int world, unsigned int seq,
                                                              long long spin_limit, float* __restrict__ out_cost,
                                                              long long* __restrict__ out_idx, int* __restrict__ err) {
    // launched with programmatic stream serialization: this grid may be resident before the score kernel
    // in front of it has drained; the wait below orders the read of my_cost / my_idx after its completion
    asm volatile("griddepcontrol.wait;" ::: "memory");
    MgbExchange x;
    x.peer_slots = peer_slots; x.rank = rank; x.world = world; x.seq = seq; x.spin_limit = spin_limit; x.err = err;
    float c;
    long long i;
    mgb_exchange_warp(x, my_cost[0], my_idx[0], c, i);
    if (threadIdx.x == 0) {
        out_cost[0] = c;
        out_idx[0] = i;
    }
}

}  // namespace

extern "C" size_t mgb_exchange_slot_bytes(int32_t world) { return (size_t)2 * (world > 0 ? world : 1) * sizeof(MgbBestSlot); }

extern "C" int mgb_exchange_best(const float* d_my_cost, const int64_t* d_my_idx, void* const* d_peer_slots, int32_t rank, int32_t world,
                                 uint32_t seq, int64_t spin_limit, float* d_out_cost, int64_t* d_out_idx, int32_t* d_err, void* stream) {
    MGB_REQUIRE(d_my_cost && d_my_idx && d_peer_slots && d_out_cost && d_out_idx && d_err, "null argument");
    MGB_REQUIRE(world >= 1 && world <= 32 && rank >= 0 && rank < world, "rank / world out of range (one box: world <= 32)");
    MGB_REQUIRE(seq != 0, "seq 0 is the cleared state of the slots; start at 1");
    if (spin_limit <= 0) spin_limit = 20000000;       // ~1 s of 40 ns sleeps, then the error flag is raised
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(32);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MGB_CUDA(cudaLaunchKernelEx(&cfg, mgb_exchange_best_kernel, d_my_cost, reinterpret_cast<const long long*>(d_my_idx), d_peer_slots, (int)rank,
                                (int)world, (unsigned int)seq, (long long)spin_limit, d_out_cost, reinterpret_cast<long long*>(d_out_idx),
                                (int*)d_err));
    mgb_count_launch();
    MGB_CUDA(cudaGetLastError());
    return MGB_OK;
}
