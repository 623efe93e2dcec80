// Internal declarations shared by the translation units of libmgb200.so.
// Not part of the ABI (include/mgb200.h is).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/mgb200.h"
#include "mgb_exchange.cuh"

// ---------------------------------------------------------------------------
// error handling: thread-local message, no exceptions across the ABI
// ---------------------------------------------------------------------------
void mgb_set_error(const char* fmt, ...);
void mgb_count_launch(int n = 1);

#define MGB_CUDA(expr)                                                                  \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            mgb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return MGB_E_CUDA;                                                          \
        }                                                                               \
    } while (0)

#define MGB_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            mgb_set_error(__VA_ARGS__);   \
            return MGB_E_INVALID;         \
        }                                 \
    } while (0)

// ---------------------------------------------------------------------------
// per-DEVICE one-time kernel configuration: cudaFuncSetAttribute applies to the current device only, so a
// process that drives several GPUs must opt every one of them in (a per-process flag would leave the second
// device at the 48 KB default)
// ---------------------------------------------------------------------------
#define MGB_MAX_DEVICES 64
template <typename Kern>
inline int mgb_opt_in_smem(Kern kern, bool (&done)[MGB_MAX_DEVICES], int bytes) {
    int dev = 0;
    MGB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MGB_MAX_DEVICES || !done[dev]) {
        MGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        if (dev >= 0 && dev < MGB_MAX_DEVICES) done[dev] = true;
    }
    return MGB_OK;
}
// SM count of the CURRENT device (cached per device)
inline int mgb_num_sms() {
    static int cache[MGB_MAX_DEVICES] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MGB_MAX_DEVICES) return 148;
    if (cache[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cache[dev] = n > 0 ? n : 148;
    }
    return cache[dev];
}

// ---------------------------------------------------------------------------
// scoring program (built by mgb_constraints_pack, interpreted by the score kernels)
// ---------------------------------------------------------------------------
struct MgbOp {
    int32_t code;
    int32_t a;  // quad index / term index / scratch slot (see op list)
    int32_t b;  // second quad index / mask / count
    int32_t c;  // offset into the float constant pool
};

enum MgbOpCode : int32_t {
    OP_ROOT = 0,      // a = translation quad, b = rotation quad, c -> root offset[3]
    OP_JOINT = 1,     // a = rotation quad, c -> offset[3]
    OP_CONST = 2,     // a = 1 if rotation is identity, c -> unit quaternion[4], offset[3]
    OP_PUSH = 3,
    OP_POP = 4,
    OP_ALIGN = 5,     // a = translation quad of control point 0, c -> ref_dir[3], prev heading(rad), prev x, prev z
    OP_POS = 6,       // a = term, b = axis mask, c -> weight, target[3]
    OP_ORIENT = 7,    // a = term, c -> weight, v2[3] (unit)
    OP_REL = 8,       // a = term, c -> weight, offset[4], target[3]
    OP_DIR2D = 9,     // a = term, c -> weight, ref_dir[3], target_dir[2], target_len
    OP_LOOKAT = 10,   // a = term, c -> weight, target[3]
    OP_STORE = 11,    // a = scratch float offset: scratch[a..a+2] = position
    OP_TWOHANDS = 12, // a = term, b = scratch offset (left, right), c -> weight, center[3], p0[3], p1[3]
    OP_FEET = 13,     // a = term, b = scratch offset (left, right), c -> weight, left[3], right[3], inner weight
    OP_POSE = 14,     // a = term, b = scratch offset: points [n*3] then root2[3]; c -> weight, n, has_vel, vel[3], W, sax, saz, (target[3], w) * n
    OP_CA_BEGIN = 15, // a = scratch offset (1 float)
    OP_CA_ACC = 16,   // a = scratch offset, b = axis mask, c -> target[3]
    OP_CA_END = 17,   // a = term, b = scratch offset, c -> weight
    OP_RESET = 18,    // G = identity (between slots; not required, kept for clarity)
    OP_JROT = 19,     // a = term, b = rotation quad of the joint (raw channels, no FK), c -> weight, inner weight, target R[9], is_root
    OP_LT_BEGIN = 20, // a = scratch offset (6 floats: sum, arc length, last position[3], started), c -> start_t
    OP_LT_ACC = 21,   // a = scratch offset, b = residual index (residual program), c -> 1 / step, n_table, full arc length, weight,
                      // table (x, z) * (n_table + 2) [entry n_table + 1 = last control point]
    OP_LT_END = 22,   // a = term, b = scratch offset, c -> weight
    OP_SCR_INIT = 23, // a = scratch offset, b = count, c -> values[count]: scratch[a + i] = values[i]
    OP_TR_ACC = 24,   // TrajectoryConstraint, one frame: a = scratch offset (2 floats: sum, min_u), b = residual index (residual
                      // program), c -> n_table, weight, 0, 0, table (x, y, z) * (n_table + 1) [the spline at parameter i / n_table]
    OP_TS_ACC = 25,   // TrajectorySetConstraint, one joint of one frame: a = scratch offset (3 + 5 J floats: sum, frame
                      // accumulator, any-active; per joint arc length, last position[3], started), b = residual index << 8 |
                      // joint << 1 | closes-the-frame, c -> 1 / step, n_table, full arc length, range start, range end,
                      // 1 / (3 J), weight, 0, table (x, y, z) * (n_table + 2) [by arc length; last entry = last control point]
    OP_MEAN_END = 26, // a = term, b = scratch offset, c -> weight, 1 / n_frames: term = scratch[b] / n_frames
};

#define MGB_SCRATCH_FLOATS 160
#define MGB_STACK_DEPTH 8

// One reusable (device block, pinned mirror) pair for the tables of a constraint program: mgb_constraints_pack runs once per
// motion-primitive step on the planner's thread, so it must not cudaMalloc / cudaFree / block on a copy.  A block is filled
// through its pinned mirror with ONE asynchronous copy on the model's pack stream; launches wait for `copied` on their own
// stream and record `used` after themselves, and a later pack that reuses the block waits for `used` on the pack stream.
struct MgbPackBlock {
    unsigned char* dev = nullptr;
    unsigned char* pin = nullptr;
    size_t bytes = 0;
    cudaEvent_t copied = nullptr, used = nullptr;
    bool busy = false;
};

struct mgb_model {
    int device = 0;
    int F = 0, Nb = 0, P = 0, K = 0, D = 0, G = 0, M = 0;  // M = Nb * P
    int n_nodes = 0;
    int aligning_node = 0;
    double aligning_dir[3] = {0, 0, 1};
    // host fp64 (translation maxima already folded into E / mean)
    std::vector<double> E;     // [K][M]
    std::vector<double> mean;  // [M]
    std::vector<double> knots; // [Nb+4]
    std::vector<int> parent, type, channel;
    std::vector<double> offset;    // [n_nodes][3]
    std::vector<double> fixed_rot; // [n_nodes][9] rotation matrix of fixed joints (identity otherwise)
    std::vector<double> fixed_quat; // [n_nodes][4] the same as a unit quaternion (identity below eps)
    std::vector<std::vector<int>> children;
    // device fp32
    float* d_E = nullptr;     // [K][M]
    float* d_mean = nullptr;  // [M]
    // GMM
    int NQ = 0;               // quads per component = ceil(D/4)
    int DP = 0;               // D rounded up to a multiple of 4
    float4* d_gmm_tab = nullptr;  // [(g*NQ + q) * (DP+1) + r]: r = 0 -> -(mu_g P_g)[4q..4q+3], r = 1+k -> P_g[k][4q..4q+3] (zero for k >= D)
    float* d_gmm_c = nullptr;     // [G]  log w_g + log det - D/2 log 2 pi
    int LS = 0;                   // row stride of d_chol = D rounded up to a multiple of 4
    float* d_chol = nullptr;      // [G][D][LS] lower Cholesky of the covariances (sampling), rows 16-byte aligned
    float* d_means = nullptr;     // [G][D]
    size_t gmm_tab_bytes = 0;
    // skeleton on device for the FK kernel
    int* d_parent = nullptr;
    int* d_type = nullptr;
    int* d_channel = nullptr;
    float* d_offset = nullptr;     // [n_nodes][3]
    float* d_fixed_rot = nullptr;  // [n_nodes][9]
    // tensor-core back-projection operands (TF32 hi/lo split), built lazily
    float* d_Bsplit = nullptr;  // packed stage images of E' (see mgb_tc.cu: TcB)
    int tc_Kp = 0, tc_Np = 0, tc_NC = 0, tc_KC = 0, tc_nst = 0, tc_W = 0, tc_nbuf = 1;
    // tensor-core score path (D > 32): GMM precision factors as a packed 3xTF32 operand
    std::vector<double> prec_chol;  // [G][D][D] upper factors P_g (y = (s - mu_g) . P_g), fp64
    std::vector<double> gmm_bias;   // [G][D]   -mu_g . P_g
    float* d_Bgmm = nullptr;
    int tcg_Kp = 0, tcg_Np = 0, tcg_NC = 0, tcg_KC = 0, tcg_CW = 0;
    std::vector<double> chol64;     // [G][D][D] lower Cholesky factors of the covariances, fp64
    std::vector<double> means64;    // [G][D]
    float* d_Bsamp = nullptr;       // sampling operand: per component [L_g^T ; mu_g], one N-chunk each
    int tcs_Kp = 0, tcs_Np = 0, tcs_NC = 0, tcs_KC = 0;
    // private streams / staging for the host-buffer entries; calls on one handle are serialised by host_mutex
    std::mutex host_mutex;
    cudaStream_t copy_stream = nullptr, compute_stream = nullptr;
    void* pinned[2] = {nullptr, nullptr};
    void* d_stage[2] = {nullptr, nullptr};
    size_t stage_bytes = 0;   // size of each d_stage buffer
    size_t pinned_bytes = 0;  // size of each pinned buffer (allocated only for pageable inputs)
    void* d_ws = nullptr;
    size_t ws_bytes = 0;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    // constraint-table blocks (see MgbPackBlock); mutable state of an otherwise immutable handle, guarded by pool_mutex
    mutable std::mutex pool_mutex;
    mutable std::vector<MgbPackBlock*> pool;
    mutable cudaStream_t pack_stream = nullptr;
};

struct mgb_constraints {
    const mgb_model* model = nullptr;
    int n_terms = 0;
    int n_times = 0;
    int n_quads = 0;
    int n_ops = 0;
    int n_cst = 0;
    int KP = 0;  // K rounded up to a multiple of 4: quad q occupies tab[q * (KP + 1) ...], bias first
    bool needs_s = false;  // program itself never needs s beyond K; logpdf/magnitude decided per call
    std::vector<MgbOp> ops_fused;   // quads index the contracted table
    std::vector<MgbOp> ops_resid;   // the same program in residual-vector form: term fields hold residual indices
    int n_resid = 0;                // length of the residual vector (before the reference's zero padding to len(s))
    MgbOp* d_ops_resid = nullptr;
    std::vector<MgbOp> ops_frames;  // quads index channels of materialised frames: slot * P + channel
    std::vector<float> cst;
    std::vector<float> tab;  // [n_quads][KP+1][4]
    std::vector<double> tab64;  // same in fp64 (source of the hi/lo split of the tensor-core operand)
    std::vector<int> quad_slot;  // evaluation slot of every quad (quads of a slot are contiguous)
    float* d_Bq = nullptr;      // packed quad operand of the tensor-core score path (models with D > 32): every N-chunk
                                // holds whole slots, so the program can run chunk by chunk
    int tcq_Kp = 0, tcq_Np = 0, tcq_NC = 0, tcq_KC = 0;
    std::vector<MgbOp> ops_tc;   // program with quad fields = position inside their chunk
    std::vector<int> phase_tc;   // [n_chunks + 1] op index where the ops of chunk c begin
    MgbOp* d_ops_tc = nullptr;
    int* d_phase_tc = nullptr;
    std::vector<double> times;
    std::vector<int> spans;
    std::vector<double> weights;
    float4* d_tab = nullptr;      // the four tables below live in `block` (one allocation, one copy)
    MgbOp* d_ops_fused = nullptr;
    MgbOp* d_ops_frames = nullptr;
    float* d_cst = nullptr;
    MgbPackBlock* block = nullptr;
};

// stream ordering of a constraint handle's tables around a launch on `stream` (no-ops for c == NULL)
int mgb_constraints_before_launch(const mgb_constraints* c, cudaStream_t stream);
int mgb_constraints_after_launch(const mgb_constraints* c, cudaStream_t stream);

#define MGB_MAX_GMM 64
struct MgbGmmPrefix {  // job-wide prefix sums of the per-component sample counts (component-ordered stacking)
    long long v[MGB_MAX_GMM + 1];
};

// device-side view handed to the score kernels by value
struct ScoreParams {
    const float4* tab;  // contracted quads [n_quads][1 + KP]
    const MgbOp* ops;
    const float* cst;
    int n_ops, n_cst, n_quads;
    int K, KP, D, DP, P, n_times;
    int n_terms;
    // GMM log-density
    const float4* gmm_tab;  // [(g * NQ + q)][1 + DP]
    const float* gmm_c;
    int G, NQ;
    int gmm_in_smem;
    int tab_in_smem;
    // GMM sampling (MODE_SAMPLE)
    const float* chol;   // [G][D][LS]
    const float* means;  // [G][D]
    int LS;
    int chol_in_smem;
    unsigned long long seed;
    MgbGmmPrefix prefix;
    // inputs
    const void* s;        // [n, D] float or double
    const float* frames;  // frames mode: [n, n_times, P]
    long long n;
    long long index_base;
    int s_aligned;        // s is 16-byte aligned: full tiles travel by TMA bulk copy
    int nbuf;             // tile buffers in shared memory (2 = double-buffered)
    float err_scale, quality_scale, magnitude_scale;
    float* cost;
    float* terms;
    // selection
    int want_best;
    float* part_cost;          // [grid]
    long long* part_idx;       // [grid]
    unsigned int* done_count;  // [1]
    float* best_cost;
    long long* best_idx;
    MgbExchange xch;           // world > 1 (xch.peer_slots != NULL): the folding CTA exchanges this rank's best with its peers
    float* best_row;           // MODE_SAMPLE: [D] the winning latent vector, re-drawn by the CTA that folds the partials (or NULL)
    const float* cost_base;    // [n] or NULL: added to every candidate's cost (the naturalness / magnitude part computed by another launch)
};

// launchers implemented in the kernel translation units
// mode: 0 = candidates from d_s, 1 = from materialised frames, 2 = sampled in-kernel (h_counts, seed)
int mgb_launch_score(const mgb_model* m, const mgb_constraints* c, int mode, const void* d_s, int s_is_f64,
                     const float* d_frames, const int64_t* h_counts, uint64_t seed, int64_t n, int64_t index_base, float err_scale,
                     float quality_scale, float magnitude_scale, float* d_cost, float* d_terms, int k,
                     float* d_best_cost, int64_t* d_best_idx, void* d_ws, size_t ws_bytes,
                     cudaStream_t stream, float* d_best_row = nullptr, const MgbExchange* xch = nullptr);
// one fused sample + score + arg-min job of a multi-job launch (mgb_sample_score_multi)
struct MgbJob {
    const mgb_model* m;
    const mgb_constraints* c;
    int64_t n, index_base;
    const int64_t* h_counts;
    uint64_t seed;
    float err_scale, quality_scale, magnitude_scale;
    float* d_best_cost;
    int64_t* d_best_idx;
    float* d_best_row;
};
int mgb_launch_score_multi(const MgbJob* jobs, int n_jobs, ScoreParams* h_params, int* h_first, ScoreParams* d_params, int* d_first,
                           unsigned char* d_ws, size_t ws_stride, cudaStream_t stream);
int mgb_backproject_tc(const mgb_model* m, const float* d_s, int64_t n, float* d_coeffs, cudaStream_t stream);
int mgb_backproject_tc_prepare(mgb_model* m);
// tensor-core score path (mgb_tc.cu): operands are packed at handle creation for models with D > 32
int mgb_score_tc_prepare_model(mgb_model* m);
int mgb_score_tc_prepare_constraints(const mgb_model* m, mgb_constraints* c);
bool mgb_score_tc_usable(const mgb_model* m, const mgb_constraints* c, float quality_scale, int sample);
int mgb_launch_score_tc(const mgb_model* m, const mgb_constraints* c, const ScoreParams& p, int s_is_f64, float quality_scale,
                        int sample, cudaStream_t stream);
