/*
 * mgb200.h -- C ABI of libmgb200.so: the B200 (sm_100a) engine for the
 * sample-and-evaluate hot path of dfki-asr/morphablegraphs.
 *
 * The reference has no FFI of its own (it is pure Python); the boundary it
 * already exercises is the duck-typed batched API towards the absent `mgrd`
 * package (reference morphablegraphs/motion_generator/mgrd_sample_filter.py:62-75,
 * motion_primitive_generator.py:167-177).  Every entry point below names the
 * reference function(s) it replaces.  INTEGRATION.md shows the ctypes stub a
 * maintainer adds on the reference side.
 *
 * Conventions
 *  - return 0 on success, a negative MGB_E_* code otherwise; the message is
 *    available from mgb_last_error() (thread-local).  No C++ exception
 *    crosses the ABI.
 *  - `d_*` pointers are CALLER-OWNED DEVICE memory on the handle's device
 *    (e.g. torch tensor data_ptr()), fp32 row-major unless stated otherwise.
 *    `h_*` pointers are host memory.
 *  - calls taking `stream` are asynchronous on that cudaStream_t (passed as
 *    void*); NULL is the legacy default stream.
 *  - the tables of a model / constraints handle are immutable after creation, and
 *    the entries that take caller-owned device pointers + a stream (mgb_sample,
 *    mgb_score, mgb_sample_score, mgb_backproject, ...) may be called on one
 *    handle from several host threads, each with its own workspace and stream.
 *    The HOST-BUFFER entries use buffers owned by the library: mgb_score_host stages
 *    through buffers and streams of the model handle (calls on one model handle are
 *    serialised by a per-handle mutex, different handles run concurrently);
 *    mgb_sample_score_best[_sharded] and mgb_sample_score_multi return their results
 *    through one pinned record block per DEVICE (calls on one device are serialised);
 *    mgb_constraints_pack takes its table block from a pool of the model handle
 *    (mutex-protected) and uploads it on the handle's own stream - launches order
 *    themselves against that upload with events.
 *  - there is no CPU fallback: every compute entry fails with MGB_E_CUDA when
 *    no sm_100-class device is usable.
 */
#ifndef MGB200_H
#define MGB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGB_OK 0
#define MGB_E_INVALID (-1)     /* bad argument / unsupported shape            */
#define MGB_E_CUDA (-2)        /* CUDA runtime error (message has the detail) */
#define MGB_E_UNSUPPORTED (-3) /* constraint type outside the device path     */
#define MGB_E_WORKSPACE (-4)   /* workspace too small                         */

typedef struct mgb_model mgb_model;
typedef struct mgb_constraints mgb_constraints;

/* Skeleton node types (anim_utils SKELETON_NODE_TYPE_*). */
#define MGB_NODE_ROOT 0
#define MGB_NODE_JOINT 1
#define MGB_NODE_END_SITE 2

/* One motion primitive: spatial fPCA basis + GMM + skeleton.  All arrays are
 * host float64, exactly the numbers of the model JSON.
 * Replaces: MotionPrimitive._initialize_from_json
 *   (reference motion_model/motion_primitive.py:93-180) and the skeleton
 *   fields the reference reads (utilities/__init__.py:13-16). */
typedef struct {
    int32_t n_frames;   /* F   n_canonical_frames                            */
    int32_t n_basis;    /* Nb  n_basis_spatial                               */
    int32_t n_dim;      /* P   n_dim_spatial = 3 + 4 * animated joints       */
    int32_t n_spatial;  /* K   spatial latent dims                           */
    int32_t n_latent;   /* D   GMM dimension (K + K_t)                       */
    int32_t n_gmm;      /* G   mixture components                            */
    const double* eigen;              /* [K, Nb*P]  eigen_vectors_spatial    */
    const double* mean;               /* [Nb*P]     mean_spatial_vector      */
    const double* knots;              /* [Nb+4]     b_spline_knots_spatial   */
    const double* translation_maxima; /* [3]                                 */
    const double* gmm_weights;        /* [G]                                 */
    const double* gmm_means;          /* [G, D]                              */
    const double* gmm_covars;         /* [G, D, D]                           */
    int32_t n_nodes;
    const int32_t* node_parent;  /* [n_nodes] parent index, -1 for the root; parents precede children */
    const int32_t* node_type;    /* [n_nodes] MGB_NODE_*                     */
    const int32_t* node_channel; /* [n_nodes] quaternion_frame_index, -1 if fixed */
    const double* node_offset;   /* [n_nodes, 3]                             */
    const double* node_rotation; /* [n_nodes, 4] (w,x,y,z), used for fixed joints */
    int32_t aligning_node;       /* skeleton.aligning_root_node              */
    double aligning_dir[3];      /* skeleton.aligning_root_dir               */
} mgb_model_desc;

/* Constraint term types (reference constraints/spatial_constraints/__init__.py:23-32). */
#define MGB_CT_POSITION 0   /* keyframe_position: GlobalTransformConstraint          */
#define MGB_CT_RELATIVE 1   /* keyframe_relative_position                            */
#define MGB_CT_DIR2D 2      /* keyframe_2d_direction                                 */
#define MGB_CT_POSE 3       /* keyframe_pose: PoseConstraint (transition continuity) */
#define MGB_CT_TWO_HANDS 4  /* keyframe_two_hands                                    */
#define MGB_CT_FEET 5       /* keyframe_feet                                         */
#define MGB_CT_LOOK_AT 6    /* keyframe_look_at                                      */
#define MGB_CT_CA 7         /* ca_constraint: min over frames                        */
#define MGB_CT_JOINT_ROT 8  /* JointRotationConstraint (joint_rotation_constraint.py:62-75)            */
#define MGB_CT_QUAT_FRAME 9 /* keyframe_pose given as a quaternion frame: PoseConstraintQuatFrame      */
#define MGB_CT_LOCAL_TRAJ 10 /* local_trajectory: LocalTrajectoryConstraint, sum over all frames       */
#define MGB_CT_TRAJ 11       /* trajectory: TrajectoryConstraint, closest point per frame, mean over frames */
#define MGB_CT_TRAJ_SET 12   /* trajectory_set: TrajectorySetConstraint, arc-length walk of several joints  */

#define MGB_MAX_POSE_POINTS 32

/* One constraint of a MotionPrimitiveConstraints list, flattened.
 * Fields used per type:
 *  POSITION : node[0], time_slot, has_position, axis_mask (bit i set = axis i constrained),
 *             target, has_orientation, orient_q   (global_transform_constraint.py:71-143)
 *  RELATIVE : as POSITION (all axes) + offset[4]  (relative_transform_constraint.py:46-50)
 *  DIR2D    : time_slot, dir[2] = normalised (x,z) target (direction_2d_constraint.py:42-52)
 *  POSE     : time_slot (t), time_slot2 (t+1), points_offset/n_points into the pose point
 *             array, has_velocity, velocity[3]    (pose_constraint.py:48-67)
 *  TWO_HANDS: node[0]=left, node[1]=right, target, target2, center (two_hand_constraint.py:68-93)
 *  FEET     : node[0]=LeftFoot, node[1]=RightFoot, target=left, target2=right; the term is
 *             multiplied by `weight` once more (quirk Q6, feet_constraint.py:49-50)
 *  LOOK_AT  : node[0]=head joint, target          (look_at_constraint.py:58-68)
 *  CA       : node[0], time_slot = first of n_slots consecutive slots, axis_mask, target
 *             (global_transform_ca_constraint.py:34-40)
 *  JOINT_ROT: node[0] (an animated joint or the root), time_slot (frame_index), orient_q = the
 *             target rotation as a unit quaternion; error = |R_target - R(q/|q|)|_F of the joint's
 *             own frame channels, no forward kinematics (joint_rotation_constraint.py:62-75)
 *  QUAT_FRAME: time_slot (frame 0), points_offset/n_points: one pose point per compared joint with
 *             node, weight and target = (x, y, z) of the target UNIT quaternion taken with w >= 0;
 *             error = sum_i weight_i |R(target_i) - R(q_i/|q_i|)|_F
 *             (pose_constraint_quat_frame.py:47-52 through anim_utils' calculate_weighted_frame_distance_quat)
 *  LOCAL_TRAJ: node[0], time_slot = first of n_slots consecutive slots (the integer frames),
 *             offset[0] = start_t, offset[1] = arc-length step of the trajectory table,
 *             points_offset/n_points: the trajectory sampled at arc length i * step, i = 0 .. n_points - 2,
 *             (n_points - 2) * step = full arc length, last entry = the last control point
 *             (returned beyond the end); only target x and z are read
 *             (local_trajectory_constraint.py:58-83, parameterized_spline.py:131-148)
 *  TRAJ     : node[0], time_slot = first of n_slots consecutive slots = the frames of the motion vector
 *             (times linspace(0, F, F), motion_primitive.py:233), offset[0] = the parameter the search of the
 *             first frame starts from (min_arc_length / full_arc_length), points_offset/n_points: the spline
 *             at parameter i / (n_points - 1).  Per frame the closest point of the curve at a parameter >= the
 *             previous frame's (the reference minimises |S(u) - p| with L-BFGS-B from that bound,
 *             parameterized_spline.py:303-322; here: the first local minimum of the distance to the tabulated
 *             polyline walking up from the bound); error = mean over the frames of the distances
 *             (trajectory_constraint.py:78-116)
 *  TRAJ_SET : node[1] = number of joint trajectories J (<= 8), time_slot / n_slots as TRAJ; points: per
 *             trajectory a block of 2 header points + n + 2 table points, all with node = the joint:
 *             header 0 target = (arc-length step, initial arc length, n), header 1 target = (range_start,
 *             range_end, 1 if a range is set else 0); table = the trajectory at arc length i * step, i = 0..n,
 *             then the last control point.  Per frame |mean of all joint coordinates - mean of all target
 *             coordinates| if any trajectory is active at its walked arc length, the walk advancing by the
 *             joint's own path AFTER the frame is scored; error = mean over the frames
 *             (trajectory_set_constraint.py:83-105)
 */
typedef struct {
    int32_t type;
    int32_t node[2];
    int32_t time_slot;
    int32_t time_slot2;
    int32_t n_slots;
    int32_t points_offset;
    int32_t n_points;
    int32_t has_position;
    int32_t has_orientation;
    int32_t has_velocity;
    uint32_t axis_mask;
    double weight; /* weight_factor */
    double target[3];
    double target2[3];
    double center[3];
    double orient_q[4]; /* target orientation quaternion (w,x,y,z) */
    double offset[4];
    double dir[2];
    double velocity[3];
} mgb_constraint_desc;

typedef struct {
    int32_t node;
    double target[3];
    double weight;
} mgb_pose_point;

/* Optional frame alignment to the previous motion (non-local path,
 * reference motion_primitive_constraints.py:111-114 with prev_frames given):
 * heading (degrees, from +z towards +x) and root translation of the last
 * previous frame. mode 0 = none (is_local), 1 = align to prev frame. */
typedef struct {
    int32_t mode;
    double prev_heading_deg;
    double prev_root[3];
} mgb_align_desc;

/* ---- handles ---------------------------------------------------------- */

/* Copies bases, GMM (precision Cholesky computed here in fp64, as
 * _compute_precision_cholesky at motion_primitive.py:141-142) and skeleton to
 * `device`. */
int mgb_model_create(const mgb_model_desc* desc, int device, mgb_model** out);
void mgb_model_destroy(mgb_model* m);

/* Cubic B-spline basis at `t`, FITPACK splev semantics incl. extrapolation
 * (what motion_spline.py:89-92 evaluates per channel): value = sum_i w[i]*c[span+i]. */
int mgb_spline_basis(const double* knots, int32_t n_knots, double t, int32_t* span, double w[4]);

/* Compile a constraint list into a device program.
 * times[n_times]: the distinct evaluation times; if spans/weights are given
 * (4 weights per time, from the caller's own splev) they are used verbatim,
 * otherwise computed with mgb_spline_basis.
 * Replaces the per-candidate body of MotionPrimitiveConstraints.evaluate
 * (motion_primitive_constraints.py:100-122). */
int mgb_constraints_pack(const mgb_model* m, const mgb_constraint_desc* c, int32_t n,
                         const mgb_pose_point* points, int32_t n_points,
                         const double* times, int32_t n_times,
                         const int32_t* spans /*[n_times] or NULL*/,
                         const double* weights /*[n_times,4] or NULL*/,
                         const mgb_align_desc* align /* or NULL */,
                         mgb_constraints** out);
void mgb_constraints_destroy(mgb_constraints* c);
int32_t mgb_constraints_n_terms(const mgb_constraints* c);
/* size of the compiled program: contracted channel quads, interpreter ops, evaluation times
 * (used by bench.py to state the algorithmic flop of a launch) */
int mgb_constraints_info(const mgb_constraints* c, int32_t* n_quads, int32_t* n_ops, int32_t* n_times);

/* ---- kernels ---------------------------------------------------------- */

/* K1: GMM sampling, s = mu_g + L_g z, samples stacked in component order as
 * sklearn GaussianMixture.sample does (motion_primitive.py:182-189).
 * h_counts[G]: samples per component over the WHOLE job (host multinomial
 * draw); this call fills global indices [index_base, index_base + n).
 * Philox4x32-10 keyed by (seed, global index): invariant to sharding. */
int mgb_sample(const mgb_model* m, int64_t n, int64_t index_base, const int64_t* h_counts,
               uint64_t seed, float* d_s /*[n,D]*/, void* stream);

/* K6: fused score.  cost[i] = err_scale * sum_j w_j c_j(s_i)
 *        + quality_scale * (-log p(s_i)) + magnitude_scale * |s_i|
 * Replaces evaluate_samples_using_constraints' loop body
 * (motion_primitive_generator.py:251-257 -> objective_functions.py:142-160 ->
 * motion_primitive_constraints.py:100-122), the naturalness term
 * (objective_functions.py:256) and `+ 2|s|` (mg_state_planner.py:110-129).
 * s_is_f64 != 0: d_s holds float64 (what the reference's sampler returns).
 * d_cost [n] and d_terms [n_terms, n] (unweighted per-constraint errors) may be
 * NULL.  d_best_cost[k] / d_best_idx[k] (global index = index_base + i)
 * receive the k smallest costs, ties -> lowest index (first-minimum rule,
 * motion_primitive_generator.py:253-257); k == 0 skips selection.
 * k > 1 requires d_cost.  Workspace: mgb_score_workspace_bytes(n, k) bytes (it grows with k > 1: the
 * one-pass top-k keeps its per-CTA partials there; size it with the largest k you will pass), zero-filled ONCE
 * by the caller when it is allocated (it holds the cross-CTA completion counter, which the
 * kernel re-arms itself); one workspace must not be shared by concurrent streams. */
size_t mgb_score_workspace_bytes(int64_t n, int32_t k);
int mgb_score(const mgb_model* m, const mgb_constraints* c, const void* d_s, int32_t s_is_f64,
              int64_t n, int64_t index_base, float err_scale, float quality_scale,
              float magnitude_scale, float* d_cost, float* d_terms, int32_t k,
              float* d_best_cost, int64_t* d_best_idx, void* d_workspace,
              size_t workspace_bytes, void* stream);

/* Fused K1 + K6: candidates [index_base, index_base + n) of a job drawn in-kernel (exactly the
 * rows mgb_sample writes for the same h_counts / seed) and scored without ever being stored:
 * _get_best_fit_sample_using_gmm (motion_primitive_generator.py:192-206) in one launch.  The
 * caller regenerates the winning row with mgb_sample(n = 1, index_base = best_idx). */
int mgb_sample_score(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base,
                     const int64_t* h_counts, uint64_t seed, float err_scale, float quality_scale,
                     float magnitude_scale, float* d_cost /*[n] or NULL*/, int32_t k, float* d_best_cost,
                     int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream);

/* The planner-facing form of mgb_sample_score: ONE call per motion-primitive step that returns, in host
 * memory, the best cost, its global index and the winning latent vector itself (re-drawn on the device by the
 * CTA that folds the per-CTA minima - no second launch, one D2H copy, one synchronisation).  Replaces the
 * whole of _get_best_fit_sample_using_gmm (motion_primitive_generator.py:192-206).  h_best_row [D] or NULL.
 * Serialised per device (library-owned result buffers). */
int mgb_sample_score_best(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base,
                          const int64_t* h_counts, uint64_t seed, float err_scale, float quality_scale,
                          float magnitude_scale, float* h_best_cost, int64_t* h_best_idx,
                          float* h_best_row, void* stream);

/* The same with the candidates of the step SHARDED over the GPUs of one box (this rank: global indices
 * [index_base, index_base + n) of the job whose per-component counts are h_counts): the CTA that folds this
 * rank's minima also exchanges the 16-byte (cost, index) record with every peer through NVLink peer memory and
 * reduces the records arriving in its own slot array - the step stays ONE launch per rank, every rank returns the
 * job-wide winner and re-draws the same winning row.  Slot arrays: mgb_exchange_slot_bytes(world) bytes per rank,
 * zero-filled, peer-mapped (e.g. torch symmetric memory); seq: a step counter > 0 shared by the ranks. */
typedef struct {
    void* const* d_peer_slots; /* device array [world] of the ranks' slot arrays (this rank's own included) */
    int32_t rank, world;
    uint32_t seq;
    int64_t spin_limit;        /* polls before the error flag is raised; <= 0: default (~1 s)             */
    int32_t* d_err;            /* device int: 1 + peer that never arrived                                 */
} mgb_exchange_desc;
/* device-resident form (k = 1; results stay in d_best_cost / d_best_idx on every rank) */
int mgb_sample_score_sharded(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base,
                             const int64_t* h_counts, uint64_t seed, float err_scale, float quality_scale,
                             float magnitude_scale, const mgb_exchange_desc* x, float* d_best_cost,
                             int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream);
int mgb_sample_score_best_sharded(const mgb_model* m, const mgb_constraints* c, int64_t n, int64_t index_base,
                                  const int64_t* h_counts, uint64_t seed, float err_scale, float quality_scale,
                                  float magnitude_scale, const mgb_exchange_desc* x, float* h_best_cost,
                                  int64_t* h_best_idx, float* h_best_row, void* stream);

/* Cross-option batching: the outgoing options of a planner step - each its own (model, constraint
 * program, sample count, seed) - sampled, scored and reduced in ONE launch, one arg-min per option.
 * The reference scores them one after another (GraphWalkPlanner._evaluate_options,
 * graph_walk_planner.py:216-226; MGStatePlanner.select_best_option, mg_state_planner.py:946-968).
 * Results per option in host memory: h_best_cost[i], h_best_idx[i], h_best_rows[i * row_stride ...] (or NULL).
 * All models on one device and of one kernel class (latent dimension <= 16, or <= 32); otherwise
 * MGB_E_UNSUPPORTED and the caller issues one mgb_sample_score_best per option.  At most 64 options. */
typedef struct {
    const mgb_model* model;
    const mgb_constraints* constraints;
    int64_t n;            /* candidates of this option                                   */
    int64_t index_base;   /* first global index (0 unless the option itself is sharded)   */
    const int64_t* counts; /* [G] per-component counts of the option's whole job           */
    uint64_t seed;
    float err_scale, quality_scale, magnitude_scale;
} mgb_option;
int mgb_sample_score_multi(const mgb_option* options, int32_t n_options, float* h_best_cost,
                           int64_t* h_best_idx, float* h_best_rows, int32_t row_stride, void* stream);

/* K11: d_C[d, d] (float64, row-major) += (X - mean)^T (X - mean) for the UPPER triangle (i <= j) of the feature covariance,
 * X [n, d] float32 with row stride ld, mean [d] or NULL: the covariance of sklearn's PCA as the reference's
 * map_motions_to_euclidean_pca calls it (space_partitioning/features.py:125-130), as a 3xTF32 tcgen05 contraction
 * (2 n d^2 flop; 4.5 TFLOP at n = 1M, d = 1500).  The caller mirrors the triangle and divides by n - 1. */
int mgb_syrk_upper(const float* d_X, int64_t n, int32_t d, int32_t ld, const float* d_mean, double* d_C, void* stream);

/* ---- residual vectors and the local optimiser --------------------------------------
 * K9: the weighted residual vector of every candidate, d_out[n, R] with R = mgb_constraints_n_residuals(c):
 * MotionPrimitiveConstraints.get_residual_vector (motion_primitive_constraints.py:124-146) - per constraint
 * what its class returns from get_residual_vector_spline (1 value; two hands 3; pose one per point; local
 * trajectory one per frame), times weight_factor, in list order; no zero padding. */
int32_t mgb_constraints_n_residuals(const mgb_constraints* c);
int mgb_residuals(const mgb_model* m, const mgb_constraints* c, const float* d_s, int64_t n, float* d_out,
                  void* stream);
/* K10: Levenberg-Marquardt on the residual objective
 *   r_i(s) = (err_scale * residual_i(s) + quality_scale * (-log p(s))) / init_error_sum
 * (obj_spatial_error_residual_vector_and_naturalness, objective_functions.py:239-267), one CTA per seed,
 * the whole minimisation in ONE launch: replaces LeastSquares.run -> scipy leastsq (least_squares.py:35-64)
 * as called by _optimize_parameters_numerically (motion_primitive_generator.py:179-190), for n_seeds start
 * points at once (e.g. the top-k of mgb_sample_score).  init_error_sum <= 0: max(|sum r(s0)|, 1) per seed,
 * as the reference computes it.  max_iter = Jacobian evaluations.  d_out_s[n_seeds, D]; d_out_err[n_seeds, 4]
 * = sum of the weighted residuals at the result (the reference's min_error), 0.5|r|^2 (un-normalised) at the
 * seed and at the result, iterations used. */
int mgb_lm_refine(const mgb_model* m, const mgb_constraints* c, const float* d_seeds, int32_t n_seeds,
                  float err_scale, float quality_scale, float init_error_sum, int32_t max_iter,
                  float* d_out_s, float* d_out_err, void* stream);

/* Same through HOST buffers (pageable or pinned): chunked H2D on a copy stream
 * overlapped with scoring, D2H of the k results; synchronous.  This is the
 * call the reference-side shim makes with the numpy array it already holds. */
int mgb_score_host(const mgb_model* m, const mgb_constraints* c, const void* h_s, int32_t s_is_f64,
                   int64_t n, int64_t index_base, float err_scale, float quality_scale,
                   float magnitude_scale, float* h_cost /*[n] or NULL*/, int32_t k,
                   float* h_best_cost, int64_t* h_best_idx);

/* K2: back-projection  coeffs[n, Nb*P] = s[:, :K] . E' + m'  (translation
 * maxima folded in), motion_primitive.py:236-256.
 * impl: 0 = auto, 1 = fp32 SIMT, 2 = tcgen05 TF32x3 tensor-core path. */
int mgb_backproject(const mgb_model* m, const float* d_s, int64_t n, float* d_coeffs,
                    int32_t impl, void* stream);

/* K3: spline evaluation  frames[n, T, P] = B[T, Nb] . coeffs[n, Nb, P]
 * (motion_spline.py:71-92).  spans/weights as in mgb_constraints_pack; device
 * copies are made internally (T small). */
int mgb_eval_frames(const mgb_model* m, const float* d_coeffs, int64_t n, const double* times,
                    int32_t n_times, const int32_t* spans, const double* weights,
                    float* d_frames, void* stream);

/* K4: forward kinematics  pos[nT, nj, 3] of the listed nodes for frames[nT, P]
 * (anim_utils SkeletonNode.get_global_position; call sites
 * global_transform_constraint.py:103, pose_constraint.py:51, features.py:146). */
int mgb_fk(const mgb_model* m, const float* d_frames, int64_t n_frames, const int32_t* h_nodes,
           int32_t n_nodes, float* d_pos, void* stream);

/* K5a: scoring from materialised frames (staged pipeline): d_frames is
 * [n, n_times, P] at the constraint set's own times, in pack order. */
int mgb_score_frames(const mgb_model* m, const mgb_constraints* c, const float* d_frames,
                     const float* d_s /* [n,D] for logpdf / magnitude, may be NULL if unused */,
                     int64_t n, int64_t index_base, float err_scale, float quality_scale,
                     float magnitude_scale, float* d_cost, float* d_terms, int32_t k,
                     float* d_best_cost, int64_t* d_best_idx, void* d_workspace,
                     size_t workspace_bytes, void* stream);

/* K5b: GMM log-density  logpdf[i] = logsumexp_g(log w_g + log N(s_i; mu_g, Sigma_g))
 * (sklearn score_samples via precisions_cholesky_, objective_functions.py:110,198,256). */
int mgb_logpdf(const mgb_model* m, const float* d_s, int64_t n, float* d_logpdf, void* stream);

/* K5c: k smallest of a cost array, ties -> lowest index. */
int mgb_topk(const float* d_cost, int64_t n, int64_t index_base, int32_t k, float* d_best_cost,
             int64_t* d_best_idx, void* d_workspace, size_t workspace_bytes, void* stream);

/* K5c without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same result as mgb_topk (bit-identical costs and indices), but the
 * scratch (per-CTA partial selections) is a file-scope device array inside the library instead of a caller
 * workspace: MGB_TOPK_NOALLOC_SCRATCH_BYTES of device memory on every GPU that loads the library, held until
 * exit.  The call makes no cudaMalloc / cudaFree / allocating library call; inputs and outputs are the caller's.
 * Limits: n <= MGB_TOPK_NOALLOC_MAX_N and 1 <= k <= 64, beyond them MGB_E_* with a message (use mgb_topk).
 * The array is shared by every call on a device and keeps the last call's contents (nothing in it has
 * to start at zero, so nothing is cleared): calls that overlap on the GPU - two streams, two host
 * threads - race; serialise them on one stream or with events. */
#define MGB_TOPK_NOALLOC_MAX_N 16777216            /* costs per call */
#define MGB_TOPK_NOALLOC_SCRATCH_BYTES 454656      /* 592 CTAs x 64 picks x (int64 index + float cost) */
int mgb_topk_noalloc(const float* d_cost, int64_t n, int64_t index_base, int32_t k, float* d_best_cost,
                     int64_t* d_best_idx, void* stream);

/* k-means++ seeding step for every node of a level at once: d2[i] = min(d2[i], |x[i, :dim] - centers[node[i], :dim]|^2), or
 * the squared distance itself when first != 0; d2 is float64.  centers[n_nodes, dim] = the centre chosen last in every node.
 * Replaces the closest_dist_sq update inside the k-means++ initialisation of sklearn KMeans as the reference calls it
 * (space_partitioning/clustering.py:74-80); the D^2 sampling on top of it (prefix sums, picks) stays index bookkeeping. */
int mgb_min_sqdist(const float* d_x, int64_t n, int32_t dim, int32_t ld, const int32_t* d_node, const float* d_centers,
                   double* d_d2, int32_t first, void* stream);

/* k-means++ seeding, the pick, for every node of a level at once: node i owns rows [starts[i], starts[i] + counts[i]) of x;
 * with prefix == NULL the row start + min(floor(u counts), counts - 1) is chosen (first centre), else the first row of the node
 * whose GLOBAL inclusive prefix sum of the closest squared distances (prefix[n], float64) exceeds
 * prefix[start - 1] + u * (mass of the node) (D^2 sampling, one trial per centre); u[i * u_stride] in [0, 1).  The chosen row's
 * first dim columns go to centers[i * center_stride ...].  Same reference lines as mgb_min_sqdist. */
int mgb_kmeanspp_pick(const float* d_x, int32_t dim, int32_t ld, const int64_t* d_starts, const int64_t* d_counts, int32_t n_nodes,
                      const double* d_u, int32_t u_stride, const double* d_prefix, float* d_centers, int32_t center_stride,
                      void* stream);

/* K7: one Lloyd iteration (E-step + M-step sums) for every node of a cluster-tree level at once.
 * x[n, dim] (row stride ld) with the samples of one node stored contiguously (any order is
 * accepted; runs of equal node ids are what makes the accumulation cheap), node[n] = tree node of
 * each sample in this level (NULL = one node), centroids[n_nodes, n_sub, dim].  Writes label[n];
 * accumulates sums[n_nodes, n_sub, dim] / counts[n_nodes, n_sub] / inertia[1] (zeroed by the
 * caller).  n_sub <= 8, dim <= 126.
 * Replaces the E/M step inside sklearn KMeans(n_clusters=4).fit_predict per node
 * (space_partitioning/cluster_tree_node_builder.py:83-103, clustering.py:74-80). */
int mgb_kmeans_assign_update(const float* d_x, int64_t n, int32_t dim, int32_t ld,
                             const int32_t* d_node, const float* d_centroids, int32_t n_sub,
                             int32_t* d_label, float* d_sums, int32_t* d_counts,
                             float* d_inertia /* [1], accumulated */, void* stream);

/* K7 driver: Lloyd iterations for a whole tree level until every node has converged
 * (squared centre shift <= d_tol[node], sklearn's rule with d_tol = 1e-4 * mean variance of the
 * node's samples) or max_iter is reached; the convergence count is read back every check_every
 * iterations.  d_centroids[n_nodes, n_sub, dim]: seeds in, final centroids out (empty clusters
 * keep their seed).  Distances use the first `dim` columns of a row; the sums cover the first
 * dim_acc >= dim columns, so the caller gets the children's means over the full latent vector
 * (cluster_tree_node_builder.py:62-69 takes np.mean over all columns) from the same pass:
 * on return label / sums[n_nodes, n_sub, dim_acc] / counts[n_nodes, n_sub] /
 * node_inertia[n_nodes] (may be NULL) belong to the FINAL centroids.  d_converged[n_nodes]: in =
 * nodes to leave untouched (1), out = converged flags.  max_iter = 0 runs the E-step only (used
 * with n_sub = 1 for the per-node mean and variance).  iters_done: host int, may be NULL. */
int mgb_kmeans_level(const float* d_x, int64_t n, int32_t dim, int32_t dim_acc, int32_t ld,
                     const int32_t* d_node, int32_t n_nodes, int32_t n_sub, float* d_centroids,
                     const float* d_tol, int32_t max_iter, int32_t check_every, int32_t* d_label,
                     float* d_sums, int32_t* d_counts, float* d_node_inertia,
                     int32_t* d_converged, int32_t* iters_done, void* stream);

/* K8: cost of every point of a cluster tree (node means and kd-tree points, one [n, D] table) under
 * the step's constraints in one launch; the host walk of the tree then only looks costs up.
 * Same arithmetic as mgb_score with d_cost set and no arg-min.  Replaces the obj(child.mean, data)
 * / obj(node.point, data) calls of cluster_tree_node.py:113-138 and kdtree.py:138-161,233-250. */
int mgb_node_costs(const mgb_model* m, const mgb_constraints* c, const void* d_points, int32_t is_f64,
                   int64_t n, float error_scale, float quality_scale, float magnitude_scale,
                   float* d_cost, void* stream);

/* Multi-GPU (one box): exchange of every rank's best (cost, global index) through NVLink peer memory and
 * reduction with the first-minimum rule (lowest cost, lowest index on ties, NaN never wins;
 * motion_primitive_generator.py:251-257 applied across ranks).  d_peer_slots: DEVICE array of `world`
 * pointers, entry r = base of rank r's slot array (mgb_exchange_slot_bytes(world) bytes, zero-filled once,
 * mapped into this process - e.g. torch symmetric memory or cudaIpc); the call stores this rank's record in
 * every peer's array, waits for all peers' records of step `seq` in its own array and writes the winner to
 * d_out_cost[1] / d_out_idx[1].  seq: 1, 2, 3, ... identical on all ranks, incremented per call.
 * spin_limit <= 0: default (~1 s); on expiry d_err[0] = 1 + missing rank and the kernel returns.
 * Asynchronous on `stream`; replaces a 16-byte ncclAllGather + host-side selection. */
size_t mgb_exchange_slot_bytes(int32_t world);
int mgb_exchange_best(const float* d_my_cost, const int64_t* d_my_idx, void* const* d_peer_slots,
                      int32_t rank, int32_t world, uint32_t seq, int64_t spin_limit, float* d_out_cost,
                      int64_t* d_out_idx, int32_t* d_err, void* stream);

/* Diagnostic: measured fp32 FMA throughput (TFLOP/s) of `device`; the denominator of the
 * compute-bound fused score kernel's roofline in bench.py (no reference counterpart). */
int mgb_fp32_fma_peak(int32_t device, double* tflops);

/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
int64_t mgb_launch_count(void);
const char* mgb_last_error(void);
const char* mgb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MGB200_H */
