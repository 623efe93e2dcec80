"""Generate tests/golden/*.npz by running the REFERENCE's own modules.

TEST INFRASTRUCTURE ONLY; runs only in the build container (needs
/root/reference).  Usage:  python -m oracle.make_golden

What runs unmodified from /root/reference (through oracle/refload.py):
  MotionStateGraphNode/MotionPrimitiveModelWrapper -> MotionPrimitive.back_project,
  MotionSpline.evaluate / get_motion_vector, sklearn GMM sample / score_samples,
  MotionPrimitiveConstraints.evaluate + every spatial constraint class,
  obj_spatial_error_sum, MotionPrimitiveGenerator.evaluate_samples_using_constraints,
  ClusterTree / FeatureClusterTree search.
What they call into and is NOT the reference (oracle/standin): anim_utils FK,
alignment, point clouds; transformations.  PARITY UNPINNED for those.

Each fixture stores the inputs that cannot be regenerated from seeds (the
sklearn-drawn samples, the constraint specs as JSON) and the reference outputs.
"""
import io
import json
import os
import sys
import contextlib
import warnings

import numpy as np

from oracle import refload
from morphablegraphs_b200 import synthetic as syn

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _quiet():
    return contextlib.redirect_stdout(io.StringIO())


def reference_factory(ns):
    sc = ns.spatial_constraints
    return {
        "GlobalTransformConstraint": sc.GlobalTransformConstraint,
        "RelativeTransformConstraint": sc.RelativeTransformConstraint,
        "Direction2DConstraint": sc.Direction2DConstraint,
        "PoseConstraint": sc.PoseConstraint,
        "TwoHandConstraintSet": sc.TwoHandConstraintSet,
        "FeetConstraint": sc.FeetConstraint,
        "LookAtConstraint": sc.LookAtConstraint,
        "GlobalTransformCAConstraint": ns.ca_constraint.GlobalTransformCAConstraint,
        "JointRotationConstraint": ns.joint_rotation_constraint.JointRotationConstraint,
        "PoseConstraintQuatFrame": ns.pose_constraint_quat_frame.PoseConstraintQuatFrame,
        "LocalTrajectoryConstraint": ns.local_trajectory_constraint.LocalTrajectoryConstraint,
        "TrajectoryConstraint": ns.trajectory_constraint.TrajectoryConstraint,
        "TrajectorySetConstraint": ns.trajectory_set_constraint.TrajectorySetConstraint,
    }


def reference_trajectory(ns):
    """desc["trajectory"] dict -> the reference's own arc-length parameterised Catmull-Rom spline
    (constraints/spatial_constraints/splines/parameterized_spline.py)."""
    def make(t):
        with _quiet():
            return ns.parameterized_spline.ParameterizedSpline([np.array(p, dtype=np.float64) for p in t["control_points"]])
    return make


def build_mp_constraints(ns, skeleton, specs, is_local=True):
    mpc = ns.motion_primitive_constraints.MotionPrimitiveConstraints()
    mpc.skeleton = skeleton
    mpc.is_local = is_local
    mpc.constraints = syn.build_constraints(reference_factory(ns), skeleton, specs, reference_trajectory(ns))
    return mpc


def make_specs(ns, node, skeleton, F, rng, which, prev_frames=None):
    """Constraint specs with targets taken from held-out draws of the same model
    (so errors are in a realistic range), SURVEY.md section 8(d).  prev_frames: the held-out motions are aligned to
    them first (targets in the global frame of a non-local step)."""
    held = node.sample_low_dimensional_vectors(3)
    spl = [node.back_project(h, use_time_parameters=False) for h in held]
    if prev_frames is not None:
        from anim_utils.animation_data.motion_concatenation import align_quaternion_frames
        for sp_ in spl:
            sp_.coeffs = align_quaternion_frames(skeleton, skeleton.aligning_root_node, sp_.coeffs, prev_frames, None)

    def pos(k, joint, t):
        return skeleton.nodes[joint].get_global_position(spl[k].evaluate(t)).tolist()

    specs = []
    if which == "config1":
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("RightFoot", pos(0, "RightFoot", F - 1), F - 1, F), 1.0, 1.0))
        return specs
    if which == "rot":
        # round 2: the constraint types of SURVEY.md rows a12 / a13 that read joint rotations or walk all frames
        from morphablegraphs_b200.constraints import tabulate_trajectory
        joints = list(skeleton.node_name_frame_map.keys())
        jn = joints[5]
        ji = joints.index(jn)
        q = np.array(spl[0].evaluate(F // 2)[3 + 4 * ji:7 + 4 * ji], dtype=np.float64)
        specs.append(syn.spec("JointRotationConstraint",
                              {"joint_name": jn, "rotation_type": "quaternion", "rotation_constraint": (q * 1.7).tolist(),
                               "frame_index": F // 2, "canonical_keyframe": F // 2, "semanticAnnotation": syn.sem("mid")}, 1.0, 1.3))
        specs.append(syn.spec("JointRotationConstraint",
                              {"joint_name": joints[0], "rotation_type": "euler", "rotation_constraint": [10.0, 35.0, -5.0],
                               "frame_index": F - 1, "canonical_keyframe": F - 1, "semanticAnnotation": syn.sem()}, 1.0, 0.6))
        specs.append(syn.spec("PoseConstraintQuatFrame",
                              {"canonical_keyframe": 0, "frame_constraint": np.asarray(spl[1].evaluate(0)).tolist(),
                               "semanticAnnotation": syn.sem("start")}, 1.0, 2.0))
        cps = [[0.0, 0.0, 0.0], [5.0, 0.0, 15.0], [12.0, 0.0, 30.0], [10.0, 0.0, 45.0], [0.0, 0.0, 60.0], [-8.0, 0.0, 80.0]]
        traj = reference_trajectory(ns)({"control_points": cps})
        tab = tabulate_trajectory(traj)
        specs.append(syn.spec("LocalTrajectoryConstraint",
                              {"canonical_keyframe": None, "joint_name": "Hips", "start_t": 5.0, "n_canonical_frames": F,
                               "semanticAnnotation": syn.sem("path"),
                               "trajectory": {"control_points": cps, "points": tab.points.tolist(), "step": tab.step,
                                              "last_control_point": tab.last_control_point.tolist()}}, 1.0, 0.01))
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("RightFoot", pos(0, "RightFoot", F - 1), F - 1, F), 1.0, 1.0))
        return specs
    if which == "traj":
        # f2: the constraints that read the whole motion vector - TrajectoryConstraint (scipy closest point per frame) on the
        # root, TrajectorySetConstraint on both hands (one trajectory with an active range, one without), plus one keyframe
        def path(k, joint, n):
            return [pos(k, joint, t) for t in np.linspace(0, F - 1, n)]
        specs.append(syn.spec("TrajectoryConstraint",
                              {"joint_name": "Hips", "control_points": path(0, "Hips", 6), "min_arc_length": 3.0,
                               "n_canonical_frames": F}, 1.0, 0.7))
        specs.append(syn.spec("TrajectorySetConstraint",
                              {"joint_names": ["LeftHand", "RightHand"], "joint_arc_lengths": [0.0, 4.0], "n_canonical_frames": F,
                               "trajectories": [{"control_points": path(1, "LeftHand", 7), "range": [5.0, 40.0]},
                                                {"control_points": path(1, "RightHand", 7), "range": None}]}, 1.0, 1.3))
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("RightFoot", pos(0, "RightFoot", F - 1), F - 1, F), 1.0, 1.0))
        return specs
    if which == "cfg4":
        # right-hand position at F-1, root position + 2-D direction at F-1, transition
        # PoseConstraint (5 joints) at keyframes 0/1 (pose at 0, root velocity from 1); naturalness is part of the objective
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("RightHand", pos(0, "RightHand", F - 1), F - 1, F), 1.0, 1.0))
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("Hips", pos(1, "Hips", F - 1), F - 1, F), 1.0, 1.0))
        specs.append(syn.spec("Direction2DConstraint",
                              {"canonical_keyframe": F - 1, "dir_vector": [0.3, 0.0, 1.0],
                               "semanticAnnotation": syn.sem()}, 1.0, 0.05))
        names = ["Hips", "LeftHand", "RightHand", "LeftFoot", "RightFoot"]
        f0 = spl[2].evaluate(0)
        f1 = spl[2].evaluate(1)
        cloud = [skeleton.nodes[n].get_global_position(f0).tolist() for n in names]
        vel = (skeleton.nodes["Hips"].get_global_position(f1) - np.array(cloud[0])).tolist()
        specs.append(syn.spec("PoseConstraint",
                              {"canonical_keyframe": 0, "frame_constraint": cloud, "velocity_constraint": vel,
                               "semanticAnnotation": syn.sem("start"), "node_names": names,
                               "weights": [skeleton.joint_weight_map[n] for n in names]}, 1.0, 1.0))
        return specs
    # mixed set: every keyframe constraint type of SURVEY.md Appendix C
    p = pos(0, "RightHand", F - 1)
    q = skeleton.nodes["RightHand"].get_global_orientation_quaternion(spl[1].evaluate(F - 1)).tolist()
    specs.append(syn.spec("GlobalTransformConstraint",
                          syn.position_constraint_desc("RightHand", p, F - 1, F, orientation_q=q), 1.0, 1.0))
    p = pos(1, "Hips", (F - 1) / 2.0)
    p[1] = None   # height unconstrained; fractional keyframe (graph_walk_planner.py:203)
    specs.append(syn.spec("GlobalTransformConstraint",
                          syn.position_constraint_desc("Hips", p, (F - 1) / 2.0, F), 1.0, 0.5))
    d = syn.position_constraint_desc("LeftHand", pos(2, "Bip01_L_Finger2", F - 1), F - 1, F)
    d["offset"] = [3.0, -1.0, 2.0, 1.0]
    specs.append(syn.spec("RelativeTransformConstraint", d, 1.0, 0.8))
    specs.append(syn.spec("Direction2DConstraint",
                          {"canonical_keyframe": F - 1, "dir_vector": [0.3, 0.0, 1.0],
                           "semanticAnnotation": syn.sem()}, 1.0, 0.02))
    names = ["Hips", "LeftHand", "RightHand", "LeftFoot", "RightFoot"]
    f0 = spl[2].evaluate(0)
    f1 = spl[2].evaluate(1)
    cloud = [skeleton.nodes[n].get_global_position(f0).tolist() for n in names]
    vel = (skeleton.nodes["Hips"].get_global_position(f1) - np.array(cloud[0])).tolist()
    specs.append(syn.spec("PoseConstraint",
                          {"canonical_keyframe": 0, "frame_constraint": cloud, "velocity_constraint": vel,
                           "semanticAnnotation": syn.sem("start"), "node_names": names,
                           "weights": [skeleton.joint_weight_map[n] for n in names]}, 1.0, 1.5))
    specs.append(syn.spec("TwoHandConstraintSet",
                          {"joint": ["LeftHand", "RightHand"],
                           "positions": [pos(0, "LeftHand", F // 2), pos(0, "RightHand", F // 2)],
                           "orientations": [None, None], "n_canonical_frames": F,
                           "canonical_keyframe": F // 2, "semanticAnnotation": syn.sem("mid")}, 1.0, 1.0))
    specs.append(syn.spec("FeetConstraint",
                          {"canonical_keyframe": F - 1, "left": pos(1, "LeftFoot", F - 1),
                           "right": pos(1, "RightFoot", F - 1), "semanticAnnotation": syn.sem()}, 1.0, 0.7))
    specs.append(syn.spec("LookAtConstraint",
                          {"canonical_keyframe": F // 3, "position": [40.0, 150.0, 200.0],
                           "semanticAnnotation": syn.sem("look")}, 1.0, 2.0))
    d = syn.position_constraint_desc("LeftFoot", pos(2, "LeftFoot", F // 2), None, F)
    d["ca_constraint"] = True
    specs.append(syn.spec("GlobalTransformCAConstraint", d, 1.0, 1.0))
    return specs


def run_case(ns, shape, which, n_samples, seed, translation_maxima=(1.0, 1.0, 1.0), non_local=False, n_joints=None, n_full=None):
    """n_full: number of leading candidates for which the per-stage outputs (coefficients, key frames, FK positions,
    motion vectors) are stored; the costs / terms / arg-min always cover all n_samples."""
    cfg = syn.SHAPES[shape]
    F = cfg["n_frames"]
    mj = syn.make_model_json(shape, translation_maxima=translation_maxima)
    skel_desc = syn.make_skeleton_desc(cfg["n_joints"])
    skeleton = refload.make_skeleton(skel_desc)
    node = refload.make_graph_node(ns, mj)
    np.random.seed(seed)          # sklearn's GaussianMixture.sample uses the global state
    rng = np.random.RandomState(syn.SEED_CONSTRAINTS)
    specs = make_specs(ns, node, skeleton, F, rng, which)
    S = node.sample_low_dimensional_vectors(n_samples)
    mpc = build_mp_constraints(ns, skeleton, specs, is_local=not non_local)
    prev_frames = None
    if non_local:
        prev = node.back_project(node.sample_low_dimensional_vectors(1)[0], use_time_parameters=False)
        with _quiet():
            pf = prev.get_motion_vector()
        # turn and shift the previous motion so that alignment has something to do
        from anim_utils.animation_data.motion_concatenation import transform_quaternion_frames
        from transformations import quaternion_matrix, quaternion_about_axis
        m = quaternion_matrix(quaternion_about_axis(0.7, [0, 1, 0]))
        m[:3, 3] = [120.0, 0.0, -40.0]
        prev_frames = transform_quaternion_frames(pf, m)[-3:]
        if which == "traj":
            # the paths of a non-local step lie where the aligned motion goes
            specs = make_specs(ns, node, skeleton, F, rng, which, prev_frames=prev_frames)
            mpc = build_mp_constraints(ns, skeleton, specs, is_local=False)
    out = {}
    out["samples"] = S
    out["specs_json"] = np.array(json.dumps(specs))
    out["shape"] = np.array(shape)
    out["translation_maxima"] = np.array(translation_maxima, dtype=np.float64)
    out["is_local"] = np.array(not non_local)
    if prev_frames is not None:
        out["prev_frames"] = np.asarray(prev_frames)
    # --- reference outputs -------------------------------------------------
    n_full = n_samples if n_full is None else min(n_full, n_samples)
    out["n_full"] = np.array(n_full)
    splines = [node.back_project(s, use_time_parameters=False) for s in S[:n_full]]
    out["coeffs"] = np.array([sp.coeffs.copy() for sp in splines])
    key_times = sorted({float(sp_["desc"]["canonical_keyframe"]) for sp_ in specs
                        if sp_["desc"].get("canonical_keyframe") is not None})
    out["key_times"] = np.array(key_times)
    out["key_frames"] = np.array([[sp.evaluate(t) for t in key_times] for sp in splines])
    n_mv = min(4, n_full)
    with _quiet():
        out["motion_vectors"] = np.array([splines[i].get_motion_vector() for i in range(n_mv)])
    fk_names = list(skel_desc["nodes"].keys())
    out["fk_names_json"] = np.array(json.dumps(fk_names))
    out["fk_positions"] = np.array([[skeleton.nodes[n].get_global_position(out["key_frames"][i, -1])
                                     for n in fk_names] for i in range(n_full)])
    # per-term errors and total cost through the reference's own loop
    terms = np.zeros((len(mpc.constraints), n_samples))
    costs = np.zeros(n_samples)
    obj = ns.objective_functions.obj_spatial_error_sum
    with _quiet(), refload.scalar_spline_parameter(ns), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, s in enumerate(S):
            sp = node.back_project(s, use_time_parameters=False)
            if non_local:
                from anim_utils.animation_data.motion_concatenation import align_quaternion_frames
                sp.coeffs = align_quaternion_frames(skeleton, skeleton.aligning_root_node, sp.coeffs,
                                                    prev_frames, None)
            for j, c in enumerate(mpc.constraints):
                terms[j, i] = c.evaluate_motion_spline(sp)
            costs[i] = obj(s, (node, mpc, prev_frames))
    out["terms"] = terms
    out["costs"] = costs
    if which == "traj":
        # the residual-vector form of the same constraints (motion_primitive_constraints.py:124-146): one value per frame
        with _quiet(), refload.scalar_spline_parameter(ns), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out["residual_vectors"] = np.array([mpc.get_residual_vector(node, s, prev_frames) for s in S[:16]])
    out["evaluations"] = np.array(mpc.evaluations)
    # the reference's own arg-min loop
    gen = ns.motion_primitive_generator.MotionPrimitiveGenerator.__new__(
        ns.motion_primitive_generator.MotionPrimitiveGenerator)
    gen.objective = obj
    with _quiet(), refload.scalar_spline_parameter(ns), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        best, min_error = gen.evaluate_samples_using_constraints(S, node, mpc, prev_frames)
    out["best_idx"] = np.array(int(np.where((S == best).all(axis=1))[0][0]))
    out["min_error"] = np.array(min_error)
    gmm = node.get_gaussian_mixture_model()
    out["logpdf"] = gmm.score_samples(S)
    out["naturalness_cost"] = costs * 1.0 + 0.1 * (-out["logpdf"])   # error_scale 1, quality_scale 0.1
    out["magnitude_cost"] = costs + 2.0 * np.linalg.norm(S, axis=1)   # mg_state_planner.py:129
    return out, (node, skeleton, mpc, S, prev_frames)


def run_local_case(ns, shape="S", n_samples=64, seed=2031):
    """a8: a GLOBAL constraint set (every constraint type, non-trivial weights) with a non-identity
    `aligning_transform`, brought to the primitive's local frame by the reference's own
    `transform_constraints_to_local_cos` (motion_primitive_constraints.py:268-378, incl. quirk Q5: position /
    relative / CA constraints come back with weight 1.0 and without their orientation), then scored by the
    reference's loop - what `generate_constrained_sample` does on the default `use_local_coordinates=True` path."""
    from tests.common import describe_constraint
    from transformations import quaternion_matrix, quaternion_about_axis
    cfg = syn.SHAPES[shape]
    F = cfg["n_frames"]
    mj = syn.make_model_json(shape)
    skeleton = refload.make_skeleton(syn.make_skeleton_desc(cfg["n_joints"]))
    node = refload.make_graph_node(ns, mj)
    np.random.seed(seed)
    local_specs = make_specs(ns, node, skeleton, F, None, "mixed")
    for i, sp_ in enumerate(local_specs):          # weights that Q5 must reset (or keep) visibly
        sp_["weight"] = [2.5, 0.5, 1.7, 0.02, 1.5, 1.3, 0.7, 2.0, 3.0][i]
    S = node.sample_low_dimensional_vectors(n_samples)
    T = quaternion_matrix(quaternion_about_axis(0.9, [0, 1, 0]))
    T[:3, 3] = [250.0, 0.0, -130.0]
    # express the targets in the GLOBAL frame: x_global = T x_local (directions without the translation)
    def gp(p, w=1.0):
        if p is None:
            return None
        idx = [i for i in range(3) if p[i] is None]  # masked axis (the rotation is about y, so a masked height stays masked)
        q = np.dot(T, [0.0 if x is None else x for x in p] + [w])[:3].tolist()
        for i in idx:
            q[i] = None
        return q
    specs = json.loads(json.dumps(local_specs))
    for sp_ in specs:
        d = sp_["desc"]
        if sp_["class"] in ("GlobalTransformConstraint", "RelativeTransformConstraint", "GlobalTransformCAConstraint", "LookAtConstraint"):
            d["position"] = gp(d["position"])
        elif sp_["class"] == "Direction2DConstraint":
            d["dir_vector"] = gp(d["dir_vector"], 0.0)
        elif sp_["class"] == "PoseConstraint":
            d["frame_constraint"] = [gp(p) for p in d["frame_constraint"]]
        elif sp_["class"] == "TwoHandConstraintSet":
            d["positions"] = [gp(p) for p in d["positions"]]
        elif sp_["class"] == "FeetConstraint":
            d["left"], d["right"] = gp(d["left"]), gp(d["right"])
    mpc = build_mp_constraints(ns, skeleton, specs, is_local=False)
    mpc.aligning_transform = T
    mpc.use_local_optimization = True
    with _quiet():
        local = mpc.transform_constraints_to_local_cos()
    assert local is not mpc and local.is_local
    out = {"samples": S, "specs_json": np.array(json.dumps(specs)), "shape": np.array(shape),
           "translation_maxima": np.array([1.0, 1.0, 1.0]), "is_local": np.array(False), "aligning_transform": T,
           "local_desc_json": np.array(json.dumps([describe_constraint(c) for c in local.constraints])),
           "local_use_local_optimization": np.array(bool(local.use_local_optimization))}
    obj = ns.objective_functions.obj_spatial_error_sum
    terms = np.zeros((len(local.constraints), n_samples))
    costs = np.zeros(n_samples)
    with _quiet():
        for i, s_ in enumerate(S):
            sp = node.back_project(s_, use_time_parameters=False)
            for j, c in enumerate(local.constraints):
                terms[j, i] = c.evaluate_motion_spline(sp)
            costs[i] = obj(s_, (node, local, None))
        gen = ns.motion_primitive_generator.MotionPrimitiveGenerator.__new__(ns.motion_primitive_generator.MotionPrimitiveGenerator)
        gen.objective = obj
        best, min_error = gen.evaluate_samples_using_constraints(S, node, local, None)
    out["terms"] = terms
    out["costs"] = costs
    if which == "traj":
        # the residual-vector form of the same constraints (motion_primitive_constraints.py:124-146): one value per frame
        with _quiet(), refload.scalar_spline_parameter(ns), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out["residual_vectors"] = np.array([mpc.get_residual_vector(node, s, prev_frames) for s in S[:16]])
    out["best_idx"] = np.array(int(np.where((S == best).all(axis=1))[0][0]))
    out["min_error"] = np.array(min_error)
    return out


@contextlib.contextmanager
def _linspace_int_num():
    """numpy >= 1.18 refuses the FLOAT sample count the reference hands to np.linspace in
    _invert_canonical_to_sample_time_function (motion_primitive.py:314, quirk Q3); older numpy truncated it with int().
    This restores exactly that truncation for the duration of the call - the reference module itself is not touched."""
    orig = np.linspace

    def linspace(start, stop, num=50, *a, **k):
        return orig(start, stop, int(num), *a, **k)
    np.linspace = linspace
    try:
        yield
    finally:
        np.linspace = orig


def run_timewarp_case(ns, shape="S", n_samples=6, seed=2051):
    """f2: the reference's own time-warp back-projection (motion_primitive.py:268-318) and the motion it yields
    (MotionSpline.get_motion_vector at the warped times, motion_spline.py:71-92), for sklearn-drawn samples."""
    node = refload.make_graph_node(ns, syn.make_model_json(shape))
    np.random.seed(seed)
    S = node.sample_low_dimensional_vectors(n_samples)
    S[:, -3:] *= 8.0        # time parameters large enough that the warp is visible
    out = {"samples": S, "shape": np.array(shape)}
    mp = node.motion_primitive
    for i, s_ in enumerate(S):
        with _linspace_int_num(), _quiet():
            spline = node.back_project(s_, use_time_parameters=True)
            frames = spline.get_motion_vector()
            canonical = mp._back_transform_gamma_to_canonical_time_function(s_[mp.s_pca["n_components"]:])
        out["tw%d_canonical" % i] = np.asarray(canonical)
        out["tw%d_time" % i] = np.asarray(spline.time_function)
        out["tw%d_frames" % i] = np.asarray(frames)
    return out


def run_config3_case(ns, n_samples=10000, seed=4100, max_steps=None):
    """BASELINE config 3 at the parity size SURVEY.md 8(d) row 3 states: the action list of examples/example_input.json
    (walk / pickRight / carryRight / placeRight) over the synthetic graph, every motion-primitive step decided by the
    reference's OWN generator loop (motion_primitive_generator.py:230-261) on 10 000 candidates drawn by the numpy
    restatement of sklearn's sampler from a per-step seed (oracle.np_oracle.gmm_sample: the test redraws them), candidates
    aligned to the walk so far (non-local path).  The step list and the constraint specs come from the product's driver
    (morphablegraphs_b200/graph_synthesis.py) - they are INPUTS; decisions, costs and the frames carried to the next
    step are the reference's."""
    from anim_utils.animation_data.motion_concatenation import align_quaternion_frames
    from morphablegraphs_b200 import graph_synthesis as GS
    from morphablegraphs_b200.motion_state_graph import MotionStateGraphLoader
    from oracle import np_oracle as O
    gd = syn.make_graph_data()
    ld = MotionStateGraphLoader()
    ld.set_graph_data(gd)
    graph = ld.build()                                            # product graph: planning only
    skeleton = refload.make_skeleton(gd["skeleton"])              # reference-side skeleton (stand-in for anim_utils)
    ref_nodes = {}
    for action, entry in gd["subgraphs"].items():
        for mp, nd in entry["nodes"].items():
            ref_nodes[(action, mp)] = refload.make_graph_node(ns, nd["mm"], name=mp, action=action)
    plan_input = GS.read_input(os.path.join(refload.SOURCE_ROOT, "examples", "example_input.json"))
    steps = GS.plan_steps(graph, plan_input)
    if max_steps:
        steps = steps[:max_steps]
    prev = GS.initial_frames(graph, plan_input, steps[0]["node_key"])
    gen = ns.motion_primitive_generator.MotionPrimitiveGenerator.__new__(ns.motion_primitive_generator.MotionPrimitiveGenerator)
    gen.objective = ns.objective_functions.obj_spatial_error_sum
    out = {"n_steps": np.array(len(steps)), "n_samples": np.array(n_samples), "node_keys_json": np.array(json.dumps([list(s_["node_key"]) for s_ in steps]))}
    for i, step in enumerate(steps):
        node = ref_nodes[step["node_key"]]
        specs = GS.step_specs(graph, step, prev)
        mpc = build_mp_constraints(ns, skeleton, specs, is_local=False)
        S, _ = O.gmm_sample(O.ModelArrays(node), n_samples, np.random.RandomState(seed + i))
        costs = np.zeros(n_samples)
        with _quiet():
            for k, s_ in enumerate(S):
                costs[k] = gen.objective(s_, (node, mpc, prev))
            best, min_error = gen.evaluate_samples_using_constraints(S, node, mpc, prev)
            frames = node.back_project(best, use_time_parameters=False).get_motion_vector()
        frames = np.asarray(align_quaternion_frames(skeleton, skeleton.aligning_root_node, frames, prev, None))
        bi = int(np.where((S == best).all(axis=1))[0][0])
        out["step%d_seed" % i] = np.array(seed + i)
        out["step%d_specs_json" % i] = np.array(json.dumps(specs))
        out["step%d_prev_frames" % i] = np.asarray(prev)
        out["step%d_costs" % i] = costs.astype(np.float32)
        out["step%d_best_idx" % i] = np.array(bi)
        out["step%d_best_sample" % i] = S[bi]
        out["step%d_min_error" % i] = np.array(min_error)
        out["step%d_last_frames" % i] = frames[-8:]
        print("config3 step", i, step["node_key"], "best", bi, "min_error %.4f" % min_error, flush=True)
        prev = frames[-8:]
    return out


def run_resid_case(ns, name_of_set="mixed", shape="S", n_samples=24, seed=2041, n_opt=3):
    """f1: the reference's own residual-vector objective (motion_primitive_constraints.py:124-146,
    objective_functions.py:239-267) on reference-drawn samples, and the reference's own LeastSquares
    (least_squares.py:35-64 -> scipy leastsq) from the best of them, as _optimize_parameters_numerically runs it
    (motion_primitive_generator.py:179-190: pose constraints dropped, residuals normalised by the start error)."""
    import importlib
    cfg = syn.SHAPES[shape]
    F = cfg["n_frames"]
    skeleton = refload.make_skeleton(syn.make_skeleton_desc(cfg["n_joints"]))
    node = refload.make_graph_node(ns, syn.make_model_json(shape))
    np.random.seed(seed)
    specs = make_specs(ns, node, skeleton, F, None, name_of_set)
    S = node.sample_low_dimensional_vectors(n_samples)
    mpc = build_mp_constraints(ns, skeleton, specs, is_local=True)
    obj = ns.objective_functions.obj_spatial_error_residual_vector_and_naturalness
    out = {"samples": S, "specs_json": np.array(json.dumps(specs)), "shape": np.array(shape),
           "translation_maxima": np.array([1.0, 1.0, 1.0]), "is_local": np.array(True)}
    with _quiet():
        out["residual_vectors"] = np.array([mpc.get_residual_vector(node, s_, None) for s_ in S])
        out["objective_q01"] = np.array([obj(s_, (node, mpc, None, 1.0, 0.1, 1.0)) for s_ in S])
    # the optimiser, exactly as the generator drives it
    ls_mod = importlib.import_module("morphablegraphs.motion_generator.optimization.least_squares")
    mpc_opt = build_mp_constraints(ns, skeleton, specs, is_local=True)
    mpc_opt.constraints = [c for c in mpc_opt.constraints if c.constraint_type != "keyframe_pose"]
    out["opt_specs_json"] = np.array(json.dumps([sp_ for sp_ in specs if sp_["class"] not in ("PoseConstraint", "PoseConstraintQuatFrame")]))
    costs = np.array([ns.objective_functions.obj_spatial_error_sum(s_, (node, mpc_opt, None)) for s_ in S])
    order = np.argsort(costs)[:n_opt]
    seeds, results, start_cost, final_cost, final_min_error = [], [], [], [], []
    for qs in (0.0, 0.1):
        for i in order:
            ls = ls_mod.LeastSquares({"max_iterations": 500, "verbose": False})
            ls.set_objective_function(obj)
            data = (node, mpc_opt, None, 1.0, qs, 1.0)
            with _quiet():
                error_sum = max(abs(np.sum(obj(S[i], data))), 1.0)
                data = (node, mpc_opt, None, 1.0, qs, error_sum)
                ls.set_objective_function_parameters(data)
                x = ls.run(initial_guess=S[i].copy())
                r0 = obj(S[i], (node, mpc_opt, None, 1.0, qs, 1.0))
                r1 = obj(x, (node, mpc_opt, None, 1.0, qs, 1.0))
            seeds.append(S[i]); results.append(x)
            start_cost.append(0.5 * np.sum(r0 ** 2)); final_cost.append(0.5 * np.sum(r1 ** 2))
            final_min_error.append(mpc_opt.min_error)
    out["opt_quality_scales"] = np.array([0.0] * n_opt + [0.1] * n_opt)
    out["opt_seeds"] = np.array(seeds)
    out["opt_results"] = np.array(results)
    out["opt_start_cost"] = np.array(start_cost)
    out["opt_final_cost"] = np.array(final_cost)
    out["opt_final_min_error"] = np.array(final_min_error)
    return out


def run_walk_case(ns, n_steps=6, n_samples=48, seed=31):
    """BASELINE config 3 in miniature: a walk over several primitives (one S-shaped model per graph
    node, its own model seed), every step picked by the reference's own
    `evaluate_samples_using_constraints` loop on the non-local path (candidates aligned to the previous
    step's last frames, motion_primitive_constraints.py:110-115), the chosen motion aligned and appended
    like graph_walk.py does.  Stores per step what the engine needs to replay the decision."""
    from anim_utils.animation_data.motion_concatenation import align_quaternion_frames
    shape = "S"
    cfg = syn.SHAPES[shape]
    F = cfg["n_frames"]
    skeleton = refload.make_skeleton(syn.make_skeleton_desc(cfg["n_joints"]))
    names = ["walk_beginLeftStance", "walk_rightStance", "walk_leftStance", "walk_endRightStance"]
    seeds = [syn.SEED_MODEL + 17 * k for k in range(len(names))]
    nodes = [refload.make_graph_node(ns, syn.make_model_json(shape, seed=sd, name=nm), name=nm) for nm, sd in zip(names, seeds)]
    order = [0, 1, 2, 1, 2, 3][:n_steps]
    np.random.seed(seed)
    rng = np.random.RandomState(syn.SEED_CONSTRAINTS)
    obj = ns.objective_functions.obj_spatial_error_sum
    out = {"node_names_json": np.array(json.dumps(names)), "model_seeds": np.array(seeds), "order": np.array(order)}
    prev_frames = None
    for step, k in enumerate(order):
        node = nodes[k]
        specs = make_specs(ns, node, skeleton, F, rng, "config1" if step % 2 == 0 else "mixed")
        if prev_frames is not None:
            # the transition term compares with the end of the previous step (pose_constraint.py:48-67)
            specs = [sp_ for sp_ in specs if sp_["class"] != "PoseConstraint"]
        S = node.sample_low_dimensional_vectors(n_samples)
        mpc = build_mp_constraints(ns, skeleton, specs, is_local=prev_frames is None)
        gen = ns.motion_primitive_generator.MotionPrimitiveGenerator.__new__(ns.motion_primitive_generator.MotionPrimitiveGenerator)
        gen.objective = obj
        with _quiet():
            costs = np.array([obj(s_, (node, mpc, prev_frames)) for s_ in S])
            best, min_error = gen.evaluate_samples_using_constraints(S, node, mpc, prev_frames)
        spline = node.back_project(best, use_time_parameters=False)
        if prev_frames is not None:
            spline.coeffs = align_quaternion_frames(skeleton, skeleton.aligning_root_node, spline.coeffs, prev_frames, None)
        with _quiet():
            frames = np.asarray(spline.get_motion_vector())
        out["step%d_samples" % step] = S
        out["step%d_specs_json" % step] = np.array(json.dumps(specs))
        out["step%d_costs" % step] = costs
        out["step%d_best_idx" % step] = np.array(int(np.where((S == best).all(axis=1))[0][0]))
        out["step%d_min_error" % step] = np.array(min_error)
        if prev_frames is not None:
            out["step%d_prev_frames" % step] = np.asarray(prev_frames)
        out["step%d_last_frames" % step] = frames[-3:]
        prev_frames = frames[-3:]
    return out


def run_tree_case(ns, n_data=400, n_candidates_list=(1, 2, 5), seed=11):
    """ClusterTree (k-means + kd leaves) and FeatureClusterTree built by the
    reference (KMeans seeded through numpy's global state) and searched with
    the config-1 objective.  Stores the tree topology so the device-side search
    can be checked on the *same* tree."""
    shape = "S"
    cfg = syn.SHAPES[shape]
    F = cfg["n_frames"]
    mj = syn.make_model_json(shape)
    skeleton = refload.make_skeleton(syn.make_skeleton_desc(cfg["n_joints"]))
    node = refload.make_graph_node(ns, mj)
    np.random.seed(seed)
    specs = make_specs(ns, node, skeleton, F, None, "config1")
    mpc = build_mp_constraints(ns, skeleton, specs)
    data = node.sample_low_dimensional_vectors(n_data)
    K = node.get_n_spatial_components()
    obj = ns.objective_functions.obj_spatial_error_sum
    out = {"data": data, "specs_json": np.array(json.dumps(specs))}
    with _quiet():
        # max_level 6: a node that still holds > n_subdivisions samples at depth == max_level gets
        # kd-tree children but leaf=False and the reference's own search then dies on
        # `KDTreeWrapper.mean` (cluster_tree_node_builder.py:139-154 vs cluster_tree_node.py:133-134);
        # deep enough that every branch ends in a proper leaf
        tree = ns.cluster_tree.ClusterTree(4, 6, K, store_indices=True, use_kd_tree=True)
        tree.construct(data)
    out["cluster_tree_json"] = np.array(json.dumps(serialize_cluster_tree(tree.root)))
    res = []
    for nc in n_candidates_list:
        with _quiet():
            v, s = tree.find_best_example_excluding_search_candidates(obj, (node, mpc, None), nc)
        res.append(np.concatenate([[v], np.asarray(s, dtype=np.float64)]))
    out["cluster_tree_results"] = np.array(res)
    out["n_candidates"] = np.array(n_candidates_list)
    with _quiet():
        opts = {"n_subdivisions": 4, "clustering_method": 0, "use_feature_mean": False}
        ftree = ns.feature_cluster_tree.FeatureClusterTree(data[:, :K], data, None, opts, [])
    out["feature_tree_json"] = np.array(json.dumps(serialize_feature_tree(ftree)))
    res = []
    for nc in n_candidates_list:
        with _quiet():
            v, s = ftree.find_best_example_excluding_search_candidates(obj, (node, mpc, None), nc)
        res.append(np.concatenate([[v], np.asarray(s, dtype=np.float64)]))
    out["feature_tree_results"] = np.array(res)
    # ---- builds with INJECTED seeds: the reference's own builder code, its
    # sklearn KMeans handed fixed initial centres (members at positions len*j//k of the node's member
    # list) so that the subdivision is reproducible by the oracle restatement and the device build
    import sklearn.cluster

    class SeededKMeans(object):
        def __init__(self, n_clusters=4):
            self.k = n_clusters

        def fit_predict(self, X):
            X = np.asarray(X)
            init = X[[(len(X) * j) // self.k for j in range(self.k)]]
            return sklearn.cluster.KMeans(n_clusters=self.k, init=init, n_init=1, algorithm="lloyd").fit_predict(X)

    import types
    shim = types.SimpleNamespace(KMeans=SeededKMeans)
    builder_mod = sys.modules[ns.cluster_tree.ClusterTreeNodeBuilder.__module__]
    clustering_mod = sys.modules[ns.feature_cluster_tree.find_clusters.__module__]
    saved = builder_mod.cluster, clustering_mod.cluster
    builder_mod.cluster = clustering_mod.cluster = shim
    try:
        with _quiet():
            for tag, max_level in (("seeded6", 6), ("seeded2", 2)):
                tree = ns.cluster_tree.ClusterTree(4, max_level, K, store_indices=True, use_kd_tree=True)
                tree.construct(data)
                out["cluster_tree_%s_json" % tag] = np.array(json.dumps(serialize_cluster_tree(tree.root)))
                if max_level == 6:
                    res = []
                    for nc in n_candidates_list:
                        v, s = tree.find_best_example_excluding_search_candidates(obj, (node, mpc, None), nc)
                        res.append(np.concatenate([[v], np.asarray(s, dtype=np.float64)]))
                    out["cluster_tree_seeded6_results"] = np.array(res)
            ftree = ns.feature_cluster_tree.FeatureClusterTree(data[:, :K], data, None, opts, [])
            out["feature_tree_seeded_json"] = np.array(json.dumps(serialize_feature_tree(ftree)))
            res = []
            for nc in n_candidates_list:
                v, s = ftree.find_best_example_excluding_search_candidates(obj, (node, mpc, None), nc)
                res.append(np.concatenate([[v], np.asarray(s, dtype=np.float64)]))
            out["feature_tree_seeded_results"] = np.array(res)
    finally:
        builder_mod.cluster, clustering_mod.cluster = saved
    # ---- euclidean PCA features (a19): the reference's features.map_motions_to_euclidean_pca on motions
    # evaluated like construction/cluster_tree_builder.py:293-301 (every integer canonical frame), with
    # use_time_parameters=False (the time-warp branch raises under numpy 2, quirk Q3)
    n_feat = 120
    motions = []
    for sv in data[:n_feat]:
        spline = node.back_project(sv, use_time_parameters=False)
        motions.append([spline.evaluate(fi) for fi in range(int(node.get_n_canonical_frames()))])
    motions = np.array(motions)
    with _quiet():
        feats, pca = ns.features.map_motions_to_euclidean_pca(motions, skeleton, ns.features.END_EFFECTORS2, step=1)
        clouds = np.array(ns.features.map_motions_to_euclidean_space(motions[:4], skeleton, ns.features.END_EFFECTORS2, step=1))
    out["euclid_n_samples"] = np.array(n_feat)
    out["euclid_point_clouds"] = clouds
    out["euclid_features"] = np.asarray(feats)
    out["euclid_explained_variance_ratio"] = np.asarray(pca.explained_variance_ratio_)
    return out


def serialize_kd(n):
    if n is None:
        return None
    return {"point": list(map(float, n.point)), "type": n.type, "left": serialize_kd(n.left),
            "right": serialize_kd(n.right)}


def serialize_cluster_tree(node):
    d = {"mean": node.mean.tolist(), "leaf": bool(node.leaf), "children": []}
    for c in node.clusters:
        if c.type == "kdtree":
            d["children"].append({"kdtree": serialize_kd(c.kdtree.root)})
        else:
            d["children"].append(serialize_cluster_tree(c))
    return d


def serialize_feature_tree(node):
    return {"mean": np.asarray(node._mean).tolist(),
            "indices": None if node._indices is None else [int(i) for i in node._indices],
            "children": [serialize_feature_tree(c) for c in node._children]}


def main():
    ns = refload.load()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    if "--only-walk" in sys.argv:
        out = run_walk_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_walk.npz"), **out)
        print("s_walk", [int(out["step%d_best_idx" % i]) for i in range(len(out["order"]))],
              [float(out["step%d_min_error" % i]) for i in range(len(out["order"]))])
        return 0
    if "--only-timewarp" in sys.argv:
        out = run_timewarp_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_timewarp.npz"), **out)
        print("s_timewarp", [len(out["tw%d_time" % i]) for i in range(len(out["samples"]))], out["tw0_time"][:4], out["tw0_time"][-3:])
        return 0
    if "--only-config3" in sys.argv:
        out = run_config3_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_config3.npz"), **out)
        return 0
    if "--only-resid" in sys.argv:
        out = run_resid_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_resid.npz"), **out)
        print("s_resid R =", out["residual_vectors"].shape, "start", out["opt_start_cost"], "final", out["opt_final_cost"])
        return 0
    if "--only-round2" in sys.argv:
        # fixtures added in round 2: reference-made goldens at the sizes SURVEY.md 8(d) states
        out = run_local_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_local.npz"), **out)
        print("s_local", out["costs"][:3], int(out["best_idx"]), float(out["min_error"]))
        for name, kw in (("s_config1_10k", dict(shape="S", which="config1", n_samples=10000, seed=2024, n_full=16)),
                         ("l_cfg4", dict(shape="L", which="cfg4", n_samples=1000, seed=2032, n_full=8)),
                         ("l_mixed_1k", dict(shape="L", which="mixed", n_samples=1000, seed=2029, n_full=4))):
            if "--skip-" + name in sys.argv:
                continue
            out, _ = run_case(ns, **kw)
            np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), **out)
            print(name, "costs[:3]", out["costs"][:3], "best", int(out["best_idx"]), float(out["min_error"]), flush=True)
        return 0
    if "--only-tree" in sys.argv:
        out = run_tree_case(ns)
        np.savez_compressed(os.path.join(GOLDEN_DIR, "s_tree.npz"), **out)
        print("s_tree", out["cluster_tree_results"][:, 0], out["cluster_tree_seeded6_results"][:, 0], out["feature_tree_seeded_results"][:, 0])
        return 0
    cases = [
        ("s_config1", dict(shape="S", which="config1", n_samples=96, seed=2024)),
        ("s_mixed", dict(shape="S", which="mixed", n_samples=64, seed=2025)),
        ("s_mixed_tm", dict(shape="S", which="mixed", n_samples=32, seed=2026, translation_maxima=(50.0, 10.0, 80.0))),
        ("s_nonlocal", dict(shape="S", which="mixed", n_samples=32, seed=2027, non_local=True)),
        ("xs_mixed", dict(shape="XS", which="mixed", n_samples=32, seed=2028)),
        ("l_mixed", dict(shape="L", which="mixed", n_samples=12, seed=2029)),
    ]
    if "--only-traj" in sys.argv:
        cases = [("s_traj", dict(shape="S", which="traj", n_samples=48, seed=2061)),
                 ("s_traj_nonlocal", dict(shape="S", which="traj", n_samples=24, seed=2062, non_local=True))]
    if "--only-rot" in sys.argv:
        cases = [("s_rot", dict(shape="S", which="rot", n_samples=64, seed=2033)),
                 ("s_rot_nonlocal", dict(shape="S", which="rot", n_samples=32, seed=2034, non_local=True))]
    for name, kw in cases:
        out, _ = run_case(ns, **kw)
        # keep fixtures small: coefficients only for the first 8 candidates
        out["coeffs"] = out["coeffs"][:8]
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), **out)
        print(name, "costs[:3]", out["costs"][:3], "best", int(out["best_idx"]), float(out["min_error"]))
    if "--only-rot" in sys.argv or "--only-traj" in sys.argv:
        return 0
    out = run_tree_case(ns)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "s_tree.npz"), **out)
    print("s_tree", out["cluster_tree_results"][:, 0], out["feature_tree_results"][:, 0])


if __name__ == "__main__":
    sys.exit(main())
