"""Synthetic motion-primitive models, skeletons and constraint sets.

The reference ships no model data (its example configs point at absolute
Windows paths, reference `examples/config/service.config:2-5`), so every
benchmark and parity input is generated here, following the shapes the
reference's training code writes:

* model JSON keys: reference `morphablegraphs/motion_model/motion_primitive.py:93-180`
  (legacy format, writer `construction/motion_model_constructor.py:479-494`);
* `Nb = int(F * 0.2)` spatial B-spline coefficients
  (`construction/motion_model_constructor.py:357`);
* clamped uniform cubic knots (`construction/utils.py:187-198`);
* frame layout `[tx, ty, tz, (w, x, y, z) * J]` (`motion_generator/utils.py:39-45`).

All draws use `numpy.random.RandomState(seed)` with the fixed seeds of
SURVEY.md section 8(d): model 1234, skeleton 7, constraints 99, samples 2024.

Everything here is plain host-side data generation (dicts of lists / numpy
arrays).  No scoring arithmetic lives in this module.
"""
import collections

import numpy as np

SEED_MODEL = 1234
SEED_SKELETON = 7
SEED_CONSTRAINTS = 99
SEED_SAMPLES = 2024

# name -> (F, J, K, K_t, G); Nb = int(0.2 F), P = 3 + 4 J
SHAPES = {
    "S": dict(n_frames=50, n_joints=19, n_spatial=10, n_time=3, n_gmm=8),
    "L": dict(n_frames=100, n_joints=70, n_spatial=100, n_time=3, n_gmm=32),
    # tiny shape used by CPU-only unit tests of host logic
    "XS": dict(n_frames=20, n_joints=19, n_spatial=4, n_time=2, n_gmm=3),
}

NODE_ROOT = "root"
NODE_JOINT = "joint"
NODE_END_SITE = "end_site"

# the 19 animated joints of the small skeleton (root included); the count
# matches the joint list in reference `mg_state_planner.py:54-73` plus root
_BODY_19 = [
    # name, parent, direction (unit-ish axis the bone points along)
    ("Hips", None, (0.0, 1.0, 0.0)),
    ("Spine", "Hips", (0.0, 1.0, 0.0)),
    ("Spine_1", "Spine", (0.0, 1.0, 0.0)),
    ("Neck", "Spine_1", (0.0, 1.0, 0.0)),
    ("Head", "Neck", (0.0, 1.0, 0.0)),
    ("LeftShoulder", "Spine_1", (1.0, 0.3, 0.0)),
    ("LeftArm", "LeftShoulder", (1.0, 0.0, 0.0)),
    ("LeftForeArm", "LeftArm", (1.0, 0.0, 0.0)),
    ("LeftHand", "LeftForeArm", (1.0, 0.0, 0.0)),
    ("RightShoulder", "Spine_1", (-1.0, 0.3, 0.0)),
    ("RightArm", "RightShoulder", (-1.0, 0.0, 0.0)),
    ("RightForeArm", "RightArm", (-1.0, 0.0, 0.0)),
    ("RightHand", "RightForeArm", (-1.0, 0.0, 0.0)),
    ("LeftUpLeg", "Hips", (0.4, -1.0, 0.0)),
    ("LeftLeg", "LeftUpLeg", (0.0, -1.0, 0.0)),
    ("LeftFoot", "LeftLeg", (0.0, -1.0, 0.0)),
    ("RightUpLeg", "Hips", (-0.4, -1.0, 0.0)),
    ("RightLeg", "RightUpLeg", (0.0, -1.0, 0.0)),
    ("RightFoot", "RightLeg", (0.0, -1.0, 0.0)),
]

# fixed (non-animated) joints and end sites hanging off the body; the
# Bip01_* names are the ones reference `space_partitioning/features.py:32`
# (END_EFFECTORS2) looks up
_FIXED = [
    ("Bip01_L_Toe0", "LeftFoot", (0.0, -0.3, 1.0)),
    ("Bip01_R_Toe0", "RightFoot", (0.0, -0.3, 1.0)),
    ("Bip01_L_Finger2", "LeftHand", (1.0, 0.0, 0.0)),
    ("Bip01_R_Finger2", "RightHand", (-1.0, 0.0, 0.0)),
]
_END_SITES = [
    ("Head_EndSite", "Head", (0.0, 1.0, 0.0)),
    ("Bip01_L_Toe0_EndSite", "Bip01_L_Toe0", (0.0, 0.0, 1.0)),
    ("Bip01_R_Toe0_EndSite", "Bip01_R_Toe0", (0.0, 0.0, 1.0)),
    ("Bip01_L_Finger2_EndSite", "Bip01_L_Finger2", (1.0, 0.0, 0.0)),
    ("Bip01_R_Finger2_EndSite", "Bip01_R_Finger2", (-1.0, 0.0, 0.0)),
]

END_EFFECTORS2 = ["LeftUpLeg", "RightUpLeg", "RightHand", "Bip01_R_Finger2", "LeftHand",
                  "Bip01_L_Finger2", "RightFoot", "Bip01_R_Toe0", "LeftFoot", "Bip01_L_Toe0"]


def _extra_animated(n_extra):
    """Extra animated joints for the large skeleton (fingers, twist bones, ...)."""
    extra = []
    for side, sx in (("Left", 1.0), ("Right", -1.0)):
        for f in range(5):
            parent = side + "Hand"
            for seg in range(3):
                name = "%sFinger%d_%d" % (side, f, seg)
                extra.append((name, parent, (sx, 0.0, 0.2 * (f - 2))))
                parent = name
    for side in ("Left", "Right"):
        extra.append((side + "ToeBase", side + "Foot", (0.0, -0.2, 1.0)))
        extra.append((side + "ToeTip", side + "ToeBase", (0.0, 0.0, 1.0)))
    for side, sx in (("Left", 1.0), ("Right", -1.0)):
        extra.append((side + "ArmTwist", side + "Arm", (sx, 0.0, 0.0)))
        extra.append((side + "ForeArmTwist", side + "ForeArm", (sx, 0.0, 0.0)))
        extra.append((side + "UpLegTwist", side + "UpLeg", (0.0, -1.0, 0.0)))
        extra.append((side + "LegTwist", side + "Leg", (0.0, -1.0, 0.0)))
    parent = "Head"
    i = 0
    while len(extra) < n_extra:
        name = "HeadTop_%d" % i
        extra.append((name, parent, (0.0, 1.0, 0.0)))
        parent = name
        i += 1
    return extra[:n_extra]


def make_skeleton_desc(n_joints=19, seed=SEED_SKELETON):
    """Skeleton description (plain dict) with `n_joints` animated joints.

    Fields per node mirror what anim_utils' SkeletonNode carries and what the
    reference reads from it (`utilities/__init__.py:13-16`,
    `construction/cluster_tree_builder.py:82-91`): `offset`, `rotation`
    (w,x,y,z of a fixed joint), `fixed`, `quaternion_frame_index`, `parent`,
    `children`, `node_type`.
    """
    rng = np.random.RandomState(seed)
    assert n_joints >= len(_BODY_19)
    animated = list(_BODY_19) + _extra_animated(n_joints - len(_BODY_19))
    nodes = collections.OrderedDict()
    animated_joints = []
    for idx, (name, parent, axis) in enumerate(animated):
        axis = np.asarray(axis, dtype=np.float64)
        axis = axis / np.linalg.norm(axis)
        length = rng.uniform(5.0, 45.0)
        if parent is None:
            offset = np.zeros(3)
        elif "Finger" in name or "Toe" in name or "HeadTop" in name or "Twist" in name:
            offset = axis * rng.uniform(2.0, 8.0)
        else:
            offset = axis * length
        nodes[name] = dict(name=name, parent=parent, children=[], offset=offset.tolist(),
                           rotation=[1.0, 0.0, 0.0, 0.0], fixed=False,
                           quaternion_frame_index=idx,
                           node_type=NODE_ROOT if parent is None else NODE_JOINT)
        animated_joints.append(name)
    for name, parent, axis in _FIXED:
        axis = np.asarray(axis, dtype=np.float64)
        axis = axis / np.linalg.norm(axis)
        q = np.array([1.0, 0.0, 0.0, 0.0]) + rng.normal(0.0, 0.2, 4)
        q = q / np.linalg.norm(q)
        nodes[name] = dict(name=name, parent=parent, children=[],
                           offset=(axis * rng.uniform(5.0, 15.0)).tolist(),
                           rotation=q.tolist(), fixed=True, quaternion_frame_index=-1,
                           node_type=NODE_JOINT)
    for name, parent, axis in _END_SITES:
        axis = np.asarray(axis, dtype=np.float64)
        axis = axis / np.linalg.norm(axis)
        nodes[name] = dict(name=name, parent=parent, children=[],
                           offset=(axis * rng.uniform(3.0, 10.0)).tolist(),
                           rotation=[1.0, 0.0, 0.0, 0.0], fixed=True, quaternion_frame_index=-1,
                           node_type=NODE_END_SITE)
    for name, node in nodes.items():
        if node["parent"] is not None:
            nodes[node["parent"]]["children"].append(name)
    weights = collections.OrderedDict()
    for name in animated_joints:
        depth = 0
        p = nodes[name]["parent"]
        while p is not None:
            depth += 1
            p = nodes[p]["parent"]
        weights[name] = float(1.0 / (1.0 + depth))
    return dict(root="Hips", nodes=nodes, animated_joints=animated_joints,
                aligning_root_node="Hips", aligning_root_dir=[0.0, 0.0, 1.0],
                head_joint="Head", joint_weight_map=weights)


def cubic_knots(n_frames, n_basis):
    """Clamped uniform cubic knot vector, reference `construction/utils.py:187-198`."""
    knots = np.zeros(n_basis + 4)
    knots[3:-3] = np.linspace(0.0, n_frames - 1.0, n_basis - 2)
    knots[-3:] = n_frames - 1.0
    return knots


def make_model_json(shape="S", seed=SEED_MODEL, translation_maxima=(1.0, 1.0, 1.0), name=None,
                    animated_joints=None, root_travel=60.0):
    """Random-init motion primitive in the legacy JSON layout."""
    cfg = SHAPES[shape] if isinstance(shape, str) else dict(shape)
    F, J, K, Kt, G = cfg["n_frames"], cfg["n_joints"], cfg["n_spatial"], cfg["n_time"], cfg["n_gmm"]
    Nb = int(F * 0.2)
    P = 3 + 4 * J
    D = K + Kt
    rng = np.random.RandomState(seed)
    eigen = rng.normal(0.0, 0.05, (K, Nb * P))
    mean = np.zeros((Nb, P))
    # root path: z 0 -> 60 across control points, y = 90, x = 0 (stored divided by the maxima,
    # which back-projection multiplies back in: reference motion_primitive.py:255)
    tm = np.asarray(translation_maxima, dtype=np.float64)
    mean[:, 0] = 0.0
    mean[:, 1] = 90.0 / tm[1]
    mean[:, 2] = np.linspace(0.0, root_travel, Nb) / tm[2]
    quats = np.zeros((Nb, J, 4))
    quats[..., 0] = 1.0
    quats += rng.normal(0.0, 0.1, (Nb, J, 4))
    quats /= np.linalg.norm(quats, axis=-1, keepdims=True)
    mean[:, 3:] = quats.reshape(Nb, 4 * J)
    weights = rng.dirichlet(np.ones(G))
    means = rng.normal(0.0, 1.0, (G, D))
    covars = np.zeros((G, D, D))
    for g in range(G):
        A = rng.normal(0.0, 0.3, (D, D))
        covars[g] = A.dot(A.T) + 0.1 * np.eye(D)
    n_basis_time = 8
    data = {
        "name": name or ("walk_leftStance_%s" % (shape if isinstance(shape, str) else "custom")),
        "n_canonical_frames": F,
        "gmm_weights": weights.tolist(),
        "gmm_means": means.tolist(),
        "gmm_covars": covars.tolist(),
        "eigen_vectors_spatial": eigen.tolist(),
        "mean_spatial_vector": mean.reshape(-1).tolist(),
        "n_basis_spatial": Nb,
        "n_dim_spatial": P,
        "b_spline_knots_spatial": cubic_knots(F, Nb).tolist(),
        "translation_maxima": list(map(float, translation_maxima)),
        "eigen_vectors_time": rng.normal(0.0, 0.02, (n_basis_time, Kt)).tolist(),
        "mean_time_vector": np.zeros(n_basis_time).tolist(),
        "n_basis_time": n_basis_time,
        "b_spline_knots_time": cubic_knots(F, n_basis_time).tolist(),
        "keyframes": {"start_contact": 0, "end_contact": F - 1},
    }
    if animated_joints is not None:
        data["animated_joints"] = list(animated_joints)
    return data


def sem(label="end"):
    return {"keyframeLabel": label}


def position_constraint_desc(joint, position, keyframe, n_frames, label="end", orientation_q=None):
    """Descriptor dict in the form reference `GlobalTransformConstraint.__init__`
    takes (`global_transform_constraint.py:44-64`)."""
    desc = {"joint": joint, "position": list(position), "canonical_keyframe": keyframe,
            "n_canonical_frames": n_frames, "semanticAnnotation": sem(label)}
    if orientation_q is not None:
        desc["qOrientation"] = list(orientation_q)
    return desc


def spec(cls_name, desc, precision=1.0, weight=1.0):
    """One constraint as plain data: class name + the descriptor dict the
    reference's constraint constructors take + precision + weight factor."""
    return {"class": cls_name, "desc": desc, "precision": precision, "weight": weight}


def build_constraints(factory, skeleton, specs, trajectories=None):
    """Instantiate constraint objects from specs.  `factory` maps class name ->
    constructor with the reference's signature
    `(skeleton, constraint_desc, precision, weight_factor)`; the reference's
    classes and this package's mirror classes both fit.  A descriptor's
    `"trajectory"` entry, when it is a plain dict (specs are JSON), is replaced by
    `trajectories(dict)` - default: an `ArcLengthTable` from its `points` / `step` /
    `last_control_point`."""
    out = []
    for sp in specs:
        desc = dict(sp["desc"])
        if sp["class"] in ("TrajectoryConstraint", "TrajectorySetConstraint"):
            # constructor signatures of trajectory_constraint.py:33-35 / trajectory_set_constraint.py:30
            def trajectory(t, joint_name):
                c = factory["TrajectoryConstraint"](joint_name, [np.array(p, dtype=np.float64) for p in t["control_points"]], None, 0,
                                                    t.get("min_arc_length", 0.0), [], skeleton, sp["precision"], sp["weight"])
                if t.get("range") is not None:
                    c.set_active_range(t["range"][0], t["range"][1])
                return c
            if sp["class"] == "TrajectoryConstraint":
                c = trajectory(desc, desc["joint_name"])
            else:
                trajs = [trajectory(t, jn) for t, jn in zip(desc["trajectories"], desc["joint_names"])]
                c = factory["TrajectorySetConstraint"](trajs, list(desc["joint_names"]), skeleton, sp["precision"], sp["weight"])
                c.joint_arc_lengths = np.array(desc.get("joint_arc_lengths", [0.0] * len(trajs)), dtype=np.float64)
            c.set_number_of_canonical_frames(desc["n_canonical_frames"])
            out.append(c)
            continue
        if isinstance(desc.get("trajectory"), dict):
            if trajectories is not None:
                desc["trajectory"] = trajectories(desc["trajectory"])
            else:
                from .constraints import ArcLengthTable
                t = desc["trajectory"]
                desc["trajectory"] = ArcLengthTable(t["points"], t["step"], t.get("last_control_point"))
        out.append(factory[sp["class"]](skeleton, desc, sp["precision"], sp["weight"]))
    return out


# ---------------------------------------------------------------------------
# BASELINE config 3: a synthetic walk / pick / carry / place graph
# ---------------------------------------------------------------------------
LOCOMOTION_NODES = ["beginLeftStance", "beginRightStance", "leftStance", "rightStance", "endLeftStance", "endRightStance"]
STEP_LENGTH = 60.0     # root travel of one synthetic step primitive (make_model_json)


def make_graph_data(shape="S", seed=SEED_MODEL):
    """The dictionary `ZipReader.get_graph_data` returns (reference utilities/zip_io.py:64-95) for a synthetic graph with the
    actions of examples/example_input.json: `walk` and `carryRight` with the stance primitives of
    motion_primitive_grounding.py:45-99, `pickRight` / `placeRight` as single-primitive actions; every node an S-shaped
    random-init model with its own seed; meta information with start_states / end_states (action_meta_info.py:58-63) and the
    keyframe labels start_contact / end_contact (examples/example_input.json:39,44,85,90)."""
    cfg = SHAPES[shape]
    F = cfg["n_frames"]
    sub = collections.OrderedDict()
    k = 0
    for action in ("walk", "carryRight"):
        nodes = collections.OrderedDict()
        for mp in LOCOMOTION_NODES:
            mm = make_model_json(shape, seed=seed + 17 * k, name="%s_%s" % (action, mp))
            mm["keyframes"] = {"contact": F - 1}
            nodes[mp] = {"name": "%s_%s" % (action, mp), "mm": mm}
            k += 1
        sub[action] = {"name": action, "nodes": nodes,
                       "info": {"start_states": ["beginLeftStance", "beginRightStance"], "end_states": ["endLeftStance", "endRightStance"]}}
    for action in ("pickRight", "placeRight"):
        mm = make_model_json(shape, seed=seed + 17 * k, name="%s_reach" % action, root_travel=0.0)
        mm["keyframes"] = {"start_contact": F // 2, "end_contact": F // 2}
        sub[action] = {"name": action, "nodes": collections.OrderedDict([("reach", {"name": "%s_reach" % action, "mm": mm})]),
                       "info": {"start_states": ["reach"], "end_states": ["reach"]}}
        k += 1
    transitions = collections.OrderedDict()
    for action in ("walk", "carryRight"):
        t = transitions
        t["%s:beginLeftStance" % action] = ["%s:rightStance" % action, "%s:endRightStance" % action]
        t["%s:beginRightStance" % action] = ["%s:leftStance" % action, "%s:endLeftStance" % action]
        t["%s:leftStance" % action] = ["%s:rightStance" % action, "%s:endRightStance" % action]
        t["%s:rightStance" % action] = ["%s:leftStance" % action, "%s:endLeftStance" % action]
        nxt = "pickRight:reach" if action == "walk" else "placeRight:reach"
        t["%s:endLeftStance" % action] = [nxt]
        t["%s:endRightStance" % action] = [nxt]
    transitions["pickRight:reach"] = ["carryRight:beginLeftStance", "carryRight:beginRightStance"]
    transitions["placeRight:reach"] = ["walk:beginLeftStance", "walk:beginRightStance"]
    return {"formatVersion": "4.0", "transitions": transitions, "startNode": ["walk", "beginLeftStance"],
            "skeleton": make_skeleton_desc(cfg["n_joints"]), "subgraphs": sub}


# reference examples/example_input.json, abridged to what graph_synthesis.read_input reads (positions verbatim)
EXAMPLE_INPUT = {
    "startPose": {"position": [530.110290527344, 268.851318359375, 0.0], "orientation": [0.0, 0.0, 91.4786896560991]},
    "elementaryActions": [
        {"action": "walk", "constraints": [{"joint": "Hips", "trajectoryConstraints": [
            {"position": [530.110290527344, 268.851318359375, None]}, {"position": [355.0, 255.0, None]}, {"position": [195.0, 255.0, None]},
            {"position": [147.011029052734, 206.885131835938, None]}, {"position": [60.0, 200.0, None]}, {"position": [30.0, 200.0, None]}]}]},
        {"action": "pickRight", "constraints": [{"joint": "RightHand", "keyframeConstraints": [
            {"position": [-30.0000009536743, 200.0, 130.0], "semanticAnnotation": {"start_contact": True}}]}]},
        {"action": "carryRight", "constraints": [{"joint": "Hips", "trajectoryConstraints": [
            {"position": [30.0000000000001, 200.0, None]}, {"position": [35.0, 215.0, None]}, {"position": [105.0, 285.0, None]},
            {"position": [108.39608, 436.52072, None]}, {"position": [161.5512, 462.8008, None]}, {"position": [191.5512, 462.8008, None]}]}]},
        {"action": "placeRight", "constraints": [{"joint": "RightHand", "keyframeConstraints": [
            {"position": [251.5512, 462.8008, 129.296196185303], "semanticAnnotation": {"end_contact": True}}]}]}]}
