"""BASELINE config 3: synthesis of a whole action list, every motion-primitive step decided on the GPU
(SURVEY.md 8(d) row 3).

Input: the reference's input file format (examples/example_input.json: `startPose` and a list of
`elementaryActions`, each with `trajectoryConstraints` - a path for the root - or `keyframeConstraints` - a
joint position at a labelled keyframe), read with the CAD -> OpenGL axis change the reference applies
(constraints/mg_input_format_reader/utils.py:27-34: [x, y, z] -> [x, z, -y], `null` = unconstrained).

What this module is NOT: the reference's planners and constraint builders (graph_walk_planner.py,
elementary_action_constraints_builder.py, ... - control plane, out of scope).  It is the smallest driver that turns
the input file into the sequence of motion-primitive steps the hot path sees - walk along the path in steps of one
stance primitive, reach at the labelled keyframe - so that the generator can be exercised, measured and checked
step by step against the reference's own sample-and-evaluate loop (tests/golden/s_config3.npz).
Every step is scored on the non-local path: candidates aligned to the last frames of the walk so far
(motion_primitive_constraints.py:110-115).
"""
import json

import numpy as np

from . import synthetic as syn
from .constraints import MotionPrimitiveConstraints, CONSTRAINT_FACTORY
from .skeleton import align_frames

POSE_JOINTS = ["Hips", "LeftHand", "RightHand", "LeftFoot", "RightFoot"]


def cad_to_gl(p):
    """constraints/mg_input_format_reader/utils.py:27-34 with `null` entries kept as None."""
    x, y, z = p
    return [x, z, None if y is None else -y]


def read_input(mg_input):
    """-> dict(start_position [x, y, z] GL, start_heading_deg, actions [dict(action, path | joint/position/label)])"""
    if isinstance(mg_input, str):
        with open(mg_input) as f:
            mg_input = json.load(f)
    start = mg_input["startPose"]
    actions = []
    for a in mg_input["elementaryActions"]:
        for c in a["constraints"]:
            if "trajectoryConstraints" in c:
                pts = [cad_to_gl(t["position"]) for t in c["trajectoryConstraints"]]
                actions.append({"action": a["action"], "joint": c["joint"], "path": [[p[0], p[2]] for p in pts]})
            for kc in c.get("keyframeConstraints", []):
                label = [k for k, v in kc.get("semanticAnnotation", {}).items() if v is True][0]
                actions.append({"action": a["action"], "joint": c["joint"], "position": cad_to_gl(kc["position"]), "label": label})
    return {"start_position": cad_to_gl(start["position"]), "start_orientation": start["orientation"], "actions": actions}


def _path_point(path, s):
    """Point and unit tangent of the polyline `path` ([x, z] vertices) at arc length s (clamped)."""
    p = np.asarray(path, dtype=np.float64)
    seg = np.linalg.norm(p[1:] - p[:-1], axis=1)
    cum = np.concatenate([[0.0], np.cumsum(seg)])
    s = min(max(s, 0.0), cum[-1])
    i = min(int(np.searchsorted(cum, s, side="right")) - 1, len(seg) - 1)
    while seg[i] == 0.0 and i + 1 < len(seg):
        i += 1
    t = (p[i + 1] - p[i]) / max(seg[i], 1e-12)
    return p[i] + (s - cum[i]) * t, t, cum[-1]


def plan_steps(graph, plan_input):
    """The motion-primitive steps of the action list: locomotion actions alternate the stance primitives along the path in
    steps of syn.STEP_LENGTH (begin..., left/right..., end...), reach actions are one primitive."""
    steps = []
    for a in plan_input["actions"]:
        action = a["action"]
        if "path" in a:
            total = _path_point(a["path"], 0.0)[2]
            n = max(2, int(np.ceil(total / syn.STEP_LENGTH)))
            for i in range(n):
                left = (i % 2 == 0)
                kind = "begin" if i == 0 else ("end" if i == n - 1 else "")
                mp = (kind + ("LeftStance" if left else "RightStance")) if kind else ("leftStance" if left else "rightStance")
                goal, tangent, _ = _path_point(a["path"], total * (i + 1) / n)
                steps.append({"node_key": (action, mp), "goal_xz": goal.tolist(), "dir_xz": tangent.tolist()})
        else:
            mp = graph.node_groups[action].start_states[0]
            frame = graph.node_groups[action].labeled_frames[mp][a["label"]]
            steps.append({"node_key": (action, mp), "joint": a["joint"], "position": a["position"], "keyframe": frame})
    return steps


def step_specs(graph, step, prev_frames, pose_points=None):
    """Constraint specs (class name + descriptor, synthetic.spec) of one step, in GLOBAL coordinates.
    pose_points: optional [2, len(POSE_JOINTS), 3] global positions of POSE_JOINTS in the last two previous frames (the
    synthesizer gets them from K4 on the device; without them the host skeleton is walked)."""
    node = graph.nodes[step["node_key"]]
    F = node.get_n_canonical_frames()
    skel = graph.skeleton
    specs = []
    if "goal_xz" in step:
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc("Hips", [step["goal_xz"][0], None, step["goal_xz"][1]], F - 1, F), 1.0, 1.0))
        specs.append(syn.spec("Direction2DConstraint", {"canonical_keyframe": F - 1, "dir_vector": [step["dir_xz"][0], 0.0, step["dir_xz"][1]],
                                                        "semanticAnnotation": syn.sem()}, 1.0, 0.05))
    else:
        specs.append(syn.spec("GlobalTransformConstraint",
                              syn.position_constraint_desc(step["joint"], step["position"], step["keyframe"], F, label="contact"), 1.0, 1.0))
    # transition continuity (motion_primitive_constraints_builder.py:286-306): pose of the last frame so far, root velocity
    if pose_points is not None:
        cloud = np.asarray(pose_points[1], dtype=np.float64).tolist()
        vel = (np.asarray(pose_points[1][0], dtype=np.float64) - np.asarray(pose_points[0][0], dtype=np.float64)).tolist()
    else:
        last, before = np.asarray(prev_frames[-1]), np.asarray(prev_frames[-2] if len(prev_frames) > 1 else prev_frames[-1])
        cloud = [skel.nodes[n].get_global_position(last).tolist() for n in POSE_JOINTS]
        vel = (np.asarray(cloud[0]) - skel.nodes["Hips"].get_global_position(before)).tolist()
    specs.append(syn.spec("PoseConstraint", {"canonical_keyframe": 0, "frame_constraint": cloud, "velocity_constraint": vel,
                                             "semanticAnnotation": syn.sem("start"), "node_names": POSE_JOINTS,
                                             "weights": [skel.joint_weight_map[n] for n in POSE_JOINTS]}, 1.0, 1.0))
    return specs


def build_constraints(graph, specs, factory=None):
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = graph.skeleton
    mpc.is_local = False
    mpc.constraints = syn.build_constraints(factory or CONSTRAINT_FACTORY, graph.skeleton, specs)
    return mpc


def initial_frames(graph, plan_input, first_node_key):
    """Two copies of the first primitive's mean pose at t = 0, placed at the start position and turned towards the first
    action's path (the reference aligns the first step to `startPose`, a host-side transform the engine refuses to repeat
    per candidate - quirk Q2; giving the walk an explicit first frame does the same once)."""
    mp = graph.nodes[first_node_key].motion_primitive
    P = mp.s_pca["n_dim"]
    frame = np.array(mp.s_pca["mean_vector"][:P], dtype=np.float64)
    frame[:3] *= mp.translation_maxima
    path = next(a["path"] for a in plan_input["actions"] if "path" in a)
    d = np.asarray(path[1]) - np.asarray(path[0])
    ang = np.arctan2(d[0], d[1])
    frame[3:7] = [np.cos(0.5 * ang), 0.0, np.sin(0.5 * ang), 0.0]
    frame[0], frame[2] = plan_input["start_position"][0], plan_input["start_position"][2]
    return np.stack([frame, frame])


class GraphWalkSynthesizer(object):
    """Runs the step list with a MotionPrimitiveGenerator (this package's or the reference's: same interface)."""

    def __init__(self, graph, generator, keep_frames=8, device_fk=False):
        self.graph, self.generator, self.keep_frames, self.device_fk = graph, generator, keep_frames, device_fk

    def run(self, mg_input, max_steps=None, on_step=None):
        plan_input = read_input(mg_input)
        steps = plan_steps(self.graph, plan_input)
        if max_steps is not None:
            steps = steps[:max_steps]
        prev = initial_frames(self.graph, plan_input, steps[0]["node_key"])
        out = []
        pose_points = None
        for i, step in enumerate(steps):
            node = self.graph.nodes[step["node_key"]]
            specs = step_specs(self.graph, step, prev, pose_points)
            mpc = build_constraints(self.graph, specs)
            s = self.generator.generate_constrained_sample(node, mpc, "", prev, None)
            frames = node.back_project(s, use_time_parameters=False).get_motion_vector()
            frames = align_frames(self.graph.skeleton, frames, prev)
            rec = {"node_key": step["node_key"], "parameters": s, "min_error": mpc.min_error, "specs": specs, "prev_frames": prev,
                   "last_frames": frames[-self.keep_frames:]}
            out.append(rec)
            if on_step is not None:
                on_step(i, rec)
            prev = frames[-self.keep_frames:]
            if self.device_fk:
                # the transition constraint of the next step needs the end-effector positions of the last two frames; K4 can
                # supply them, but for TWO frames its per-call plan upload costs more (0.5 ms) than walking the host skeleton
                # (0.2 ms, tools/c3_cost.py), so this is off by default
                dm = node.motion_primitive.device_model(self.graph.skeleton)
                pose_points = dm.fk(np.ascontiguousarray(frames[-2:], dtype=np.float32), POSE_JOINTS).cpu().numpy()
        return out
