"""Host-side boundary logic that needs no GPU: the content key of a compiled constraint program,
the mapping of objective callables, and the generator entry points the reference's callers use
(generate_constrained_motion_spline / generate_random_sample,
reference motion_generator/motion_primitive_generator.py:83-124,208-213)."""
import types

import numpy as np
import pytest

from tests.common import Case
from morphablegraphs_b200 import engine, MotionPrimitiveGenerator
from morphablegraphs_b200 import objective_functions as OF
from morphablegraphs_b200._lib import MgbError


class _StubModel(object):
    """what flatten_constraints reads of a DeviceModel (no device needed)"""

    def __init__(self, skeleton):
        sk = engine.pack_skeleton(skeleton)
        self.node_index = sk["index"]
        self.skeleton = skeleton


@pytest.fixture(scope="module")
def case(lib):
    return Case("s_mixed")


def _key(case, dm, prev=None):
    flat = engine.flatten_constraints(dm, case.constraints)
    return OF.constraints_key(dm, flat, case.constraints, prev)


def test_constraints_key_follows_in_place_mutation(case):
    dm = _StubModel(case.skeleton)
    k0 = _key(case, dm)
    assert _key(case, dm) == k0                      # unchanged content -> same key, whatever the object ids
    c = case.constraints.constraints[0]
    old = c.canonical_keyframe
    c.canonical_keyframe = old / 2.0                 # graph_walk_planner.py:199-203 does exactly this
    k1 = _key(case, dm)
    assert k1 != k0
    c.canonical_keyframe = old
    assert _key(case, dm) == k0
    c.position[0] += 1.0                             # a target edited in place
    assert _key(case, dm) != k0
    c.position[0] -= 1.0
    c.weight_factor *= 2.0                           # a weight edited in place
    assert _key(case, dm) != k0
    c.weight_factor /= 2.0
    assert _key(case, dm) == k0
    # swapping a constraint for another one at the same list length
    saved = case.constraints.constraints[1]
    case.constraints.constraints[1] = case.constraints.constraints[2]
    assert _key(case, dm) != k0
    case.constraints.constraints[1] = saved
    assert _key(case, dm) == k0


def test_constraints_key_reads_prev_frames_content():
    case = Case("s_nonlocal")
    dm = _StubModel(case.skeleton)
    prev = np.array(case.prev_frames, dtype=np.float64)
    k0 = _key(case, dm, prev)
    assert _key(case, dm, prev.copy()) == k0         # a different array object with the same last frame
    prev2 = prev.copy()
    prev2[-1, 0] += 5.0                              # the frame the alignment reads changed in place
    assert _key(case, dm, prev2) != k0


def test_objective_scales_maps_reference_functions_and_refuses_others():
    assert OF.objective_scales(OF.obj_spatial_error_sum) == (1.0, 0.0, 0.0)
    assert OF.objective_scales(OF.obj_spatial_error_sum_and_magnitude) == (1.0, 0.0, 2.0)
    assert OF.objective_scales(OF.ObjSpatialErrorAndNaturalness(1.0, 0.1)) == (1.0, 0.1, 0.0)

    # the reference's planner assigns a PLAIN FUNCTION (mg_state_planner.py:250): mapped by name
    def obj_spatial_error_sum_and_magnitude(s, data):
        raise AssertionError("never called on the device path")

    def obj_spatial_error_sum(s, data):
        raise AssertionError("never called on the device path")

    assert OF.objective_scales(obj_spatial_error_sum_and_magnitude) == (1.0, 0.0, 2.0)
    assert OF.objective_scales(obj_spatial_error_sum) == (1.0, 0.0, 0.0)

    def my_objective(s, data):
        return 0.0

    with pytest.raises(MgbError):
        OF.objective_scales(my_objective)
    with pytest.raises(MgbError):
        OF.objective_scales(lambda s, d: 0.0)
    with pytest.raises(MgbError):
        OF.objective_scales(None)


def test_reference_objective_functions_map_by_name():
    """the reference's own function objects (loaded unmodified) resolve to the same scales"""
    from oracle import refload
    if not refload.available():
        pytest.skip("reference tree not present")
    ns = refload.load()
    assert OF.objective_scales(ns.objective_functions.obj_spatial_error_sum) == (1.0, 0.0, 0.0)


class _Node(object):
    def __init__(self, tag, D=4):
        self.tag, self.D = tag, D
        self.calls = []

    def sample_low_dimensional_vector(self):
        self.calls.append("sample")
        return np.full(self.D, 1.0)

    def back_project(self, parameters, use_time_parameters=True):
        self.calls.append(("back_project", use_time_parameters))
        return ("spline", self.tag, tuple(parameters))

    def has_transition_model(self, to_key):
        return True

    def predict_parameters(self, to_key, prev_parameters):
        self.calls.append(("predict", to_key))
        return np.asarray(prev_parameters) * 2.0


def _generator(use_transition_model=False):
    nodes = {("walk", "leftStance"): _Node("l"), ("walk", "rightStance"): _Node("r")}
    graph = types.SimpleNamespace(nodes=nodes, skeleton=None)
    ac = types.SimpleNamespace(action_name="walk", prev_action_name="walk", motion_state_graph=graph)
    gen = MotionPrimitiveGenerator(ac, {"n_random_samples": 8, "use_transition_model": use_transition_model}, seed=1)
    return gen, nodes


def test_generate_random_sample():
    gen, nodes = _generator()
    s = gen.generate_random_sample(("walk", "leftStance"))
    assert np.all(s == 1.0) and nodes[("walk", "leftStance")].calls == ["sample"]
    gen, nodes = _generator(use_transition_model=True)
    s = gen.generate_random_sample(("walk", "leftStance"), "rightStance", np.array([1.0, 2.0]))
    assert np.all(s == [2.0, 4.0]) and nodes[("walk", "rightStance")].calls == [("predict", ("walk", "leftStance"))]


def test_generate_constrained_motion_spline_bookkeeping():
    gen, nodes = _generator()
    seen = {}

    def fake_sample(graph_node, mpc, prev_mp_name="", prev_frames=None, prev_parameters=None):
        seen["args"] = (graph_node.tag, prev_mp_name, prev_frames, prev_parameters)
        return np.arange(4.0)

    gen.generate_constrained_sample = fake_sample
    mpc = types.SimpleNamespace(motion_primitive_name="leftStance", constraints=[object()], time=None)
    step = types.SimpleNamespace(node_key=("walk", "rightStance"), parameters=np.ones(4))
    walk = types.SimpleNamespace(steps=[step], get_quat_frames=lambda: "frames")
    spline, params = gen.generate_constrained_motion_spline(mpc, walk)
    assert seen["args"][0] == "l" and seen["args"][1] == "rightStance" and seen["args"][2] == "frames"
    assert np.all(seen["args"][3] == 1.0)
    assert spline == ("spline", "l", (0.0, 1.0, 2.0, 3.0)) and np.all(params == np.arange(4.0))
    assert nodes[("walk", "leftStance")].calls == [("back_project", False)]
    assert mpc.time is not None and mpc.time >= 0.0
    # no constraints: random sample, time still set
    mpc = types.SimpleNamespace(motion_primitive_name="leftStance", constraints=[], time=None)
    walk = types.SimpleNamespace(steps=[], get_quat_frames=lambda: None)
    spline, params = gen.generate_constrained_motion_spline(mpc, walk)
    assert np.all(params == 1.0) and mpc.time is not None
