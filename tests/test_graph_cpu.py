"""f3 / config 3 host logic (CPU): the graph archive round trip, the loader, the step plan of the reference's example input,
the host-side frame alignment against the stand-in of anim_utils' align_quaternion_frames."""
import io
import json
import os

import numpy as np
import pytest

from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import graph_synthesis as GS
from morphablegraphs_b200.motion_state_graph import MotionStateGraphLoader, ZipReader, graph_to_bytes, write_graph_zip
from morphablegraphs_b200.skeleton import align_frames

EXAMPLE_INPUT = syn.EXAMPLE_INPUT


@pytest.fixture(scope="module")
def graph_data():
    return syn.make_graph_data()


def test_archive_round_trip(graph_data, tmp_path):
    path = os.path.join(str(tmp_path), "graph.zip")
    write_graph_zip(path, graph_data)
    back = ZipReader(path).get_graph_data()
    assert float(back["formatVersion"]) == 4.0
    assert back["transitions"] == json.loads(json.dumps(graph_data["transitions"])) and back["startNode"] == graph_data["startNode"]
    assert back["skeleton"] == json.loads(json.dumps(graph_data["skeleton"]))
    assert list(back["subgraphs"].keys()) == list(graph_data["subgraphs"].keys())
    for action, entry in graph_data["subgraphs"].items():
        assert back["subgraphs"][action]["info"] == entry["info"]
        assert list(back["subgraphs"][action]["nodes"].keys()) == list(entry["nodes"].keys())
        for mp, nd in entry["nodes"].items():
            assert back["subgraphs"][action]["nodes"][mp]["mm"] == json.loads(json.dumps(nd["mm"]))
            assert back["subgraphs"][action]["nodes"][mp]["name"] == "%s_%s" % (action, mp)     # zip_io.py:186 naming
    # the loader takes the reference's path-without-suffix form (motion_state_graph_loader.py:184-185)
    ld = MotionStateGraphLoader()
    ld.set_data_source(path[:-4])
    g = ld.build()
    assert len(g.nodes) == 14


def test_loader_builds_the_graph(graph_data):
    ld = MotionStateGraphLoader()
    ld.set_data_source(io.BytesIO(graph_to_bytes(graph_data)))
    g = ld.build()
    assert g.start_node == ("walk", "beginLeftStance")
    assert g.nodes[("walk", "beginLeftStance")].node_type == "start" and g.nodes[("walk", "leftStance")].node_type == "standard"
    assert g.nodes[("walk", "endRightStance")].node_type == "end" and g.nodes[("pickRight", "reach")].node_type == "single_primitive"
    assert g.node_groups["walk"].start_states == ["beginLeftStance", "beginRightStance"]
    assert g.node_groups["pickRight"].labeled_frames["reach"]["start_contact"] == 25
    assert g.node_groups["placeRight"].label_to_motion_primitive_map["end_contact"] == ["reach"]
    assert g.get_outgoing(("walk", "leftStance"), "standard") == [("walk", "rightStance")]
    assert g.get_outgoing(("walk", "leftStance"), "end") == [("walk", "endRightStance")]
    assert g.get_outgoing(("walk", "endLeftStance")) == [("pickRight", "reach")]
    assert g.nodes[("pickRight", "reach")].outgoing_edges[("carryRight", "beginLeftStance")].transition_type == "action_transition"
    # the primitive a node carries is the model that went in
    mm = graph_data["subgraphs"]["carryRight"]["nodes"]["rightStance"]["mm"]
    mp = g.nodes[("carryRight", "rightStance")].motion_primitive
    assert np.array_equal(mp.s_pca["mean_vector"], np.asarray(mm["mean_spatial_vector"])) and mp.n_canonical_frames == mm["n_canonical_frames"]


def test_step_plan_of_the_example_input(graph_data):
    ld = MotionStateGraphLoader()
    ld.set_graph_data(graph_data)
    g = ld.build()
    pi = GS.read_input(EXAMPLE_INPUT)
    assert pi["start_position"] == [530.110290527344, 0.0, -268.851318359375]           # CAD -> GL: [x, z, -y]
    assert pi["actions"][1]["position"] == [-30.0000009536743, 130.0, -200.0] and pi["actions"][1]["label"] == "start_contact"
    steps = GS.plan_steps(g, pi)
    keys = [s["node_key"] for s in steps]
    assert len(steps) == 17 and keys[0] == ("walk", "beginLeftStance") and keys[8] == ("walk", "endLeftStance")
    assert keys[9] == ("pickRight", "reach") and keys[-1] == ("placeRight", "reach") and steps[9]["keyframe"] == 25
    for a, b in zip(keys[:8], keys[1:9]):
        assert g.nodes[a].outgoing_edges.get(b) is not None                               # the plan follows graph edges
    assert np.allclose(steps[8]["goal_xz"], [30.0, -200.0]) and np.allclose(steps[15]["goal_xz"], [191.5512, -462.8008])
    prev = GS.initial_frames(g, pi, keys[0])
    specs = GS.step_specs(g, steps[0], prev)
    assert [s["class"] for s in specs] == ["GlobalTransformConstraint", "Direction2DConstraint", "PoseConstraint"]
    assert specs[0]["desc"]["position"][1] is None
    mpc = GS.build_constraints(g, specs)
    assert not mpc.is_local and len(mpc.constraints) == 3


def test_align_frames_matches_the_anim_utils_stand_in(graph_data):
    from oracle import refload
    if not refload.available():
        pytest.skip("reference modules not present")
    refload.load()
    from anim_utils.animation_data.motion_concatenation import align_quaternion_frames
    ld = MotionStateGraphLoader()
    ld.set_graph_data(graph_data)
    g = ld.build()
    ref_skel = refload.make_skeleton(graph_data["skeleton"])
    rng = np.random.RandomState(5)
    P = 79
    def frames(n):
        f = rng.normal(0, 1, (n, P))
        f[:, 3:] = (f[:, 3:].reshape(n, -1, 4) / np.linalg.norm(f[:, 3:].reshape(n, -1, 4), axis=-1, keepdims=True)).reshape(n, -1)
        return f
    new, prev = frames(7), frames(3)
    want = np.asarray(align_quaternion_frames(ref_skel, ref_skel.aligning_root_node, new.copy(), prev, None))
    got = align_frames(g.skeleton, new, prev)
    assert np.abs(got - want).max() < 1e-10
