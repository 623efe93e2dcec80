"""Round-2 fixtures, CPU side: the reference-made goldens at the sizes SURVEY.md 8(d) states pin the oracle
(`s_config1_10k`: config 1's 10 000 reference-drawn samples; `l_cfg4`: config 4's constraint set on the L model;
`l_mixed_1k`), the local-frame transform a8 (`s_local`: the reference's own transform_constraints_to_local_cos
incl. quirk Q5) pins the product's host mirror, and the reference's OWN objects (loaded unmodified through
oracle/refload.py) flatten to the same descriptor tables as the mirror classes."""
import ctypes as C
import json

import numpy as np
import pytest

from oracle import np_oracle as O
from tests.common import Case, rel_err, describe_constraint, assert_same_description
from morphablegraphs_b200 import engine, _lib
from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import objective_functions as OF

BIG = ["s_config1_10k", "l_cfg4", "l_mixed_1k"]


@pytest.mark.parametrize("name", BIG)
def test_oracle_against_reference_costs_at_size(name):
    c = Case(name)
    g = c.g
    n = len(c.samples)
    assert n >= 1000
    cost, terms = O.candidate_costs(c.node, c.skeleton, c.constraints, c.samples, None, return_terms=True)
    assert rel_err(terms, g["terms"]).max() < 1e-8
    assert rel_err(cost, g["costs"]).max() < 1e-8
    idx, err = O.first_argmin(cost)
    assert idx == int(g["best_idx"]) and abs(err - float(g["min_error"])) < 1e-8 * max(1.0, abs(err))
    ma = O.ModelArrays(c.node)
    assert rel_err(O.gmm_logpdf(ma, c.samples), g["logpdf"]).max() < 1e-9
    c_nat = O.candidate_costs(c.node, c.skeleton, c.constraints, c.samples, None, error_scale=1.0, quality_scale=0.1)
    assert rel_err(c_nat, g["naturalness_cost"]).max() < 1e-8
    nf = int(g["n_full"])
    coeffs = O.back_project_coeffs(ma, c.samples[:nf])
    assert rel_err(coeffs, g["coeffs"]).max() < 1e-9
    assert rel_err(O.eval_frames(ma, coeffs, g["key_times"]), g["key_frames"]).max() < 1e-9


def test_config4_constraint_set_is_the_surveys():
    """right-hand position at F-1, root position + 2-D direction at F-1, 5-joint PoseConstraint
    at keyframes 0 / 1."""
    c = Case("l_cfg4")
    F = syn.SHAPES["L"]["n_frames"]
    kinds = [(s["class"], s["desc"].get("joint"), s["desc"]["canonical_keyframe"]) for s in c.specs]
    assert kinds == [("GlobalTransformConstraint", "RightHand", F - 1), ("GlobalTransformConstraint", "Hips", F - 1),
                     ("Direction2DConstraint", None, F - 1), ("PoseConstraint", None, 0)]
    pose = c.specs[3]["desc"]
    assert len(pose["node_names"]) == 5 and pose["velocity_constraint"] is not None


# ---------------------------------------------------------------------------
# a8: transform_constraints_to_local_cos
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def local_case():
    c = Case("s_local")
    c.constraints.aligning_transform = np.array(c.g["aligning_transform"])
    c.constraints.use_local_optimization = True
    return c


def test_local_transform_matches_reference(local_case):
    c = local_case
    ref = json.loads(str(c.g["local_desc_json"]))
    assert not c.constraints.is_local
    local = c.constraints.transform_constraints_to_local_cos()
    assert local is not c.constraints and local.is_local and local.skeleton is c.skeleton
    assert local.use_local_optimization == bool(c.g["local_use_local_optimization"])
    assert local.start_pose == {"orientation": [0, 0, 0], "position": [0, 0, 0]}
    mine = [describe_constraint(x) for x in local.constraints]
    assert len(mine) == len(ref)
    for a, b in zip(mine, ref):
        assert_same_description(a, b)
    # quirk Q5, spelled out: position / relative / CA come back with weight 1.0 and no orientation; the others keep theirs
    by_type = {d["type"]: d for d in mine}
    src = {s_["class"]: s_["weight"] for s_ in c.specs}
    assert by_type["keyframe_relative_position"]["weight_factor"] == 1.0 and src["RelativeTransformConstraint"] != 1.0
    assert by_type["ca_constraint"]["weight_factor"] == 1.0 and src["GlobalTransformCAConstraint"] != 1.0
    assert all(d["weight_factor"] == 1.0 and d["orientation"] is None for d in mine if d["type"] == "keyframe_position")
    for cls, typ in (("Direction2DConstraint", "keyframe_2d_direction"), ("PoseConstraint", "keyframe_pose"),
                     ("TwoHandConstraintSet", "keyframe_two_hands"), ("FeetConstraint", "keyframe_feet"),
                     ("LookAtConstraint", "keyframe_look_at")):
        assert by_type[typ]["weight_factor"] == src[cls] != 1.0
    # already local / no transform: returned unchanged (motion_primitive_constraints.py:269-270)
    assert local.transform_constraints_to_local_cos() is local


def test_oracle_scores_the_local_set_like_the_reference(local_case):
    c = local_case
    local = c.constraints.transform_constraints_to_local_cos()
    cost, terms = O.candidate_costs(c.node, c.skeleton, local, c.samples, None, return_terms=True)
    assert rel_err(terms, c.g["terms"]).max() < 1e-8
    assert rel_err(cost, c.g["costs"]).max() < 1e-8
    assert O.first_argmin(cost)[0] == int(c.g["best_idx"])


# ---------------------------------------------------------------------------
# "accepts the reference's own objects": reference classes -> the same flat tables as the mirror classes
# ---------------------------------------------------------------------------
class _StubModel(object):
    def __init__(self, skeleton, n_frames=None):
        sk = engine.pack_skeleton(skeleton)
        self.F = n_frames
        self.node_index = sk["index"]
        self.channel_node = {int(ch): i for i, ch in enumerate(sk["channel"]) if ch >= 0}
        self.skeleton = skeleton


def _flat_bytes(flat):
    descs, n, parr, n_points, times = flat
    return (n, n_points, bytes(descs)[:n * C.sizeof(_lib.ConstraintDesc)], bytes(parr)[:n_points * C.sizeof(_lib.PosePoint)],
            times.tobytes())


def _struct_values(st):
    out = []
    for fname, ftype in st._fields_:
        v = getattr(st, fname)
        out.extend(list(v) if hasattr(v, "__len__") else [v])
    return np.asarray(out, dtype=np.float64)


def _flat_close(fa, fb, tol):
    """Field by field, to `tol`: a target rotation reaches the table as a unit quaternion derived from the constraint's
    rotation MATRIX, which the reference builds with transformations.euler_matrix and the mirror with its own product of
    axis rotations (last-bit differences)."""
    (da, na, pa, npa, ta), (db, nb, pb, npb, tb) = fa, fb
    assert na == nb and npa == npb and np.array_equal(ta, tb)
    for i in range(na):
        assert np.allclose(_struct_values(da[i]), _struct_values(db[i]), rtol=0.0, atol=tol), i
    for i in range(npa):
        assert np.allclose(_struct_values(pa[i]), _struct_values(pb[i]), rtol=0.0, atol=tol), i


@pytest.mark.parametrize("name", ["s_mixed", "s_local", "l_cfg4", "s_rot", "s_traj"])
def test_reference_objects_flatten_like_the_mirror(name):
    from oracle import refload, make_golden
    if not refload.available():
        pytest.skip("reference modules not present (neither /root/reference nor oracle/_ref)")
    ns = refload.load()
    c = Case(name)
    ref_skel = refload.make_skeleton(c.skeleton_desc)                   # anim_utils stand-in Skeleton
    ref_node = refload.make_graph_node(ns, c.model_json)                # the reference's MotionStateGraphNode
    ref_mpc = make_golden.build_mp_constraints(ns, ref_skel, c.specs, is_local=bool(c.g["is_local"]))   # reference classes
    assert type(ref_mpc).__module__.startswith("morphablegraphs.") and type(ref_mpc.constraints[0]).__module__.startswith("morphablegraphs.")
    if name == "s_local":
        ref_mpc.aligning_transform = np.array(c.g["aligning_transform"])
        c.constraints.aligning_transform = np.array(c.g["aligning_transform"])
        ref_mpc, mine = ref_mpc.transform_constraints_to_local_cos(), c.constraints.transform_constraints_to_local_cos()
    else:
        mine = c.constraints
    # skeleton packing reads the same fields of either skeleton class
    a, b = engine.pack_skeleton(ref_skel), engine.pack_skeleton(c.skeleton)
    assert a["names"] == b["names"] and a["aligning_node"] == b["aligning_node"]
    for k in ("parent", "type", "channel", "offset", "rotation"):
        assert np.array_equal(a[k], b[k]), k
    # constraint tables
    fa = engine.flatten_constraints(_StubModel(ref_skel, c.F), ref_mpc)
    fb = engine.flatten_constraints(_StubModel(c.skeleton, c.F), mine)
    if name == "s_rot":
        _flat_close(fa, fb, 1e-14)
    elif name == "s_traj":
        # the trajectory tables are queried from the reference's ParameterizedSpline on one side and from the mirror's own
        # Catmull-Rom restatement on the other
        _flat_close(fa, fb, 1e-10)
    else:
        assert _flat_bytes(fa) == _flat_bytes(fb)
        # and therefore the same cache key content
        ka = OF.constraints_key(None, fa, ref_mpc, None)
        kb = OF.constraints_key(None, fb, mine, None)
        assert ka[1:] == kb[1:]
    # model arrays the packer reads (DeviceModel.__init__): identical numbers from the reference's model object
    mp_ref, mp_mine = ref_node.motion_primitive, c.node.motion_primitive
    for k in ("eigen_vectors", "mean_vector", "knots"):
        assert np.array_equal(np.asarray(mp_ref.s_pca[k]), np.asarray(mp_mine.s_pca[k])), k
    assert np.array_equal(mp_ref.translation_maxima, mp_mine.translation_maxima)
    gr, gm = mp_ref.gaussian_mixture_model, mp_mine.gaussian_mixture_model
    assert np.array_equal(gr.weights_, gm.weights_) and np.array_equal(gr.means_, gm.means_) and np.array_equal(gr.covariances_, gm.covariances_)


# ---------------------------------------------------------------------------
# f1: residual-vector objective, pinned by the reference's own functions (tests/golden/s_resid.npz)
# ---------------------------------------------------------------------------
def test_oracle_residual_vectors_match_the_reference():
    c = Case("s_resid")
    r = O.residual_vectors(c.node, c.skeleton, c.constraints, c.samples)
    assert r.shape == c.g["residual_vectors"].shape
    assert rel_err(r, c.g["residual_vectors"]).max() < 1e-9
    obj = O.residual_objective(c.node, c.skeleton, c.constraints, c.samples, quality_scale=0.1)
    assert rel_err(obj, c.g["objective_q01"]).max() < 1e-9
    # the sum of the residual vector is the scalar objective for every class but the pose term (mean vs per point) ...
    cost, terms = O.candidate_costs(c.node, c.skeleton, c.constraints, c.samples, return_terms=True)
    keep = [i for i, k in enumerate(c.constraints.constraints) if k.constraint_type != "keyframe_pose"]
    mpc2 = type(c.constraints)()
    mpc2.skeleton, mpc2.is_local = c.skeleton, True
    mpc2.constraints = [c.constraints.constraints[i] for i in keep]
    assert rel_err(O.residual_vectors(c.node, c.skeleton, mpc2, c.samples).sum(axis=1),
                   sum(c.constraints.constraints[i].weight_factor * terms[i] for i in keep)).max() < 1e-9


def test_quirk_q8_reference_closest_point_needs_a_scalar_parameter():
    """as written, ParameterizedSpline.find_closest_point_fast (parameterized_spline.py:303-322) hands the
    Catmull-Rom spline scipy's 1-element array; numpy >= 1.24 refuses the ragged list built from it (catmull_rom_spline.py:141-165).
    With the parameter handed over as a scalar (oracle/refload.py:scalar_spline_parameter) the reference's own class runs and finds
    the closest point of the curve."""
    import warnings
    from oracle import refload
    if not refload.available():
        pytest.skip("reference modules not present (neither /root/reference nor oracle/_ref)")
    ns = refload.load()
    cps = [np.array(p) for p in ([0.0, 0.0, 0.0], [5.0, 2.0, 15.0], [12.0, 0.0, 30.0], [10.0, 3.0, 45.0], [0.0, 0.0, 60.0])]
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        spline = ns.parameterized_spline.ParameterizedSpline(cps)
    p = np.asarray(spline.query_point_by_parameter(0.4)) + np.array([0.3, -0.2, 0.1])
    if tuple(int(x) for x in np.__version__.split(".")[:2]) >= (1, 24):
        with pytest.raises(ValueError), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            spline.find_closest_point_fast(p, 0.1)
    with refload.scalar_spline_parameter(ns), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        target, u = spline.find_closest_point_fast(p, 0.1)
    assert abs(float(np.ravel(u)[0]) - 0.4) < 0.02 and np.linalg.norm(np.asarray(target) - p) < 0.4
    # and the mirror's own Catmull-Rom restatement is the same curve
    from morphablegraphs_b200.constraints import CatmullRomTrajectory
    mine = CatmullRomTrajectory(cps)
    assert abs(mine.full_arc_length - spline.full_arc_length) < 1e-9
    for x in np.linspace(0.0, 1.0, 41):
        assert np.abs(mine.query_point_by_parameter(x) - np.asarray(spline.query_point_by_parameter(x))).max() < 1e-9
    for a_len in np.linspace(0.0, 1.1 * spline.full_arc_length, 37):
        assert np.abs(mine.query_point_by_absolute_arc_length(a_len) - np.asarray(spline.query_point_by_absolute_arc_length(a_len))).max() < 1e-9


@pytest.mark.parametrize("name", ["s_traj", "s_traj_nonlocal"])
def test_oracle_trajectory_residuals_match_the_reference(name):
    """f2: TrajectoryConstraint / TrajectorySetConstraint, one residual per frame of the motion vector, made by the reference's own
    classes (quirk Q8: with the spline parameter handed over as a scalar, oracle/refload.py:scalar_spline_parameter)."""
    import warnings
    c = Case(name)
    want = c.g["residual_vectors"]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = O.residual_vectors(c.node, c.skeleton, c.constraints, c.samples[:len(want)], c.prev_frames)
    assert r.shape == want.shape == (len(want), 2 * c.F + 1)
    assert rel_err(r, want).max() < 1e-9


# ---------------------------------------------------------------------------
# f2: time-warp back-projection (host part), pinned by the reference's own functions (tests/golden/s_timewarp.npz)
# ---------------------------------------------------------------------------
def test_time_warp_functions_match_the_reference():
    from tests.common import load_golden
    from morphablegraphs_b200 import MotionStateGraphNode, synthetic as syn
    g = load_golden("s_timewarp")
    node = MotionStateGraphNode()
    node._initialize_from_json(None, syn.make_model_json(str(g["shape"])))
    mp = node.motion_primitive
    K = mp.s_pca["n_components"]
    for i, s in enumerate(g["samples"]):
        can = mp._back_transform_gamma_to_canonical_time_function(s[K:])
        assert rel_err(can, g["tw%d_canonical" % i]).max() < 1e-12
        tf = mp.back_project_time_function(s[K:])
        assert tf.shape == g["tw%d_time" % i].shape            # the warp changes the NUMBER of frames
        assert rel_err(tf, g["tw%d_time" % i]).max() < 1e-10
        # the oracle's spline evaluation at the warped times gives the reference's motion
        ma = O.ModelArrays(node)
        fr = O.eval_frames(ma, O.back_project_coeffs(ma, s[None]), tf)[0]
        assert rel_err(fr, g["tw%d_frames" % i]).max() < 1e-9
