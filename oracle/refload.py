"""Load the reference's own hot-path modules, unmodified, from /root/reference.

TEST INFRASTRUCTURE ONLY, and only usable in the build container:
/root/reference does not exist on the GPU box, so nothing under `-m gpu`,
`smoke()` or `bench.py` may import this module.  It is used by
`oracle/make_golden.py` (fixture generation) and by the `not gpu` tests that
pin `oracle/np_oracle.py` against the live reference when it is present.

The reference cannot be imported as a package:
`morphablegraphs/__init__.py:3` pulls the whole generator, which needs
`anim_utils`, `transformations`, `motion_db_interface`.  So the package
`__init__`s that drag those in are replaced by empty namespace stubs carrying
only the constants the hot-path modules read; every *module* on the path is
then executed from its original file.  `oracle/standin/` supplies
`anim_utils` / `transformations`.
"""
import importlib
import os
import sys
import types

_PKG = "morphablegraphs"
_HERE = os.path.dirname(os.path.abspath(__file__))
_STANDIN = os.path.join(_HERE, "standin")
# The reference tree where it lies (build container), else the staged copy of its hot-path modules that
# `stage_reference()` (called by __graft_entry__.build()) puts under the git-ignored oracle/_ref/ so that the
# reference's OWN loop can be timed on the GPU box, where /root/reference does not exist.
SOURCE_ROOT = "/root/reference"
STAGED_ROOT = os.path.join(_HERE, "_ref")
REFERENCE_ROOT = SOURCE_ROOT if os.path.isdir(os.path.join(SOURCE_ROOT, _PKG)) else STAGED_ROOT


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, _PKG))


def stage_reference():
    """Copy the reference modules this loader executes (and nothing else) from /root/reference into
    oracle/_ref/, unmodified, keeping their relative paths.  oracle/_ref/ is listed in .gitignore (never
    committed) but not in .gpurunignore (it travels to the GPU box like a built .so).  Returns the number
    of files staged; a no-op when /root/reference is absent."""
    import shutil
    if not os.path.isdir(os.path.join(SOURCE_ROOT, _PKG)) or REFERENCE_ROOT != SOURCE_ROOT:
        return 0
    load()
    n = 0
    for name, mod in list(sys.modules.items()):
        f = getattr(mod, "__file__", None)
        if not (name == _PKG or name.startswith(_PKG + ".")) or not f or not f.startswith(SOURCE_ROOT + os.sep):
            continue
        dst = os.path.join(STAGED_ROOT, os.path.relpath(f, SOURCE_ROOT))
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(f, dst)
        n += 1
    with open(os.path.join(STAGED_ROOT, "README"), "w") as fh:
        fh.write("Unmodified copies of %d reference modules (dfki-asr/morphablegraphs) staged by oracle/refload.py:stage_reference()\n"
                 "for the CPU baseline on the GPU box.  Git-ignored; never edit, never commit.\n" % n)
    return n


def _stub(name, relpath, **attrs):
    mod = types.ModuleType(name)
    mod.__path__ = [os.path.join(REFERENCE_ROOT, relpath)]
    mod.__package__ = name
    for k, v in attrs.items():
        setattr(mod, k, v)
    sys.modules[name] = mod
    return mod


_loaded = False


def load():
    """Install the stubs; afterwards `importlib.import_module("morphablegraphs.<...>")`
    executes reference files from where they lie.  Returns a namespace of the
    modules the oracle uses."""
    global _loaded
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    if _STANDIN not in sys.path:
        sys.path.insert(0, _STANDIN)
    if not _loaded:
        # scipy>=1.8 dropped the private module path the reference imports at
        # objective_functions.py:30; alias it to the public one.
        try:
            importlib.import_module("scipy.optimize.optimize")
        except Exception:  # pragma: no cover
            import scipy.optimize as so
            shim = types.ModuleType("scipy.optimize.optimize")
            shim.approx_fprime = so.approx_fprime
            sys.modules["scipy.optimize.optimize"] = shim
        # numpy>=2 removed np.alltrue, which feature_cluster_tree.py:92 calls
        import numpy as _np
        if not hasattr(_np, "alltrue"):
            _np.alltrue = _np.all
        # plotting / DTW packages the spline helpers import at module level
        # (reference splines/b_spline.py:38) but never call on this path
        for missing in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d", "fastdtw"):
            try:
                importlib.import_module(missing)
            except Exception:
                m = types.ModuleType(missing)
                m.__path__ = []
                m.Axes3D = object
                sys.modules[missing] = m
        _stub(_PKG, _PKG)
        # constants: reference motion_model/__init__.py:1-13
        _stub(_PKG + ".motion_model", _PKG + "/motion_model",
              NODE_TYPE_START="start", NODE_TYPE_STANDARD="standard", NODE_TYPE_END="end",
              NODE_TYPE_IDLE="idle", NODE_TYPE_SINGLE="single_primitive",
              NODE_TYPE_CYCLE_END="cycle_end", B_SPLINE_DEGREE=3,
              ELEMENTARY_ACTION_DIRECTORY_NAME="elementary_action_models",
              TRANSITION_MODEL_DIRECTORY_NAME="transition_models",
              META_INFORMATION_FILE_NAME="meta_information.json",
              TRANSITION_DEFINITION_FILE_NAME="graph_definition.json",
              TRANSITION_MODEL_FILE_ENDING=".GPM")
        # constants: reference constraints/__init__.py
        _stub(_PKG + ".constraints", _PKG + "/constraints",
              OPTIMIZATION_MODE_ALL="all", OPTIMIZATION_MODE_KEYFRAMES="keyframes",
              OPTIMIZATION_MODE_TWO_HANDS="two_hands", OPTIMIZATION_MODE_NONE="none",
              CA_CONSTRAINTS_MODE_NONE="none", CA_CONSTRAINTS_MODE_INDIVIDUAL="individual",
              CA_CONSTRAINTS_MODE_SET="create_constraint_set",
              CA_CONSTRAINTS_MODE_DIRECT_CONNECTION="direct_connection")
        _stub(_PKG + ".motion_generator", _PKG + "/motion_generator")
        _stub(_PKG + ".construction", _PKG + "/construction")
        _loaded = True
    ns = types.SimpleNamespace()
    imp = importlib.import_module
    ns.motion_primitive = imp(_PKG + ".motion_model.motion_primitive")
    ns.motion_spline = imp(_PKG + ".motion_model.motion_spline")
    ns.motion_primitive_wrapper = imp(_PKG + ".motion_model.motion_primitive_wrapper")
    ns.motion_state_graph_node = imp(_PKG + ".motion_model.motion_state_graph_node")
    ns.spatial_constraints = imp(_PKG + ".constraints.spatial_constraints")
    ns.ca_constraint = imp(_PKG + ".constraints.spatial_constraints.keyframe_constraints.global_transform_ca_constraint")
    ns.joint_rotation_constraint = imp(_PKG + ".constraints.spatial_constraints.keyframe_constraints.joint_rotation_constraint")
    ns.pose_constraint_quat_frame = imp(_PKG + ".constraints.spatial_constraints.keyframe_constraints.pose_constraint_quat_frame")
    ns.local_trajectory_constraint = imp(_PKG + ".constraints.spatial_constraints.keyframe_constraints.local_trajectory_constraint")
    ns.trajectory_constraint = imp(_PKG + ".constraints.spatial_constraints.trajectory_constraint")
    ns.trajectory_set_constraint = imp(_PKG + ".constraints.spatial_constraints.trajectory_set_constraint")
    ns.catmull_rom_spline = imp(_PKG + ".constraints.spatial_constraints.splines.catmull_rom_spline")
    ns.parameterized_spline = imp(_PKG + ".constraints.spatial_constraints.splines.parameterized_spline")
    ns.motion_primitive_constraints = imp(_PKG + ".constraints.motion_primitive_constraints")
    ns.objective_functions = imp(_PKG + ".motion_generator.optimization.objective_functions")
    ns.motion_primitive_generator = imp(_PKG + ".motion_generator.motion_primitive_generator")
    ns.algorithm_configuration = imp(_PKG + ".motion_generator.algorithm_configuration")
    ns.space_partitioning = imp(_PKG + ".space_partitioning")
    ns.cluster_tree = imp(_PKG + ".space_partitioning.cluster_tree")
    ns.feature_cluster_tree = imp(_PKG + ".space_partitioning.feature_cluster_tree")
    ns.kdtree = imp(_PKG + ".space_partitioning.kdtree")
    ns.features = imp(_PKG + ".space_partitioning.features")
    try:  # needs anim_utils modules far off the path; only used to pin the knot formula
        ns.construction_utils = imp(_PKG + ".construction.utils")
    except Exception:
        ns.construction_utils = None
    return ns


def scalar_spline_parameter(ns):
    """Context manager for quirk Q8.  scipy's minimize hands `ParameterizedSpline.find_closest_point_fast`'s objective
    (parameterized_spline.py:304-322) a 1-element ARRAY u; `CatmullRomSpline.query_point_by_parameter`
    (catmull_rom_spline.py:132-153) then builds the ragged list [u**3, u**2, u, 1] and hands it to np.dot, which raises since
    numpy 1.24 (and gave a (3, 1) point before: a 3 x 3 broadcast against the joint position).  While this context is active
    the spline receives u as the Python float it stands for, so the reference's own classes run and compute the distance
    the code states.  The class attribute is wrapped and restored; the reference's source is not touched."""
    import contextlib

    @contextlib.contextmanager
    def cm():
        import numpy
        cls = ns.catmull_rom_spline.CatmullRomSpline
        orig = cls.query_point_by_parameter

        def query_point_by_parameter(self, u):
            return orig(self, float(numpy.ravel(u)[0]))
        cls.query_point_by_parameter = query_point_by_parameter
        try:
            yield
        finally:
            cls.query_point_by_parameter = orig
    return cm()


def make_wrapper(ns, model_json):
    """reference MotionPrimitiveModelWrapper over the legacy model
    (`motion_primitive_wrapper.py:63-75`; mgrd absent -> legacy path)."""
    w = ns.motion_primitive_wrapper.MotionPrimitiveModelWrapper()
    w._initialize_from_json(None, model_json)
    return w


def make_graph_node(ns, model_json, name="node", action="walk"):
    """reference MotionStateGraphNode (wrapper + cluster_tree slot)."""
    node = ns.motion_state_graph_node.MotionStateGraphNode(None)
    node._initialize_from_json(None, model_json)
    node.name = name
    node.action_name = action
    return node


def make_skeleton(desc):
    if _STANDIN not in sys.path:
        sys.path.insert(0, _STANDIN)
    from anim_utils.animation_data.skeleton import Skeleton
    return Skeleton.from_desc(desc)
