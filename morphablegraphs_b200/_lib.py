"""ctypes binding of libmgb200.so (include/mgb200.h).

The library is built in-tree by `__graft_entry__.build()` (explicit nvcc,
sm_100a) so that the .so travels with the repository snapshot.  Loading fails
loudly: there is no fallback implementation of any entry point.
"""
import ctypes as C
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libmgb200.so")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include", "mgb200.h")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def build(force=False, verbose=False):
    """Compile csrc/*.cu into csrc/libmgb200.so for sm_100a (cross-compiles without a GPU):
    one object per translation unit (compiled in parallel, only when stale), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = sources()
    headers = glob.glob(os.path.join(CSRC, "*.cuh")) + [INCLUDE]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(CSRC, "build")
    os.makedirs(objdir, exist_ok=True)
    hdr_time = max(os.path.getmtime(p) for p in headers)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + os.environ.get("MGB_NVCC_EXTRA", "").split()   # e.g. -DMGB_TC_PROFILE (diagnostic builds)
    jobs = []
    objs = []
    for src in srcs:
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time)
        if stale:
            jobs.append([nvcc] + flags + ["-c", "-o", obj, src])

    def run(cmd):
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)

    if jobs:
        with ThreadPoolExecutor(max_workers=len(jobs)) as ex:
            list(ex.map(run, jobs))
    if jobs or not os.path.exists(LIB_PATH):
        run([nvcc] + NVCC_FLAGS + ["-o", LIB_PATH] + objs)
    return LIB_PATH


class ModelDesc(C.Structure):
    _fields_ = [("n_frames", C.c_int32), ("n_basis", C.c_int32), ("n_dim", C.c_int32), ("n_spatial", C.c_int32),
                ("n_latent", C.c_int32), ("n_gmm", C.c_int32),
                ("eigen", C.POINTER(C.c_double)), ("mean", C.POINTER(C.c_double)), ("knots", C.POINTER(C.c_double)),
                ("translation_maxima", C.POINTER(C.c_double)), ("gmm_weights", C.POINTER(C.c_double)),
                ("gmm_means", C.POINTER(C.c_double)), ("gmm_covars", C.POINTER(C.c_double)),
                ("n_nodes", C.c_int32),
                ("node_parent", C.POINTER(C.c_int32)), ("node_type", C.POINTER(C.c_int32)),
                ("node_channel", C.POINTER(C.c_int32)), ("node_offset", C.POINTER(C.c_double)),
                ("node_rotation", C.POINTER(C.c_double)),
                ("aligning_node", C.c_int32), ("aligning_dir", C.c_double * 3)]


class ConstraintDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("node", C.c_int32 * 2), ("time_slot", C.c_int32), ("time_slot2", C.c_int32),
                ("n_slots", C.c_int32), ("points_offset", C.c_int32), ("n_points", C.c_int32),
                ("has_position", C.c_int32), ("has_orientation", C.c_int32), ("has_velocity", C.c_int32),
                ("axis_mask", C.c_uint32), ("weight", C.c_double),
                ("target", C.c_double * 3), ("target2", C.c_double * 3), ("center", C.c_double * 3),
                ("orient_q", C.c_double * 4), ("offset", C.c_double * 4), ("dir", C.c_double * 2),
                ("velocity", C.c_double * 3)]


class PosePoint(C.Structure):
    _fields_ = [("node", C.c_int32), ("target", C.c_double * 3), ("weight", C.c_double)]


class AlignDesc(C.Structure):
    _fields_ = [("mode", C.c_int32), ("prev_heading_deg", C.c_double), ("prev_root", C.c_double * 3)]


class ExchangeDesc(C.Structure):
    """mgb_exchange_desc (include/mgb200.h)."""
    _fields_ = [("d_peer_slots", C.c_void_p), ("rank", C.c_int32), ("world", C.c_int32), ("seq", C.c_uint32),
                ("spin_limit", C.c_int64), ("d_err", C.c_void_p)]


class Option(C.Structure):
    """mgb_option (include/mgb200.h): one outgoing option of a planner step."""
    _fields_ = [("model", C.c_void_p), ("constraints", C.c_void_p), ("n", C.c_int64), ("index_base", C.c_int64),
                ("counts", C.POINTER(C.c_int64)), ("seed", C.c_uint64), ("err_scale", C.c_float), ("quality_scale", C.c_float),
                ("magnitude_scale", C.c_float)]


CT_POSITION, CT_RELATIVE, CT_DIR2D, CT_POSE, CT_TWO_HANDS, CT_FEET, CT_LOOK_AT, CT_CA, CT_JOINT_ROT, CT_QUAT_FRAME, CT_LOCAL_TRAJ, CT_TRAJ, CT_TRAJ_SET = range(13)
MAX_POSE_POINTS = 32
TOPK_NOALLOC_MAX_N = 16777216          # include/mgb200.h MGB_TOPK_NOALLOC_MAX_N

# every symbol include/mgb200.h declares: (restype, argtypes)
_vp, _i32, _i64, _f, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t
_pd, _pi32 = C.POINTER(C.c_double), C.POINTER(C.c_int32)
SYMBOLS = {
    "mgb_model_create": (C.c_int, [C.POINTER(ModelDesc), C.c_int, C.POINTER(_vp)]),
    "mgb_model_destroy": (None, [_vp]),
    "mgb_spline_basis": (C.c_int, [_pd, _i32, C.c_double, _pi32, _pd]),
    "mgb_constraints_pack": (C.c_int, [_vp, C.POINTER(ConstraintDesc), _i32, C.POINTER(PosePoint), _i32, _pd, _i32,
                                       _pi32, _pd, C.POINTER(AlignDesc), C.POINTER(_vp)]),
    "mgb_constraints_destroy": (None, [_vp]),
    "mgb_constraints_n_terms": (_i32, [_vp]),
    "mgb_constraints_info": (C.c_int, [_vp, _pi32, _pi32, _pi32]),
    "mgb_sample": (C.c_int, [_vp, _i64, _i64, C.POINTER(C.c_int64), C.c_uint64, _vp, _vp]),
    "mgb_score_workspace_bytes": (_sz, [_i64, _i32]),
    "mgb_score": (C.c_int, [_vp, _vp, _vp, _i32, _i64, _i64, _f, _f, _f, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mgb_sample_score": (C.c_int, [_vp, _vp, _i64, _i64, C.POINTER(C.c_int64), C.c_uint64, _f, _f, _f, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mgb_sample_score_best": (C.c_int, [_vp, _vp, _i64, _i64, C.POINTER(C.c_int64), C.c_uint64, _f, _f, _f, C.POINTER(C.c_float),
                                        C.POINTER(C.c_int64), _vp, _vp]),
    "mgb_sample_score_sharded": (C.c_int, [_vp, _vp, _i64, _i64, C.POINTER(C.c_int64), C.c_uint64, _f, _f, _f, C.POINTER(ExchangeDesc),
                                           _vp, _vp, _vp, _sz, _vp]),
    "mgb_sample_score_best_sharded": (C.c_int, [_vp, _vp, _i64, _i64, C.POINTER(C.c_int64), C.c_uint64, _f, _f, _f, C.POINTER(ExchangeDesc),
                                                C.POINTER(C.c_float), C.POINTER(C.c_int64), _vp, _vp]),
    "mgb_sample_score_multi": (C.c_int, [C.POINTER(Option), _i32, _vp, _vp, _vp, _i32, _vp]),
    "mgb_syrk_upper": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "mgb_constraints_n_residuals": (_i32, [_vp]),
    "mgb_residuals": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "mgb_lm_refine": (C.c_int, [_vp, _vp, _vp, _i32, _f, _f, _f, _i32, _vp, _vp, _vp]),
    "mgb_score_host": (C.c_int, [_vp, _vp, _vp, _i32, _i64, _i64, _f, _f, _f, _vp, _i32, _vp, _vp]),
    "mgb_backproject": (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp]),
    "mgb_eval_frames": (C.c_int, [_vp, _vp, _i64, _pd, _i32, _pi32, _pd, _vp, _vp]),
    "mgb_fk": (C.c_int, [_vp, _vp, _i64, _pi32, _i32, _vp, _vp]),
    "mgb_score_frames": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _f, _f, _f, _vp, _vp, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mgb_logpdf": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "mgb_topk": (C.c_int, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _sz, _vp]),
    "mgb_topk_noalloc": (C.c_int, [_vp, _i64, _i64, _i32, _vp, _vp, _vp]),
    "mgb_kmeans_assign_update": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp]),
    "mgb_min_sqdist": (C.c_int, [_vp, _i64, _i32, _i32, _vp, _vp, _vp, _i32, _vp]),
    "mgb_kmeanspp_pick": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _i32, _vp, _i32, _vp, _vp, _i32, _vp]),
    "mgb_kmeans_level": (C.c_int, [_vp, _i64, _i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp,
                                   C.POINTER(C.c_int32), _vp]),
    "mgb_exchange_slot_bytes": (C.c_size_t, [_i32]),
    "mgb_exchange_best": (C.c_int, [_vp, _vp, _vp, _i32, _i32, C.c_uint32, _i64, _vp, _vp, _vp, _vp]),
    "mgb_node_costs": (C.c_int, [_vp, _vp, _vp, _i32, _i64, _f, _f, _f, _vp, _vp]),
    "mgb_fp32_fma_peak": (C.c_int, [_i32, _pd]),
    "mgb_launch_count": (C.c_int64, []),
    "mgb_last_error": (C.c_char_p, []),
    "mgb_version": (C.c_char_p, []),
}

_lib = None


class MgbError(RuntimeError):
    pass


def load():
    """dlopen libmgb200.so and bind every declared symbol; raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MgbError("libmgb200.so not built (%s missing): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                       "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)     # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise MgbError("libmgb200: %s (code %d)" % (load().mgb_last_error().decode("utf-8", "replace"), rc))
