"""Thin Python layer over libmgb200.so: packs reference-shaped objects into the
C-ABI descriptors and launches the kernels on torch-owned device memory.

Duck-typed on the reference's attribute names, so `DeviceModel` /
`DeviceConstraints` accept either the reference's own objects
(`MotionPrimitiveModelWrapper`, anim_utils `Skeleton`,
`MotionPrimitiveConstraints`) or this package's mirrors.

PyTorch is plumbing only: device allocations, streams, `torch.distributed`.
All arithmetic of the path runs in the CUDA library; nothing here falls back
to the CPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import MgbError  # noqa: F401


class UnsupportedConstraintError(MgbError):
    """A constraint outside the device path (trajectory constraints that query a host
    spline per frame, quaternion-frame pose constraints, start-pose alignment with its
    call-order dependent mutation - SURVEY.md Q2).  Callers decide what to do; the engine
    never drops a term silently and never evaluates it on the CPU itself."""


def _torch():
    import torch
    return torch


def _motion_primitive(model):
    return getattr(model, "motion_primitive", model)


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _iptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


# ---------------------------------------------------------------------------
# skeleton packing
# ---------------------------------------------------------------------------
def _node_type_int(node):
    t = node.node_type
    if isinstance(t, str):
        return {"root": 0, "joint": 1, "end_site": 2}[t]
    return int(t)


def pack_skeleton(skeleton):
    """Topologically ordered flat arrays + name -> index map.
    Fields read: reference utilities/__init__.py:13-16 (node_name, offset, rotation,
    fixed, children), quaternion_frame_index (mg_state_planner.py:288)."""
    if skeleton is None:
        # root-only skeleton: enough for sampling / back-projection / spline evaluation
        names = ["root"]
        return dict(names=names, index={"root": 0}, parent=np.array([-1], dtype=np.int32),
                    type=np.array([0], dtype=np.int32), channel=np.array([0], dtype=np.int32),
                    offset=np.zeros((1, 3)), rotation=np.array([[1.0, 0, 0, 0]]), aligning_node=0,
                    aligning_dir=[0.0, 0.0, 1.0])
    root_name = skeleton.root if isinstance(skeleton.root, str) else skeleton.root.node_name
    order = []
    stack = [skeleton.nodes[root_name]]
    while stack:                      # pre-order: parents precede children
        node = stack.pop()
        order.append(node)
        stack.extend(reversed(list(node.children)))
    index = {n.node_name: i for i, n in enumerate(order)}
    n = len(order)
    parent = np.full(n, -1, dtype=np.int32)
    ntype = np.zeros(n, dtype=np.int32)
    channel = np.full(n, -1, dtype=np.int32)
    offset = np.zeros((n, 3))
    rotation = np.zeros((n, 4))
    rotation[:, 0] = 1.0
    for i, node in enumerate(order):
        if node.parent is not None:
            parent[i] = index[node.parent.node_name]
        ntype[i] = _node_type_int(node)
        if ntype[i] == 0:
            channel[i] = 0
        elif ntype[i] == 1 and not node.fixed:
            channel[i] = int(node.quaternion_frame_index)
        offset[i] = np.asarray(node.offset, dtype=np.float64)
        if node.rotation is not None:
            rotation[i] = np.asarray(node.rotation, dtype=np.float64)
    aligning = getattr(skeleton, "aligning_root_node", None) or root_name
    return dict(names=[n.node_name for n in order], index=index, parent=parent, type=ntype, channel=channel,
                offset=offset, rotation=rotation, aligning_node=index[aligning],
                aligning_dir=list(getattr(skeleton, "aligning_root_dir", [0.0, 0.0, 1.0])))


# ---------------------------------------------------------------------------
# spline basis: the same FITPACK call the reference makes (motion_spline.py:89-92)
# ---------------------------------------------------------------------------
def spline_basis(knots, t, degree=3):
    """(span, w[4]) with value = sum_i w[i] * coeffs[span+i], from scipy.splev on identity
    coefficients; falls back to the library's own FITPACK-compatible recurrence."""
    knots = np.ascontiguousarray(knots, dtype=np.float64)
    nb = len(knots) - degree - 1
    nk1 = nb
    l = degree
    while l + 1 < nk1 and t >= knots[l + 1]:
        l += 1
    span = l - degree
    try:
        import scipy.interpolate as si
        eye = np.eye(nb)
        w = np.array([float(si.splev(t, (knots, eye[span + i], degree))) for i in range(degree + 1)])
    except ImportError:  # pragma: no cover
        lib = _lib.load()
        sp = C.c_int32()
        w = np.zeros(4)
        _lib.check(lib.mgb_spline_basis(_dptr(knots), len(knots), float(t), C.byref(sp), _dptr(w)))
        span = sp.value
    return span, w


def lib_spline_basis(knots, t):
    """The library's own basis (host code in csrc/mgb_api.cu); no GPU needed."""
    lib = _lib.load()
    knots = np.ascontiguousarray(knots, dtype=np.float64)
    sp = C.c_int32()
    w = np.zeros(4)
    _lib.check(lib.mgb_spline_basis(_dptr(knots), len(knots), float(t), C.byref(sp), _dptr(w)))
    return sp.value, w


# ---------------------------------------------------------------------------
# model
# ---------------------------------------------------------------------------
_BASIS_CACHE = {}


class DeviceModel(object):
    """One motion primitive + skeleton resident on one GPU (mgb_model handle)."""

    def __init__(self, model, skeleton=None, device=None):
        torch = _torch()
        if not torch.cuda.is_available():
            raise MgbError("no CUDA device: morphablegraphs_b200 has no CPU fallback")
        self.lib = _lib.load()
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(device)
        self.torch_device = torch.device("cuda", self.device)
        mp = _motion_primitive(model)
        s = mp.s_pca
        # reference stores eigen vectors transposed at load (motion_primitive.py:158): [Nb*P, K]
        E = np.ascontiguousarray(np.asarray(s["eigen_vectors"], dtype=np.float64).T)
        mean = np.ascontiguousarray(s["mean_vector"], dtype=np.float64)
        self.knots = np.ascontiguousarray(s["knots"], dtype=np.float64)
        self.Nb, self.P, self.K = int(s["n_basis"]), int(s["n_dim"]), int(s["n_components"])
        self.F = int(mp.n_canonical_frames)
        tm = np.ascontiguousarray(mp.translation_maxima, dtype=np.float64)
        gmm = mp.gaussian_mixture_model
        self.gmm_weights = np.ascontiguousarray(gmm.weights_, dtype=np.float64)
        means = np.ascontiguousarray(gmm.means_, dtype=np.float64)
        covars = np.ascontiguousarray(gmm.covariances_, dtype=np.float64)
        self.G, self.D = means.shape
        assert E.shape == (self.K, self.Nb * self.P), E.shape
        self.skeleton = skeleton
        sk = pack_skeleton(skeleton)
        self.node_index = sk["index"]
        self.node_names = sk["names"]
        self.channel_node = {int(ch): i for i, ch in enumerate(sk["channel"]) if ch >= 0}
        offset = np.ascontiguousarray(sk["offset"], dtype=np.float64)
        rotation = np.ascontiguousarray(sk["rotation"], dtype=np.float64)
        d = _lib.ModelDesc()
        d.n_frames, d.n_basis, d.n_dim, d.n_spatial, d.n_latent, d.n_gmm = self.F, self.Nb, self.P, self.K, self.D, self.G
        d.eigen, d.mean, d.knots, d.translation_maxima = _dptr(E), _dptr(mean), _dptr(self.knots), _dptr(tm)
        d.gmm_weights, d.gmm_means, d.gmm_covars = _dptr(self.gmm_weights), _dptr(means), _dptr(covars)
        d.n_nodes = len(sk["names"])
        d.node_parent, d.node_type, d.node_channel = _iptr(sk["parent"]), _iptr(sk["type"]), _iptr(sk["channel"])
        d.node_offset, d.node_rotation = _dptr(offset), _dptr(rotation)
        d.aligning_node = sk["aligning_node"]
        d.aligning_dir = (C.c_double * 3)(*sk["aligning_dir"])
        h = C.c_void_p()
        _lib.check(self.lib.mgb_model_create(C.byref(d), self.device, C.byref(h)))
        self.handle = h
        self._ws = None

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.mgb_model_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    # -- helpers ---------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(_torch().cuda.current_stream(self.torch_device).cuda_stream)

    def workspace(self):
        torch = _torch()
        nbytes = int(self.lib.mgb_score_workspace_bytes(0, 64))      # large enough for every k (top-k keeps its partials in it)
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.zeros(nbytes, dtype=torch.uint8, device=self.torch_device)
        return self._ws

    def _f32(self, t):
        torch = _torch()
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(np.ascontiguousarray(t, dtype=np.float32), device=self.torch_device)
        if t.device != self.torch_device or t.dtype != torch.float32 or not t.is_contiguous():
            t = t.to(device=self.torch_device, dtype=torch.float32).contiguous()
        return t

    # -- K1 ----------------------------------------------------------------------
    def component_counts(self, n_total, seed):
        """Multinomial split over components, drawn on the host like sklearn's sampler does
        (GaussianMixture.sample: rng.multinomial(n, weights_))."""
        if seed is None:
            return np.random.RandomState(None).multinomial(int(n_total), self.gmm_weights).astype(np.int64)
        # re-seeding a kept RandomState gives the stream of RandomState(seed) at 1 % of the cost of constructing one
        # (the constructor alone is ~0.3 ms: four times the 1M-candidate kernel of the S primitive)
        rng = self.__dict__.get("_count_rng")
        if rng is None:
            rng = self.__dict__["_count_rng"] = np.random.RandomState(0)
        rng.seed(int(seed) & 0x7FFFFFFF)
        return rng.multinomial(int(n_total), self.gmm_weights).astype(np.int64)

    def sample(self, n, seed=0, index_base=0, counts=None, n_total=None, out=None):
        """[n, D] float32 device tensor: global indices [index_base, index_base+n) of a job of
        n_total samples stacked in component order (motion_primitive.py:182-189)."""
        torch = _torch()
        n_total = int(n if n_total is None else n_total)
        if counts is None:
            counts = self.component_counts(n_total, seed)
        counts = np.ascontiguousarray(counts, dtype=np.int64)
        if out is None:
            out = torch.empty((int(n), self.D), dtype=torch.float32, device=self.torch_device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_sample(self.handle, int(n), int(index_base), counts.ctypes.data_as(C.POINTER(C.c_int64)),
                                           C.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF), C.c_void_p(out.data_ptr()), self._stream()))
        return out

    # -- K2 / K3 / K4 / K5b ---------------------------------------------------
    def backproject(self, s, impl=0, out=None):
        torch = _torch()
        s = self._f32(s)
        n = s.shape[0]
        if out is None:
            out = torch.empty((n, self.Nb, self.P), dtype=torch.float32, device=self.torch_device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_backproject(self.handle, C.c_void_p(s.data_ptr()), n, C.c_void_p(out.data_ptr()), int(impl),
                                                self._stream()))
        return out

    def basis(self, times):
        times = np.ascontiguousarray(np.atleast_1d(times), dtype=np.float64)
        spans = np.zeros(len(times), dtype=np.int32)
        weights = np.zeros((len(times), 4), dtype=np.float64)
        # shared by every model with the same knot vector (the primitives of a graph usually share it)
        kk = self.__dict__.get("_knots_key")
        if kk is None:
            kk = self.__dict__["_knots_key"] = self.knots.tobytes()
        cache = _BASIS_CACHE.get(kk)
        if cache is None:
            cache = _BASIS_CACHE[kk] = {}
        for i, t in enumerate(times):
            key = float(t)
            hit = cache.get(key)
            if hit is None:         # the same FITPACK call the reference makes, once per distinct (knot vector, time)
                if len(cache) > 4096:
                    cache.clear()
                hit = cache[key] = spline_basis(self.knots, key)
            spans[i], weights[i] = hit
        return times, spans, weights

    def eval_frames(self, coeffs, times, out=None):
        torch = _torch()
        coeffs = self._f32(coeffs)
        n = coeffs.shape[0]
        times, spans, weights = self.basis(times)
        if out is None:
            out = torch.empty((n, len(times), self.P), dtype=torch.float32, device=self.torch_device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_eval_frames(self.handle, C.c_void_p(coeffs.data_ptr()), n, _dptr(times), len(times), _iptr(spans),
                                                _dptr(weights), C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def fk(self, frames, node_names, out=None):
        torch = _torch()
        frames = self._f32(frames)
        lead = frames.shape[:-1]
        nf = int(np.prod(lead)) if len(lead) else 1
        nodes = np.ascontiguousarray([self.node_index[nm] for nm in node_names], dtype=np.int32)
        if out is None:
            out = torch.empty(tuple(lead) + (len(nodes), 3), dtype=torch.float32, device=self.torch_device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_fk(self.handle, C.c_void_p(frames.data_ptr()), nf, _iptr(nodes), len(nodes),
                                       C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def logpdf(self, s):
        torch = _torch()
        s = self._f32(s)
        out = torch.empty((s.shape[0],), dtype=torch.float32, device=self.torch_device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.mgb_logpdf(self.handle, C.c_void_p(s.data_ptr()), s.shape[0], C.c_void_p(out.data_ptr()), self._stream()))
        return out


# ---------------------------------------------------------------------------
# constraints
# ---------------------------------------------------------------------------
def _mask_and_target(position):
    mask, tgt = 0, [0.0, 0.0, 0.0]
    for i in range(3):
        if position[i] is not None:
            mask |= 1 << i
            tgt[i] = float(position[i])
    return mask, tgt


def _quaternion_from_rotmat(R):
    """Unit quaternion (w, x, y, z), w >= 0, of a rotation matrix (Shepperd's method, fp64)."""
    R = np.asarray(R, dtype=np.float64)[:3, :3]
    t = np.trace(R)
    if t > 0.0:
        s = np.sqrt(t + 1.0) * 2.0
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2.0
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2.0
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2.0
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    q = np.asarray(q, dtype=np.float64)
    q /= np.linalg.norm(q)
    return -q if q[0] < 0.0 else q


def flatten_constraints(dmodel, mp_constraints):
    """MotionPrimitiveConstraints -> (ConstraintDesc[], PosePoint[], times).
    Field names per type: SURVEY.md Appendix C / the reference classes cited in
    include/mgb200.h."""
    skeleton = mp_constraints.skeleton if getattr(mp_constraints, "skeleton", None) is not None else dmodel.skeleton
    idx = dmodel.node_index
    times = []
    slot_of = {}

    def slot(t):
        t = float(t)
        if t not in slot_of:
            slot_of[t] = len(times)
            times.append(t)
        return slot_of[t]

    clist = list(mp_constraints.constraints)
    # CA / local-trajectory constraints need consecutive slots for the integer frames 0..F-1: allocate them first
    ca_frames = 0
    for c in clist:
        if c.constraint_type in ("ca_constraint", "local_trajectory"):
            ca_frames = max(ca_frames, int(c.n_canonical_frames))
    for f in range(ca_frames):
        slot(float(f))
    # trajectory / trajectory-set constraints read the motion vector: its frames sit at linspace(0, F, F)
    # (motion_primitive.py:233, quirk Q1), in consecutive slots of their own (t = 0 may repeat an integer-frame slot)
    mv_first = -1
    if any(c.constraint_type in ("trajectory", "trajectory_set") for c in clist):
        mv_first = len(times)
        for t in np.linspace(0, dmodel.F, dmodel.F):
            slot_of.setdefault(float(t), len(times))
            times.append(float(t))
    descs = (_lib.ConstraintDesc * max(len(clist), 1))()
    points = []

    def node(name):
        if name not in idx:
            raise UnsupportedConstraintError("joint %r is not in the packed skeleton" % (name,))
        return idx[name]

    def channel_node(channel):
        # node whose quaternion occupies frame channels [3 + 4 channel, 7 + 4 channel) (mg_state_planner.py:288)
        if channel not in dmodel.channel_node:
            raise UnsupportedConstraintError("no joint with quaternion frame index %d" % channel)
        return dmodel.channel_node[channel]

    for j, c in enumerate(clist):
        d = descs[j]
        ct = c.constraint_type
        d.weight = float(c.weight_factor)
        if ct is None and hasattr(c, "constraint_rotmat"):
            # JointRotationConstraint (joint_rotation_constraint.py:62-75): channel = position of the joint in node_name_frame_map
            d.type = _lib.CT_JOINT_ROT
            d.node[0] = channel_node(list(skeleton.node_name_frame_map.keys()).index(c.joint_name))
            d.time_slot = slot(c.frame_idx)
            d.orient_q = (C.c_double * 4)(*_quaternion_from_rotmat(c.constraint_rotmat))
        elif ct == "keyframe_pose" and not hasattr(c, "node_names"):
            # PoseConstraintQuatFrame (pose_constraint_quat_frame.py:47-52): frame 0, joints 1.. against the target frame,
            # weight i of skeleton.get_joint_weights() on joint i + 1 (anim_utils calculate_weighted_frame_distance_quat)
            weights = list(skeleton.get_joint_weights())
            target = np.asarray(c.pose_constraint, dtype=np.float64)
            d.type = _lib.CT_QUAT_FRAME
            d.time_slot = slot(0.0)
            d.points_offset = len(points)
            d.n_points = len(weights) - 1
            for i in range(len(weights) - 1):
                q = target[3 + 4 * (i + 1):7 + 4 * (i + 1)]
                q = q / np.linalg.norm(q)
                if q[0] < 0.0:
                    q = -q
                pp = _lib.PosePoint()
                pp.node = channel_node(i + 1)
                pp.target = (C.c_double * 3)(float(q[1]), float(q[2]), float(q[3]))
                pp.weight = float(weights[i])
                points.append(pp)
        elif ct == "local_trajectory":
            from .constraints import tabulate_trajectory
            table = tabulate_trajectory(c.trajectory)
            d.type = _lib.CT_LOCAL_TRAJ
            d.node[0] = node(c.joint_name)
            d.time_slot = 0
            d.n_slots = int(c.n_canonical_frames)
            d.offset = (C.c_double * 4)(float(c.start_t), float(table.step), 0.0, 0.0)
            d.points_offset = len(points)
            d.n_points = len(table.points) + 1
            for pt in list(table.points) + [table.last_control_point]:
                pp = _lib.PosePoint()
                pp.node = 0
                pp.target = (C.c_double * 3)(float(pt[0]), float(pt[1]), float(pt[2]))
                pp.weight = 0.0
                points.append(pp)
        elif ct == "trajectory":
            # TrajectoryConstraint (trajectory_constraint.py:78-116): every frame of the motion vector against the closest
            # point of the spline at a parameter >= the previous frame's, the first search starting at
            # min_arc_length / full_arc_length (:104) read as a spline PARAMETER (parameterized_spline.py:313-320)
            from .constraints import tabulate_by_parameter
            table = tabulate_by_parameter(c)
            d.type = _lib.CT_TRAJ
            d.node[0] = node(c.joint_name)
            d.time_slot = mv_first
            d.n_slots = dmodel.F
            d.offset = (C.c_double * 4)(float(c.min_arc_length) / float(c.full_arc_length), 0.0, 0.0, 0.0)
            d.points_offset = len(points)
            d.n_points = len(table.points)
            for pt in table.points:
                pp = _lib.PosePoint()
                pp.node = 0
                pp.target = (C.c_double * 3)(float(pt[0]), float(pt[1]), float(pt[2]))
                pp.weight = 0.0
                points.append(pp)
        elif ct == "trajectory_set":
            # TrajectorySetConstraint (trajectory_set_constraint.py:83-105)
            from .constraints import tabulate_trajectory
            if int(c.n_canonical_frames) > dmodel.F or int(c.n_canonical_frames) < 1:
                raise UnsupportedConstraintError("trajectory set over %d frames of a %d-frame motion" % (c.n_canonical_frames, dmodel.F))
            d.type = _lib.CT_TRAJ_SET
            d.node[1] = len(c.joint_trajectories)
            d.time_slot = mv_first
            d.n_slots = int(c.n_canonical_frames)
            d.points_offset = len(points)

            def header(nd, a, b, cc):
                pp = _lib.PosePoint()
                pp.node = nd
                pp.target = (C.c_double * 3)(float(a), float(b), float(cc))
                pp.weight = 0.0
                points.append(pp)

            for traj, jn, arc0 in zip(c.joint_trajectories, c.joint_names, c.joint_arc_lengths):
                table = tabulate_trajectory(traj, 1024)
                nd = node(jn)
                header(nd, table.step, arc0, len(table.points) - 1)
                ranged = getattr(traj, "range_start", None) is not None
                header(nd, traj.range_start if ranged else 0.0, traj.range_end if ranged else 0.0, 1.0 if ranged else 0.0)
                for pt in list(table.points) + [table.last_control_point]:
                    header(nd, pt[0], pt[1], pt[2])
            d.n_points = len(points) - d.points_offset
        elif ct in ("keyframe_position", "keyframe_relative_position"):
            d.type = _lib.CT_POSITION if ct == "keyframe_position" else _lib.CT_RELATIVE
            d.node[0] = node(c.joint_name)
            d.time_slot = slot(c.canonical_keyframe)
            if c.position is not None:
                d.has_position = 1
                if ct == "keyframe_relative_position":
                    if any(p is None for p in c.position):
                        raise UnsupportedConstraintError("relative constraint with an unconstrained axis")
                    d.axis_mask = 7
                    d.target = (C.c_double * 3)(*[float(p) for p in c.position])
                    d.offset = (C.c_double * 4)(*[float(o) for o in c.offset])
                else:
                    mask, tgt = _mask_and_target(c.position)
                    d.axis_mask = mask
                    d.target = (C.c_double * 3)(*tgt)
            if getattr(c, "orientation", None) is not None:
                d.has_orientation = 1
                d.orient_q = (C.c_double * 4)(*[float(q) for q in c.orientation])
        elif ct == "keyframe_2d_direction":
            d.type = _lib.CT_DIR2D
            d.time_slot = slot(c.canonical_keyframe)
            d.dir = (C.c_double * 2)(float(c.target_dir[0]), float(c.target_dir[1]))
        elif ct == "keyframe_pose":
            n = len(c.node_names)
            if n > _lib.MAX_POSE_POINTS:
                raise UnsupportedConstraintError("pose constraint with %d points (max %d)" % (n, _lib.MAX_POSE_POINTS))
            d.type = _lib.CT_POSE
            d.time_slot = slot(c.canonical_keyframe)
            d.points_offset = len(points)
            d.n_points = n
            for nm, tgt, w in zip(c.node_names, c.pose_constraint, c.weights):
                pp = _lib.PosePoint()
                pp.node = node(nm)
                pp.target = (C.c_double * 3)(*[float(x) for x in tgt])
                pp.weight = float(w)
                points.append(pp)
            if c.velocity_constraint is not None:
                d.has_velocity = 1
                d.velocity = (C.c_double * 3)(*[float(x) for x in c.velocity_constraint])
                d.time_slot2 = slot(c.canonical_keyframe + 1)
        elif ct == "keyframe_two_hands":
            d.type = _lib.CT_TWO_HANDS
            d.node[0], d.node[1] = node(c.joint_names[0]), node(c.joint_names[1])
            d.time_slot = slot(c.canonical_keyframe)
            d.target = (C.c_double * 3)(*[float(x) for x in c.positions[0]])
            d.target2 = (C.c_double * 3)(*[float(x) for x in c.positions[1]])
            d.center = (C.c_double * 3)(*[float(x) for x in c.target_center])
        elif ct == "keyframe_feet":
            d.type = _lib.CT_FEET
            d.node[0], d.node[1] = node("LeftFoot"), node("RightFoot")   # names fixed by feet_constraint.py:49-50
            d.time_slot = slot(c.canonical_keyframe)
            d.target = (C.c_double * 3)(*[float(x) for x in c.left])
            d.target2 = (C.c_double * 3)(*[float(x) for x in c.right])
        elif ct == "keyframe_look_at":
            d.type = _lib.CT_LOOK_AT
            d.node[0] = node(skeleton.head_joint)
            d.time_slot = slot(c.canonical_keyframe)
            d.target = (C.c_double * 3)(*[float(x) for x in c.target_position])
        elif ct == "ca_constraint":
            d.type = _lib.CT_CA
            d.node[0] = node(c.joint_name)
            d.time_slot = 0
            d.n_slots = int(c.n_canonical_frames)
            mask, tgt = _mask_and_target(c.position)
            d.axis_mask = mask
            d.target = (C.c_double * 3)(*tgt)
        else:
            raise UnsupportedConstraintError("constraint type %r is outside the device path" % (ct,))
    parr = (_lib.PosePoint * max(len(points), 1))(*points)
    return descs, len(clist), parr, len(points), np.asarray(times, dtype=np.float64)


class DeviceConstraints(object):
    """A MotionPrimitiveConstraints list compiled for the GPU (mgb_constraints handle)."""

    def __init__(self, dmodel, mp_constraints, prev_frames=None, flat=None):
        self.dmodel = dmodel
        self.lib = dmodel.lib
        descs, n, parr, n_points, times = flat if flat is not None else flatten_constraints(dmodel, mp_constraints)
        align = _lib.AlignDesc()
        align.mode = 0
        is_local = bool(getattr(mp_constraints, "is_local", False))
        if not is_local:
            if prev_frames is not None:
                from .skeleton import heading_degrees
                skeleton = mp_constraints.skeleton if getattr(mp_constraints, "skeleton", None) is not None else dmodel.skeleton
                last = np.asarray(prev_frames[-1], dtype=np.float64)
                align.mode = 1
                align.prev_heading_deg = heading_degrees(skeleton, last)
                align.prev_root = (C.c_double * 3)(*last[:3])
                if 0.0 not in times:            # control point 0 == frame at t = 0 (clamped spline)
                    times = np.concatenate([times, [0.0]])
            elif getattr(mp_constraints, "start_pose", None) is not None:
                raise UnsupportedConstraintError("start-pose alignment mutates start_pose per call (SURVEY.md Q2)")
        self.times, self.spans, self.weights = dmodel.basis(times) if len(times) else (np.zeros(0), np.zeros(0, np.int32), np.zeros((0, 4)))
        self.n_terms = n
        h = C.c_void_p()
        _lib.check(self.lib.mgb_constraints_pack(dmodel.handle, descs, n, parr, n_points, _dptr(self.times), len(self.times),
                                                 _iptr(self.spans), _dptr(self.weights), C.byref(align), C.byref(h)))
        self.handle = h
        nq, no, nt = C.c_int32(), C.c_int32(), C.c_int32()
        _lib.check(self.lib.mgb_constraints_info(h, C.byref(nq), C.byref(no), C.byref(nt)))
        self.n_quads, self.n_ops = nq.value, no.value

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.mgb_constraints_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def score(self, s, k=1, error_scale=1.0, quality_scale=0.0, magnitude_scale=0.0, return_cost=False,
              return_terms=False, index_base=0, frames=None):
        """Score device-resident candidates.  s: [n, D] float32 or float64 CUDA tensor.
        Returns dict(best_cost[k], best_idx[k] (device tensors), cost[n]?, terms[n_terms, n]?)."""
        torch = _torch()
        dm = self.dmodel
        if not isinstance(s, torch.Tensor):
            s = torch.as_tensor(np.ascontiguousarray(s), device=dm.torch_device)
        if s.device != dm.torch_device:
            s = s.to(dm.torch_device)
        if s.dtype not in (torch.float32, torch.float64):
            s = s.float()
        s = s.contiguous()
        n = int(s.shape[0])
        assert s.dim() == 2 and s.shape[1] == dm.D, "samples must be [n, %d]" % dm.D
        # frames mode at a large latent dimension: the density part is computed by a tensor-core launch INTO the cost array
        need_cost = return_cost or k > 1 or (frames is not None and dm.D > 32 and (quality_scale != 0.0 or magnitude_scale != 0.0))
        cost = torch.empty((n,), dtype=torch.float32, device=dm.torch_device) if need_cost else None
        terms = torch.empty((self.n_terms, n), dtype=torch.float32, device=dm.torch_device) if return_terms else None
        best_cost = torch.empty((max(k, 1),), dtype=torch.float32, device=dm.torch_device)
        best_idx = torch.empty((max(k, 1),), dtype=torch.int64, device=dm.torch_device)
        ws = dm.workspace()
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        with torch.cuda.device(dm.device):
            if frames is None:
                _lib.check(self.lib.mgb_score(dm.handle, self.handle, ptr(s), 1 if s.dtype == torch.float64 else 0, n, int(index_base),
                                              float(error_scale), float(quality_scale), float(magnitude_scale), ptr(cost), ptr(terms),
                                              int(k), ptr(best_cost), ptr(best_idx), ptr(ws), ws.numel(), dm._stream()))
            else:
                frames = dm._f32(frames)
                s32 = s if s.dtype == torch.float32 else s.float()
                _lib.check(self.lib.mgb_score_frames(dm.handle, self.handle, ptr(frames), ptr(s32), n, int(index_base),
                                                     float(error_scale), float(quality_scale), float(magnitude_scale), ptr(cost),
                                                     ptr(terms), int(k), ptr(best_cost), ptr(best_idx), ptr(ws), ws.numel(), dm._stream()))
        out = {"best_cost": best_cost[:k], "best_idx": best_idx[:k]}
        if return_cost:
            out["cost"] = cost
        if return_terms:
            out["terms"] = terms
        return out

    def node_costs(self, points, error_scale=1.0, quality_scale=0.0, magnitude_scale=0.0, out=None):
        """K8 (mgb_node_costs): cost of every row of a cluster tree's point table [M, D]
        (device tensor, float32 or float64) -> float32 device tensor [M]; no selection."""
        torch = _torch()
        dm = self.dmodel
        assert isinstance(points, torch.Tensor) and points.device == dm.torch_device and points.is_contiguous()
        assert points.dtype in (torch.float32, torch.float64) and points.dim() == 2 and points.shape[1] == dm.D
        n = int(points.shape[0])
        if out is None:
            out = torch.empty((n,), dtype=torch.float32, device=dm.torch_device)
        with torch.cuda.device(dm.device):
            _lib.check(self.lib.mgb_node_costs(dm.handle, self.handle, C.c_void_p(points.data_ptr()),
                                               1 if points.dtype == torch.float64 else 0, n, float(error_scale),
                                               float(quality_scale), float(magnitude_scale), C.c_void_p(out.data_ptr()), dm._stream()))
        return out

    def sample_score(self, n, seed=0, counts=None, n_total=None, index_base=0, k=1, error_scale=1.0, quality_scale=0.0,
                     magnitude_scale=0.0, return_cost=False):
        """Fused K1 + K6 (mgb_sample_score): draw candidates [index_base, index_base + n) of a job of
        n_total samples in-kernel and score them; nothing but the k best (cost, global index) is
        written.  `DeviceModel.sample(1, seed, index_base=idx, counts=counts, n_total=n_total)`
        regenerates a winning row bit-exactly."""
        torch = _torch()
        dm = self.dmodel
        n_total = int(n if n_total is None else n_total)
        if counts is None:
            counts = dm.component_counts(n_total, seed)
        counts = np.ascontiguousarray(counts, dtype=np.int64)
        need_cost = return_cost or k > 1
        cost = torch.empty((int(n),), dtype=torch.float32, device=dm.torch_device) if need_cost else None
        best_cost = torch.empty((max(k, 1),), dtype=torch.float32, device=dm.torch_device)
        best_idx = torch.empty((max(k, 1),), dtype=torch.int64, device=dm.torch_device)
        ws = dm.workspace()
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        with torch.cuda.device(dm.device):
            _lib.check(self.lib.mgb_sample_score(dm.handle, self.handle, int(n), int(index_base),
                                                 counts.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF),
                                                 float(error_scale), float(quality_scale), float(magnitude_scale), ptr(cost), int(k),
                                                 ptr(best_cost), ptr(best_idx), ptr(ws), ws.numel(), dm._stream()))
        out = {"best_cost": best_cost[:k], "best_idx": best_idx[:k], "counts": counts}
        if return_cost:
            out["cost"] = cost
        return out

    def sample_score_sharded(self, n, peer, seed=0, counts=None, n_total=None, index_base=0, error_scale=1.0, quality_scale=0.0,
                             magnitude_scale=0.0):
        """This rank's shard of a step with the exchange of the per-rank best inside the kernel's epilogue
        (mgb_sample_score_sharded): one launch, returns device tensors (best_cost[1], best_idx[1]) holding the JOB-WIDE winner."""
        torch = _torch()
        dm = self.dmodel
        counts = np.ascontiguousarray(counts if counts is not None else dm.component_counts(int(n_total), seed), dtype=np.int64)
        best_cost = torch.empty((1,), dtype=torch.float32, device=dm.torch_device)
        best_idx = torch.empty((1,), dtype=torch.int64, device=dm.torch_device)
        ws = dm.workspace()
        desc = peer.next_desc()
        with torch.cuda.device(dm.device):
            _lib.check(self.lib.mgb_sample_score_sharded(dm.handle, self.handle, int(n), int(index_base), counts.ctypes.data_as(C.POINTER(C.c_int64)),
                                                         C.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF), float(error_scale), float(quality_scale),
                                                         float(magnitude_scale), C.byref(desc), C.c_void_p(best_cost.data_ptr()),
                                                         C.c_void_p(best_idx.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel(), dm._stream()))
        return best_cost, best_idx

    def sample_score_best(self, n, seed=0, counts=None, n_total=None, index_base=0, error_scale=1.0, quality_scale=0.0,
                          magnitude_scale=0.0, peer=None):
        """The planner-facing step (mgb_sample_score_best): draw + score + arg-min in one launch and get
        (best_cost, best_global_index, winning latent vector as float64 numpy [D]) back with one copy and one
        synchronisation.  What `_get_best_fit_sample_using_gmm` needs, nothing else."""
        dm = self.dmodel
        n_total = int(n if n_total is None else n_total)
        if counts is None:
            counts = dm.component_counts(n_total, seed)
        counts = np.ascontiguousarray(counts, dtype=np.int64)
        cost, idx = C.c_float(), C.c_int64()
        row = np.empty(dm.D, dtype=np.float32)
        desc = peer.next_desc() if peer is not None else None     # world > 1: the exchange runs in the kernel's epilogue
        with _torch().cuda.device(dm.device):
            _lib.check(self.lib.mgb_sample_score_best_sharded(dm.handle, self.handle, int(n), int(index_base),
                                                              counts.ctypes.data_as(C.POINTER(C.c_int64)), C.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF),
                                                              float(error_scale), float(quality_scale), float(magnitude_scale),
                                                              C.byref(desc) if desc is not None else None, C.byref(cost), C.byref(idx),
                                                              C.c_void_p(row.ctypes.data), dm._stream()))
        return float(cost.value), int(idx.value), row.astype(np.float64)

    # -- K9 / K10: residual vectors and the local optimiser ------------------------------------
    def n_residuals(self):
        return int(self.lib.mgb_constraints_n_residuals(self.handle))

    def residuals(self, s):
        """Weighted residual vectors [n, R] (float32 CUDA tensor) of candidates s [n, D]:
        MotionPrimitiveConstraints.get_residual_vector (motion_primitive_constraints.py:124-146), batched."""
        torch = _torch()
        dm = self.dmodel
        s = dm._f32(s)
        n = int(s.shape[0])
        out = torch.empty((n, self.n_residuals()), dtype=torch.float32, device=dm.torch_device)
        with torch.cuda.device(dm.device):
            _lib.check(self.lib.mgb_residuals(dm.handle, self.handle, C.c_void_p(s.data_ptr()), n, C.c_void_p(out.data_ptr()), dm._stream()))
        return out

    def lm_refine(self, seeds, error_scale=1.0, quality_scale=0.0, init_error_sum=0.0, max_iter=50):
        """Levenberg-Marquardt from every row of seeds [k, D] in one launch (mgb_lm_refine).  Returns
        dict(s [k, D] float64 numpy, min_error [k] = sum of the weighted residuals at the result, cost_start [k],
        cost [k] = 0.5 |r|^2 of the un-normalised residual objective, iterations [k])."""
        torch = _torch()
        dm = self.dmodel
        seeds = dm._f32(np.atleast_2d(seeds) if not isinstance(seeds, torch.Tensor) else seeds)
        k = int(seeds.shape[0])
        out_s = torch.empty((k, dm.D), dtype=torch.float32, device=dm.torch_device)
        out_e = torch.empty((k, 4), dtype=torch.float32, device=dm.torch_device)
        with torch.cuda.device(dm.device):
            _lib.check(self.lib.mgb_lm_refine(dm.handle, self.handle, C.c_void_p(seeds.data_ptr()), k, float(error_scale), float(quality_scale),
                                              float(init_error_sum), int(max_iter), C.c_void_p(out_s.data_ptr()), C.c_void_p(out_e.data_ptr()),
                                              dm._stream()))
        e = out_e.cpu().numpy().astype(np.float64)
        return {"s": out_s.cpu().numpy().astype(np.float64), "min_error": e[:, 0], "cost_start": e[:, 1], "cost": e[:, 2],
                "iterations": e[:, 3].astype(np.int64)}

    def score_host(self, samples, k=1, error_scale=1.0, quality_scale=0.0, magnitude_scale=0.0, return_cost=False,
                   index_base=0):
        """Score HOST candidates (numpy float32/float64 [n, D], or a pinned CPU torch tensor):
        chunked H2D overlapped with scoring, only the k results come back."""
        torch = _torch()
        if isinstance(samples, torch.Tensor):
            assert samples.device.type == "cpu" and samples.is_contiguous()
            is64 = samples.dtype == torch.float64
            assert is64 or samples.dtype == torch.float32
            n, D = samples.shape
            sptr = samples.data_ptr()
        else:
            if samples.dtype not in (np.float32, np.float64) or not samples.flags["C_CONTIGUOUS"]:
                samples = np.ascontiguousarray(samples, dtype=np.float64)
            is64 = samples.dtype == np.float64
            n, D = samples.shape
            sptr = samples.ctypes.data
        assert D == self.dmodel.D
        cost = np.empty(n, dtype=np.float32) if return_cost else None
        bc = np.empty(max(k, 1), dtype=np.float32)
        bi = np.empty(max(k, 1), dtype=np.int64)
        _lib.check(self.lib.mgb_score_host(self.dmodel.handle, self.handle, C.c_void_p(sptr), 1 if is64 else 0, int(n), int(index_base),
                                           float(error_scale), float(quality_scale), float(magnitude_scale),
                                           C.c_void_p(cost.ctypes.data if cost is not None else 0), int(k),
                                           C.c_void_p(bc.ctypes.data), C.c_void_p(bi.ctypes.data)))
        out = {"best_cost": bc[:k], "best_idx": bi[:k]}
        if return_cost:
            out["cost"] = cost
        return out


def sample_score_multi(options):
    """Cross-option batching (mgb_sample_score_multi): `options` is a list of dicts with keys `dc` (DeviceConstraints),
    `n`, `seed` and optionally `counts`, `error_scale`, `quality_scale`, `magnitude_scale`.  All options are drawn,
    scored and reduced in ONE launch; returns a list of (best_cost, best_index, latent vector float64 [D]) per option.
    Raises MgbError (code MGB_E_UNSUPPORTED) when the options do not share a kernel class - the caller then issues one
    `sample_score_best` per option."""
    lib = _lib.load()
    m = len(options)
    arr = (_lib.Option * m)()
    keep = []
    dmax = 0
    for j, o in enumerate(options):
        dc = o["dc"]
        dm = dc.dmodel
        n = int(o["n"])
        seed = int(o.get("seed", 0))
        counts = o.get("counts")
        if counts is None:
            counts = dm.component_counts(n, seed)
        counts = np.ascontiguousarray(counts, dtype=np.int64)
        keep.append(counts)
        a = arr[j]
        a.model, a.constraints, a.n, a.index_base = dm.handle, dc.handle, n, int(o.get("index_base", 0))
        a.counts = counts.ctypes.data_as(C.POINTER(C.c_int64))
        a.seed = seed & 0xFFFFFFFFFFFFFFFF
        a.err_scale, a.quality_scale, a.magnitude_scale = float(o.get("error_scale", 1.0)), float(o.get("quality_scale", 0.0)), float(o.get("magnitude_scale", 0.0))
        dmax = max(dmax, dm.D)
    cost = np.empty(m, dtype=np.float32)
    idx = np.empty(m, dtype=np.int64)
    rows = np.zeros((m, dmax), dtype=np.float32)
    dm0 = options[0]["dc"].dmodel
    with _torch().cuda.device(dm0.device):
        _lib.check(lib.mgb_sample_score_multi(arr, m, C.c_void_p(cost.ctypes.data), C.c_void_p(idx.ctypes.data), C.c_void_p(rows.ctypes.data),
                                              int(dmax), dm0._stream()))
    return [(float(cost[j]), int(idx[j]), rows[j, :options[j]["dc"].dmodel.D].astype(np.float64)) for j in range(m)]


def kmeans_level(x, dim, node, n_nodes, n_sub, centroids, tol=None, max_iter=300, check_every=4, dim_acc=None,
                 frozen=None):
    """K7 driver (mgb_kmeans_level) on torch CUDA tensors.  x [n, ld] float32 with the samples of a
    node stored contiguously, node [n] int32 (or None = one node), centroids [n_nodes, n_sub, dim]
    float32 (seeds in, final centroids out, in place).  Returns dict(label [n] int32,
    sums [n_nodes, n_sub, dim_acc], counts [n_nodes, n_sub], inertia [n_nodes], converged [n_nodes],
    iters) for the final centroids."""
    torch = _torch()
    lib = _lib.load()
    if not x.is_cuda:
        raise MgbError("kmeans_level needs CUDA tensors: morphablegraphs_b200 has no CPU fallback")
    dev = x.device
    assert x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
    n, ld = int(x.shape[0]), int(x.shape[1])
    dim_acc = int(dim if dim_acc is None else dim_acc)
    assert centroids.dtype == torch.float32 and centroids.is_contiguous() and tuple(centroids.shape) == (n_nodes, n_sub, dim)
    if node is not None:
        assert node.dtype == torch.int32 and node.is_contiguous() and node.shape[0] == n
    alloc = torch.zeros if n == 0 else torch.empty          # the library returns early on an empty sample set
    label = torch.empty((n,), dtype=torch.int32, device=dev)
    sums = alloc((n_nodes, n_sub, dim_acc), dtype=torch.float32, device=dev)
    counts = alloc((n_nodes, n_sub), dtype=torch.int32, device=dev)
    inertia = alloc((n_nodes,), dtype=torch.float32, device=dev)
    converged = torch.zeros((n_nodes,), dtype=torch.int32, device=dev) if frozen is None else frozen.to(torch.int32).contiguous()
    if tol is not None:
        assert tol.dtype == torch.float32 and tol.is_contiguous() and tol.shape[0] == n_nodes
    iters = C.c_int32(0)
    ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
    with torch.cuda.device(dev):
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.mgb_kmeans_level(ptr(x), n, int(dim), dim_acc, ld, ptr(node), int(n_nodes), int(n_sub), ptr(centroids), ptr(tol),
                                        int(max_iter), int(check_every), ptr(label), ptr(sums), ptr(counts), ptr(inertia),
                                        ptr(converged), C.byref(iters), stream))
    return {"label": label, "sums": sums, "counts": counts, "inertia": inertia, "converged": converged, "iters": int(iters.value)}


def launch_count():
    return int(_lib.load().mgb_launch_count())
