"""Vectorised numpy (float64) restatement of the reference's sample-and-evaluate path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  This is the checker the
GPU parity tests compare against, and the CPU port timed by bench.py's
`cpu_baseline` / `--impl reference` legs.  Every function cites the reference
lines it restates; `tests/test_oracle_golden.py` pins it against fixtures
produced by the reference's own modules (`oracle/make_golden.py`).

Inputs are duck-typed on the reference's attribute names so the same call
works on reference objects (build container) and on the product's host-side
mirror classes (GPU box):
  model  : has .s_pca / .translation_maxima / .gaussian_mixture_model /
           .n_canonical_frames (reference motion_primitive.py:137-163), or a
           wrapper exposing it as .motion_primitive
  skeleton: .nodes[name] with .parent/.offset/.rotation/.fixed/
           .quaternion_frame_index/.node_type, .aligning_root_node, ...
  constraints: MotionPrimitiveConstraints-like with .constraints list
"""
import heapq
import math

import numpy as np
import scipy.interpolate as si
from scipy.special import logsumexp

_EPS = np.finfo(float).eps * 4.0
NODE_ROOT, NODE_JOINT, NODE_END_SITE = 0, 1, 2


# --------------------------------------------------------------------------
# model access
# --------------------------------------------------------------------------
def _mp(model):
    return getattr(model, "motion_primitive", model)


class ModelArrays(object):
    """Flat float64 arrays of one motion primitive.
    reference motion_primitive.py:126-180 (GMM, spatial PCA)."""

    def __init__(self, model):
        mp = _mp(model)
        s = mp.s_pca
        self.E = np.asarray(s["eigen_vectors"], dtype=np.float64)      # [Nb*P, K]
        self.mean = np.asarray(s["mean_vector"], dtype=np.float64)     # [Nb*P]
        self.Nb = int(s["n_basis"])
        self.P = int(s["n_dim"])
        self.K = int(s["n_components"])
        self.knots = np.asarray(s["knots"], dtype=np.float64)
        self.tm = np.asarray(mp.translation_maxima, dtype=np.float64)
        self.F = int(mp.n_canonical_frames)
        gmm = mp.gaussian_mixture_model
        self.weights = np.asarray(gmm.weights_, dtype=np.float64)
        self.means = np.asarray(gmm.means_, dtype=np.float64)
        self.covars = np.asarray(gmm.covariances_, dtype=np.float64)
        self.prec_chol = np.asarray(gmm.precisions_cholesky_, dtype=np.float64)
        self.D = self.means.shape[1]
        self.G = self.means.shape[0]


def precision_cholesky(covars):
    """sklearn `_compute_precision_cholesky(full)`, called at reference
    motion_primitive.py:141-142: P = solve(L, I).T with L = chol(cov) lower."""
    from scipy import linalg
    G, D, _ = covars.shape
    out = np.empty((G, D, D))
    for g in range(G):
        L = linalg.cholesky(covars[g], lower=True)
        out[g] = linalg.solve_triangular(L, np.eye(D), lower=True).T
    return out


# --------------------------------------------------------------------------
# B-spline basis (a5, Q1)
# --------------------------------------------------------------------------
def spline_basis_fitpack(knots, t, degree=3):
    """Non-zero basis weights at `t` exactly as scipy.splev evaluates them
    (reference motion_spline.py:89-92 calls splev per channel): evaluate splev
    on identity coefficients.  Returns (span, w[4]) with
    value = sum_i w[i] * coeffs[span + i].  Extrapolates outside the knot
    range like FITPACK (ext=0), which matters for Q1."""
    knots = np.asarray(knots, dtype=np.float64)
    nb = len(knots) - degree - 1
    row = np.array([float(si.splev(t, (knots, np.eye(nb)[i], degree))) for i in range(nb)])
    span = fitpack_span(knots, t, degree)
    w = row[span:span + degree + 1].copy()
    rest = row.copy()
    rest[span:span + degree + 1] = 0.0
    assert np.all(np.abs(rest) < 1e-12), "basis support mismatch"
    return span, w


def fitpack_span(knots, t, degree=3):
    """Index of the first of the 4 active coefficients: FITPACK splev interval
    search, clamped to the first/last polynomial piece."""
    n = len(knots)
    nk1 = n - degree - 1
    l = degree  # 0-based: knots[l] <= t < knots[l+1]
    while l + 1 < nk1 and t >= knots[l + 1]:
        l += 1
    return l - degree


def spline_basis_deboor(knots, t, degree=3):
    """Own restatement of FITPACK fpbspl (de Boor-Cox recurrence on the clamped
    interval); the C library implements the same recurrence."""
    knots = np.asarray(knots, dtype=np.float64)
    span = fitpack_span(knots, t, degree)
    l = span + degree
    h = np.zeros(degree + 1)
    h[0] = 1.0
    for j in range(1, degree + 1):
        hh = h[:j].copy()
        h[0] = 0.0
        for i in range(j):
            li = l + i + 1
            lj = li - j
            f = hh[i] / (knots[li] - knots[lj])
            h[i] += f * (knots[li] - t)
            h[i + 1] = f * (t - knots[lj])
    return span, h


def basis_matrix(knots, times, nb, degree=3):
    B = np.zeros((len(times), nb))
    for r, t in enumerate(times):
        span, w = spline_basis_fitpack(knots, t, degree)
        B[r, span:span + degree + 1] = w
    return B


def motion_vector_times(F, speed=1.0):
    """reference motion_primitive.py:233 (Q1: last sample at t=F, outside the knots)."""
    return np.linspace(0, F, int(F * (1.0 / speed)))


# --------------------------------------------------------------------------
# back-projection and frames (a3-a5)
# --------------------------------------------------------------------------
def back_project_coeffs(ma, S):
    """reference motion_primitive.py:251-255: coefs = E.alpha + mean; reshape;
    translation scaled after the mean is added."""
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    alpha = S[:, :ma.K]
    coefs = alpha.dot(ma.E.T) + ma.mean
    coefs = coefs.reshape(-1, ma.Nb, ma.P).copy()
    coefs[:, :, :3] *= ma.tm
    return coefs


def eval_frames(ma, coeffs, times):
    """reference motion_spline.py:84-92 for a list of times -> [N, T, P]."""
    B = basis_matrix(ma.knots, times, ma.Nb)
    return np.einsum("tb,nbp->ntp", B, coeffs)


# --------------------------------------------------------------------------
# forward kinematics (a14)  -- PARITY UNPINNED (anim_utils)
# --------------------------------------------------------------------------
def quaternion_matrix_batch(q):
    """transformations.quaternion_matrix over [..., 4] -> [..., 3, 3], including
    the |q|^2 < 4 eps identity branch."""
    q = np.asarray(q, dtype=np.float64)
    n = np.sum(q * q, axis=-1)
    small = n < _EPS
    scale = np.sqrt(2.0 / np.where(small, 1.0, n))
    qs = q * scale[..., None]
    w, x, y, z = qs[..., 0], qs[..., 1], qs[..., 2], qs[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1.0 - y * y - z * z
    R[..., 0, 1] = x * y - z * w
    R[..., 0, 2] = x * z + y * w
    R[..., 1, 0] = x * y + z * w
    R[..., 1, 1] = 1.0 - x * x - z * z
    R[..., 1, 2] = y * z - x * w
    R[..., 2, 0] = x * z - y * w
    R[..., 2, 1] = y * z + x * w
    R[..., 2, 2] = 1.0 - x * x - y * y
    R[small] = np.eye(3)
    return R


def _node_type(node):
    t = node.node_type
    if isinstance(t, str):
        return {"root": NODE_ROOT, "joint": NODE_JOINT, "end_site": NODE_END_SITE}[t]
    return int(t)


def _chain(skeleton, node_name):
    chain = []
    node = skeleton.nodes[node_name]
    while node is not None:
        chain.append(node)
        node = node.parent
    return chain[::-1]


def fk_global(skeleton, frames, node_name):
    """Global rotation [..., 3, 3] and position [..., 3] of `node_name` for
    frames [..., P]: G(node) = G(parent) . L(node)."""
    frames = np.asarray(frames, dtype=np.float64)
    R = None
    t = None
    for node in _chain(skeleton, node_name):
        nt = _node_type(node)
        off = np.asarray(node.offset, dtype=np.float64)
        if nt == NODE_ROOT:
            Rl = quaternion_matrix_batch(frames[..., 3:7])
            tl = frames[..., :3] + off
        elif nt == NODE_END_SITE:
            Rl = None
            tl = off
        else:
            if not node.fixed:
                i = node.quaternion_frame_index * 4 + 3
                Rl = quaternion_matrix_batch(frames[..., i:i + 4])
            else:
                Rl = quaternion_matrix_batch(np.asarray(node.rotation, dtype=np.float64))
            tl = off
        if R is None:
            R = Rl if Rl is not None else np.broadcast_to(np.eye(3), frames.shape[:-1] + (3, 3))
            t = np.broadcast_to(tl, frames.shape[:-1] + (3,))
        else:
            t = t + np.einsum("...ij,...j->...i", R, np.broadcast_to(tl, frames.shape[:-1] + (3,)))
            if Rl is not None:
                R = np.einsum("...ij,...jk->...ik", R, np.broadcast_to(Rl, frames.shape[:-1] + (3, 3)))
    return R, t


def fk_positions(skeleton, frames, node_names):
    """[..., len(node_names), 3] global positions (reference call sites
    global_transform_constraint.py:103, pose_constraint.py:51, features.py:146)."""
    return np.stack([fk_global(skeleton, frames, n)[1] for n in node_names], axis=-2)


# --------------------------------------------------------------------------
# constraint terms (a9-a13), vectorised over candidates
# --------------------------------------------------------------------------
def _frames_at(ma, coeffs, t):
    return eval_frames(ma, coeffs, [t])[:, 0, :]


def _position_term(c, skeleton, frame):
    """reference global_transform_constraint.py:102-104,136-143."""
    p = fk_global(skeleton, frame, c.joint_name)[1]
    d = np.zeros(frame.shape[0])
    for i in range(3):
        if c.position[i] is not None:
            d += (c.position[i] - p[:, i]) ** 2
    return np.sqrt(d)


def _rotate_z_by_quat(q):
    """anim_utils quaternion_rotate_vector(q, [0,0,0,1]) for a general (not
    necessarily unit) q: (q (x) z (x) q*)[1:]."""
    w, x, y, z = q
    return np.array([2 * (x * z + w * y), 2 * (y * z - w * x), w * w - x * x - y * y + z * z])


def _orientation_term(c, skeleton, frame):
    """reference global_transform_constraint.py:106-120: angle between +z
    rotated by the joint's global quaternion and by the target quaternion."""
    R = fk_global(skeleton, frame, c.joint_name)[0]
    v1 = R[:, :, 2]
    v2 = _rotate_z_by_quat(np.asarray(c.orientation, dtype=np.float64))
    dot = v1.dot(v2) / (np.linalg.norm(v1, axis=1) * np.linalg.norm(v2))
    return np.arccos(np.clip(dot, -1.0, 1.0))


def term_errors(c, ma, skeleton, coeffs):
    """Unweighted error of one constraint for every candidate (what
    `c.evaluate_motion_spline` returns)."""
    ct = c.constraint_type
    N = coeffs.shape[0]
    if ct is None and hasattr(c, "constraint_rotmat"):
        # reference joint_rotation_constraint.py:62-75: |R_target - R(q / |q|)|_F of the joint's own channels
        frame = _frames_at(ma, coeffs, c.frame_idx)
        ji = list(skeleton.node_name_frame_map.keys()).index(c.joint_name)
        q = frame[:, 3 + 4 * ji:7 + 4 * ji]
        R = quaternion_matrix_batch(q / np.linalg.norm(q, axis=1, keepdims=True))
        Rt = np.asarray(c.constraint_rotmat, dtype=np.float64)[:3, :3]
        return np.sqrt(np.sum((Rt[None] - R) ** 2, axis=(1, 2)))
    if ct == "keyframe_pose" and not hasattr(c, "node_names"):
        # reference pose_constraint_quat_frame.py:47-52 through anim_utils' calculate_weighted_frame_distance_quat
        # (PARITY UNPINNED, see oracle/standin): joints 1.., weight i on joint i + 1, rotation-matrix distance
        frame = _frames_at(ma, coeffs, 0)
        w = list(skeleton.get_joint_weights())
        tgt = np.asarray(c.pose_constraint, dtype=np.float64)
        err = np.zeros(N)
        for i in range(len(w) - 1):
            q = frame[:, 3 + 4 * (i + 1):7 + 4 * (i + 1)]
            R = quaternion_matrix_batch(q / np.linalg.norm(q, axis=1, keepdims=True))
            qt = tgt[3 + 4 * (i + 1):7 + 4 * (i + 1)]
            Rt = quaternion_matrix_batch((qt / np.linalg.norm(qt))[None])[0]
            err += w[i] * np.sqrt(np.sum((Rt[None] - R) ** 2, axis=(1, 2)))
        return err
    if ct == "local_trajectory":
        # reference local_trajectory_constraint.py:58-83: walk the frames, the arc length covered so far picks the target
        times = list(range(int(c.n_canonical_frames)))
        frames = eval_frames(ma, coeffs, times)
        p = fk_global(skeleton, frames, c.joint_name)[1]         # [N, F, 3]
        err = np.zeros(N)
        for n in range(N):
            t = float(c.start_t)
            for f in range(p.shape[1]):
                if f > 0:
                    t += np.linalg.norm(p[n, f - 1] - p[n, f])
                target = c.trajectory.query_point_by_absolute_arc_length(t)
                err[n] += (target[0] - p[n, f, 0]) ** 2 + (target[2] - p[n, f, 2]) ** 2
        return err
    if ct in ("trajectory", "trajectory_set"):
        # trajectory_constraint.py:78-85 / trajectory_set_constraint.py:41-47: the average of the per-frame errors
        return np.average(term_residuals(c, ma, skeleton, coeffs), axis=1)
    if ct in ("keyframe_position", "keyframe_relative_position"):
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        err = np.zeros(N)
        if ct == "keyframe_relative_position":
            # reference relative_transform_constraint.py:46-50
            if c.position is not None:
                R, t = fk_global(skeleton, frame, c.joint_name)
                off = np.asarray(c.offset, dtype=np.float64)
                p = np.einsum("nij,j->ni", R, off[:3]) + t * off[3]
                err += np.linalg.norm(np.asarray(c.position, dtype=np.float64) - p, axis=1)
        elif c.position is not None:
            err += _position_term(c, skeleton, frame)
        if getattr(c, "orientation", None) is not None:
            err += _orientation_term(c, skeleton, frame)
        return err
    if ct == "keyframe_2d_direction":
        # reference direction_2d_constraint.py:42-52 (degrees)
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        R = fk_global(skeleton, frame, skeleton.aligning_root_node)[0]
        p = np.einsum("nij,j->ni", R, np.asarray(skeleton.aligning_root_dir, dtype=np.float64))
        d = np.stack([p[:, 0], p[:, 2]], axis=1)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        td = np.asarray(c.target_dir, dtype=np.float64)
        mag = np.linalg.norm(td) * np.linalg.norm(d, axis=1)
        cosang = np.clip(d.dot(td) / mag, -1.0, 1.0)
        return np.abs(np.degrees(np.arccos(cosang)))
    if ct == "keyframe_pose":
        # reference pose_constraint.py:48-67
        f1 = _frames_at(ma, coeffs, c.canonical_keyframe)
        f2 = _frames_at(ma, coeffs, c.canonical_keyframe + 1)
        pc = fk_positions(skeleton, f1, c.node_names)           # [N, n, 3]
        root2 = fk_global(skeleton, f2, c.node_names[0])[1]
        if c.velocity_constraint is not None:
            vel = root2 - pc[:, 0]
            vel_err = np.linalg.norm(np.asarray(c.velocity_constraint, dtype=np.float64) - vel, axis=1)
        else:
            vel_err = 0.0
        a = np.asarray(c.pose_constraint, dtype=np.float64)     # [n, 3] target
        w = np.asarray(c.weights, dtype=np.float64)
        W = w.sum()
        num_l = np.sum(w * (a[:, 0] * pc[:, :, 2] - pc[:, :, 0] * a[:, 2]), axis=1)
        den_l = np.sum(w * (a[:, 0] * pc[:, :, 0] + a[:, 2] * pc[:, :, 2]), axis=1)
        sax, saz = np.sum(w * a[:, 0]), np.sum(w * a[:, 2])
        sbx, sbz = np.sum(w * pc[:, :, 0], axis=1), np.sum(w * pc[:, :, 2], axis=1)
        num = num_l - (sax * sbz - sbx * saz) / W
        den = den_l - (sax * sbx + saz * sbz) / W
        theta = np.arctan2(num, den)
        cth, sth = np.cos(theta), np.sin(theta)
        ox = (sax - sbx * cth - sbz * sth) / W
        oz = (saz + sbx * sth - sbz * cth) / W
        tx = cth[:, None] * pc[:, :, 0] + sth[:, None] * pc[:, :, 2] + ox[:, None]
        tz = -sth[:, None] * pc[:, :, 0] + cth[:, None] * pc[:, :, 2] + oz[:, None]
        d = np.sqrt((a[:, 0] - tx) ** 2 + (a[:, 1] - pc[:, :, 1]) ** 2 + (a[:, 2] - tz) ** 2)
        return d.mean(axis=1) + vel_err
    if ct == "keyframe_two_hands":
        # reference two_hand_constraint.py:68-93 (sum of three distances)
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        lp = fk_global(skeleton, frame, c.joint_names[0])[1]
        rp = fk_global(skeleton, frame, c.joint_names[1])[1]
        delta = rp - lp
        center = np.asarray(c.target_center, dtype=np.float64)
        r0 = np.linalg.norm(center - (lp + 0.5 * delta), axis=1)
        r1 = np.linalg.norm(np.asarray(c.positions[0], dtype=np.float64) - lp, axis=1)
        r2 = np.linalg.norm(np.asarray(c.positions[1], dtype=np.float64) - rp, axis=1)
        return r0 + r1 + r2
    if ct == "keyframe_feet":
        # reference feet_constraint.py:39-52; weight applied here AND by the caller (Q6)
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        le = np.linalg.norm(np.asarray(c.left, dtype=np.float64) - fk_global(skeleton, frame, "LeftFoot")[1], axis=1)
        re = np.linalg.norm(np.asarray(c.right, dtype=np.float64) - fk_global(skeleton, frame, "RightFoot")[1], axis=1)
        return (le + re) * c.weight_factor
    if ct == "keyframe_look_at":
        # reference look_at_constraint.py:58-68: |angle| between the head's +z and
        # the direction to the target
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        R, hp = fk_global(skeleton, frame, skeleton.head_joint)
        td = np.asarray(c.target_position, dtype=np.float64) - hp
        td /= np.linalg.norm(td, axis=1, keepdims=True)
        hd = R[:, :, 2]
        hd = hd / np.linalg.norm(hd, axis=1, keepdims=True)
        cr = np.linalg.norm(np.cross(hd, td), axis=1)
        return np.abs(np.arctan2(cr, np.sum(hd * td, axis=1)))
    if ct == "ca_constraint":
        # reference global_transform_ca_constraint.py:34-40: min over integer frames
        times = list(range(int(c.n_canonical_frames)))
        frames = eval_frames(ma, coeffs, times)                  # [N, F, P]
        p = fk_global(skeleton, frames, c.joint_name)[1]         # [N, F, 3]
        d = np.zeros(p.shape[:2])
        for i in range(3):
            if c.position[i] is not None:
                d += (c.position[i] - p[..., i]) ** 2
        return np.sqrt(d).min(axis=1)
    raise NotImplementedError("constraint type %r is outside the device path" % ct)


# --------------------------------------------------------------------------
# frame alignment (a7) -- PARITY UNPINNED (anim_utils)
# --------------------------------------------------------------------------
def heading_angle_deg(skeleton, frame_batch):
    """Heading (deg, from +z towards +x) of the aligning root node."""
    R = fk_global(skeleton, frame_batch, skeleton.aligning_root_node)[0]
    p = np.einsum("...ij,j->...i", R, np.asarray(skeleton.aligning_root_dir, dtype=np.float64))
    return np.degrees(np.arctan2(p[..., 0], p[..., 2]))


def quaternion_multiply_batch(q1, q0):
    w1, x1, y1, z1 = q1[..., 0], q1[..., 1], q1[..., 2], q1[..., 3]
    w0, x0, y0, z0 = q0[..., 0], q0[..., 1], q0[..., 2], q0[..., 3]
    return np.stack([-x1 * x0 - y1 * y0 - z1 * z0 + w1 * w0,
                     x1 * w0 + y1 * z0 - z1 * y0 + w1 * x0,
                     -x1 * z0 + y1 * w0 + z1 * x0 + w1 * y0,
                     x1 * y0 - y1 * x0 + z1 * w0 + w1 * z0], axis=-1)


def align_coeffs_to_prev_frames(ma, skeleton, coeffs, prev_frames):
    """reference motion_primitive_constraints.py:111-114 with prev_frames given
    (anim_utils align_quaternion_frames_automatically): per candidate, rotate
    about +y so the heading of control point 0 matches the last previous frame,
    translate x,z so its root lands on the previous root."""
    prev_last = np.asarray(prev_frames[-1], dtype=np.float64)
    ang_a = heading_angle_deg(skeleton, prev_last[None, :])[0]
    ang_b = heading_angle_deg(skeleton, coeffs[:, 0, :])
    ang = ang_a - ang_b
    ang = np.where(ang > 180.0, ang - 360.0, ang)
    ang = np.where(ang <= -180.0, ang + 360.0, ang)
    half = np.radians(ang) / 2.0
    q = np.stack([np.cos(half), np.zeros_like(half), np.sin(half), np.zeros_like(half)], axis=-1)
    Ry = quaternion_matrix_batch(q)                                    # [N,3,3]
    p0 = np.einsum("nij,nj->ni", Ry, coeffs[:, 0, :3])
    delta = prev_last[:3] - p0
    out = coeffs.copy()
    tr = np.einsum("nij,nbj->nbi", Ry, coeffs[:, :, :3])
    tr[:, :, 0] += delta[:, None, 0]
    tr[:, :, 2] += delta[:, None, 2]
    out[:, :, :3] = tr
    # the reference derives q back from the matrix (quaternion_from_matrix, w>=0)
    qm = np.where(q[:, :1] < 0.0, -q, q)
    out[:, :, 3:7] = quaternion_multiply_batch(qm[:, None, :], coeffs[:, :, 3:7])
    return out


# --------------------------------------------------------------------------
# GMM (a1, a2, a15)
# --------------------------------------------------------------------------
def gmm_component_logprob(ma, S):
    """sklearn `_estimate_log_gaussian_prob(full)` + log weights -> [N, G]."""
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    N, D = S.shape
    out = np.empty((N, ma.G))
    for g in range(ma.G):
        Pc = ma.prec_chol[g]
        y = S.dot(Pc) - ma.means[g].dot(Pc)
        logdet = np.sum(np.log(np.diag(Pc)))
        out[:, g] = -0.5 * (D * np.log(2 * np.pi) + np.sum(y * y, axis=1)) + logdet + np.log(ma.weights[g])
    return out


def gmm_logpdf(ma, S):
    """`gmm.score_samples(S)`; naturalness = -logpdf * quality_scale
    (reference objective_functions.py:256)."""
    return logsumexp(gmm_component_logprob(ma, S), axis=1)


def gmm_sample(ma, n, rng):
    """sklearn GaussianMixture.sample: multinomial split, per-component
    multivariate normal, stacked in component order (reference
    motion_primitive.py:182-189)."""
    counts = rng.multinomial(n, ma.weights)
    X = np.vstack([rng.multivariate_normal(ma.means[g], ma.covars[g], int(c))
                   for g, c in enumerate(counts)])
    return X, counts


# --------------------------------------------------------------------------
# the loop (a6, a15, a16)
# --------------------------------------------------------------------------
def candidate_costs(model, skeleton, constraints, S, prev_frames=None, error_scale=1.0,
                    quality_scale=0.0, magnitude_scale=0.0, return_terms=False):
    """cost[n] = error_scale * sum_i w_i c_i(s_n) + quality_scale * (-log p(s_n))
    + magnitude_scale * |s_n|.

    error part: reference motion_primitive_constraints.py:100-122;
    magnitude: mg_state_planner.py:110-129 (scale 2); naturalness term:
    objective_functions.py:256."""
    ma = model if isinstance(model, ModelArrays) else ModelArrays(model)
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    coeffs = back_project_coeffs(ma, S)
    if not constraints.is_local and prev_frames is not None:
        coeffs = align_coeffs_to_prev_frames(ma, skeleton, coeffs, prev_frames)
    total = np.zeros(S.shape[0])
    terms = []
    for c in constraints.constraints:
        e = term_errors(c, ma, skeleton, coeffs)
        terms.append(e)
        total += c.weight_factor * e
    cost = error_scale * total
    if quality_scale != 0.0:
        cost = cost + quality_scale * (-gmm_logpdf(ma, S))
    if magnitude_scale != 0.0:
        cost = cost + magnitude_scale * np.linalg.norm(S, axis=1)
    if return_terms:
        return cost, np.array(terms)
    return cost


# --------------------------------------------------------------------------
# residual-vector form (f1)
# --------------------------------------------------------------------------
def term_residuals(c, ma, skeleton, coeffs):
    """[N, n_res]: what `c.get_residual_vector_spline` returns per candidate (unweighted).  One value for most classes
    (their evaluate); two_hand_constraint.py:68-80 three distances; pose_constraint.py:107-121 one distance per point
    (no mean, no velocity term); pose_constraint_quat_frame.py:60-70 one weighted rotation distance per joint;
    local_trajectory_constraint.py:66-90 one error per frame."""
    ct = c.constraint_type
    N = coeffs.shape[0]
    if ct == "keyframe_two_hands":
        frame = _frames_at(ma, coeffs, c.canonical_keyframe)
        lp = fk_global(skeleton, frame, c.joint_names[0])[1]
        rp = fk_global(skeleton, frame, c.joint_names[1])[1]
        center = np.asarray(c.target_center, dtype=np.float64)
        return np.stack([np.linalg.norm(center - (lp + 0.5 * (rp - lp)), axis=1),
                         np.linalg.norm(np.asarray(c.positions[0], dtype=np.float64) - lp, axis=1),
                         np.linalg.norm(np.asarray(c.positions[1], dtype=np.float64) - rp, axis=1)], axis=1)
    if ct == "keyframe_pose" and hasattr(c, "node_names"):
        f1 = _frames_at(ma, coeffs, c.canonical_keyframe)
        pc = fk_positions(skeleton, f1, c.node_names)
        a = np.asarray(c.pose_constraint, dtype=np.float64)
        w = np.asarray(c.weights, dtype=np.float64)
        W = w.sum()
        num_l = np.sum(w * (a[:, 0] * pc[:, :, 2] - pc[:, :, 0] * a[:, 2]), axis=1)
        den_l = np.sum(w * (a[:, 0] * pc[:, :, 0] + a[:, 2] * pc[:, :, 2]), axis=1)
        sax, saz = np.sum(w * a[:, 0]), np.sum(w * a[:, 2])
        sbx, sbz = np.sum(w * pc[:, :, 0], axis=1), np.sum(w * pc[:, :, 2], axis=1)
        theta = np.arctan2(num_l - (sax * sbz - sbx * saz) / W, den_l - (sax * sbx + saz * sbz) / W)
        cth, sth = np.cos(theta), np.sin(theta)
        ox = (sax - sbx * cth - sbz * sth) / W
        oz = (saz + sbx * sth - sbz * cth) / W
        tx = cth[:, None] * pc[:, :, 0] + sth[:, None] * pc[:, :, 2] + ox[:, None]
        tz = -sth[:, None] * pc[:, :, 0] + cth[:, None] * pc[:, :, 2] + oz[:, None]
        return np.sqrt((a[:, 0] - tx) ** 2 + (a[:, 1] - pc[:, :, 1]) ** 2 + (a[:, 2] - tz) ** 2)
    if ct == "keyframe_pose":
        frame = _frames_at(ma, coeffs, 0)
        w = list(skeleton.get_joint_weights())
        tgt = np.asarray(c.pose_constraint, dtype=np.float64)
        out = np.zeros((N, len(w) - 1))
        for i in range(len(w) - 1):
            q = frame[:, 3 + 4 * (i + 1):7 + 4 * (i + 1)]
            R = quaternion_matrix_batch(q / np.linalg.norm(q, axis=1, keepdims=True))
            qt = tgt[3 + 4 * (i + 1):7 + 4 * (i + 1)]
            Rt = quaternion_matrix_batch((qt / np.linalg.norm(qt))[None])[0]
            out[:, i] = w[i] * np.sqrt(np.sum((Rt[None] - R) ** 2, axis=(1, 2)))
        return out
    if ct == "trajectory":
        return trajectory_frame_errors(c, ma, skeleton, coeffs)
    if ct == "trajectory_set":
        return trajectory_set_frame_errors(c, ma, skeleton, coeffs)
    if ct == "local_trajectory":
        times = list(range(int(c.n_canonical_frames)))
        p = fk_global(skeleton, eval_frames(ma, coeffs, times), c.joint_name)[1]
        out = np.zeros((N, len(times)))
        for n in range(N):
            t = float(c.start_t)
            for f in range(p.shape[1]):
                if f > 0:
                    t += np.linalg.norm(p[n, f - 1] - p[n, f])
                target = c.trajectory.query_point_by_absolute_arc_length(t)
                out[n, f] = (target[0] - p[n, f, 0]) ** 2 + (target[2] - p[n, f, 2]) ** 2
        return out
    return term_errors(c, ma, skeleton, coeffs)[:, None]


def trajectory_frame_errors(c, ma, skeleton, coeffs):
    """[N, F]: TrajectoryConstraint.get_residual_vector (trajectory_constraint.py:95-116) with find_closest_point_fast
    (splines/parameterized_spline.py:303-322): per frame a bounded L-BFGS-B minimisation of the distance between the
    spline at parameter u and the joint position over [u of the previous frame, 1], started at the lower bound; the first
    bound is min_arc_length / full_arc_length (a relative arc length used as a spline parameter, as the reference does).
    Quirk Q8: scipy hands the objective a 1-element ARRAY u, and catmull_rom_spline.py:141-151 builds the ragged list
    [u^3, u^2, u, 1] from it - numpy >= 1.24 raises there, older numpy returned a (3, 1) point whose difference with the
    (3,) joint position broadcasts to a 3 x 3 matrix (and a (3,) point at u = 1, so the optimiser usually ends on that
    bound).  Restated here is the distance the code states, with u read as the scalar it is."""
    from scipy.optimize import minimize
    p = fk_global(skeleton, eval_frames(ma, coeffs, motion_vector_times(ma.F)), c.joint_name)[1]      # [N, F, 3]
    N, F = p.shape[0], p.shape[1]
    out = np.zeros((N, F))

    def dist(u, q):
        return np.linalg.norm(np.asarray(c.query_point_by_parameter(float(np.ravel(u)[0])), dtype=np.float64).reshape(3) - q)

    for n in range(N):
        min_u = float(c.min_arc_length) / float(c.full_arc_length)
        for f in range(F):
            res = minimize(dist, np.array([min_u]), (p[n, f],), method="L-BFGS-B", bounds=[(min_u, 1.0)])
            min_u = float(res["x"][0])
            out[n, f] = dist(min_u, p[n, f])
    return out


def trajectory_set_frame_errors(c, ma, skeleton, coeffs):
    """[N, n_canonical_frames]: TrajectorySetConstraint.get_residual_vector (trajectory_set_constraint.py:83-105).  Every
    joint walks its own trajectory by the arc length of its own path; a frame is scored with the arc lengths reached BEFORE
    it (the step from the previous frame is added after the frame's residual); `np.average` of the list of positions is
    the mean over all joints AND coordinates, so the residual is |mean(positions) - mean(targets)|, and it is taken only if
    some trajectory is active at its arc length."""
    nf = int(c.n_canonical_frames)
    frames = eval_frames(ma, coeffs, motion_vector_times(ma.F)[:nf])
    pos = np.stack([fk_global(skeleton, frames, jn)[1] for jn in c.joint_names], axis=2)     # [N, nf, J, 3]
    N = pos.shape[0]
    out = np.zeros((N, nf))
    for n in range(N):
        arcs = [float(a) for a in c.joint_arc_lengths]
        last = None
        for i in range(nf):
            jp = pos[n, i]
            active = [getattr(t, "range_start", None) is not None and t.range_start <= a <= t.range_end
                      for t, a in zip(c.joint_trajectories, arcs)]
            if np.any(active):
                targets = [np.asarray(t.query_point_by_absolute_arc_length(a), dtype=np.float64)[:3]
                           for t, a in zip(c.joint_trajectories, arcs)]
                out[n, i] = abs(np.average(jp) - np.average(targets))
            if last is not None:
                arcs = [a + np.linalg.norm(jp[j] - last[j]) for j, a in enumerate(arcs)]
            last = jp
    return out


def residual_vectors(model, skeleton, constraints, S, prev_frames=None):
    """[N, R] weighted residual vectors: reference motion_primitive_constraints.py:124-146."""
    ma = model if isinstance(model, ModelArrays) else ModelArrays(model)
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    coeffs = back_project_coeffs(ma, S)
    if not constraints.is_local and prev_frames is not None:
        coeffs = align_coeffs_to_prev_frames(ma, skeleton, coeffs, prev_frames)
    return np.concatenate([c.weight_factor * term_residuals(c, ma, skeleton, coeffs) for c in constraints.constraints], axis=1)


def residual_objective(model, skeleton, constraints, S, prev_frames=None, error_scale=1.0, quality_scale=0.0, init_error_sum=1.0):
    """[N, max(R, D)]: reference objective_functions.py:239-267 (obj_spatial_error_residual_vector_and_naturalness)."""
    ma = model if isinstance(model, ModelArrays) else ModelArrays(model)
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    r = residual_vectors(ma, skeleton, constraints, S, prev_frames) * error_scale
    r = r + (-gmm_logpdf(ma, S) * quality_scale)[:, None]
    if r.shape[1] < S.shape[1]:
        r = np.concatenate([r, np.zeros((S.shape[0], S.shape[1] - r.shape[1]))], axis=1)
    return r / init_error_sum


def first_argmin(cost):
    """reference motion_primitive_generator.py:251-257: strict `>` so the first
    minimum wins; NaN never wins."""
    best_idx = 0
    min_error = np.inf
    c = np.where(np.isnan(cost), np.inf, cost)
    i = int(np.argmin(c)) if len(c) else 0
    if len(c) and c[i] < min_error:
        best_idx, min_error = i, float(c[i])
    return best_idx, min_error


def evaluate_samples_using_constraints(model, skeleton, constraints, S, prev_frames=None, chunk=65536, **kw):
    """Batched equivalent of reference motion_primitive_generator.py:230-261."""
    ma = model if isinstance(model, ModelArrays) else ModelArrays(model)
    S = np.asarray(S, dtype=np.float64)
    best_idx, min_error = 0, np.inf
    for lo in range(0, len(S), chunk):
        cost = candidate_costs(ma, skeleton, constraints, S[lo:lo + chunk], prev_frames, **kw)
        i, e = first_argmin(cost)
        if e < min_error:
            best_idx, min_error = lo + i, e
    return best_idx, min_error


# --------------------------------------------------------------------------
# per-candidate loop port (reference-shaped; timed as the faithful CPU baseline)
# --------------------------------------------------------------------------
def loop_evaluate_one_position(ma, skeleton, constraint, S):
    """Scalar restatement of what the reference executes per candidate for one
    joint-position constraint: mat-vec back-projection (motion_primitive.py:251-255),
    P separate `splev` calls for one frame (motion_spline.py:89-92), a chain of
    4x4 products for FK, Euclidean distance; first-minimum arg-min
    (motion_primitive_generator.py:253-257)."""
    best_idx, min_error = 0, np.inf
    chain = _chain(skeleton, constraint.joint_name)
    target = constraint.position
    t = constraint.canonical_keyframe
    for idx in range(len(S)):
        alpha = S[idx][:ma.K]
        coefs = np.dot(ma.E, alpha)
        coefs += ma.mean
        coefs = coefs.reshape((ma.Nb, ma.P))
        coefs[:, :3] *= ma.tm
        cT = coefs.T
        frame = np.asarray([si.splev(t, (ma.knots, cT[p], 3)) for p in range(ma.P)]).T
        G = None
        for node in chain:
            L = np.identity(4)
            nt = _node_type(node)
            if nt == NODE_ROOT:
                L[:3, :3] = quaternion_matrix_batch(frame[3:7])
                L[:3, 3] = frame[:3] + node.offset
            elif nt == NODE_END_SITE:
                L[:3, 3] = node.offset
            else:
                if not node.fixed:
                    i = node.quaternion_frame_index * 4 + 3
                    L[:3, :3] = quaternion_matrix_batch(frame[i:i + 4])
                else:
                    L[:3, :3] = quaternion_matrix_batch(np.asarray(node.rotation))
                L[:3, 3] = node.offset
            G = L if G is None else np.dot(G, L)
        p = G[:3, 3]
        d = 0.0
        for i in range(3):
            if target[i] is not None:
                d += (target[i] - p[i]) ** 2
        error = math.sqrt(d)
        if min_error > error:
            min_error = error
            best_idx = idx
    return best_idx, min_error


# --------------------------------------------------------------------------
# cluster tree (a18, a20)
# --------------------------------------------------------------------------
def kmeans_lloyd(X, init_centroids, max_iter=300, tol=1e-4):
    """Lloyd iterations from given initial centroids (what sklearn KMeans(lloyd)
    runs after its k-means++ seeding; reference cluster_tree_node_builder.py:87-91
    constructs KMeans with no random_state, so label parity is only defined for
    injected seeds - SURVEY.md section 7).  Convergence: centre shift^2 <= tol *
    mean(var(X)) like sklearn, or labels unchanged.  Empty clusters keep their
    previous centroid."""
    X = np.asarray(X, dtype=np.float64)
    C = np.array(init_centroids, dtype=np.float64, copy=True)
    tol_abs = tol * np.mean(np.var(X, axis=0))
    labels = np.full(len(X), -1)
    for it in range(max_iter):
        d = ((X[:, None, :] - C[None, :, :]) ** 2).sum(axis=2)
        new_labels = np.argmin(d, axis=1)
        newC = C.copy()
        for k in range(len(C)):
            m = new_labels == k
            if m.any():
                newC[k] = X[m].mean(axis=0)
        shift = ((newC - C) ** 2).sum()
        same = np.array_equal(new_labels, labels)
        C, labels = newC, new_labels
        if same or shift <= tol_abs:
            break
    d = ((X[:, None, :] - C[None, :, :]) ** 2).sum(axis=2)
    labels = np.argmin(d, axis=1)
    inertia = float(d[np.arange(len(X)), labels].sum())
    return labels, C, inertia


def heap_prefix(values_with_payload, n):
    """Q4: push everything on a heap, take a prefix of the heap ARRAY
    (reference cluster_tree_node.py:133-138)."""
    q = []
    for item in values_with_payload:
        heapq.heappush(q, item)
    return q[:n]


def injected_seed_rows(n_members, n_sub):
    """The injected-seed rule of the seeded golden builds (oracle/make_golden.py: SeededKMeans):
    initial centres = members at positions len * j // k of the node's member list."""
    return [(n_members * j) // n_sub for j in range(n_sub)]


def _kd_dict(points, dim, depth=0):
    """kdtree.py:44-73: sort by axis depth % dim (list.sort: stable), median len // 2."""
    if len(points) > 1:
        axis = depth % dim
        points = sorted(points, key=lambda p: p[axis])
        median = len(points) // 2
        left, right = points[:median], points[median + 1:]
        return {"point": list(map(float, points[median])), "type": "inner",
                "left": _kd_dict(left, dim, depth + 1) if len(left) > 0 else None,
                "right": _kd_dict(right, dim, depth + 1) if len(right) > 0 else None}
    return {"point": list(map(float, points[0])), "type": "leaf", "left": None, "right": None}


def build_cluster_tree_dict(data, n_sub=4, max_level=4, dim=10, indices=None, depth=0):
    """cluster_tree_node_builder.py:105-170 (use_kd_tree=True) with injected seeds; nested dicts in the
    layout of make_golden.serialize_cluster_tree."""
    data = np.asarray(data, dtype=np.float64)
    dim = min(dim, data.shape[1])
    idx = np.arange(len(data)) if indices is None else np.asarray(indices)
    mean = data[idx].mean(axis=0)
    n = len(idx)
    children = []
    if n > n_sub and n_sub > 1:
        leaf = False
        X = data[idx, :dim]
        labels, _, _ = kmeans_lloyd(X, X[injected_seed_rows(n, n_sub)])
        for j in range(n_sub):
            members = idx[labels == j]
            if len(members) == 0:
                continue
            if depth < max_level:
                children.append(build_cluster_tree_dict(data, n_sub, max_level, dim, members, depth + 1))
            else:
                children.append({"kdtree": _kd_dict(data[members].tolist(), dim)})
    else:
        leaf = True
        children.append({"kdtree": _kd_dict(data[idx].tolist(), dim)})
    return {"mean": mean.tolist(), "leaf": leaf, "children": children}


def build_feature_tree_dict(features, data, n_sub=4, indices=None):
    """feature_cluster_tree.py:61-98 + clustering.py:105-129 with injected seeds; nested dicts in the
    layout of make_golden.serialize_feature_tree."""
    features = np.asarray(features, dtype=np.float64)
    data = np.asarray(data, dtype=np.float64)
    idx = list(range(len(features))) if indices is None else list(indices)
    node = {"mean": data[idx].mean(axis=0).tolist(), "indices": None if indices is None else [int(i) for i in idx], "children": []}
    if len(idx) > 1:
        all_eq = len(idx) < 10 and all(np.array_equal(features[a], features[idx[0]]) for a in idx)
        if all_eq:
            clusters = [[i] for i in idx]
        elif len(idx) > n_sub:
            X = features[idx]
            labels, _, _ = kmeans_lloyd(X, X[injected_seed_rows(len(idx), n_sub)])
            clusters = [[i for i, l in zip(idx, labels) if l == j] for j in range(n_sub)]
            clusters = [c for c in clusters if len(c) > 0]
            if len(clusters) == 1:
                clusters = [[i] for i in idx]
        else:
            clusters = [[i] for i in idx]
        for c in clusters:
            node["children"].append(build_feature_tree_dict(features, data, n_sub, c))
    return node


def _kd_search_dict(kd, cost):
    """kdtree.py:233-250, k = 1."""
    pts = [kd["point"]]
    q = [(cost(kd["point"]), 0)]
    node, depth = kd, 0
    while node is not None and node["type"] == "inner":
        depth += 1
        left, right = node["left"], node["right"]
        if left is not None and right is not None:
            r_d, l_d = cost(right["point"]), cost(left["point"])
            node, least = (left, l_d) if l_d < r_d else (right, r_d)
        elif right is not None:
            node, least = right, cost(right["point"])
        elif left is not None:
            node, least = left, cost(left["point"])
        else:
            node, least = None, np.inf
        if node is not None:
            heapq.heappush(q, (least, depth))
            pts.append(node["point"])
    return q[0][0], pts[q[0][1]]


def cluster_tree_search_dict(root, cost, n_candidates=1):
    """cluster_tree.py:117-149 + cluster_tree_node.py:63-76,113-138 on the nested-dict tree;
    cost(point list) -> float."""
    results, candidates, serial = [], [(np.inf, 0, 0, root)], 0
    while candidates:
        new = []
        for c_idx, (value, _t, _s, node) in enumerate(candidates):
            if not node["leaf"]:
                items = []
                for ci, c in enumerate(node["children"]):
                    serial += 1
                    items.append((cost(c["mean"]), ci, serial, c))
                for i, c in enumerate(heap_prefix(items, n_candidates)):
                    serial += 1
                    heapq.heappush(new, (c[0], i, serial, c[3]))
            else:
                q = []
                for c in node["children"]:
                    heapq.heappush(q, _kd_search_dict(c["kdtree"], cost))
                v, sample = heapq.heappop(q)
                heapq.heappush(results, (v, c_idx, sample))
        candidates = new[:n_candidates]
    if results:
        r = heapq.heappop(results)
        return r[0], r[2]
    return np.inf, root["mean"]


def feature_tree_search_dict(root, data, cost, n_candidates=1):
    """feature_cluster_tree.py:154-187."""
    results, candidates, serial = [], [(np.inf, 0, root)], 0
    while candidates:
        new = []
        for value, _s, node in candidates:
            if len(node["children"]) > 0:
                items = []
                for c in node["children"]:
                    serial += 1
                    items.append((cost(c["mean"]), serial, c))
                for c in heap_prefix(items, n_candidates):
                    heapq.heappush(new, c)
            else:
                serial += 1
                heapq.heappush(results, (value, serial, node))
        candidates = new[:n_candidates]
    if results:
        value, _, node = heapq.heappop(results)
        return value, np.asarray(data)[node["indices"][0]]
    return np.inf, np.asarray(data)[0]


# --------------------------------------------------------------------------
# euclidean PCA features of the cluster-tree build (a19)
# --------------------------------------------------------------------------
END_EFFECTORS2 = ["LeftUpLeg", "RightUpLeg", "RightHand", "Bip01_R_Finger2", "LeftHand", "Bip01_L_Finger2", "RightFoot",
                  "Bip01_R_Toe0", "LeftFoot", "Bip01_L_Toe0"]        # space_partitioning/features.py:32


def euclidean_point_clouds(model, skeleton, S, joint_names=END_EFFECTORS2):
    """construction/cluster_tree_builder.py:293-301 (every integer canonical frame of the back-projected
    spline) + space_partitioning/features.py:133-153 with step = 1: [M, F * len(joints) * 3], frame-major,
    joints in list order."""
    ma = model if isinstance(model, ModelArrays) else ModelArrays(model)
    S = np.atleast_2d(np.asarray(S, dtype=np.float64))
    frames = eval_frames(ma, back_project_coeffs(ma, S), np.arange(ma.F, dtype=np.float64))     # [M, F, P]
    M, F, P = frames.shape
    pos = fk_positions(skeleton, frames.reshape(M * F, P), joint_names)                      # [M*F, J, 3]
    return pos.reshape(M, F * len(joint_names) * 3)


def pca_by_variance(X, variance=0.95):
    """sklearn PCA(n_components=0.95).fit_transform (features.py:125-130): centre, SVD, keep the
    smallest number of components whose explained-variance ratio sums to more than `variance`.
    Component signs are a convention (sklearn's svd_flip); callers must compare sign-invariantly."""
    X = np.asarray(X, dtype=np.float64)
    Xc = X - X.mean(axis=0)
    U, s, Vt = np.linalg.svd(Xc, full_matrices=False)
    ev = s ** 2 / (len(X) - 1)
    ratio = ev / ev.sum()
    k = int(np.searchsorted(np.cumsum(ratio), variance, side="right") + 1)
    return Xc @ Vt[:k].T, ratio[:k]
