"""Features of the cluster-tree build on the GPU.

Reference: `ClusterTreeBuilder._extract_features` (construction/cluster_tree_builder.py:266-291):
either the spatial part of the latent vectors (`data[:, :K]`) or "euclidean PCA" features - every
sample back-projected, every integer canonical frame evaluated (`_back_project`, :293-301), the
global positions of ten end-effector joints per frame (space_partitioning/features.py:32,133-153)
flattened to one row per sample and reduced by `sklearn.decomposition.PCA(n_components=0.95)`
(features.py:125-130).

Here the point clouds come from the staged kernels K2 -> K3 -> K4 chunk by chunk (the motion
[M, F, P] is never held as a whole: 15.8 GB at M = 1M for the S model), the PCA is a covariance
eigen-decomposition on the device: the covariance (the 4.5 TFLOP part at M = 1M) by the hand-written tcgen05 kernel K11
(csrc/mgb_syrk.cu), the small d x d symmetric eigen-decomposition by cuSOLVER in float64, the projection onto the k kept
components by a library GEMM (2 M d k flop, bound by the one read of X).  Component SIGNS are sklearn's svd_flip convention in the reference
and unspecified here; everything downstream (k-means on the features) only sees distances, which do
not depend on them.
"""
import numpy as np

from . import engine

END_EFFECTORS = ["Hips", "RightHand", "LeftHand", "RightFoot", "LeftFoot"]                      # features.py:31
END_EFFECTORS2 = ["LeftUpLeg", "RightUpLeg", "RightHand", "Bip01_R_Finger2", "LeftHand", "Bip01_L_Finger2", "RightFoot",
                  "Bip01_R_Toe0", "LeftFoot", "Bip01_L_Toe0"]                                   # features.py:32
LEN_END_EFFECTOR = 3


def map_samples_to_euclidean_space(dmodel, samples, joint_names=END_EFFECTORS2, chunk=32768, out=None):
    """latent samples [M, D] (device tensor or array) -> point clouds [M, F * len(joints) * 3] float32 on
    the device: K2 back-projection, K3 spline evaluation at t = 0 .. F-1, K4 forward kinematics."""
    torch = engine._torch()
    S = dmodel._f32(samples)
    M = int(S.shape[0])
    F, J = dmodel.F, len(joint_names)
    if out is None:
        out = torch.empty((M, F * J * LEN_END_EFFECTOR), dtype=torch.float32, device=dmodel.torch_device)
    times = np.arange(F, dtype=np.float64)
    chunk = max(1, min(chunk, M))
    coeffs = torch.empty((chunk, dmodel.Nb, dmodel.P), dtype=torch.float32, device=dmodel.torch_device)
    frames = torch.empty((chunk, F, dmodel.P), dtype=torch.float32, device=dmodel.torch_device)
    for a in range(0, M, chunk):
        b = min(M, a + chunk)
        m = b - a
        dmodel.backproject(S[a:b], out=coeffs[:m])
        dmodel.eval_frames(coeffs[:m], times, out=frames[:m])
        dmodel.fk(frames[:m], joint_names, out=out[a:b].view(m, F, J, LEN_END_EFFECTOR))
    return out


def pca_by_variance(X, variance=0.95, sharded=False, group=None):
    """PCA(n_components=variance).fit_transform on the device: X [M, d] float32 -> (features [M, k] float32,
    explained-variance ratios [k] float64).  Covariance accumulated in float64 from fp32 chunk products,
    symmetric eigen-decomposition in float64.  sharded=True: X holds this rank's rows only; the row count,
    the column sums and the d x d covariance are all-reduced, every rank gets the
    same basis and projects its own rows."""
    torch = engine._torch()
    M, d = int(X.shape[0]), int(X.shape[1])
    total = torch.tensor([float(M)], dtype=torch.float64, device=X.device)
    colsum = X.double().sum(dim=0)
    if sharded:
        import torch.distributed as dist
        dist.all_reduce(total, group=group)
        dist.all_reduce(colsum, group=group)
    mean = colsum / total
    # K11 (mgb_syrk_upper): the 2 M d^2 contraction on the tensor cores (3xTF32, float64 result), upper triangle
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    mean32 = mean.float().contiguous()
    assert X.dtype == torch.float32 and X.stride(1) == 1
    upper = torch.zeros((d, d), dtype=torch.float64, device=X.device)
    with torch.cuda.device(X.device):
        _lib.check(lib.mgb_syrk_upper(C.c_void_p(X.data_ptr()), M, d, int(X.stride(0)), C.c_void_p(mean32.data_ptr()),
                                      C.c_void_p(upper.data_ptr()), C.c_void_p(torch.cuda.current_stream(X.device).cuda_stream)))
    cov = upper + upper.t() - torch.diag(torch.diagonal(upper))
    step = 131072
    if sharded:
        dist.all_reduce(cov, group=group)
    cov /= max(float(total.item()) - 1.0, 1.0)
    ev, V = torch.linalg.eigh(cov)
    ev, V = ev.flip(0).clamp(min=0), V.flip(1)
    ratio = ev / ev.sum()
    k = int(torch.searchsorted(torch.cumsum(ratio, 0), torch.tensor(variance, dtype=torch.float64, device=X.device), right=True).item()) + 1
    k = min(k, d)
    W = V[:, :k].float().contiguous()
    feats = torch.empty((M, k), dtype=torch.float32, device=X.device)
    for a in range(0, M, step):
        feats[a:a + step] = (X[a:a + step] - mean32) @ W
    return feats, ratio[:k].cpu().numpy()


def map_to_euclidean_pca(dmodel, samples, joint_names=END_EFFECTORS2, variance=0.95):
    """features.map_motions_to_euclidean_pca (features.py:125-130) from latent samples."""
    clouds = map_samples_to_euclidean_space(dmodel, samples, joint_names)
    return pca_by_variance(clouds, variance)
