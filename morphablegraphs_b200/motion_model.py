"""Host-side mirror of the reference's model classes for the hot path:
`MotionPrimitive` (reference motion_model/motion_primitive.py), `MotionSpline`
(motion_spline.py), `MotionPrimitiveModelWrapper` (motion_primitive_wrapper.py,
legacy path - the only one loadable without mgrd) and `MotionStateGraphNode`
(motion_state_graph_node.py).  Same method names and argument meaning; the
arithmetic runs on the GPU through `engine.DeviceModel`.
"""
import json

import numpy as np

from . import engine

B_SPLINE_DEGREE = 3


class GaussianMixtureParams(object):
    """The attributes the reference sets on sklearn's GaussianMixture
    (motion_primitive.py:135-143).  `precisions_cholesky_` is computed on first use
    for host-side consumers; the device copy is computed inside the library."""

    def __init__(self, weights, means, covars):
        self.weights_ = np.array(weights, dtype=np.float64)
        self.means_ = np.array(means, dtype=np.float64)
        self.covariances_ = np.array(covars, dtype=np.float64)
        self.converged_ = True
        self.covariance_type = "full"
        self.n_components = len(self.weights_)
        self.n_dims = self.means_.shape[1]
        self._prec_chol = None
        self._owner = None

    @property
    def precisions_cholesky_(self):
        if self._prec_chol is None:
            out = np.empty_like(self.covariances_)
            eye = np.eye(self.n_dims)
            for g in range(self.n_components):
                L = np.linalg.cholesky(self.covariances_[g])
                out[g] = np.linalg.solve(L, eye).T
            self._prec_chol = out
        return self._prec_chol

    def sample(self, n_samples=1):
        s = self._owner.sample_low_dimensional_vector(n_samples)
        return s, None

    def score_samples(self, X):
        dm = self._owner.device_model()
        return dm.logpdf(np.atleast_2d(np.asarray(X, dtype=np.float32))).cpu().numpy().astype(np.float64)

    def score(self, X):
        return float(np.mean(self.score_samples(X)))


class MotionSpline(object):
    """reference motion_spline.py:32-107: coefficients [Nb, P] + knots + time function."""

    def __init__(self, canonical_motion_coeffs, time_function, knots, semantic_annotation=None,
                 low_dimensional_parameters=None, owner=None):
        self.low_dimensional_parameters = low_dimensional_parameters
        self.time_function = time_function
        self.buffered_frames = None
        self.coeffs = canonical_motion_coeffs
        self.knots = knots
        self.semantic_annotation = semantic_annotation
        self.n_pose_parameters = len(canonical_motion_coeffs[0])
        self.n_max_frame = knots[-1]
        self._owner = owner

    def _eval(self, times):
        dm = self._owner.device_model()
        fr = dm.eval_frames(np.asarray(self.coeffs, dtype=np.float32)[None], times)
        return fr[0].cpu().numpy().astype(np.float64)

    def evaluate(self, canonical_t):
        """motion_spline.py:89-92"""
        if np.ndim(canonical_t) == 0:
            return self._eval([float(canonical_t)])[0]
        return self._eval(canonical_t)

    def get_motion_vector(self, step_size=None):
        """motion_spline.py:71-86 (step_size path raises in the reference under numpy>=1.18, Q3)"""
        if step_size is not None:
            raise NotImplementedError("step_size re-sampling is broken in the reference itself (SURVEY.md Q3)")
        return self._eval(self.time_function)

    def get_buffered_motion_vector(self):
        if self.buffered_frames is None:
            self.buffered_frames = self.get_motion_vector()
        return self.buffered_frames


class MotionPrimitive(object):
    """reference motion_primitive.py:41-380, legacy JSON."""

    def __init__(self, filename=None):
        self.filename = filename
        self.name = ""
        self.gaussian_mixture_model = None
        self.s_pca = dict()
        self.t_pca = dict()
        self.n_canonical_frames = 0
        self.translation_maxima = np.array([1.0, 1.0, 1.0])
        self.has_time_parameters = True
        self.has_semantic_parameters = False
        self.animated_joints = []
        self.skeleton = None
        self._device_models = {}
        if filename is not None:
            with open(filename, "r") as f:
                self._initialize_from_json(json.load(f))

    def _initialize_from_json(self, data):
        if "name" in data:
            self.name = data["name"]
        self.n_canonical_frames = data["n_canonical_frames"]
        self.canonical_time_range = np.arange(0, self.n_canonical_frames)
        self.gaussian_mixture_model = GaussianMixtureParams(data["gmm_weights"], data["gmm_means"], data["gmm_covars"])
        self.gaussian_mixture_model._owner = self
        self.translation_maxima = np.array(data["translation_maxima"], dtype=np.float64)
        self.s_pca = dict()
        self.s_pca["eigen_vectors"] = np.transpose(np.array(data["eigen_vectors_spatial"], dtype=np.float64))
        self.s_pca["mean_vector"] = np.array(data["mean_spatial_vector"], dtype=np.float64)
        self.s_pca["n_basis"] = int(data["n_basis_spatial"])
        self.s_pca["n_dim"] = int(data["n_dim_spatial"])
        self.s_pca["n_components"] = len(self.s_pca["eigen_vectors"].T)
        self.s_pca["knots"] = np.asarray(data["b_spline_knots_spatial"], dtype=np.float64)
        if "eigen_vectors_time" in data:
            self.t_pca = dict()
            self.t_pca["eigen_vectors"] = np.array(data["eigen_vectors_time"])
            self.t_pca["mean_vector"] = np.array(data["mean_time_vector"])
            self.t_pca["n_basis"] = int(data["n_basis_time"])
            self.t_pca["n_dim"] = 1
            self.t_pca["n_components"] = len(self.t_pca["eigen_vectors"].T)
            self.t_pca["knots"] = np.asarray(data["b_spline_knots_time"])
            self.has_time_parameters = True
        else:
            self.has_time_parameters = False
            self.t_pca = {"n_components": 0}
        if "animated_joints" in data:
            self.animated_joints = data["animated_joints"]

    # -- device residency ------------------------------------------------------
    def device_model(self, skeleton=None, device=None):
        skeleton = skeleton if skeleton is not None else self.skeleton
        import torch
        dev = torch.cuda.current_device() if device is None else int(device)
        key = (id(skeleton), dev)
        if key not in self._device_models:
            self._device_models[key] = engine.DeviceModel(self, skeleton, dev)
        return self._device_models[key]

    # -- reference API -----------------------------------------------------------
    def sample_low_dimensional_vector(self, n_samples=1, seed=None):
        """motion_primitive.py:182-189 -> [n, D] float64"""
        if seed is None:
            seed = int(np.random.randint(0, 2 ** 31 - 1))
        return self.device_model().sample(int(n_samples), seed=seed).cpu().numpy().astype(np.float64)

    def back_project_spatial_coeffs(self, alpha):
        """motion_primitive.py:236-256 -> [Nb, P]"""
        s = np.zeros((1, self.gaussian_mixture_model.n_dims), dtype=np.float32)
        s[0, :self.s_pca["n_components"]] = np.asarray(alpha, dtype=np.float32)
        return self.device_model().backproject(s)[0].cpu().numpy().astype(np.float64)

    # -- time warping: host, once per CHOSEN sample; the frames are evaluated on the device ----------
    def _mean_temporal(self):
        """motion_primitive.py:259-266"""
        import scipy.interpolate as si
        return si.splev(self.canonical_time_range, (self.t_pca["knots"], self.t_pca["mean_vector"], 3))

    def _back_transform_gamma_to_canonical_time_function(self, gamma):
        """motion_primitive.py:289-302: t(t') = cumulative sum of exp(mean + harmonics . gamma) over the canonical frames, - 1."""
        import scipy.interpolate as si
        mean_t = self._mean_temporal()
        eigen_coefs = list(zip(*self.t_pca["eigen_vectors"]))
        t_eigen_discrete = np.array([si.splev(self.canonical_time_range, (self.t_pca["knots"], c, 3)) for c in eigen_coefs]).T
        steps = np.exp(mean_t[:self.n_canonical_frames] + t_eigen_discrete[:self.n_canonical_frames].dot(np.asarray(gamma, dtype=np.float64)))
        return np.cumsum(steps) - 1.0

    def _invert_canonical_to_sample_time_function(self, canonical_time_function, speed=1.0):
        """motion_primitive.py:304-318: spline through (t, t') inverted and sampled at the integer sample frames.  The
        reference passes a FLOAT sample count to np.linspace, which numpy >= 1.18 refuses (quirk Q3); the count is the
        integer older numpy truncated it to."""
        import scipy.interpolate as si
        x_sample = np.arange(self.n_canonical_frames)
        sample_time_spline = si.splrep(canonical_time_function, x_sample, w=None, k=3)
        num = int(np.round(canonical_time_function[-2]) * (1.0 / speed))
        frames = np.linspace(1, stop=canonical_time_function[-2], num=num)
        sample_time_function = si.splev(frames, sample_time_spline)
        sample_time_function = np.insert(sample_time_function, 0, 0)
        return np.insert(sample_time_function, len(sample_time_function), self.n_canonical_frames - 1)

    def back_project_time_function(self, gamma, speed=1.0):
        """motion_primitive.py:268-287 (smooth_time_parameters is False in the reference, :72)."""
        return self._invert_canonical_to_sample_time_function(self._back_transform_gamma_to_canonical_time_function(gamma), speed)

    def back_project(self, s, use_time_parameters=True, speed=1.0):
        """motion_primitive.py:206-234; the scored path always passes use_time_parameters=False
        (objective_functions.py:159).  With time parameters the chosen sample's time function t'(t) is computed on
        the host (back_project_time_function) and the motion is evaluated at those times by K3."""
        s = np.ravel(np.asarray(s, dtype=np.float64))
        coeffs = self.back_project_spatial_coeffs(s[:self.s_pca["n_components"]])
        if use_time_parameters and self.has_time_parameters:
            time_function = self.back_project_time_function(s[self.s_pca["n_components"]:], speed)
            return MotionSpline(coeffs, time_function, self.s_pca["knots"], None, low_dimensional_parameters=s, owner=self)
        time_function = np.linspace(0, self.n_canonical_frames, int(self.n_canonical_frames * (1.0 / speed)))
        return MotionSpline(coeffs, time_function, self.s_pca["knots"], None, low_dimensional_parameters=s, owner=self)

    def get_n_canonical_frames(self):
        return self.n_canonical_frames

    def get_n_spatial_components(self):
        return self.s_pca["n_components"]

    def get_n_time_components(self):
        return self.t_pca.get("n_components", 0)


class MotionPrimitiveModelWrapper(object):
    """reference motion_primitive_wrapper.py:41-375, legacy branch (`mgrd == False`)."""

    def __init__(self):
        self.motion_primitive = None
        self.keyframes = dict()
        self.mgrd = False

    def _initialize_from_json(self, mgrd_skeleton, data, animated_joints=None, use_mgrd_mixture_model=False, scale=None):
        if "keyframes" in data:
            self.keyframes = data["keyframes"]
        if "tspm" in data:      # mgrd-format JSON converted to the legacy model (wrapper.py:87-115)
            sspm, gmm = data["sspm"], data["gmm"]
            data = {"eigen_vectors_spatial": sspm["eigen"], "mean_spatial_vector": sspm["mean"],
                    "n_basis_spatial": sspm["n_coeffs"], "n_dim_spatial": sspm["n_dims"],
                    "b_spline_knots_spatial": sspm["knots"], "animated_joints": sspm["animated_joints"],
                    "gmm_covars": gmm["covars"], "gmm_means": gmm["means"], "gmm_weights": gmm["weights"],
                    "n_canonical_frames": int(max(data["tspm"]["knots"]) + 1), "translation_maxima": [1, 1, 1]}
        mp = MotionPrimitive(None)
        mp._initialize_from_json(data)
        self.motion_primitive = mp

    def _load_from_file(self, mgrd_skeleton, file_name, animated_joints=None, use_mgrd_mixture_model=False, scale=None):
        with open(file_name, "r") as f:
            self._initialize_from_json(mgrd_skeleton, json.load(f), animated_joints, use_mgrd_mixture_model, scale)

    def set_skeleton(self, skeleton):
        self.motion_primitive.skeleton = skeleton

    def sample_low_dimensional_vector(self):
        return self.motion_primitive.sample_low_dimensional_vector(1)

    def sample_low_dimensional_vectors(self, n_samples=1, seed=None):
        return self.motion_primitive.sample_low_dimensional_vector(n_samples, seed=seed)

    def back_project(self, s_vec, use_time_parameters=True, speed=1.0):
        return self.motion_primitive.back_project(s_vec, use_time_parameters, speed)

    def get_n_canonical_frames(self):
        return self.motion_primitive.n_canonical_frames

    def get_n_spatial_components(self):
        return self.motion_primitive.get_n_spatial_components()

    def get_n_time_components(self):
        return self.motion_primitive.get_n_time_components()

    def get_gaussian_mixture_model(self):
        return self.motion_primitive.gaussian_mixture_model

    def get_spatial_eigen_vectors(self, joints=None, frame_idx=-1):
        return self.motion_primitive.s_pca["eigen_vectors"].T


class MotionStateGraphNode(MotionPrimitiveModelWrapper):
    """reference motion_state_graph_node.py:43-142: primitive + cluster tree."""

    def __init__(self, motion_state_group=None):
        super(MotionStateGraphNode, self).__init__()
        self.motion_state_group = motion_state_group
        self.outgoing_edges = dict()
        self.node_type = "standard"
        self.action_name = None
        self.name = None
        self.cluster_tree = None

    def search_best_sample(self, obj, data, n_candidates=2):
        """motion_state_graph_node.py:119-142"""
        if self.cluster_tree is not None:
            return self.cluster_tree.find_best_example_excluding_search_candidates(obj, data, n_candidates)
        return np.inf, None
