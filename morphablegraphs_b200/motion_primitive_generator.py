"""`MotionPrimitiveGenerator` with the reference's interface
(reference motion_generator/motion_primitive_generator.py:47-261) whose
sample-and-evaluate loop runs on the GPU.

Drop-in points kept: `generate_constrained_sample`, `_get_best_fit_sample_using_gmm`,
`evaluate_samples_using_constraints`, `_get_best_fit_sample_using_cluster_tree`,
the `objective` attribute, and the bookkeeping on the constraints object
(`min_error`, `evaluations`, first-minimum tie-break, returned row is a row of
the input `samples`).  The scipy local optimiser stays on the host, seeded by
the arg-min.
"""
import numpy as np

from . import engine, sharding
from .objective_functions import obj_spatial_error_sum, _device_constraints

SAMPLING_MODE_RANDOM = "random_discrete"
SAMPLING_MODE_CLUSTER_TREE_SEARCH = "cluster_tree_search"
SAMPLING_MODE_RANDOM_SPLINE = "random_spline"

# hot-path keys of reference algorithm_configuration.py:141-257
DEFAULT_ALGORITHM_CONFIG = {
    "n_random_samples": 100,
    "constrained_sampling_mode": SAMPLING_MODE_RANDOM,
    "n_cluster_search_candidates": 2,
    "use_local_coordinates": True,
    "local_optimization_mode": "none",
    "local_optimization_settings": {"start_error_threshold": 0.0, "error_scale_factor": 1.0,
                                    "quality_scale_factor": 0.1},
    "use_constraints": True,
    "use_transition_model": False,
    "verbose": False,
}


class MotionPrimitiveGenerator(object):
    def __init__(self, action_constraints=None, algorithm_config=None, skeleton=None, seed=None):
        self._action_constraints = action_constraints
        self.set_algorithm_config(dict(DEFAULT_ALGORITHM_CONFIG, **(algorithm_config or {})))
        self.skeleton = skeleton
        if action_constraints is not None:
            self.action_name = action_constraints.action_name
            self._motion_state_graph = action_constraints.motion_state_graph
            self.skeleton = self._motion_state_graph.skeleton
        self.objective = obj_spatial_error_sum
        self._rng = np.random.RandomState(seed)
        self.numerical_minimizer = None

    def set_algorithm_config(self, algorithm_config):
        self._algorithm_config = algorithm_config
        self.n_random_samples = algorithm_config["n_random_samples"]
        self.verbose = algorithm_config.get("verbose", False)
        self.use_constraints = algorithm_config.get("use_constraints", True)
        self.local_optimization_mode = algorithm_config.get("local_optimization_mode", "none")
        self._settings = algorithm_config.get("local_optimization_settings", {})
        self.optimization_start_error_threshold = self._settings.get("start_error_threshold", 0.0)
        self.use_transition_model = algorithm_config.get("use_transition_model", False)
        self.constrained_sampling_mode = algorithm_config.get("constrained_sampling_mode", SAMPLING_MODE_RANDOM)
        self.n_cluster_search_candidates = int(algorithm_config.get("n_cluster_search_candidates", 2))
        self.use_local_coordinates = algorithm_config.get("use_local_coordinates", True)

    # ------------------------------------------------------------------ :126-161
    def generate_constrained_sample(self, graph_node, in_mp_constraints, prev_mp_name="", prev_frames=None, prev_parameters=None):
        if self.use_local_coordinates:
            prev_frames_copy = None
            mp_constraints = in_mp_constraints.transform_constraints_to_local_cos()
        else:
            mp_constraints = in_mp_constraints
            prev_frames_copy = prev_frames
        if self.constrained_sampling_mode == SAMPLING_MODE_RANDOM_SPLINE:
            sample = self._get_best_fit_sample_using_mgrd(graph_node, mp_constraints)
        elif self.constrained_sampling_mode == SAMPLING_MODE_CLUSTER_TREE_SEARCH and graph_node.cluster_tree is not None:
            sample = self._get_best_fit_sample_using_cluster_tree(graph_node, mp_constraints, prev_frames_copy)
        else:
            sample = self._get_best_fit_sample_using_gmm(graph_node, mp_constraints, prev_mp_name, prev_frames_copy, prev_parameters)
        if self._is_optimization_required(mp_constraints):
            sample = self._optimize_parameters_numerically(sample, graph_node, mp_constraints, prev_frames_copy)
        in_mp_constraints.min_error = mp_constraints.min_error
        in_mp_constraints.evaluations = mp_constraints.evaluations
        return sample

    def _is_optimization_required(self, mp_constraints):
        return (mp_constraints.use_local_optimization and not self.use_transition_model and
                self.numerical_minimizer is not None and mp_constraints.min_error >= self.optimization_start_error_threshold)

    def _optimize_parameters_numerically(self, initial_guess, graph_node, mp_constraints, prev_frames):
        """:179-190 - host scipy optimiser injected by the caller (`numerical_minimizer`); untouched."""
        data = (graph_node, mp_constraints, prev_frames, self._settings["error_scale_factor"],
                self._settings["quality_scale_factor"], 1.0)
        self.numerical_minimizer.set_objective_function_parameters(data)
        return self.numerical_minimizer.run(initial_guess=initial_guess)

    # ------------------------------------------------------------------ :192-206
    def _get_best_fit_sample_using_gmm(self, graph_node, mp_constraints, prev_mp_name="", prev_frames=None, prev_parameters=None):
        """Sample N and score them in ONE fused launch (mgb_sample_score: K1 + K6, candidates never
        leave the registers), arg-min on the device; the winning row is regenerated from its
        global index (Philox is keyed by it) and only that row and its cost come back to the host."""
        dc = _device_constraints(graph_node, mp_constraints, prev_frames)
        dm = dc.dmodel
        seed = int(self._rng.randint(0, 2 ** 31 - 1))
        n = int(self.n_random_samples)
        obj = self.objective
        # one process per GPU: every rank holds the same generator state (same seed), scores its own
        # index range of the job and the per-rank bests are exchanged (sharding.py)
        rank, world = 0, 1
        if sharding.is_distributed():
            import torch.distributed as dist
            rank, world = dist.get_rank(), dist.get_world_size()
        base, n_mine = sharding.shard_range(n, rank, world)
        counts = dm.component_counts(n, seed)
        res = dc.sample_score(n_mine, seed=seed, counts=counts, n_total=n, index_base=base, k=1,
                              error_scale=getattr(obj, "error_scale", 1.0), quality_scale=getattr(obj, "quality_scale", 0.0),
                              magnitude_scale=getattr(obj, "magnitude_scale", 0.0))
        peer = sharding.peer_exchange(dm.torch_device) if world > 1 else None
        if peer is not None:
            # per-rank bests exchanged and reduced on the device over NVLink peer memory (mgb_exchange_best)
            oc, oi = peer.exchange(res["best_cost"], res["best_idx"])
            best_idx, min_error = int(oi.item()), float(oc.item())
            peer.check()
        else:
            best_idx = int(res["best_idx"][0].item())
            min_error = float(res["best_cost"][0].item())
            if world > 1:
                min_error, best_idx = sharding.exchange_best(min_error, best_idx, device=dm.torch_device)
        mp_constraints.min_error = min_error
        mp_constraints.evaluations += n
        best = dm.sample(1, seed=seed, index_base=best_idx, counts=counts, n_total=n)
        return best[0].cpu().numpy().astype(np.float64)

    # ------------------------------------------------------------------ :167-177
    def _get_best_fit_sample_using_mgrd(self, graph_node, mp_constraints):
        """The reference's batched slot (samples [N, D] -> scores [N] -> np.argmin), scored by the engine
        (mgrd_sample_filter.py of this package)."""
        from .mgrd_sample_filter import MGRDSampleFilter, motion_primitive_get_random_samples
        samples = motion_primitive_get_random_samples(graph_node, self.n_random_samples)
        scores = MGRDSampleFilter.score_samples_cartesian(graph_node, samples, mp_constraints)
        best_idx = int(np.argmin(scores))
        mp_constraints.min_error = float(scores[best_idx])
        mp_constraints.evaluations += len(samples)
        return samples[best_idx]

    # ------------------------------------------------------------------ :220-228
    def _get_best_fit_sample_using_cluster_tree(self, graph_node, constraints, prev_frames, n_candidates=-1):
        n_candidates = self.n_cluster_search_candidates if n_candidates < 1 else n_candidates
        data = graph_node, constraints, prev_frames
        distance, s = graph_node.search_best_sample(self.objective, data, n_candidates)
        constraints.min_error = distance
        return np.array(s)

    # ------------------------------------------------------------------ :230-261
    def evaluate_samples_using_constraints(self, samples, mp_node, constraints, prev_frames):
        """Picks the best of the given samples (host array [N, D], what the reference's sampler
        returned).  Returns (samples[best_idx], min_error); first minimum wins."""
        samples = np.asarray(samples)
        if samples.dtype not in (np.float32, np.float64):
            samples = samples.astype(np.float64)
        samples_c = np.ascontiguousarray(samples)
        dc = _device_constraints(mp_node, constraints, prev_frames)
        obj = self.objective
        res = dc.score_host(samples_c, k=1, error_scale=getattr(obj, "error_scale", 1.0),
                            quality_scale=getattr(obj, "quality_scale", 0.0), magnitude_scale=getattr(obj, "magnitude_scale", 0.0))
        best_idx = int(res["best_idx"][0])
        min_error = float(res["best_cost"][0])
        constraints.min_error = min_error
        constraints.evaluations += len(samples)
        return samples[best_idx], min_error
