"""`MotionPrimitiveGenerator` with the reference's interface
(reference motion_generator/motion_primitive_generator.py:47-261) whose
sample-and-evaluate loop runs on the GPU.

Drop-in points kept: `generate_constrained_motion_spline`, `generate_constrained_sample`,
`generate_random_sample`, `_get_best_fit_sample_using_gmm`,
`evaluate_samples_using_constraints`, `_get_best_fit_sample_using_cluster_tree`,
the `objective` attribute, and the bookkeeping on the constraints object
(`min_error`, `evaluations`, `time`, first-minimum tie-break, returned row is a row of
the input `samples`).  The local optimiser is whatever the caller injects as
`numerical_minimizer`: the reference's scipy LeastSquares on the host, or this
package's batched device Levenberg-Marquardt (optimization.py).
"""
import time

import numpy as np

from . import sharding
from .objective_functions import obj_spatial_error_sum, objective_scales, _device_constraints

SAMPLING_MODE_RANDOM = "random_discrete"
SAMPLING_MODE_CLUSTER_TREE_SEARCH = "cluster_tree_search"
SAMPLING_MODE_RANDOM_SPLINE = "random_spline"
SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE = "keyframe_pose"

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
        self.action_name = None
        self.prev_action_name = None
        self._motion_state_graph = None
        if action_constraints is not None:
            self.action_name = action_constraints.action_name
            self.prev_action_name = getattr(action_constraints, "prev_action_name", None)
            self._motion_state_graph = action_constraints.motion_state_graph
            self.skeleton = self._motion_state_graph.skeleton
        self.objective = obj_spatial_error_sum
        self._rng = np.random.RandomState(seed)
        self._rng_shared = seed is not None     # an explicit seed is the same on every rank by construction
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

    # ------------------------------------------------------------------ :83-124
    def generate_constrained_motion_spline(self, mp_constraints, prev_graph_walk):
        """Calls generate_constrained_sample and back-projects the result to a MotionSpline; sets
        `mp_constraints.time` (one of the bookkeeping fields a drop-in keeps)."""
        node_key = (self.action_name, mp_constraints.motion_primitive_name)
        if len(prev_graph_walk.steps) > 0:
            prev_mp_name = prev_graph_walk.steps[-1].node_key[1]
            prev_parameters = prev_graph_walk.steps[-1].parameters
        else:
            prev_mp_name = ""
            prev_parameters = None
        start = time.time()
        if self.use_constraints and len(mp_constraints.constraints) > 0:
            graph_node = self._motion_state_graph.nodes[node_key]
            parameters = self.generate_constrained_sample(graph_node, mp_constraints, prev_mp_name,
                                                          prev_graph_walk.get_quat_frames(), prev_parameters)
        else:  # no constraints were given
            parameters = self.generate_random_sample(node_key, prev_mp_name, prev_parameters)
        mp_constraints.time = time.time() - start
        motion_spline = self._motion_state_graph.nodes[node_key].back_project(parameters, use_time_parameters=False)
        return motion_spline, parameters

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

    # ------------------------------------------------------------------ :179-190
    def _optimize_parameters_numerically(self, initial_guess, graph_node, mp_constraints, prev_frames):
        """Pose constraints are dropped first and the residuals are normalised by the start error, as the
        reference does; the minimiser itself is the caller's (`numerical_minimizer`)."""
        mp_constraints.constraints = [c for c in mp_constraints.constraints if c.constraint_type != SPATIAL_CONSTRAINT_TYPE_KEYFRAME_POSE]
        if len(mp_constraints.constraints) == 0:
            return initial_guess
        data = (graph_node, mp_constraints, prev_frames, self._settings["error_scale_factor"],
                self._settings["quality_scale_factor"], 1.0)
        error_sum = max(abs(np.sum(self.numerical_minimizer._objective_function(initial_guess, data))), 1.0)
        data = (graph_node, mp_constraints, prev_frames, self._settings["error_scale_factor"],
                self._settings["quality_scale_factor"], error_sum)
        self.numerical_minimizer.set_objective_function_parameters(data)
        return self.numerical_minimizer.run(initial_guess=initial_guess)

    # ------------------------------------------------------------------ :208-218
    def generate_random_sample(self, node_key, prev_mp_name="", prev_parameters=None):
        """One unconstrained draw (what motion_generator.py:233 falls back to when a step has no constraints)."""
        nodes = self._motion_state_graph.nodes
        prev_key = (self.prev_action_name, prev_mp_name)
        if (self.use_transition_model and prev_parameters is not None and prev_key in nodes
                and nodes[prev_key].has_transition_model(node_key)):
            return nodes[prev_key].predict_parameters(node_key, prev_parameters)
        return nodes[node_key].sample_low_dimensional_vector()

    def _predict_gmm(self, mp_name, prev_mp_name, prev_parameters):
        to_key = (self.action_name, mp_name)
        return self._motion_state_graph.nodes[(self.prev_action_name, prev_mp_name)].predict_gmm(to_key, prev_parameters)

    def _step_seed(self):
        """Seed of one sample-and-evaluate step.  Under torch.distributed every rank must score the SAME job
        (same multinomial split, same Philox key): a generator built without an explicit seed synchronises its
        stream once, by broadcasting a seed from rank 0 (collective: every rank reaches its first step together)."""
        if sharding.is_distributed() and not self._rng_shared:
            self._rng = np.random.RandomState(sharding.broadcast_seed(int(self._rng.randint(0, 2 ** 31 - 1))))
            self._rng_shared = True
        return int(self._rng.randint(0, 2 ** 31 - 1))

    # ------------------------------------------------------------------ :192-206
    def _get_best_fit_sample_using_gmm(self, graph_node, mp_constraints, prev_mp_name="", prev_frames=None, prev_parameters=None):
        """Sample N and score them in ONE fused launch (mgb_sample_score: K1 + K6, candidates never
        leave the registers), arg-min on the device; the winning row is re-drawn by the kernel's last CTA from its
        global index (Philox is keyed by it) and only that row and its cost come back to the host."""
        if self.use_transition_model and prev_parameters is not None:
            # :195-197 - candidates from the transition model's predicted mixture (host sampler of the caller's
            # model object), scored by the same kernels through the host-sample entry
            samples = self._predict_gmm(mp_constraints.motion_primitive_name, prev_mp_name, prev_parameters).sample(self.n_random_samples)
            if isinstance(samples, tuple):      # sklearn >= 0.18 returns (X, component labels)
                samples = samples[0]
            return self.evaluate_samples_using_constraints(samples, graph_node, mp_constraints, prev_frames)[0]
        dc = _device_constraints(graph_node, mp_constraints, prev_frames)
        dm = dc.dmodel
        error_scale, quality_scale, magnitude_scale = objective_scales(self.objective)
        seed = self._step_seed()
        n = int(self.n_random_samples)
        # one process per GPU: every rank holds the same generator state (same seed), scores its own
        # index range of the job and the per-rank bests are exchanged (sharding.py)
        rank, world = 0, 1
        if sharding.is_distributed():
            import torch.distributed as dist
            rank, world = dist.get_rank(), dist.get_world_size()
        counts = dm.component_counts(n, seed)
        if world == 1:
            # one call: launch, winner re-drawn by the kernel, (cost, index, row) back in one copy
            min_error, best_idx, best = dc.sample_score_best(n, seed=seed, counts=counts, error_scale=error_scale,
                                                             quality_scale=quality_scale, magnitude_scale=magnitude_scale)
            mp_constraints.min_error = min_error
            mp_constraints.evaluations += n
            return best
        base, n_mine = sharding.shard_range(n, rank, world)
        peer = sharding.peer_exchange(dm.torch_device)
        if peer is not None:
            # one launch per rank: the CTA that folds this rank's minima exchanges the 16-byte record with the peers over NVLink
            # and re-draws the job-wide winner (every rank gets the same row)
            min_error, best_idx, best = dc.sample_score_best(n_mine, seed=seed, counts=counts, n_total=n, index_base=base, error_scale=error_scale,
                                                             quality_scale=quality_scale, magnitude_scale=magnitude_scale, peer=peer)
            peer.check()
            mp_constraints.min_error = min_error
            mp_constraints.evaluations += n
            return best
        res = dc.sample_score(n_mine, seed=seed, counts=counts, n_total=n, index_base=base, k=1,
                              error_scale=error_scale, quality_scale=quality_scale, magnitude_scale=magnitude_scale)
        min_error, best_idx = sharding.exchange_best(float(res["best_cost"][0].item()), int(res["best_idx"][0].item()), device=dm.torch_device)
        mp_constraints.min_error = min_error
        mp_constraints.evaluations += n
        best = dm.sample(1, seed=seed, index_base=best_idx, counts=counts, n_total=n)
        return best[0].cpu().numpy().astype(np.float64)

    # ------------------------------------------------------------------ graph_walk_planner.py:216-226, mg_state_planner.py:946-968
    def generate_constrained_samples(self, options):
        """Cross-option batching.  `options`: list of `(graph_node, in_mp_constraints, prev_frames)`
        - the outgoing edges a planner scores one after another with `generate_constrained_sample`.  All of them are
        sampled, scored and reduced in ONE launch; returns the list of best latent vectors and leaves `min_error` /
        `evaluations` on every constraints object exactly as the per-option calls would.  Options whose models do not
        share a kernel class are scored by one call each (still on the device)."""
        from . import engine
        error_scale, quality_scale, magnitude_scale = objective_scales(self.objective)
        n = int(self.n_random_samples)
        jobs, locals_ = [], []
        for graph_node, in_c, prev_frames in options:
            if self.use_local_coordinates:
                mp_c, prev = in_c.transform_constraints_to_local_cos(), None
            else:
                mp_c, prev = in_c, prev_frames
            dc = _device_constraints(graph_node, mp_c, prev)
            jobs.append(dict(dc=dc, n=n, seed=self._step_seed(), error_scale=error_scale, quality_scale=quality_scale,
                             magnitude_scale=magnitude_scale))
            locals_.append(mp_c)
        try:
            results = engine.sample_score_multi(jobs)
        except engine.MgbError as e:
            if "kernel class" not in str(e):
                raise
            results = [j["dc"].sample_score_best(j["n"], seed=j["seed"], error_scale=error_scale, quality_scale=quality_scale,
                                                 magnitude_scale=magnitude_scale) for j in jobs]
        out = []
        for (graph_node, in_c, prev_frames), mp_c, (cost, idx, row) in zip(options, locals_, results):
            mp_c.min_error = cost
            mp_c.evaluations += n
            in_c.min_error = mp_c.min_error
            in_c.evaluations = mp_c.evaluations
            out.append(row)
        return out

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
        error_scale, quality_scale, magnitude_scale = objective_scales(self.objective)
        res = dc.score_host(samples_c, k=1, error_scale=error_scale, quality_scale=quality_scale, magnitude_scale=magnitude_scale)
        best_idx = int(res["best_idx"][0])
        min_error = float(res["best_cost"][0])
        constraints.min_error = min_error
        constraints.evaluations += len(samples)
        return samples[best_idx], min_error
