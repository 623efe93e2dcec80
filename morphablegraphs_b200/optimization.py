"""The local optimiser of the sample-and-evaluate path on the device.

`DeviceLeastSquares` has the interface of the reference's `LeastSquares`
(motion_generator/optimization/least_squares.py:28-64 over optimizer_base.py:32-59): the generator assigns it as
`numerical_minimizer`, calls `_objective_function(guess, data)` once for the start error, hands the six-tuple
`data = (graph_node, mp_constraints, prev_frames, error_scale, quality_scale, init_error_sum)` to
`set_objective_function_parameters` and calls `run(initial_guess)`
(motion_primitive_generator.py:179-190).  `run` is ONE kernel launch (mgb_lm_refine); `run_batch` refines several
seeds - e.g. the top-k of the sampling step - in the same launch.
"""
import numpy as np

from .objective_functions import _device_constraints


def obj_spatial_error_residual_vector_and_naturalness(s, data):
    """reference objective_functions.py:239-267 evaluated on the device: residual vector (K9) scaled by
    error_scale, plus quality_scale * (-log p(s)) on every entry, zero-padded to len(s), divided by
    init_error_sum; sets mp_constraints.min_error to the sum of the weighted residuals."""
    mp_node, mp_constraints, prev_frames, error_scale, quality_scale, init_error_sum = data
    dc = _device_constraints(mp_node, mp_constraints, prev_frames)
    s = np.asarray(s, dtype=np.float64)
    r = dc.residuals(s[None, :])[0].cpu().numpy().astype(np.float64)
    mp_constraints.min_error = float(np.sum(r))
    mp_constraints.evaluations += 1
    nll = 0.0
    if quality_scale != 0.0:
        nll = -float(dc.dmodel.logpdf(s[None, :])[0].item()) * quality_scale
    r = r * error_scale + nll
    if len(r) < len(s):
        r = np.concatenate([r, np.zeros(len(s) - len(r))])
    return r / init_error_sum


class DeviceLeastSquares(object):
    def __init__(self, optimization_settings=None):
        self.optimization_settings = dict({"max_iterations": 500, "verbose": False}, **(optimization_settings or {}))
        self.verbose = self.optimization_settings["verbose"]
        self._objective_function = obj_spatial_error_residual_vector_and_naturalness
        self._error_func_params = None
        self._jacobian = None
        self.last_result = None

    def set_objective_function(self, obj):
        if getattr(obj, "__name__", None) != "obj_spatial_error_residual_vector_and_naturalness":
            raise ValueError("DeviceLeastSquares minimises obj_spatial_error_residual_vector_and_naturalness; %r is refused" % (obj,))

    def set_objective_function_parameters(self, data):
        self._error_func_params = data

    def set_jacobian(self, jac):
        self._jacobian = jac

    def _jacobian_evaluations(self, n_dim):
        # The reference bounds leastsq by FUNCTION evaluations (maxfev = max_iterations), i.e. ~max_iterations / (n_dim + 1)
        # Jacobians, and on its own fixtures stops on that bound well before the valley floor (tests/golden/s_resid.npz).  A
        # Jacobian here is one parallel pass of the block; the bound is applied to Jacobians, so the device run goes at least
        # as far as the reference's.
        return max(8, int(self.optimization_settings["max_iterations"]))

    def run_batch(self, seeds):
        """Refine every row of seeds [k, D]; returns the engine's result dict (see DeviceConstraints.lm_refine)."""
        if self._error_func_params is None:
            raise ValueError("no objective function parameters set")
        mp_node, mp_constraints, prev_frames, error_scale, quality_scale, init_error_sum = self._error_func_params
        dc = _device_constraints(mp_node, mp_constraints, prev_frames)
        seeds = np.atleast_2d(np.asarray(seeds, dtype=np.float64))
        res = dc.lm_refine(seeds, error_scale=error_scale, quality_scale=quality_scale, init_error_sum=init_error_sum,
                           max_iter=self._jacobian_evaluations(seeds.shape[1]))
        mp_constraints.evaluations += int(np.sum(res["iterations"])) * (seeds.shape[1] + 1)
        self.last_result = res
        return res

    def run(self, initial_guess):
        if initial_guess is None:
            return initial_guess
        res = self.run_batch(np.asarray(initial_guess, dtype=np.float64)[None, :])
        self._error_func_params[1].min_error = float(res["min_error"][0])
        return res["s"][0]
