"""Oracle replay of the reference-decided walk (config 3 in miniature): costs, chosen index and the
aligned end frames that seed the next step."""
import numpy as np

from oracle import np_oracle as O
from tests.walk_common import WalkCase


def test_oracle_replays_reference_walk():
    w = WalkCase()
    for i in range(len(w.order)):
        st = w.step(i)
        ma = O.ModelArrays(st["node"])
        cost = O.candidate_costs(ma, w.skeleton, st["constraints"], st["samples"], st["prev_frames"])
        assert (np.abs(cost - st["costs"]) / np.maximum(np.abs(st["costs"]), 1.0)).max() < 1e-9
        idx, err = O.first_argmin(cost)
        assert idx == st["best_idx"] and abs(err - st["min_error"]) <= 1e-9 * max(1.0, abs(err))
        coeffs = O.back_project_coeffs(ma, st["samples"][idx][None])
        if st["prev_frames"] is not None:
            coeffs = O.align_coeffs_to_prev_frames(ma, w.skeleton, coeffs, st["prev_frames"])
        frames = O.eval_frames(ma, coeffs, O.motion_vector_times(ma.F))[0]
        assert np.abs(frames[-3:] - st["last_frames"]).max() < 1e-9
