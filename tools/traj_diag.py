"""GPU diagnosis: per-residual deviation of the trajectory cases from the reference-made golden."""
import sys
import numpy as np
sys.path.insert(0, ".")
from tests.common import Case, rel_err
from morphablegraphs_b200 import engine

for name in sys.argv[1:] or ["s_traj_nonlocal"]:
    c = Case(name)
    want = c.g["residual_vectors"]
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints, c.prev_frames)
    r = dc.residuals(c.samples[:len(want)]).cpu().numpy()
    e = rel_err(r, want)
    F = c.F
    print(name, "trajectory part max", e[:, :F].max(), "set part max", e[:, F:2 * F].max(), "keyframe", e[:, 2 * F:].max())
    for (i, j) in np.argwhere(e > 1e-4)[:20]:
        print("  cand", i, "res", j, "got", r[i, j], "want", want[i, j], "neighbours want", want[i, max(0, j - 2):j + 3], "got", r[i, max(0, j - 2):j + 3])
    res = dc.score(c.samples, k=1, return_cost=True, return_terms=True)
    print("  terms err", rel_err(res["terms"].cpu().numpy(), c.g["terms"]).max(axis=1))
