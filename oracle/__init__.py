"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (numpy, float64) of the reference's sample-and-evaluate hot
path plus a stand-in for the third-party packages the reference needs but
does not ship (`anim_utils`, `transformations`).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` /
`--impl reference` legs may import anything from here, and there only as the
checker or the timed CPU baseline -- never as part of the shipped path.
`morphablegraphs_b200` never imports `oracle`.

Parity status (see DESIGN.md section "Oracle"):
* back-projection, B-spline evaluation, GMM sampling / log-density, arg-min
  and the cluster-tree search: PINNED against the reference's own modules run
  in the build container (`oracle/make_golden.py` -> `tests/golden/*.npz`).
* forward kinematics, quaternion->matrix, heading, frame alignment and
  point-cloud alignment live in `anim_utils` / `transformations`, which are
  NOT under /root/reference and are not pinned by any reference test:
  PARITY UNPINNED for those functions.  `oracle/standin/` restates their
  published algorithms; the golden vectors run the reference's *own*
  constraint classes on top of that stand-in.
"""
