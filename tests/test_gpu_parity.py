"""GPU parity tests: the CUDA path (through the C ABI) against the golden vectors
produced by the reference's own modules and against oracle/np_oracle.py on the
same inputs.  Tolerance: |gpu - ref| <= 1e-4 * max(|ref|, 1) (BASELINE.json
north_star, fp32 vs the float64 reference); indices identical whenever the gap
between best and runner-up exceeds that tolerance.
"""
import json

import numpy as np
import pytest

from tests.common import Case, CASES, RTOL, rel_err

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(lib):
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device (there is no CPU fallback to test)")


@pytest.fixture(scope="module", params=CASES)
def case(request):
    from morphablegraphs_b200 import engine
    c = Case(request.param)
    c.dm = engine.DeviceModel(c.node, c.skeleton)
    c.dc = engine.DeviceConstraints(c.dm, c.constraints, c.prev_frames)
    return c


def _check_argmin(cost_ref, idx_gpu, cost_gpu_best):
    order = np.argsort(cost_ref, kind="stable")
    best, second = cost_ref[order[0]], cost_ref[order[1]] if len(order) > 1 else np.inf
    tol = RTOL * max(abs(best), 1.0)
    if second - best > 2 * tol:
        assert idx_gpu == order[0]
    else:   # near tie: any index within tolerance of the minimum is acceptable
        assert cost_ref[idx_gpu] - best <= 2 * tol
    assert abs(cost_gpu_best - cost_ref[idx_gpu]) <= tol


def test_fused_score_terms_cost_argmin(case):
    g = case.g
    S = torch.as_tensor(case.samples, device="cuda")          # float64, what the reference sampler returns
    res = case.dc.score(S, k=1, return_cost=True, return_terms=True)
    terms = res["terms"].cpu().numpy()
    cost = res["cost"].cpu().numpy()
    assert rel_err(terms, g["terms"]).max() < RTOL, rel_err(terms, g["terms"]).max(axis=1)
    assert rel_err(cost, g["costs"]).max() < RTOL
    _check_argmin(g["costs"], int(res["best_idx"][0]), float(res["best_cost"][0]))
    # float32 input gives the same thing
    res32 = case.dc.score(S.float(), k=1, return_cost=True)
    assert rel_err(res32["cost"].cpu().numpy(), g["costs"]).max() < RTOL


def test_fused_score_against_oracle(case):
    from oracle import np_oracle as O
    rng = np.random.RandomState(5)
    ma = O.ModelArrays(case.node)
    S, _ = O.gmm_sample(ma, 1000, rng)
    ref = O.candidate_costs(ma, case.skeleton, case.constraints, S, case.prev_frames)
    res = case.dc.score(torch.as_tensor(S, device="cuda"), k=1, return_cost=True)
    assert rel_err(res["cost"].cpu().numpy(), ref).max() < RTOL
    _check_argmin(ref, int(res["best_idx"][0]), float(res["best_cost"][0]))


def test_naturalness_and_magnitude(case):
    g = case.g
    S = torch.as_tensor(case.samples, device="cuda")
    lp = case.dm.logpdf(S.float()).cpu().numpy()
    assert rel_err(lp, g["logpdf"]).max() < RTOL
    nat = case.dc.score(S, k=0, error_scale=1.0, quality_scale=0.1, return_cost=True)["cost"].cpu().numpy()
    assert rel_err(nat, g["naturalness_cost"]).max() < RTOL
    mag = case.dc.score(S, k=0, magnitude_scale=2.0, return_cost=True)["cost"].cpu().numpy()
    assert rel_err(mag, g["magnitude_cost"]).max() < RTOL


def test_staged_pipeline(case):
    g = case.g
    S = torch.as_tensor(case.samples, device="cuda").float()
    coeffs = case.dm.backproject(S, impl=1)
    assert rel_err(coeffs[:8].cpu().numpy(), g["coeffs"]).max() < RTOL
    # K2 on the tensor cores (tcgen05, 3xTF32): same parity bound, and ~fp32 agreement with the SIMT kernel
    coeffs_tc = case.dm.backproject(S, impl=2)
    assert rel_err(coeffs_tc[:8].cpu().numpy(), g["coeffs"]).max() < RTOL
    assert rel_err(coeffs_tc.cpu().numpy(), coeffs.cpu().numpy()).max() < 2e-5
    for n_ragged in (1, 127, 129, 1000):       # ragged last tile / fewer rows than a tile
        c_r = case.dm.backproject(S[:n_ragged], impl=2)
        assert rel_err(c_r.cpu().numpy(), coeffs[:n_ragged].cpu().numpy()).max() < 2e-5
    frames = case.dm.eval_frames(coeffs, g["key_times"])
    assert rel_err(frames.cpu().numpy(), g["key_frames"]).max() < RTOL
    n_mv = g["motion_vectors"].shape[0]
    mv = case.dm.eval_frames(coeffs[:n_mv], np.linspace(0, case.F, case.F))       # Q1: last sample extrapolates
    assert rel_err(mv.cpu().numpy(), g["motion_vectors"]).max() < RTOL
    names = json.loads(str(g["fk_names_json"]))
    pos = case.dm.fk(frames[:, -1, :].contiguous(), names)
    assert rel_err(pos.cpu().numpy(), g["fk_positions"]).max() < RTOL
    # K5a: score from materialised frames at the constraint set's own times
    fr = case.dm.eval_frames(coeffs, case.dc.times)
    res = case.dc.score(S, k=1, return_cost=True, return_terms=True, frames=fr)
    assert rel_err(res["terms"].cpu().numpy(), g["terms"]).max() < RTOL
    assert rel_err(res["cost"].cpu().numpy(), g["costs"]).max() < RTOL
    _check_argmin(g["costs"], int(res["best_idx"][0]), float(res["best_cost"][0]))


def test_host_entry_and_topk(case):
    g = case.g
    for arr in (case.samples, case.samples.astype(np.float32)):
        res = case.dc.score_host(arr, k=1, return_cost=True)
        assert rel_err(res["cost"], g["costs"]).max() < RTOL
        _check_argmin(g["costs"], int(res["best_idx"][0]), float(res["best_cost"][0]))
    k = 5
    res = case.dc.score_host(case.samples, k=k, return_cost=True)
    cost = res["cost"]
    order = np.lexsort((np.arange(len(cost)), cost))[:k]      # ties -> lowest index
    assert list(res["best_idx"]) == list(order)
    assert np.allclose(res["best_cost"], cost[order], rtol=0, atol=0)
    res_d = case.dc.score(torch.as_tensor(case.samples, device="cuda"), k=k, return_cost=True)
    assert list(res_d["best_idx"].cpu().numpy()) == list(order)


def test_generator_api(case):
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 512}, skeleton=case.skeleton, seed=3)
    case.constraints.evaluations = 0
    best, min_error = gen.evaluate_samples_using_constraints(case.samples, case.node, case.constraints, case.prev_frames)
    g = case.g
    idx = int(np.where((case.samples == best).all(axis=1))[0][0])      # a row of the *input* samples
    _check_argmin(g["costs"], idx, min_error)
    assert case.constraints.evaluations == len(case.samples)
    assert abs(case.constraints.min_error - min_error) == 0.0
    if case.prev_frames is None:
        s = gen._get_best_fit_sample_using_gmm(case.node, case.constraints)
        assert s.shape == (case.dm.D,)
        from oracle import np_oracle as O
        ref = O.candidate_costs(case.node, case.skeleton, case.constraints, s[None])[0]
        assert abs(ref - case.constraints.min_error) <= RTOL * max(1.0, abs(ref))


def test_mgrd_shaped_batch_slot(case):
    """a17: the reference's batched slot (samples [N, D] -> scores [N] -> argmin,
    motion_primitive_generator.py:167-177, mgrd_sample_filter.py:52-75) served by the engine: scores of the
    position constraints the reference hands to mgrd (unset axes -> 0, orientation dropped)."""
    if case.prev_frames is not None:
        pytest.skip("local constraints only")
    from morphablegraphs_b200 import MotionPrimitiveGenerator, MotionPrimitiveConstraints
    from morphablegraphs_b200.mgrd_sample_filter import MGRDSampleFilter
    from oracle import np_oracle as O
    cart = MGRDSampleFilter.extract_cartesian_constraints(case.constraints)
    scores = MGRDSampleFilter.score_samples_cartesian(case.node, case.samples, case.constraints)
    assert scores.shape == (len(case.samples),)
    if len(cart) == 0:
        assert not scores.any()
        return
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton, mpc.is_local, mpc.constraints = case.skeleton, True, cart
    ref = O.candidate_costs(case.node, case.skeleton, mpc, case.samples)
    assert rel_err(scores, ref).max() < RTOL
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 256, "constrained_sampling_mode": "random_spline"},
                                   skeleton=case.skeleton, seed=3)
    s = gen.generate_constrained_sample(case.node, case.constraints)
    got = O.candidate_costs(case.node, case.skeleton, mpc, np.asarray(s)[None])[0]
    assert abs(got - case.constraints.min_error) <= RTOL * max(1.0, abs(got))


# ---------------------------------------------------------------------------
# size-independent properties at the benchmark's full size
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def big():
    from morphablegraphs_b200 import engine
    c = Case("s_config1")
    c.dm = engine.DeviceModel(c.node, c.skeleton)
    c.dc = engine.DeviceConstraints(c.dm, c.constraints)
    return c


def test_full_size_properties(big):
    n = 1_000_000
    S = big.dm.sample(n, seed=2024)
    assert S.shape == (n, big.dm.D) and torch.isfinite(S).all()
    res = big.dc.score(S, k=1, return_cost=True)
    cost = res["cost"]
    # arg-min == min of the materialised costs, first index on ties
    m = cost.min()
    first = int((cost == m).nonzero()[0, 0])
    assert int(res["best_idx"][0]) == first and float(res["best_cost"][0]) == float(m)
    # invariance to sharding: 4 shards with index_base give the same global winner and costs
    shard = n // 4
    bests = []
    for r in range(4):
        rr = big.dc.score(S[r * shard:(r + 1) * shard], k=1, return_cost=True, index_base=r * shard)
        assert torch.equal(rr["cost"], cost[r * shard:(r + 1) * shard])
        bests.append((float(rr["best_cost"][0]), int(rr["best_idx"][0])))
    assert min(bests) == (float(m), first)
    # host entry == device entry (bit-identical costs: same kernel, same inputs)
    Sh = S[:300_001].cpu().numpy()
    rh = big.dc.score_host(Sh, k=1, return_cost=True)
    assert np.array_equal(rh["cost"], cost[:300_001].cpu().numpy())
    # staged pipeline == fused kernel within tolerance on a slice
    sl = S[:20_000]
    fr = big.dm.eval_frames(big.dm.backproject(sl, impl=1), big.dc.times)
    rs = big.dc.score(sl, k=0, return_cost=True, frames=fr)["cost"]
    assert rel_err(rs.cpu().numpy(), cost[:20_000].cpu().numpy()).max() < RTOL
    # spot-check against the oracle
    from oracle import np_oracle as O
    pick = np.random.RandomState(1).choice(n, 2000, replace=False)
    ref = O.candidate_costs(big.node, big.skeleton, big.constraints, S[pick].cpu().numpy().astype(np.float64))
    assert rel_err(cost[pick].cpu().numpy(), ref).max() < RTOL


def test_fused_sample_score(big):
    """mgb_sample_score (K1 + K6 in one launch, candidates never stored) == mgb_sample then mgb_score."""
    dm, dc = big.dm, big.dc
    for n, base, total in ((70_001, 0, 70_001), (50_000, 123_457, 400_000)):
        counts = dm.component_counts(total, seed=11)
        S = dm.sample(n, seed=11, counts=counts, index_base=base, n_total=total)
        for kw in ({}, {"quality_scale": 0.1}, {"magnitude_scale": 2.0}):
            a = dc.score(S, k=1, return_cost=True, index_base=base, **kw)
            b = dc.sample_score(n, seed=11, counts=counts, n_total=total, index_base=base, k=1, return_cost=True, **kw)
            assert rel_err(b["cost"].cpu().numpy(), a["cost"].cpu().numpy()).max() < 1e-6
            ia, ib = int(a["best_idx"][0]), int(b["best_idx"][0])
            assert ia == ib or abs(float(a["cost"][ia - base]) - float(a["cost"][ib - base])) < 1e-6
            # the winning row regenerated from its global index is the row that was scored
            row = dm.sample(1, seed=11, counts=counts, index_base=ib, n_total=total)
            assert torch.equal(row[0], S[ib - base])
    k = 4
    r = dc.sample_score(5000, seed=3, k=k, return_cost=True)
    order = np.lexsort((np.arange(5000), r["cost"].cpu().numpy()))[:k]
    assert list(r["best_idx"].cpu().numpy()) == list(order)


def test_sampler_distribution(big):
    """The reference's RNG stream cannot be reproduced on the device, so K1 is checked
    distributionally: component frequencies, per-component mean and covariance."""
    dm = big.dm
    n = 400_000
    counts = dm.component_counts(n, seed=7)
    assert counts.sum() == n
    S = dm.sample(n, seed=7, counts=counts).cpu().numpy().astype(np.float64)
    mp = big.node.motion_primitive
    gmm = mp.gaussian_mixture_model
    assert np.abs(counts / n - gmm.weights_).max() < 5e-3
    lo = 0
    for g, c in enumerate(counts):       # stacked in component order, like sklearn
        X = S[lo:lo + c]
        lo += c
        if c < 2000:
            continue
        se = np.sqrt(np.diag(gmm.covariances_[g]) / c)
        assert np.abs(X.mean(axis=0) - gmm.means_[g]).max() < 6 * se.max()
        cov = np.cov(X.T)
        scale = np.sqrt(np.outer(np.diag(gmm.covariances_[g]), np.diag(gmm.covariances_[g])))
        assert np.abs(cov - gmm.covariances_[g]).max() < 8 * scale.max() / np.sqrt(c)
    # sharding invariance: the same global indices give the same samples
    a = dm.sample(1000, seed=7, counts=counts, index_base=12345, n_total=n).cpu().numpy()
    assert np.array_equal(a, S[12345:13345].astype(np.float32))
    # different seeds differ
    b = dm.sample(1000, seed=8, counts=counts, n_total=n).cpu().numpy()
    assert not np.array_equal(b, S[:1000].astype(np.float32))


def test_edge_cases(big):
    dc, dm = big.dc, big.dm
    D = dm.D
    # empty batch: reference returns index 0 with min_error = inf
    r = dc.score(torch.empty((0, D), device="cuda"), k=1, return_cost=True)
    assert r["cost"].numel() == 0 and int(r["best_idx"][0]) == 0 and np.isinf(float(r["best_cost"][0]))
    r = dc.score_host(np.zeros((0, D)), k=1)
    assert int(r["best_idx"][0]) == 0 and np.isinf(r["best_cost"][0])
    # ragged sizes around the tile size
    S = dm.sample(1025, seed=1)
    full = dc.score(S, k=1, return_cost=True)
    for n in (1, 2, 255, 256, 257, 1023, 1025):
        rr = dc.score(S[:n], k=1, return_cost=True)
        assert torch.equal(rr["cost"], full["cost"][:n])
        assert int(rr["best_idx"][0]) == int(torch.argmin(full["cost"][:n]))
    # NaN candidates never win (reference: `if min_error > error`)
    S2 = S.clone()
    best = int(full["best_idx"][0])
    S2[best] = float("nan")
    rr = dc.score(S2, k=1, return_cost=True)
    assert int(rr["best_idx"][0]) != best and torch.isnan(rr["cost"][best])
    # ties: identical candidates -> lowest index
    S3 = S[:1].repeat(700, 1)
    assert int(dc.score(S3, k=1)["best_idx"][0]) == 0
    assert int(dc.score(S3, k=1, index_base=1000)["best_idx"][0]) == 1000


def test_errors_are_loud(big):
    from morphablegraphs_b200 import engine, MotionPrimitiveConstraints
    from morphablegraphs_b200.constraints import SpatialConstraintBase

    class Discrete(SpatialConstraintBase):           # DiscreteTrajectoryConstraint: outside the device path
        def __init__(self):
            super(Discrete, self).__init__(1.0, 1.0)
            self.constraint_type = "discrete_trajectory"
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = big.skeleton
    mpc.is_local = True
    mpc.constraints = [Discrete()]
    with pytest.raises(engine.UnsupportedConstraintError):
        engine.DeviceConstraints(big.dm, mpc)
    with pytest.raises(engine.MgbError):
        big.dc.score(big.dm.sample(10, seed=1), k=100)     # k outside [0, 64] is refused by the library


# ---------------------------------------------------------------------------
# large latent space (D = 103, G = 32): the tensor-core kernels (tcgen05, 3xTF32)
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def large():
    from morphablegraphs_b200 import engine
    c = Case("l_mixed")
    c.dm = engine.DeviceModel(c.node, c.skeleton)
    c.dc = engine.DeviceConstraints(c.dm, c.constraints, c.prev_frames)
    return c


def test_large_tc_against_oracle(large):
    """K6-TC (contraction + GMM log-density on the tensor cores) against the numpy oracle on seeded
    draws, incl. a ragged last tile, naturalness and magnitude terms; log-pdf alone (K5b-TC)."""
    from oracle import np_oracle as O
    dm, dc = large.dm, large.dc
    n = 1000 + 37
    S = dm.sample(n, seed=5)
    S64 = S.cpu().numpy().astype(np.float64)
    ref = O.candidate_costs(large.node, large.skeleton, large.constraints, S64, prev_frames=large.prev_frames)
    lp_ref = O.gmm_logpdf(O.ModelArrays(large.node), S64)
    res = dc.score(S, k=1, return_cost=True)
    assert rel_err(res["cost"].cpu().numpy(), ref).max() < RTOL
    assert rel_err(dm.logpdf(S).cpu().numpy(), lp_ref).max() < RTOL
    resn = dc.score(S.double(), k=1, return_cost=True, quality_scale=0.1, magnitude_scale=2.0)       # float64 input path
    refn = ref + 0.1 * (-lp_ref) + 2.0 * np.linalg.norm(S64, axis=1)
    assert rel_err(resn["cost"].cpu().numpy(), refn).max() < RTOL
    _check_argmin(refn, int(resn["best_idx"][0]), float(resn["best_cost"][0]))


def test_large_rotation_and_trajectory_terms(large):
    """JointRotationConstraint (root and inner joint), the quaternion-frame pose term and a LocalTrajectoryConstraint
    on the L model: the raw-channel ops ride the tensor-core path (their quads are remapped per chunk) and agree
    with the oracle; sampled and given candidates agree with each other."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import engine, synthetic as syn, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    from morphablegraphs_b200.constraints import ArcLengthTable
    dm = large.dm
    F = large.F
    joints = list(large.skeleton.node_name_frame_map.keys())
    ma = O.ModelArrays(large.node)
    held = dm.sample(2, seed=77).cpu().numpy().astype(np.float64)
    fr = O.eval_frames(ma, O.back_project_coeffs(ma, held), [0.0, F // 2])
    ji = 23
    q = fr[0, 1, 3 + 4 * ji:7 + 4 * ji]
    pts = np.stack([np.linspace(0.0, 9.0, 65) ** 1.5, np.zeros(65), np.linspace(0.0, 70.0, 65)], axis=1)
    step = float(np.mean(np.linalg.norm(pts[1:] - pts[:-1], axis=1)))
    specs = [syn.spec("JointRotationConstraint", {"joint_name": joints[ji], "rotation_type": "quaternion", "rotation_constraint": q.tolist(),
                                                  "frame_index": F // 2, "canonical_keyframe": F // 2, "semanticAnnotation": syn.sem()}, 1.0, 1.5),
             syn.spec("JointRotationConstraint", {"joint_name": joints[0], "rotation_type": "euler", "rotation_constraint": [5.0, -20.0, 3.0],
                                                  "frame_index": F - 1, "canonical_keyframe": F - 1, "semanticAnnotation": syn.sem()}, 1.0, 0.7),
             syn.spec("LocalTrajectoryConstraint", {"canonical_keyframe": None, "joint_name": "Hips", "start_t": 2.0, "n_canonical_frames": F,
                                                    "semanticAnnotation": syn.sem(), "trajectory": {"points": pts.tolist(), "step": step}}, 1.0, 0.02),
             syn.spec("GlobalTransformConstraint", syn.position_constraint_desc("RightHand", [10.0, 120.0, 50.0], F - 1, F), 1.0, 1.0)]
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = large.skeleton
    mpc.is_local = True
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, large.skeleton, specs)
    # the quaternion-frame pose term of the reference asserts a 79-float frame (pose_constraint_quat_frame.py:38); the device op
    # is the same per-joint rotation distance, exercised here on the 283-float frame of the L skeleton through the mirror class
    mpc.constraints.append(CONSTRAINT_FACTORY["PoseConstraintQuatFrame"](large.skeleton, {"canonical_keyframe": 0, "frame_constraint": fr[1, 0].tolist(),
                                                                                        "semanticAnnotation": syn.sem()}, 1.0, 0.3))
    dc = engine.DeviceConstraints(dm, mpc)
    n = 700
    S = dm.sample(n, seed=6)
    S64 = S.cpu().numpy().astype(np.float64)
    ref, ref_terms = O.candidate_costs(large.node, large.skeleton, mpc, S64, return_terms=True)
    res = dc.score(S, k=1, return_cost=True, return_terms=True)
    assert rel_err(res["terms"].cpu().numpy(), ref_terms).max() < RTOL
    assert rel_err(res["cost"].cpu().numpy(), ref).max() < RTOL
    _check_argmin(ref, int(res["best_idx"][0]), float(res["best_cost"][0]))
    counts = dm.component_counts(n, seed=6)
    fused = dc.sample_score(n, seed=6, counts=counts, k=1, return_cost=True, quality_scale=0.1)
    staged = dc.score(dm.sample(n, seed=6, counts=counts), k=1, return_cost=True, quality_scale=0.1)
    assert rel_err(fused["cost"].cpu().numpy(), staged["cost"].cpu().numpy()).max() < RTOL


def test_large_score_from_frames_with_naturalness(large):
    """K5a at D = 103 with the naturalness and magnitude terms: the density comes from a tensor-core launch into the cost array,
    the frames kernel adds the constraint error; equal to the fused scorer on the same candidates."""
    dm, dc = large.dm, large.dc
    S = dm.sample(3000, seed=4)
    fr = dm.eval_frames(dm.backproject(S), dc.times)
    for kw in ({"quality_scale": 0.1}, {"quality_scale": 0.1, "magnitude_scale": 2.0}, {}):
        a = dc.score(S, k=1, return_cost=True, **kw)
        b = dc.score(S, k=1, return_cost=True, frames=fr, **kw)
        ca, cb = a["cost"].cpu().numpy(), b["cost"].cpu().numpy()
        assert rel_err(cb, ca).max() < RTOL
        ia, ib = int(a["best_idx"][0]), int(b["best_idx"][0])
        assert ia == ib or abs(float(ca[ia]) - float(ca[ib])) < RTOL * max(1.0, abs(float(ca[ia])))
        c = dc.score(S, k=1, frames=fr, **kw)                      # no cost array asked for: the engine supplies one
        assert int(c["best_idx"][0]) == ib


def test_large_sampler_tiled_equals_per_row(large):
    """K1 at D = 103: the register-tiled sampler (tiles of 128 candidates of one component, n >= 512) returns bit for bit the rows
    of the per-row sampler (which small calls still use), for a job of 40 003 candidates over 32 components, whole and as a
    window that starts and ends inside components."""
    import torch
    dm = large.dm
    total = 40_003
    counts = dm.component_counts(total, seed=21)
    whole = dm.sample(total, seed=21, counts=counts)
    assert torch.isfinite(whole).all()
    # per-row path: windows of 300 rows (< 4 tiles)
    for base in (0, 299, 17_000, total - 300):
        part = dm.sample(300, seed=21, counts=counts, index_base=base, n_total=total)
        assert torch.equal(part, whole[base:base + 300]), base
    # tiled path on a window
    win = dm.sample(12_345, seed=21, counts=counts, index_base=7_777, n_total=total)
    assert torch.equal(win, whole[7_777:7_777 + 12_345])


def test_large_whole_motion_trajectory_constraints(large):
    """f2 on the L model (tensor-core step kernel): TrajectoryConstraint on the root (closest point per frame of the 100-frame
    motion vector) and a TrajectorySetConstraint on both hands, against the float64 oracle (scipy L-BFGS-B per frame)."""
    import warnings
    from oracle import np_oracle as O
    from morphablegraphs_b200 import engine, synthetic as syn, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    dm, F = large.dm, large.F
    ma = O.ModelArrays(large.node)
    held = dm.sample(2, seed=91).cpu().numpy().astype(np.float64)
    fr = O.eval_frames(ma, O.back_project_coeffs(ma, held), list(np.linspace(0, F - 1, 7)))

    def path(k, joint):
        return O.fk_global(large.skeleton, fr[k:k + 1], joint)[1][0].tolist()
    specs = [syn.spec("TrajectoryConstraint", {"joint_name": "Hips", "control_points": path(0, "Hips"), "min_arc_length": 1.0,
                                               "n_canonical_frames": F}, 1.0, 0.9),
             syn.spec("TrajectorySetConstraint", {"joint_names": ["LeftHand", "RightHand"], "joint_arc_lengths": [2.0, 0.0], "n_canonical_frames": F,
                                                  "trajectories": [{"control_points": path(1, "LeftHand"), "range": [0.0, 30.0]},
                                                                   {"control_points": path(1, "RightHand"), "range": [10.0, 1e9]}]}, 1.0, 1.1),
             syn.spec("GlobalTransformConstraint", syn.position_constraint_desc("RightHand", [10.0, 120.0, 50.0], F - 1, F), 1.0, 1.0)]
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = large.skeleton
    mpc.is_local = True
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, large.skeleton, specs)
    dc = engine.DeviceConstraints(dm, mpc)
    n = 300
    S = dm.sample(n, seed=8)
    S64 = S.cpu().numpy().astype(np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref, ref_terms = O.candidate_costs(large.node, large.skeleton, mpc, S64, return_terms=True)
    res = dc.score(S, k=1, return_cost=True, return_terms=True)
    assert rel_err(res["terms"].cpu().numpy(), ref_terms).max() < RTOL
    assert rel_err(res["cost"].cpu().numpy(), ref).max() < RTOL
    _check_argmin(ref, int(res["best_idx"][0]), float(res["best_cost"][0]))
    counts = dm.component_counts(n, seed=8)
    fused = dc.sample_score(n, seed=8, counts=counts, k=1, return_cost=True, quality_scale=0.1)
    staged = dc.score(dm.sample(n, seed=8, counts=counts), k=1, return_cost=True, quality_scale=0.1)
    assert rel_err(fused["cost"].cpu().numpy(), staged["cost"].cpu().numpy()).max() < RTOL


def test_large_fused_sample_score(large):
    """mgb_sample_score on the tensor cores (z -> s = mu + L z as a GEMM, tiles aligned to mixture components)
    == mgb_sample (SIMT) then mgb_score, up to the 3xTF32 rounding of s."""
    dm, dc = large.dm, large.dc
    for n, base, total in ((20_001, 0, 20_001), (9_000, 12_345, 60_000)):
        counts = dm.component_counts(total, seed=11)
        S = dm.sample(n, seed=11, counts=counts, index_base=base, n_total=total)
        for kw in ({}, {"quality_scale": 0.1}, {"magnitude_scale": 2.0}):
            a = dc.score(S, k=1, return_cost=True, index_base=base, **kw)
            b = dc.sample_score(n, seed=11, counts=counts, n_total=total, index_base=base, k=1, return_cost=True, **kw)
            ca, cb = a["cost"].cpu().numpy(), b["cost"].cpu().numpy()
            assert rel_err(cb, ca).max() < 2e-5
            ia, ib = int(a["best_idx"][0]), int(b["best_idx"][0])
            assert ia == ib or abs(float(ca[ia - base]) - float(ca[ib - base])) < 2e-5 * max(1.0, abs(float(ca[ia - base])))


@pytest.mark.parametrize("n_gmm,n_spatial,n_constraints", [(1, 40, 1), (2, 37, 1), (3, 60, 4), (9, 48, 2), (32, 33, 1)])
def test_tc_chunk_schedule_shapes(n_gmm, n_spatial, n_constraints):
    """The tensor-core step kernel interleaves the chunks of the quad operand with the chunks of the GMM operand
    (two quad buffers + one GMM buffer in TMEM).  Models whose operand chunk counts differ in every way from the L
    fixture (one GMM chunk, one quad chunk, more GMM chunks than quad chunks and the reverse), scored in both
    modes (samples from memory / drawn in the kernel) with and without the naturalness term, against the oracle."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import engine, synthetic as syn
    from morphablegraphs_b200 import Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    shape = dict(n_frames=60, n_joints=19, n_spatial=n_spatial, n_time=3, n_gmm=n_gmm)
    skeleton = Skeleton.from_desc(syn.make_skeleton_desc(19))
    node = MotionStateGraphNode()
    node._initialize_from_json(None, syn.make_model_json(shape, seed=100 + n_gmm))
    node.set_skeleton(skeleton)
    joints = ["RightHand", "LeftFoot", "Head", "LeftHand"]
    rng = np.random.RandomState(n_gmm)
    specs = [syn.spec("GlobalTransformConstraint",
                      syn.position_constraint_desc(joints[i % 4], list(rng.uniform(-50, 100, 3)), [59, 30, 0, 45][i % 4], 60),
                      weight=1.0 + 0.5 * i) for i in range(n_constraints)]
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton, mpc.is_local = skeleton, True
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, skeleton, specs)
    dm = engine.DeviceModel(node, skeleton)
    dc = engine.DeviceConstraints(dm, mpc, None)
    assert dm.D >= 33                                           # the tensor-core path
    for n, base, total in ((128, 0, 128), (1000 + 37, 0, 1000 + 37), (2500, 777, 9000)):
        counts = dm.component_counts(total, seed=3)
        S = dm.sample(n, seed=3, counts=counts, index_base=base, n_total=total)
        S64 = S.cpu().numpy().astype(np.float64)
        ref = O.candidate_costs(node, skeleton, mpc, S64)
        lp = O.gmm_logpdf(O.ModelArrays(node), S64)
        for qs in (0.0, 0.1):
            want = ref + qs * (-lp)
            a = dc.score(S, k=1, return_cost=True, index_base=base, quality_scale=qs)
            b = dc.sample_score(n, seed=3, counts=counts, n_total=total, index_base=base, k=1, return_cost=True, quality_scale=qs)
            assert rel_err(a["cost"].cpu().numpy(), want).max() < RTOL
            assert rel_err(b["cost"].cpu().numpy(), want).max() < RTOL
            _check_argmin(want, int(a["best_idx"][0]) - base, float(a["best_cost"][0]))
            _check_argmin(want, int(b["best_idx"][0]) - base, float(b["best_cost"][0]))


def test_peer_exchange_kernel_single_process():
    """mgb_exchange_best with several 'ranks' inside one process: one slot array per rank on the same GPU,
    one stream per rank (the kernels wait for each other's records, so they must be co-resident: 1 CTA of
    32 threads each).  Winner = lowest cost, lowest global index on ties, NaN never wins; repeated steps
    exercise the parity double-buffering."""
    import ctypes as C
    from morphablegraphs_b200 import _lib, sharding
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    for world in (1, 2, 4, 8):
        nbytes = int(lib.mgb_exchange_slot_bytes(world))
        slots = [torch.zeros(nbytes // 4, dtype=torch.int32, device=dev) for _ in range(world)]
        ptrs = [int(t.data_ptr()) for t in slots]
        ranks = [sharding.PeerExchange(dev, slots=slots[r], peer_ptrs=ptrs, rank=r, world=world) for r in range(world)]
        streams = [torch.cuda.Stream(dev) for _ in range(world)]
        rng = np.random.RandomState(world)
        for step in range(6):
            costs = rng.randint(0, 3, size=world).astype(np.float32)
            if step == 2:
                costs[:] = 1.5                              # all tie -> lowest index
            if step == 3 and world > 1:
                costs[0] = np.nan                           # NaN never wins
            if step == 4:
                costs[:] = np.nan                           # nothing finite -> (inf, 0)
            idx = (rng.permutation(world) * 1000 + 2 ** 40).astype(np.int64)
            cs = [torch.tensor([costs[r]], dtype=torch.float32, device=dev) for r in range(world)]
            ix = [torch.tensor([idx[r]], dtype=torch.int64, device=dev) for r in range(world)]
            torch.cuda.synchronize()
            outs = []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    outs.append(ranks[r].exchange(cs[r], ix[r], spin_limit=2_000_000))
            torch.cuda.synchronize()
            want_c, want_i = sharding.select_best(costs, idx)
            for r in range(world):
                ranks[r].check()
                oc, oi = float(outs[r][0].item()), int(outs[r][1].item())
                assert (oc == want_c or (np.isinf(oc) and np.isinf(want_c))) and oi == want_i, (world, step, r, oc, oi, want_c, want_i)
    # a peer that never shows up raises instead of hanging
    slots = [torch.zeros(int(lib.mgb_exchange_slot_bytes(2)) // 4, dtype=torch.int32, device=dev) for _ in range(2)]
    lone = sharding.PeerExchange(dev, slots=slots[0], peer_ptrs=[int(t.data_ptr()) for t in slots], rank=0, world=2)
    lone.exchange(torch.zeros(1, device=dev), torch.zeros(1, dtype=torch.int64, device=dev), spin_limit=1000)
    with pytest.raises(_lib.MgbError):
        lone.check()


def test_walk_sequence_matches_reference_decisions():
    """Config 3 in miniature: a six-step walk over four primitives whose every decision was taken by the
    reference's generator loop (tests/golden/s_walk.npz); the engine, given the same samples, constraints
    and previous frames, must choose the same candidate at every step."""
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    from tests.walk_common import WalkCase
    w = WalkCase()
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 64}, skeleton=w.skeleton, seed=1)
    for i in range(len(w.order)):
        st = w.step(i)
        best, min_error = gen.evaluate_samples_using_constraints(st["samples"], st["node"], st["constraints"], st["prev_frames"])
        idx = int(np.where((st["samples"] == best).all(axis=1))[0][0])
        _check_argmin(st["costs"], idx, min_error)
        assert idx == st["best_idx"], "step %d" % i
        assert abs(min_error - st["min_error"]) <= RTOL * max(1.0, abs(st["min_error"]))


# ---------------------------------------------------------------------------
# boundary holes closed in round 2 (VERDICT r01 "weak" 1, 2)
# ---------------------------------------------------------------------------
def test_in_place_constraint_mutation_recompiles():
    """graph_walk_planner.py:199-203 rewrites `canonical_keyframe` on the SAME constraint objects before every
    scoring call: the cached device program must follow (content key, objective_functions.constraints_key)."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    c = Case("s_mixed")
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 64}, skeleton=c.skeleton, seed=3)
    S = np.ascontiguousarray(c.samples)
    best0, e0 = gen.evaluate_samples_using_constraints(S, c.node, c.constraints, None)
    ref0 = O.candidate_costs(c.node, c.skeleton, c.constraints, S)
    assert abs(e0 - ref0.min()) <= RTOL * max(1.0, abs(ref0.min()))
    con = c.constraints.constraints[0]
    con.canonical_keyframe = con.canonical_keyframe / 2.0      # in place, same object, same list length
    con.position[1] += 7.5
    con.weight_factor = 3.0
    best1, e1 = gen.evaluate_samples_using_constraints(S, c.node, c.constraints, None)
    ref1 = O.candidate_costs(c.node, c.skeleton, c.constraints, S)
    assert abs(ref1.min() - ref0.min()) > 1e-2                  # the mutation matters ...
    assert abs(e1 - ref1.min()) <= RTOL * max(1.0, abs(ref1.min()))   # ... and the engine saw it
    _check_argmin(ref1, int(np.where((S == best1).all(axis=1))[0][0]), e1)
    # unchanged content: the compiled program is reused (same handle object)
    dc_a = c.constraints._mgb_cache[1]
    gen.evaluate_samples_using_constraints(S, c.node, c.constraints, None)
    assert c.constraints._mgb_cache[1] is dc_a


def test_prev_frames_content_is_part_of_the_key():
    from oracle import np_oracle as O
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    c = Case("s_nonlocal")
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 32, "use_local_coordinates": False}, skeleton=c.skeleton, seed=3)
    S = np.ascontiguousarray(c.samples)
    prev = np.array(c.prev_frames, dtype=np.float64)
    _, e0 = gen.evaluate_samples_using_constraints(S, c.node, c.constraints, prev)
    prev[-1, 0] += 25.0                                         # same array object, edited in place
    prev[-1, 2] -= 10.0
    _, e1 = gen.evaluate_samples_using_constraints(S, c.node, c.constraints, prev)
    ref1 = O.candidate_costs(c.node, c.skeleton, c.constraints, S, prev)
    assert abs(e1 - e0) > 1e-2
    assert abs(e1 - ref1.min()) <= RTOL * max(1.0, abs(ref1.min()))


def test_reference_style_plain_function_objective():
    """mg_state_planner.py:250 assigns the plain function obj_spatial_error_sum_and_magnitude: the engine must
    score error + 2|s| (mapped by name), and refuse callables it does not know."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    from morphablegraphs_b200._lib import MgbError
    c = Case("s_config1")
    S = np.ascontiguousarray(c.samples)
    ref = O.candidate_costs(c.node, c.skeleton, c.constraints, S)
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 64}, skeleton=c.skeleton, seed=3)

    def obj_spatial_error_sum_and_magnitude(s, data):     # stands for the reference's function object
        raise AssertionError("the host callable is never used on the device path")

    gen.objective = obj_spatial_error_sum_and_magnitude
    best, e = gen.evaluate_samples_using_constraints(S, c.node, c.constraints, None)
    ref_mag = ref + 2.0 * np.linalg.norm(S, axis=1)
    assert abs(e - ref_mag.min()) <= RTOL * max(1.0, abs(ref_mag.min()))
    assert abs(e - c.g["magnitude_cost"].min()) <= RTOL * max(1.0, abs(e))
    s_best = gen.generate_constrained_sample(c.node, c.constraints)       # fused path takes the same scales
    assert abs(c.constraints.min_error - (O.candidate_costs(c.node, c.skeleton, c.constraints, s_best[None])[0]
                                          + 2.0 * np.linalg.norm(s_best))) <= 2 * RTOL * max(1.0, c.constraints.min_error)
    gen.objective = lambda s, data: 0.0
    with pytest.raises(MgbError):
        gen.evaluate_samples_using_constraints(S, c.node, c.constraints, None)
    with pytest.raises(MgbError):
        gen.generate_constrained_sample(c.node, c.constraints)


def test_host_entry_buffer_sequence():
    """pinned float64 input, then pageable float32, then pageable float64 with more than one
    chunk (2^18 rows) - the pinned staging buffers must grow with the chunk size."""
    from oracle import np_oracle as O
    c = Case("s_config1")
    from morphablegraphs_b200 import engine
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints)
    n = (1 << 18) + 4097
    S32 = dm.sample(n, seed=5).cpu().numpy()
    S64 = S32.astype(np.float64)
    pinned = torch.from_numpy(S64).pin_memory()
    r_a = dc.score_host(pinned, k=1, return_cost=True)
    r_b = dc.score_host(S32, k=1, return_cost=True)
    r_c = dc.score_host(S64, k=1, return_cost=True)
    assert np.array_equal(r_a["cost"], r_b["cost"]) and np.array_equal(r_a["cost"], r_c["cost"])
    assert int(r_a["best_idx"][0]) == int(r_b["best_idx"][0]) == int(r_c["best_idx"][0])
    sub = np.random.RandomState(0).choice(n, 512, replace=False)
    ref = O.candidate_costs(c.node, c.skeleton, c.constraints, S64[sub])
    assert rel_err(r_c["cost"][sub], ref).max() < RTOL


# ---------------------------------------------------------------------------
# reference-made goldens at the sizes SURVEY.md 8(d) states (round 2)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["s_config1_10k", "l_cfg4", "l_mixed_1k"])
def test_reference_goldens_at_size(name):
    """config 1's 10 000 reference-drawn samples, config 4's constraint set on the L model (1000 reference-scored
    candidates), the mixed L set at 1000: GPU costs / terms / arg-min against the reference's own loop."""
    from morphablegraphs_b200 import engine
    c = Case(name)
    g = c.g
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints)
    S = torch.as_tensor(c.samples, device="cuda")
    res = dc.score(S, k=1, return_cost=True, return_terms=True)
    assert rel_err(res["terms"].cpu().numpy(), g["terms"]).max() < RTOL
    assert rel_err(res["cost"].cpu().numpy(), g["costs"]).max() < RTOL
    _check_argmin(g["costs"], int(res["best_idx"][0]), float(res["best_cost"][0]))
    nat = dc.score(S, k=1, return_cost=True, quality_scale=0.1)
    assert rel_err(nat["cost"].cpu().numpy(), g["naturalness_cost"]).max() < RTOL
    _check_argmin(g["naturalness_cost"], int(nat["best_idx"][0]), float(nat["best_cost"][0]))
    # the reference-facing host entry on the same float64 array
    h = dc.score_host(np.ascontiguousarray(c.samples), k=1, return_cost=True, quality_scale=0.1)
    assert np.array_equal(h["cost"], nat["cost"].cpu().numpy()) and int(h["best_idx"][0]) == int(nat["best_idx"][0])


def test_local_coordinates_default_path():
    """a8 on the device path: a GLOBAL constraint set with a non-identity aligning_transform goes through
    generate_constrained_sample(use_local_coordinates=True) = transform_constraints_to_local_cos (quirk Q5) + scoring;
    costs against the reference's loop on the reference-transformed set (tests/golden/s_local.npz)."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    c = Case("s_local")
    g = c.g
    c.constraints.aligning_transform = np.array(g["aligning_transform"])
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 4096, "use_local_coordinates": True}, skeleton=c.skeleton, seed=5)
    local = c.constraints.transform_constraints_to_local_cos()
    S = np.ascontiguousarray(c.samples)
    best, e = gen.evaluate_samples_using_constraints(S, c.node, local, None)
    assert int(np.where((S == best).all(axis=1))[0][0]) == int(g["best_idx"])
    assert abs(e - float(g["min_error"])) <= RTOL * max(1.0, abs(e))
    from morphablegraphs_b200 import engine
    dc = engine.DeviceConstraints(engine.DeviceModel(c.node, c.skeleton), local)
    res = dc.score(torch.as_tensor(S, device="cuda"), k=1, return_cost=True, return_terms=True)
    assert rel_err(res["terms"].cpu().numpy(), g["terms"]).max() < RTOL
    assert rel_err(res["cost"].cpu().numpy(), g["costs"]).max() < RTOL
    # the whole generator call on the global set: the winner's error is its cost under the LOCAL set
    s = gen.generate_constrained_sample(c.node, c.constraints)
    ref = O.candidate_costs(c.node, c.skeleton, local, s[None])[0]
    assert abs(c.constraints.min_error - ref) <= 2 * RTOL * max(1.0, abs(ref))
    assert c.constraints.evaluations == 4096
    # the global set scored WITHOUT the transform is a different (much larger) number: the transform is not a no-op
    c.constraints.is_local = True
    wrong = O.candidate_costs(c.node, c.skeleton, c.constraints, s[None])[0]
    c.constraints.is_local = False
    assert abs(wrong - ref) > 10.0


# ---------------------------------------------------------------------------
# planner-facing entries: one call per step, cross-option batching
# ---------------------------------------------------------------------------
def test_sample_score_best_returns_the_row_it_scored():
    """mgb_sample_score_best == mgb_sample_score + mgb_sample(1, best_idx): same cost and index, and the row re-drawn
    by the kernel's last CTA is the row of that global index, bit for bit."""
    from morphablegraphs_b200 import engine
    c = Case("s_mixed")
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints)
    for n, seed, qs in ((10, 3, 0.0), (5000, 4, 0.1), (70001, 5, 0.0)):
        counts = dm.component_counts(n, seed)
        cost, idx, row = dc.sample_score_best(n, seed=seed, counts=counts, quality_scale=qs, magnitude_scale=0.5)
        ref = dc.sample_score(n, seed=seed, counts=counts, k=1, quality_scale=qs, magnitude_scale=0.5)
        assert idx == int(ref["best_idx"][0]) and cost == float(ref["best_cost"][0])
        want = dm.sample(1, seed=seed, index_base=idx, counts=counts, n_total=n)[0].cpu().numpy()
        assert np.array_equal(row.astype(np.float32), want)


def test_sample_score_best_large(large):
    dm, dc = large.dm, large.dc
    n, seed = 3000, 9
    counts = dm.component_counts(n, seed)
    cost, idx, row = dc.sample_score_best(n, seed=seed, counts=counts, quality_scale=0.1)
    ref = dc.sample_score(n, seed=seed, counts=counts, k=1, quality_scale=0.1)
    assert idx == int(ref["best_idx"][0]) and cost == float(ref["best_cost"][0])
    want = dm.sample(1, seed=seed, index_base=idx, counts=counts, n_total=n)[0].cpu().numpy()
    assert np.array_equal(row.astype(np.float32), want)


def test_options_scored_in_one_launch_equal_separate_calls():
    """f4: four outgoing options (four different primitives, their own constraint sets and seeds) x 10 samples - and
    x 20 000 samples - in ONE launch give exactly the per-option (cost, index, row) of four separate calls; one kernel
    launch is counted for the batch."""
    from morphablegraphs_b200 import engine
    from tests.walk_common import WalkCase
    w = WalkCase()
    dms = [engine.DeviceModel(nd, w.skeleton) for nd in w.nodes]
    opts = []
    for i in range(4):
        st = w.step(i)
        k = w.order[i]
        opts.append((dms[k], engine.DeviceConstraints(dms[k], st["constraints"], st["prev_frames"])))
    for n in (10, 20000):
        jobs = [dict(dc=dc, n=n + 3 * j, seed=100 + j, magnitude_scale=2.0 if j == 1 else 0.0) for j, (dm, dc) in enumerate(opts)]
        before = engine.launch_count()
        multi = engine.sample_score_multi(jobs)
        assert engine.launch_count() - before == 1
        for j, (dm, dc) in enumerate(opts):
            cost, idx, row = dc.sample_score_best(n + 3 * j, seed=100 + j, magnitude_scale=2.0 if j == 1 else 0.0)
            assert multi[j][0] == cost and multi[j][1] == idx and np.array_equal(multi[j][2], row), (n, j)


def test_generator_scores_all_options_at_once():
    """MotionPrimitiveGenerator.generate_constrained_samples (the planner's option loop in one launch) leaves the same
    min_error / evaluations on every constraints object and returns the same latent vectors as one
    generate_constrained_sample per option with the same seed stream."""
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    from tests.walk_common import WalkCase
    w = WalkCase()
    cfgd = {"n_random_samples": 500, "use_local_coordinates": False}
    a = MotionPrimitiveGenerator(None, cfgd, skeleton=w.skeleton, seed=5)
    b = MotionPrimitiveGenerator(None, cfgd, skeleton=w.skeleton, seed=5)
    steps_a = [w.step(i) for i in range(4)]
    steps_b = [w.step(i) for i in range(4)]
    got = a.generate_constrained_samples([(s["node"], s["constraints"], s["prev_frames"]) for s in steps_a])
    for s, row in zip(steps_b, got):
        want = b.generate_constrained_sample(s["node"], s["constraints"], "", s["prev_frames"], None)
        assert np.array_equal(row, want)
    for sa, sb in zip(steps_a, steps_b):
        assert sa["constraints"].min_error == sb["constraints"].min_error
        assert sa["constraints"].evaluations == sb["constraints"].evaluations == 500


# ---------------------------------------------------------------------------
# f1: residual vectors (K9) and the batched Levenberg-Marquardt (K10)
# ---------------------------------------------------------------------------
def _opt_case():
    import json
    from morphablegraphs_b200 import synthetic as syn, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    c = Case("s_resid")
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton, mpc.is_local = c.skeleton, True
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, c.skeleton, json.loads(str(c.g["opt_specs_json"])))
    return c, mpc


def test_residual_vectors_match_the_reference():
    from morphablegraphs_b200 import engine
    c = Case("s_resid")
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints)
    r = dc.residuals(c.samples).cpu().numpy()
    assert r.shape == c.g["residual_vectors"].shape
    assert rel_err(r, c.g["residual_vectors"]).max() < RTOL
    # the objective function with the reference's signature, evaluated on the device
    from morphablegraphs_b200.optimization import obj_spatial_error_residual_vector_and_naturalness as obj
    v = obj(c.samples[3], (c.node, c.constraints, None, 1.0, 0.1, 1.0))
    assert rel_err(v, c.g["objective_q01"][3]).max() < RTOL
    assert abs(c.constraints.min_error - c.g["residual_vectors"][3].sum()) < RTOL * abs(c.g["residual_vectors"][3].sum())


@pytest.mark.parametrize("name", ["s_traj", "s_traj_nonlocal"])
def test_trajectory_residual_vectors_match_the_reference(name):
    """f2: one residual per frame of the motion vector for TrajectoryConstraint (closest point of the path at a parameter >=
    the previous frame's) and TrajectorySetConstraint (arc-length walk of several joints), against the reference's own
    classes (tests/golden/s_traj*.npz); the L step kernel family is covered by the `case` tests through CASES."""
    from morphablegraphs_b200 import engine
    c = Case(name)
    want = c.g["residual_vectors"]
    dm = engine.DeviceModel(c.node, c.skeleton)
    dc = engine.DeviceConstraints(dm, c.constraints, c.prev_frames)
    r = dc.residuals(c.samples[:len(want)]).cpu().numpy()
    assert r.shape == want.shape
    assert rel_err(r, want).max() < RTOL


def test_device_least_squares_reaches_the_reference_optimisers_error():
    """mgb_lm_refine from the seeds the reference's own LeastSquares (scipy leastsq, maxfev 500) started from: the
    error it reaches - re-evaluated in float64 by the oracle - is not worse than the reference's result (1e-3
    relative), for quality_scale 0 and 0.1; all seeds of a quality scale are refined in ONE launch."""
    from oracle import np_oracle as O
    from morphablegraphs_b200 import engine
    from morphablegraphs_b200.optimization import DeviceLeastSquares
    c, mpc = _opt_case()
    g = c.g
    for qs in (0.0, 0.1):
        sel = np.where(g["opt_quality_scales"] == qs)[0]
        ls = DeviceLeastSquares({"max_iterations": 500})
        ls.set_objective_function_parameters((c.node, mpc, None, 1.0, qs, 0.0))
        before = engine.launch_count()
        res = ls.run_batch(g["opt_seeds"][sel])
        assert engine.launch_count() - before == 1
        r1 = O.residual_objective(c.node, c.skeleton, mpc, res["s"], quality_scale=qs)
        cost64 = 0.5 * np.sum(r1 ** 2, axis=1)
        assert rel_err(res["cost"], cost64).max() < 1e-3                      # the kernel's own account of where it ended
        assert rel_err(res["cost_start"], g["opt_start_cost"][sel]).max() < 1e-3
        assert (cost64 <= g["opt_final_cost"][sel] * (1.0 + 1e-3)).all(), (qs, cost64, g["opt_final_cost"][sel])
        assert (res["cost"] < res["cost_start"]).all()


def test_generator_runs_the_device_optimiser():
    """generate_constrained_sample with use_local_optimization: the arg-min seeds the device Levenberg-Marquardt through the
    reference's numerical_minimizer slot (motion_primitive_generator.py:179-190); the error does not go up."""
    from morphablegraphs_b200 import MotionPrimitiveGenerator
    from morphablegraphs_b200.optimization import DeviceLeastSquares
    c, mpc = _opt_case()
    cfgd = {"n_random_samples": 2000, "local_optimization_mode": "keyframes",
            "local_optimization_settings": {"start_error_threshold": 0.0, "error_scale_factor": 1.0, "quality_scale_factor": 0.1,
                                            "max_iterations": 500, "verbose": False}}
    plain = MotionPrimitiveGenerator(None, cfgd, skeleton=c.skeleton, seed=3)
    s0 = plain.generate_constrained_sample(c.node, mpc)
    e0 = mpc.min_error
    opt = MotionPrimitiveGenerator(None, cfgd, skeleton=c.skeleton, seed=3)
    opt.numerical_minimizer = DeviceLeastSquares(cfgd["local_optimization_settings"])
    mpc.use_local_optimization = True
    s1 = opt.generate_constrained_sample(c.node, mpc)
    assert s1.shape == s0.shape and not np.array_equal(s0, s1)
    assert mpc.min_error < e0


# ---------------------------------------------------------------------------
# config 3 (J1): the example_input.json action list over the synthetic graph, loaded from the archive format
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def config3_graph():
    import io
    from morphablegraphs_b200 import synthetic as syn
    from morphablegraphs_b200.motion_state_graph import MotionStateGraphLoader, graph_to_bytes
    ld = MotionStateGraphLoader()
    ld.set_data_source(io.BytesIO(graph_to_bytes(syn.make_graph_data())))     # through the zip reader (f3)
    g = ld.build(resident=True)
    assert len(g.make_resident()) == 14
    return g


def test_config3_steps_match_the_reference_decisions(config3_graph):
    """All 17 motion-primitive steps of examples/example_input.json (walk, pickRight, carryRight, placeRight): the
    reference's own generator loop chose among 10 000 candidates per step (tests/golden/s_config3.npz, candidates redrawn
    here from the stored seed by the oracle's restatement of sklearn's sampler); the engine must take the same decision
    at every step and agree on all 170 000 costs."""
    import json
    from oracle import np_oracle as O
    from morphablegraphs_b200 import MotionPrimitiveGenerator, graph_synthesis as GS
    from tests.common import load_golden
    g = load_golden("s_config3")
    graph = config3_graph
    keys = [tuple(k) for k in json.loads(str(g["node_keys_json"]))]
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": int(g["n_samples"]), "use_local_coordinates": False}, skeleton=graph.skeleton, seed=1)
    for i, key in enumerate(keys):
        node = graph.nodes[key]
        S, _ = O.gmm_sample(O.ModelArrays(node), int(g["n_samples"]), np.random.RandomState(int(g["step%d_seed" % i])))
        assert np.array_equal(S[int(g["step%d_best_idx" % i])], g["step%d_best_sample" % i])      # the redraw is the golden's draw
        mpc = GS.build_constraints(graph, json.loads(str(g["step%d_specs_json" % i])))
        prev = g["step%d_prev_frames" % i]
        best, err = gen.evaluate_samples_using_constraints(S, node, mpc, prev)
        ref = g["step%d_costs" % i].astype(np.float64)
        bi = int(np.where((S == best).all(axis=1))[0][0])
        _check_argmin(ref, bi, err)
        assert abs(err - float(g["step%d_min_error" % i])) < RTOL * max(1.0, float(g["step%d_min_error" % i]))
        from morphablegraphs_b200.objective_functions import _device_constraints
        cost = _device_constraints(node, mpc, prev).score_host(np.ascontiguousarray(S), k=0, return_cost=True)["cost"]
        assert rel_err(cost, ref).max() < 2 * RTOL        # the golden stores float32 costs


def test_config3_synthesis_runs_end_to_end(config3_graph):
    """The whole action list synthesised on the device (candidates drawn in-kernel, 100 000 per step): 17 steps, every step
    starts where the previous one ended and ends near its goal."""
    from morphablegraphs_b200 import MotionPrimitiveGenerator, graph_synthesis as GS
    from tests.test_graph_cpu import EXAMPLE_INPUT
    graph = config3_graph
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": 100000, "use_local_coordinates": False}, skeleton=graph.skeleton, seed=7)
    out = GS.GraphWalkSynthesizer(graph, gen).run(EXAMPLE_INPUT)
    assert len(out) == 17
    for rec in out:
        assert np.isfinite(rec["min_error"]) and rec["parameters"].shape == (13,)
        assert np.abs(rec["last_frames"][0][:3]).max() < 1e4
    for a, b in zip(out[:-1], out[1:]):
        assert np.array_equal(a["last_frames"], b["prev_frames"])


def test_exchange_fused_into_the_score_kernel():
    """mgb_sample_score_sharded: several "ranks" of one process (own streams, own handles and workspaces) score their shard of
    one job; the CTA that folds a rank's minima exchanges the record with the peers inside the kernel.  Every rank ends with
    the winner of the single-GPU run of the whole job, for the register-resident kernel (S) and the tensor-core kernel (L)."""
    import torch
    from morphablegraphs_b200 import engine, sharding, _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    for name, n, qs in (("s_config1", 60001, 0.0), ("l_cfg4", 3000, 0.1)):
        c = Case(name)
        dm0 = engine.DeviceModel(c.node, c.skeleton)
        dc0 = engine.DeviceConstraints(dm0, c.constraints, c.prev_frames)
        seed = 13
        counts = dm0.component_counts(n, seed)
        want_cost, want_idx, _ = dc0.sample_score_best(n, seed=seed, counts=counts, quality_scale=qs)
        for world in (2, 4):
            nbytes = int(lib.mgb_exchange_slot_bytes(world))
            slots = [torch.zeros(nbytes // 4, dtype=torch.int32, device=dev) for _ in range(world)]
            ptrs = [int(t.data_ptr()) for t in slots]
            peers = [sharding.PeerExchange(dev, slots=slots[r], peer_ptrs=ptrs, rank=r, world=world) for r in range(world)]
            dms = [engine.DeviceModel(c.node, c.skeleton) for _ in range(world)]
            dcs = [engine.DeviceConstraints(dms[r], c.constraints, c.prev_frames) for r in range(world)]
            streams = [torch.cuda.Stream(dev) for _ in range(world)]
            torch.cuda.synchronize()
            for step in range(2):
                outs = []
                before = engine.launch_count()
                for r in range(world):
                    base, n_r = sharding.shard_range(n, r, world)
                    with torch.cuda.stream(streams[r]):
                        outs.append(dcs[r].sample_score_sharded(n_r, peers[r], seed=seed, counts=counts, n_total=n, index_base=base, quality_scale=qs))
                torch.cuda.synchronize()
                assert engine.launch_count() - before == world          # one launch per rank and step
                for r in range(world):
                    peers[r].check()
                    assert float(outs[r][0].item()) == want_cost and int(outs[r][1].item()) == want_idx, (name, world, r)


def test_time_warped_motion_matches_the_reference():
    """f2: back_project(s, use_time_parameters=True): the time function t'(t) on the host (pinned on the CPU), the motion
    evaluated at the warped times by K3 - against the reference's own MotionSpline.get_motion_vector."""
    from tests.common import load_golden
    from morphablegraphs_b200 import MotionStateGraphNode, synthetic as syn
    g = load_golden("s_timewarp")
    node = MotionStateGraphNode()
    node._initialize_from_json(None, syn.make_model_json(str(g["shape"])))
    for i, s in enumerate(g["samples"]):
        spline = node.back_project(s, use_time_parameters=True)
        fr = spline.get_motion_vector()
        assert fr.shape == g["tw%d_frames" % i].shape
        assert rel_err(fr, g["tw%d_frames" % i]).max() < RTOL


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 3, 255, 4097, 100_003, 1_000_000])
def test_topk_kernel_against_numpy(n):
    """K5c through the C-ABI (mgb_topk) on raw cost arrays: order (cost, global index) ascending, ties -> lowest
    index, NaN never selected, +inf selectable, slots beyond the selectable entries = (inf, index_base)
    (the reference keeps the first minimum: motion_primitive_generator.py:245-252)."""
    import ctypes as C
    import torch
    from morphablegraphs_b200 import _lib
    lib = _lib.load()
    rs = np.random.RandomState(n + 7)
    cost = rs.randint(0, max(2, n // 3 + 1), size=n).astype(np.float32)      # many ties
    if n > 10:
        cost[rs.randint(0, n, size=n // 10)] = np.nan
        cost[rs.randint(0, n, size=5)] = np.inf
        cost[rs.randint(0, n, size=3)] = -np.inf
    base = 12345
    d_cost = torch.as_tensor(cost, device="cuda") if n else torch.zeros(1, device="cuda")
    for k in (1, 5, 64):
        ws = torch.zeros(int(lib.mgb_score_workspace_bytes(n, k)), dtype=torch.uint8, device="cuda")
        bc = torch.empty(k, dtype=torch.float32, device="cuda")
        bi = torch.empty(k, dtype=torch.int64, device="cuda")
        rc = lib.mgb_topk(C.c_void_p(d_cost.data_ptr()), n, base, k, C.c_void_p(bc.data_ptr()), C.c_void_p(bi.data_ptr()),
                          C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, lib.mgb_last_error().decode()
        torch.cuda.synchronize()
        ok = np.where(~np.isnan(cost))[0]
        order = ok[np.lexsort((ok, cost[ok]))][:k]
        want_i = np.full(k, base, dtype=np.int64)
        want_c = np.full(k, np.inf, dtype=np.float32)
        want_i[:len(order)] = order + base
        want_c[:len(order)] = cost[order]
        assert list(bi.cpu().numpy()) == list(want_i), (n, k)
        assert np.array_equal(bc.cpu().numpy(), want_c), (n, k)


@pytest.mark.gpu
def test_topk_k_pass_form_on_an_array_beyond_the_one_pass_limit():
    """n > 592 x 49 152 costs: mgb_topk falls back to k grid arg-min passes.  Dense case against numpy, and the exhausted case
    (one finite and one infinite entry among NaNs, k = 5): the slots beyond the two selectable entries are (inf, index_base)."""
    import ctypes as C
    import torch
    from morphablegraphs_b200 import _lib
    lib = _lib.load()
    n, base, k = 29_200_000, 1000, 5
    ws = torch.zeros(int(lib.mgb_score_workspace_bytes(n, k)), dtype=torch.uint8, device="cuda")
    bc = torch.empty(k, dtype=torch.float32, device="cuda")
    bi = torch.empty(k, dtype=torch.int64, device="cuda")

    def run(d_cost):
        rc = lib.mgb_topk(C.c_void_p(d_cost.data_ptr()), n, base, k, C.c_void_p(bc.data_ptr()), C.c_void_p(bi.data_ptr()),
                          C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, lib.mgb_last_error().decode()
        torch.cuda.synchronize()
        return bc.cpu().numpy(), bi.cpu().numpy()

    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    cost = torch.randint(0, 1 << 20, (n,), generator=g, device="cuda").float()           # ties are likely among the smallest
    got_c, got_i = run(cost)
    h = cost.cpu().numpy()
    kth = np.partition(h, k)[k]
    cand = np.where(h <= kth)[0]
    order = cand[np.lexsort((cand, h[cand]))][:k]
    assert list(got_i) == list(order + base) and np.array_equal(got_c, h[order])
    sparse = torch.full((n,), float("nan"), device="cuda")
    sparse[17_000_123] = 3.5
    sparse[28_999_999] = float("inf")
    got_c, got_i = run(sparse)
    assert list(got_i) == [base + 17_000_123, base + 28_999_999, base, base, base]
    assert got_c[0] == 3.5 and np.all(np.isinf(got_c[1:]))
