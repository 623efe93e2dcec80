"""Timed CPU baseline for bench.py (`cpu_baseline` leg and `--impl reference`).

TEST INFRASTRUCTURE ONLY.  Times the oracle port (oracle/np_oracle.py: the
vectorised float64 numpy restatement of the reference's sample -> back-project
-> spline -> FK -> cost -> arg-min path) on the host cores, one worker process
per core, each drawing its own GMM samples exactly like the reference's sampler
(sklearn GaussianMixture.sample algorithm, np_oracle.gmm_sample) and scoring
them in chunks.  Also times the per-candidate loop port
(np_oracle.loop_evaluate_one_position), which has the reference's own call
structure (one Python iteration, P splev calls and a chain of 4x4 products per
candidate - reference motion_primitive_generator.py:253-257) and therefore its
single-core speed.

The reference itself cannot run on the GPU box (/root/reference is absent there
and its third-party dependencies are not installed), so kind = "port".
Nothing in here imports torch or touches CUDA (workers are spawned).
"""
import multiprocessing as mp
import os
import time

import numpy as np

_STATE = {}


def build_case(golden="s_config1", shape="S"):
    """Same workload as bench.py's GPU arm: model from seeds, constraint list from the committed
    golden fixture's specs."""
    import json
    from morphablegraphs_b200 import synthetic as syn
    from morphablegraphs_b200 import Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    from oracle import np_oracle as O
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    g = np.load(os.path.join(root, "tests", "golden", golden + ".npz"), allow_pickle=False)
    cfg = syn.SHAPES[shape]
    skeleton = Skeleton.from_desc(syn.make_skeleton_desc(cfg["n_joints"]))
    node = MotionStateGraphNode()
    node._initialize_from_json(None, syn.make_model_json(shape, translation_maxima=tuple(g["translation_maxima"])))
    ma = O.ModelArrays(node)
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = skeleton
    mpc.is_local = bool(g["is_local"])
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, skeleton, json.loads(str(g["specs_json"])))
    prev = g["prev_frames"] if "prev_frames" in g.files else None
    return node, skeleton, mpc, ma, prev


def _init(golden, shape, quality_scale):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    _STATE["case"] = build_case(golden, shape)
    _STATE["quality_scale"] = quality_scale


def _work(args):
    from oracle import np_oracle as O
    seed, n, chunk = args
    node, skeleton, mpc, ma, prev = _STATE["case"]
    rng = np.random.RandomState(seed)
    best = (np.inf, -1)
    done = 0
    while done < n:
        m = min(chunk, n - done)
        S, _ = O.gmm_sample(ma, m, rng)
        cost = O.candidate_costs(ma, skeleton, mpc, S, prev, quality_scale=_STATE["quality_scale"])
        i, e = O.first_argmin(cost)
        if e < best[0]:
            best = (e, done + i)
        done += m
    return best[0], n


class Pool(object):
    """Persistent worker pool so that process start-up is outside the timed region."""

    def __init__(self, golden="s_config1", shape="S", quality_scale=0.0, workers=None, chunk=16384):
        self.workers = workers or os.cpu_count() or 1
        self.chunk = chunk
        ctx = mp.get_context("spawn")
        # one BLAS thread per worker: the variables must be in the environment BEFORE the child imports
        # numpy (the initializer runs too late), otherwise workers x cores BLAS threads spin on each other
        saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")}
        for k in saved:
            os.environ[k] = "1"
        try:
            self._start(ctx, golden, shape, quality_scale)
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v

    def _start(self, ctx, golden, shape, quality_scale):
        self.pool = ctx.Pool(self.workers, initializer=_init, initargs=(golden, shape, quality_scale))
        # make sure every worker has finished its initialiser
        self.pool.map(_work, [(i, 64, 64) for i in range(self.workers * 2)])

    def run(self, n_total, seed=0, chunk=None):
        """Sample and score n_total candidates across the workers; returns seconds."""
        chunk = chunk or self.chunk
        per = int(np.ceil(n_total / float(self.workers * 4)))
        jobs = []
        left = n_total
        j = 0
        while left > 0:
            m = min(per, left)
            jobs.append((seed * 100003 + j, m, chunk))
            left -= m
            j += 1
        t0 = time.perf_counter()
        res = self.pool.map(_work, jobs)
        dt = time.perf_counter() - t0
        assert sum(r[1] for r in res) == n_total
        return dt

    def close(self):
        self.pool.close()
        self.pool.join()


def loop_port_rate(golden="s_config1", shape="S", n=1500, seed=3):
    """candidates/s of the per-candidate loop port on one core (one position constraint)."""
    from oracle import np_oracle as O
    node, skeleton, mpc, ma, _ = build_case(golden, shape)
    S, _ = O.gmm_sample(ma, n, np.random.RandomState(seed))
    t0 = time.perf_counter()
    O.loop_evaluate_one_position(ma, skeleton, mpc.constraints[0], S)
    return n / (time.perf_counter() - t0)


def tree_build_rate(golden="s_config1", shape="S", m=10000, seed=4):
    """Config 5 on the CPU: the oracle's recursive cluster-tree build (np_oracle.build_cluster_tree_dict: numpy Lloyd per
    node, kd-tree leaves - the reference's cluster_tree_node_builder.py call structure) over m latent samples, one core."""
    from oracle import np_oracle as O
    node, skeleton, mpc, ma, _ = build_case(golden, shape)
    S, _ = O.gmm_sample(ma, m, np.random.RandomState(seed))
    t0 = time.perf_counter()
    O.build_cluster_tree_dict(S, 4, 8, ma.K if hasattr(ma, "K") else 10)
    dt = time.perf_counter() - t0
    return {"value": m / dt, "unit": "samples/s", "cores": 1, "seconds": dt,
            "sample": "%d latent samples, 4 subdivisions, max_level 8, recursive build (oracle port of cluster_tree_node_builder.py)" % m}


# ---------------------------------------------------------------------------------------------
# kind = "reference": the reference's OWN loop (MotionPrimitiveGenerator.evaluate_samples_using_constraints,
# motion_primitive_generator.py:230-261 -> obj_spatial_error_sum -> MotionPrimitiveConstraints.evaluate -> the
# reference's constraint classes, MotionPrimitive.back_project, MotionSpline.evaluate, sklearn GMM sampler), loaded
# unmodified by oracle/refload.py from /root/reference or from the staged copy oracle/_ref/ (GPU box).  Only the
# third-party anim_utils / transformations arithmetic underneath comes from oracle/standin (absent upstream packages).
# The loop is single-threaded Python; "all host cores" = one independent copy of the loop per core.
# ---------------------------------------------------------------------------------------------
def reference_available():
    from oracle import refload
    return refload.available()


def _init_ref(golden, shape, quality_scale):
    import io
    import json
    import sys
    from oracle import refload, make_golden
    from morphablegraphs_b200 import synthetic as syn
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    g = np.load(os.path.join(root, "tests", "golden", golden + ".npz"), allow_pickle=False)
    sys.stdout = io.StringIO()           # the reference prints "best sample ..." per call and "MGRD is not installed" on import
    ns = refload.load()
    cfg = syn.SHAPES[shape]
    skeleton = refload.make_skeleton(syn.make_skeleton_desc(cfg["n_joints"]))
    node = refload.make_graph_node(ns, syn.make_model_json(shape, translation_maxima=tuple(g["translation_maxima"])))
    mpc = make_golden.build_mp_constraints(ns, skeleton, json.loads(str(g["specs_json"])), is_local=bool(g["is_local"]))
    prev = g["prev_frames"] if "prev_frames" in g.files else None
    gen = ns.motion_primitive_generator.MotionPrimitiveGenerator.__new__(ns.motion_primitive_generator.MotionPrimitiveGenerator)
    ref_obj = ns.objective_functions.obj_spatial_error_sum
    if quality_scale != 0.0:
        # the reference's scalar naturalness objective never returns (quirk Q7); its usable form is the expression of
        # objective_functions.py:256 added to the spatial error, evaluated with the reference's own GMM object
        gmm = node.get_gaussian_mixture_model()

        def objective(s, data):
            return ref_obj(s, data) - gmm.score(np.asarray(s).reshape(1, -1)) * quality_scale
        gen.objective = objective
    else:
        gen.objective = ref_obj
    _STATE["ref"] = (gen, node, mpc, prev)


def _work_ref(args):
    seed, n = args
    gen, node, mpc, prev = _STATE["ref"]
    np.random.seed(seed)                  # sklearn's sampler draws from numpy's global state (motion_primitive.py:182-189)
    samples = node.sample_low_dimensional_vectors(n)
    best, err = gen.evaluate_samples_using_constraints(samples, node, mpc, prev)
    sys_out = __import__("sys").stdout
    if hasattr(sys_out, "truncate"):
        sys_out.seek(0)
        sys_out.truncate(0)
    return float(err), n


class RefPool(Pool):
    """Same interface as Pool, every worker running the reference's own per-candidate loop."""

    def _start(self, ctx, golden, shape, quality_scale):
        self.pool = ctx.Pool(self.workers, initializer=_init_ref, initargs=(golden, shape, quality_scale))
        self.pool.map(_work_ref, [(i, 2) for i in range(self.workers * 2)])

    def run(self, n_total, seed=0, chunk=None):
        per = max(1, int(np.ceil(n_total / float(self.workers))))
        jobs, left, j = [], n_total, 0
        while left > 0:
            m = min(per, left)
            jobs.append((seed * 100003 + j, m))
            left -= m
            j += 1
        t0 = time.perf_counter()
        res = self.pool.map(_work_ref, jobs)
        dt = time.perf_counter() - t0
        assert sum(r[1] for r in res) == n_total
        return dt
