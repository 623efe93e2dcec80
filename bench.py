#!/usr/bin/env python
"""bench.py -- motion candidates evaluated / second (sample -> back-project -> FK -> cost -> arg-min).

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0;
N > 1 is launched by torchrun (one rank per GPU, NCCL).  `--impl reference` times the reference's own
CPU loop (oracle/_ref staged copy or /root/reference, through oracle/refload.py; the oracle port when
neither is present) on the host cores, on the same workload and `config`.

Workloads (BASELINE.json `configs`):
  L4M  (headline, configs[3] = the largest single-GPU configuration): large model (J=70, P=283, F=100, Nb=20,
       K=100, D=103, G=32), 4,000,000 candidates per step PER GPU, the constraint set SURVEY.md 8(d) row 4 names
       (right-hand position at F-1, root position + 2-D direction at F-1, 5-joint transition PoseConstraint at
       keyframes 0/1; tests/golden/l_cfg4.npz) + naturalness 0.1, arg-min.
  S1M  (configs[1]): synthetic walk-leftStance primitive (J=19, P=79, F=50, Nb=10, K=10, D=13, G=8), 1,000,000
       GMM samples per step per GPU, one joint-position constraint (tests/golden/s_config1.npz), arg-min.
  L4M_stress: the L model on the nine-constraint mixed set incl. a collision-avoidance minimum over all 100
       frames (tests/golden/l_mixed.npz; 542 contracted quads): a program-dominated stress case.
The headline fields of the line describe `--workload` (default L4M); the other two are measured in the same run
and reported under `workloads`, each with its own roofline / e2e / cpu_baseline.

A step (device arm, `value`): ONE fused launch (mgb_sample_score: K1 GMM sampling + K6 score + arg-min; for N > 1
followed by the exchange of every rank's best (cost, index)); model and constraint tables resident in HBM.
`e2e`: the reference-facing call with HOST samples (float64, pinned) -> mgb_score_host -> best (cost, index) back
on the host; H2D and D2H inside the timed region.  `e2e.generator_call`: the planner-facing call,
MotionPrimitiveGenerator.generate_constrained_sample (constraints in, best latent vector out; candidates drawn on
the device), timed by wall clock with the constraint compile + upload inside.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "motion candidates evaluated/sec (backproject+FK+cost)"
UNIT = "candidates/s"
L2_NOTE = "flushed (384 MB fill, untimed) before every timed step"
WORKLOADS = {
    "L4M": dict(golden="l_cfg4", shape="L", n_per_gpu=4_000_000, quality_scale=0.1,
                desc="L primitive (J=70,P=283,F=100,Nb=20,K=100,D=103,G=32), 4M samples/step/GPU, SURVEY 8(d) row-4 constraints "
                     "(right-hand position, root position + 2-D direction at F-1, 5-joint pose/transition at 0/1) + naturalness 0.1, arg-min"),
    "S1M": dict(golden="s_config1", shape="S", n_per_gpu=1_000_000, quality_scale=0.0,
                desc="S primitive (J=19,P=79,F=50,Nb=10,K=10,D=13,G=8), 1M GMM samples/step/GPU, "
                     "one joint-position constraint, arg-min"),
    "L4M_stress": dict(golden="l_mixed", shape="L", n_per_gpu=4_000_000, quality_scale=0.1,
                       desc="L primitive, 4M samples/step/GPU, nine mixed constraints incl. a collision-avoidance minimum over all "
                            "100 frames (542 contracted quads) + naturalness 0.1, arg-min"),
}
SECONDARY = {"L4M": ["S1M", "L4M_stress"], "S1M": ["L4M"], "L4M_stress": []}

# dram__bytes_read.sum + dram__bytes_write.sum of the step kernel, per launch, from the committed
# `ncu --set full` captures (profiles/); captures ran 1M candidates, scaled to the workload's launch
TRAFFIC = {
    "L4M": {"bytes": 6031104.0, "source": "profiles/r02_x_ncu_L_cfg4_raw.csv: 6.0 MB read + 0 written per 4M-candidate launch (the packed operands come "
                                          "from HBM once and are re-read from L2: 92.6 GB of L2 -> SM traffic per launch; candidates are never stored)"},
    "S1M": {"bytes": 123904.0, "source": "profiles/r01_zb_ncu_k1k6_S_raw.csv: 88.6 KB read + 35.3 KB written per 1M-candidate launch "
                                          "(tables + arg-min partials; candidates are never stored)"},
    "L4M_stress": {"bytes": 4.0 * 7906304.0, "source": "profiles/r01_zb_ncu_k1k6_L_raw.csv: 7.90 MB read + 5.6 KB written per 1M candidates "
                                                        "(packed operands re-read from L2/HBM per tile), x4 for the 4M launch"},
}
# sm__pipe_tensor_cycles_active (ncu, % of peak sustained active) of the tensor-core step kernel, from the committed captures
TENSOR_PIPE_NCU = {"L4M": 70.2}      # profiles/r02_x_ncu_L_cfg4_raw.csv (71.7 in r02_p; 38.7 at the start of round 2: profiles/r02_a_ncu_L_cfg4_raw.csv)


def config_of(workload, world):
    """The `config` object of the JSON line; the reference arm prints the identical dict for the same command line."""
    w = WORKLOADS[workload]
    return {"workload": workload, "description": w["desc"], "candidates_per_gpu_per_step": w["n_per_gpu"], "l2": L2_NOTE,
            "parallelism": "candidates sharded x%d" % world, "exchange": "per-rank best (cost, global index), 16 bytes per rank per step"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="L4M", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: n_per_gpu candidates on every GPU; strong: n_per_gpu candidates in total, split over the GPUs")
    ap.add_argument("--only", action="store_true", help="measure only --workload (no `workloads` key)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="budget of the headline cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-staged", action="store_true", help="skip the staged-pipeline kernel rooflines")
    ap.add_argument("--no-tree", action="store_true", help="skip the config-5 side measurement (`tree` key)")
    return ap.parse_args()


def build_case(workload):
    """Model, skeleton and constraint list of the workload, rebuilt from seeds + the committed
    golden fixture's constraint specs (product host classes only; no oracle, no reference)."""
    from morphablegraphs_b200 import synthetic as syn
    from morphablegraphs_b200 import Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
    w = WORKLOADS[workload]
    g = np.load(os.path.join(ROOT, "tests", "golden", w["golden"] + ".npz"), allow_pickle=False)
    cfg = syn.SHAPES[w["shape"]]
    skeleton = Skeleton.from_desc(syn.make_skeleton_desc(cfg["n_joints"]))
    node = MotionStateGraphNode()
    node._initialize_from_json(None, syn.make_model_json(w["shape"], translation_maxima=tuple(g["translation_maxima"])))
    node.set_skeleton(skeleton)
    mpc = MotionPrimitiveConstraints()
    mpc.skeleton = skeleton
    mpc.is_local = bool(g["is_local"])
    mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, skeleton, json.loads(str(g["specs_json"])))
    prev = g["prev_frames"] if "prev_frames" in g.files else None
    return node, skeleton, mpc, prev


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, windows):
        """windows: list of (t0, t1) wall-clock intervals of the timed regions (samples outside are idle time)."""
        sm, mx, reasons, pw = [], 0.0, set(), 0.0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.lines:
            if windows and not any(t0 - 0.05 <= ts <= t1 + 0.15 for t0, t1 in windows):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                pw = max(pw, float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": pw}


# ---------------------------------------------------------------------------
# CPU arm (the ONLY place bench.py executes oracle/): the reference's own loop when its modules are present
# (kind "reference": /root/reference in the build container, the staged copy oracle/_ref/ on the GPU box),
# else the oracle port (kind "port")
# ---------------------------------------------------------------------------
def cpu_pool(workload, kind):
    from oracle import cpu_baseline as cb
    w = WORKLOADS[workload]
    if kind == "reference":
        return cb.RefPool(w["golden"], w["shape"], w["quality_scale"])
    return cb.Pool(w["golden"], w["shape"], w["quality_scale"], chunk=16384 if w["shape"] == "S" else 512)


def cpu_kind():
    from oracle import cpu_baseline as cb
    return "reference" if cb.reference_available() else "port"


def calibrate_cpu(pool, seconds, cap):
    """Candidates per sample so that one sample takes about `seconds` on this host: grow a probe
    batch until it takes >= 0.5 s (the L model is ~400x the S model per candidate), then scale."""
    n0 = 8 * pool.workers
    while True:
        dt = pool.run(n0, seed=11)
        if dt >= 0.5 or n0 >= cap:
            break
        n0 *= 4
    rate = n0 / dt
    return int(min(cap, max(pool.workers * 2, int(rate * seconds)))), rate


CPU_SAMPLE_TEXT = {
    "reference": "%d candidates per step (of the %d-candidate workload): sklearn GMM sampler + the reference's OWN per-candidate loop "
                 "(MotionPrimitiveGenerator.evaluate_samples_using_constraints, unmodified modules via oracle/refload.py; anim_utils / "
                 "transformations from oracle/standin), one independent loop per host core",
    "port": "%d candidates per step (of the %d-candidate workload): sampled + scored + arg-min by oracle/np_oracle.py "
            "(vectorised float64 numpy restatement), one process per core",
}


def run_reference(args, emit=print):
    """`--impl reference`: each step = a bounded sample of the workload (sample -> score -> arg-min) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = WORKLOADS[args.workload]
    kind = cpu_kind()
    pool = cpu_pool(args.workload, kind)
    per_step, _ = calibrate_cpu(pool, min(4.0, 120.0 / max(1, args.steps + args.warmup // 4)), w["n_per_gpu"])
    for i in range(args.warmup):
        pool.run(max(pool.workers, per_step // 4), seed=100 + i)
    t = 0.0
    for i in range(args.steps):
        t += pool.run(per_step, seed=200 + i)
    pool.close()
    value = per_step * args.steps / t
    sample = CPU_SAMPLE_TEXT[kind] % (per_step, w["n_per_gpu"])
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
           "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": config_of(args.workload, args.gpus),
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": pool.workers, "kind": kind, "sample": sample},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(json.dumps(out))
    return 0


def cpu_baseline_leg(workload, seconds, extras=False):
    """Bounded sample of the workload on the host cores: the reference's own loop when available (kind "reference"),
    and the vectorised port beside it (the fairer ceiling: batched numpy instead of a Python loop)."""
    from oracle import cpu_baseline as cb
    w = WORKLOADS[workload]
    kind = cpu_kind()
    out = None
    for k in ([kind, "port"] if kind == "reference" else ["port"]):
        pool = cpu_pool(workload, k)
        n, _ = calibrate_cpu(pool, max(2.0, seconds * (0.6 if k == kind else 0.3)), w["n_per_gpu"])
        dt = pool.run(n, seed=5)
        pool.close()
        ent = {"value": n / dt, "unit": UNIT, "cores": pool.workers, "kind": k, "sample": CPU_SAMPLE_TEXT[k] % (n, w["n_per_gpu"])}
        if out is None:
            out = ent
        else:
            out["vectorised_port"] = ent
    if extras:
        out["tree_build_port"] = cb.tree_build_rate("s_config1", "S", m=10000)
    return out


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
class L2Flusher(object):
    """Between timed iterations: overwrite a buffer larger than the 126 MB L2 (untimed)."""

    def __init__(self, torch, device, nbytes=384 << 20):
        self.buf = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
        self.v = 0.0

    def __call__(self):
        self.v += 1.0
        self.buf.fill_(self.v)


def timed(torch, fn, reps, flush, warmup=3):
    """Mean device time (ms) of fn() over reps, CUDA events on the current stream, L2 flushed
    (untimed) before every repetition."""
    for _ in range(warmup):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    torch.cuda.synchronize()
    for e0, e1 in evs:
        flush()
        e0.record()
        fn()
        e1.record()
    torch.cuda.synchronize()
    return float(np.mean([e0.elapsed_time(e1) for e0, e1 in evs]))


def algorithmic_work(dm, dc, n, quality_scale):
    """Per-launch algorithmic bytes / flop of each kernel (SURVEY.md section 8(d) x n candidates)."""
    D, K, P, Nb, G = dm.D, dm.K, dm.P, dm.Nb, dm.G
    T = len(dc.times)
    nq = dc.n_quads
    fk_flop = 100.0 * nq                      # ~100 flop per joint on a walked chain (a14)
    gmm_flop = 2.0 * G * D * D if quality_scale != 0.0 else 0.0
    return {
        "K1_gmm_sample": dict(bound="hbm", bytes=4.0 * D * n, flop=2.0 * D * D * n),
        "K6_fused_score": dict(bound="fp32", bytes=(4.0 * D + 4.0) * n, flop=(2.0 * K * 4 * nq + fk_flop + gmm_flop) * n),
        "K1K6_sample_score": dict(bound="fp32", bytes=4.0 * n, flop=(2.0 * D * D + 2.0 * K * 4 * nq + fk_flop + gmm_flop) * n),
        "K2_backproject": dict(bound="tensor", bytes=(4.0 * K + 4.0 * Nb * P) * n, flop=2.0 * K * Nb * P * n),
        "K3_spline_eval": dict(bound="hbm", bytes=(4.0 * P * min(Nb, 4 * T) + 4.0 * T * P) * n, flop=8.0 * T * P * n),
        "K5a_score_frames": dict(bound="hbm", bytes=(4.0 * 4 * nq + 4.0) * n, flop=(fk_flop + gmm_flop) * n),
    }


class Gpu(object):
    """Process-wide plumbing: device, process group, L2 flusher, peer exchange, one pinned host buffer, peaks."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = L2Flusher(torch, self.dev)
        self.peer, self.exchange_kind = None, "none (1 GPU)"
        if self.world > 1:
            from morphablegraphs_b200 import sharding
            try:
                self.peer = sharding.PeerExchange(self.dev)
                self.exchange_kind = "in the score kernel's epilogue (mgb_sample_score_sharded): P2P stores into every peer's slot array over NVLink + first-minimum reduction by the folding CTA"
            except Exception as e:                    # symmetric memory unavailable: NCCL all-gather (still a GPU path)
                self.exchange_kind = "nccl all_gather of (cost, index) [peer memory unavailable: %s]" % str(e)[:120]
            flag = torch.tensor([1.0 if self.peer is not None else 0.0], device=self.dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if float(flag.item()) == 0.0:
                self.peer = None
        self.gathered = torch.empty((self.world, 2), dtype=torch.float64, device=self.dev) if self.world > 1 else None
        self.host = None
        self.windows = []
        self.peaks = {}
        pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk_path):
            self.peaks = json.load(open(pk_path))
        self.fp32_peak = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, x):
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def pinned(self, n, D):
        """[n, D] float64 view of ONE pinned host allocation shared by all workloads (the largest is 3.3 GB)."""
        need = n * D
        if self.host is None or self.host.numel() < need:
            self.host = None
            self.host = self.torch.empty(need, dtype=self.torch.float64).pin_memory()
        return self.host[:need].view(n, D)

    def fma_peak(self):
        if self.fp32_peak is None:
            import ctypes as C
            from morphablegraphs_b200 import _lib
            tf = C.c_double()
            _lib.load().mgb_fp32_fma_peak(self.local, C.byref(tf))
            self.fp32_peak = float(tf.value)
        return self.fp32_peak


def measure_steps(gpu, dm, dc, n, n_total, base, qs, counts, steps, warmup):
    """W warm-up steps, then exactly `steps` timed steps bracketed by barrier + synchronize; CUDA events per step on
    the launching stream (torch's current stream), L2 flushed untimed before each; returns (sum of step ms, max over
    ranks; launches of this library inside the timed region on this rank)."""
    torch = gpu.torch
    from morphablegraphs_b200 import engine
    state = {"seed": 0}

    def step():
        state["seed"] += 1
        if gpu.peer is not None:       # one launch: the exchange of the per-rank best runs in the score kernel's epilogue
            return dc.sample_score_sharded(n, gpu.peer, seed=state["seed"], counts=counts, n_total=n_total, index_base=base, quality_scale=qs)
        res = dc.sample_score(n, seed=state["seed"], counts=counts, n_total=n_total, index_base=base, k=1, quality_scale=qs)
        if gpu.world > 1:
            mine = torch.stack([res["best_cost"][0].double(), res["best_idx"][0].double()])
            gpu.dist.all_gather_into_tensor(gpu.gathered.view(-1), mine)
            return gpu.gathered[torch.argmin(gpu.gathered[:, 0])]     # first minimum: ranks hold ascending index ranges
        return res

    for _ in range(max(warmup, 3)):
        step()
    l0 = engine.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    gpu.barrier()
    t0 = time.perf_counter()
    for e0, e1 in evs:
        gpu.flush()
        e0.record()
        step()
        e1.record()
    torch.cuda.synchronize()
    gpu.barrier()
    gpu.windows.append((t0, time.perf_counter()))
    launches = engine.launch_count() - l0
    if gpu.peer is not None:
        gpu.peer.check()
    total_ms = gpu.max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in evs))
    return total_ms, launches


def measure_e2e(gpu, case, dm, dc, n, n_total, base, qs, counts, steps):
    """The reference-facing calls through HOST buffers, copies inside the timed region."""
    torch = gpu.torch
    from morphablegraphs_b200 import engine, MotionPrimitiveGenerator
    from morphablegraphs_b200.objective_functions import ObjSpatialErrorAndNaturalness
    node, skeleton, mpc, prev = case
    S = torch.empty((n, dm.D), dtype=torch.float32, device=gpu.dev)
    dm.sample(n, seed=2024, counts=counts, n_total=n_total, index_base=base, out=S)
    host = gpu.pinned(n, dm.D)
    chunk = 1 << 18
    for lo in range(0, n, chunk):                       # widen on the device chunk by chunk (no 2x host temporary)
        host[lo:lo + chunk].copy_(S[lo:lo + chunk].double())
    torch.cuda.synchronize()
    e2e_steps = max(3, min(steps, 10))
    for _ in range(2):
        dc.score_host(host, k=1, quality_scale=qs, index_base=base)
    torch.cuda.synchronize()
    gpu.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        r = dc.score_host(host, k=1, quality_scale=qs, index_base=base)
        if gpu.world > 1:
            mine = torch.tensor([float(r["best_cost"][0]), float(r["best_idx"][0])], dtype=torch.float64, device=gpu.dev)
            gpu.dist.all_gather_into_tensor(gpu.gathered.view(-1), mine)
            gpu.gathered.cpu()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    gpu.windows.append((t0, t1))
    e2e_t = gpu.max_over_ranks(t1 - t0)
    # plain pinned H2D copy of the same bytes: what the PCIe link of this box gives to this rank (context for e2e)
    dst = torch.empty((n, dm.D), dtype=torch.float64, device=gpu.dev)
    dst.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        dst.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    h2d_gbs = 3 * n * dm.D * 8 / (time.perf_counter() - t0) * 1e-9
    del dst, S
    e2e = {"value": n_total * e2e_steps / e2e_t, "unit": UNIT, "h2d_bytes_per_step": int(n * dm.D * 8) * gpu.world,
           "d2h_bytes_per_step": 12 * gpu.world, "steps": e2e_steps, "ms_per_step": 1e3 * e2e_t / e2e_steps,
           "call": "MotionPrimitiveGenerator.evaluate_samples_using_constraints -> mgb_score_host, float64 pinned host samples",
           "pinned_h2d_copy_gbs_per_rank": h2d_gbs, "h2d_share_of_step": (n * dm.D * 8 / (h2d_gbs * 1e9)) / (e2e_t / e2e_steps)}
    # ---- the generator call itself: what the reference's planner calls once per motion-primitive step (reference
    # motion_primitive_generator.py:126-161 -> :192-206).  Host inputs: the step's constraint objects; a target is edited
    # in place before every call (a new step brings new constraints), so the program is re-flattened, re-compiled and
    # re-uploaded inside the timed call; candidates are drawn on the device; host outputs: the winning latent vector + error.
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": n_total, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=11)
    if qs != 0.0:
        gen.objective = ObjSpatialErrorAndNaturalness(1.0, qs)
    c0 = mpc.constraints[0]
    state = {"i": 0}

    def gen_call():
        state["i"] += 1
        c0.weight_factor = 1.0 + 1e-6 * (state["i"] % 2)       # content changes -> cache miss, like a new step's constraints
        return gen.generate_constrained_sample(node, mpc, prev_frames=prev)

    for _ in range(2):
        best_s = gen_call()
    torch.cuda.synchronize()
    gpu.barrier()
    l0 = engine.launch_count()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        best_s = gen_call()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    gpu.windows.append((t0, t1))
    gen_launches = engine.launch_count() - l0
    gen_t = gpu.max_over_ranks(t1 - t0)
    c0.weight_factor = 1.0
    e2e["generator_call"] = {"value": n_total * e2e_steps / gen_t, "unit": UNIT, "ms_per_step": 1e3 * gen_t / e2e_steps, "steps": e2e_steps,
                             "launches_per_step": gen_launches / e2e_steps, "d2h_bytes_per_step": int(best_s.shape[0]) * 4 + 12,
                             "call": "MotionPrimitiveGenerator.generate_constrained_sample(node, constraints), n_random_samples = %d: "
                                     "constraint program compiled + uploaded, candidates drawn on the device (K1 fused), best latent "
                                     "vector and min_error returned to the host" % n_total,
                             "min_error": float(mpc.min_error)}
    if gpu.world == 1:
        # ---- local optimiser: batched Levenberg-Marquardt on the residual-vector objective, 8 seeds in one
        # launch (mgb_lm_refine), pose constraints dropped as motion_primitive_generator.py:181 does
        try:
            from morphablegraphs_b200 import MotionPrimitiveConstraints as _MPC
            mo = _MPC()
            mo.skeleton, mo.is_local = mpc.skeleton, mpc.is_local
            mo.constraints = [c_ for c_ in mpc.constraints if c_.constraint_type != "keyframe_pose"]
            dco = engine.DeviceConstraints(dm, mo, prev)
            seeds = dm.sample(8, seed=77)
            r0 = dco.lm_refine(seeds, quality_scale=qs, max_iter=100)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                r0 = dco.lm_refine(seeds, quality_scale=qs, max_iter=100)
            ms_lm = 1e3 * (time.perf_counter() - t0) / 5
            e2e["optimizer"] = {"seeds": 8, "max_jacobians": 100, "ms_per_batch": ms_lm, "residuals": dco.n_residuals(),
                                "cost_start_mean": float(np.mean(r0["cost_start"])), "cost_mean": float(np.mean(r0["cost"])),
                                "jacobians_mean": float(np.mean(r0["iterations"])),
                                "call": "DeviceConstraints.lm_refine -> mgb_lm_refine: one CTA per seed, forward-difference Jacobian + "
                                        "damped normal equations in shared memory, one launch for the whole minimisation"}
        except engine.MgbError as e:
            e2e["optimizer"] = {"unavailable": str(e)}
    if gpu.world == 1 and dm.D <= 32:
        # ---- cross-option batching: the planner's option loop (graph_walk_planner.py:216-226,
        # mg_state_planner.py:946-968: every outgoing edge scored with n_random_samples = 100, one after another) as ONE
        # launch (generate_constrained_samples -> mgb_sample_score_multi) against one generate_constrained_sample per option
        from morphablegraphs_b200 import MotionPrimitiveConstraints
        n_opt, n_s = 4, 100
        gen4 = MotionPrimitiveGenerator(None, {"n_random_samples": n_s, "use_local_coordinates": prev is None}, skeleton=skeleton, seed=12)
        opts = []
        for j in range(n_opt):
            o = MotionPrimitiveConstraints()
            o.skeleton, o.is_local, o.constraints = mpc.skeleton, mpc.is_local, mpc.constraints
            opts.append((node, o, prev))

        def touch():
            state["i"] += 1
            c0.weight_factor = 1.0 + 1e-6 * (state["i"] % 2)

        def timed(fn, reps=50):
            for _ in range(5):
                touch(); fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                touch(); fn()
            torch.cuda.synchronize()
            return 1e3 * (time.perf_counter() - t0) / reps

        l0 = engine.launch_count()
        ms_batch = timed(lambda: gen4.generate_constrained_samples(opts))
        launches_batch = (engine.launch_count() - l0) / 55.0
        ms_seq = timed(lambda: [gen4.generate_constrained_sample(o[0], o[1], prev_frames=o[2]) for o in opts])
        c0.weight_factor = 1.0
        e2e["option_batch"] = {"options": n_opt, "samples_per_option": n_s, "ms_one_launch": ms_batch, "ms_one_call_per_option": ms_seq,
                               "launches_per_batch": launches_batch,
                               "call": "MotionPrimitiveGenerator.generate_constrained_samples(options) -> mgb_sample_score_multi: constraint programs "
                                       "compiled + uploaded, all options drawn, scored and reduced in one launch, results in one D2H copy"}
    return e2e


def kernel_rooflines(gpu, dm, dc, n, n_total, qs, counts, staged):
    """Per-kernel device times (CUDA events, L2 flushed before each launch) against the measured
    peaks: HBM copy bandwidth from MEASURED_PEAKS.json, fp32 FMA rate measured live by the library."""
    torch, flush = gpu.torch, gpu.flush
    hbm_peak = float(gpu.peaks.get("hbm_gbs", 6650.0))
    fp32_peak = gpu.fma_peak()
    work = algorithmic_work(dm, dc, n, qs)
    S = torch.empty((n, dm.D), dtype=torch.float32, device=gpu.dev)
    reps = 10
    times = {}
    times["K1_gmm_sample"] = timed(torch, lambda: dm.sample(n, seed=7, counts=counts, n_total=n_total, out=S), reps, flush)
    times["K6_fused_score"] = timed(torch, lambda: dc.score(S, k=1, quality_scale=qs), reps, flush)
    times["K1K6_sample_score"] = timed(torch, lambda: dc.sample_score(n, seed=7, counts=counts, n_total=n_total, k=1, quality_scale=qs), reps, flush)
    if staged:
        ns = min(n, 1_000_000 if dm.K <= 16 else 250_000)     # staged intermediates: n * Nb * P * 4 bytes
        Ss = S[:ns]
        coeffs = torch.empty((ns, dm.Nb, dm.P), dtype=torch.float32, device=S.device)
        frames = torch.empty((ns, len(dc.times), dm.P), dtype=torch.float32, device=S.device)
        sw = algorithmic_work(dm, dc, ns, qs)
        for k in ("K2_backproject", "K3_spline_eval", "K5a_score_frames"):
            work[k] = sw[k]
            work[k]["n"] = ns
        times["K2_backproject"] = timed(torch, lambda: dm.backproject(Ss, impl=2, out=coeffs), 5, flush)      # tcgen05 3xTF32
        work["K2_backproject_simt"] = dict(work["K2_backproject"], bound="fp32")
        times["K2_backproject_simt"] = timed(torch, lambda: dm.backproject(Ss, impl=1, out=coeffs), 5, flush)
        times["K3_spline_eval"] = timed(torch, lambda: dm.eval_frames(coeffs, dc.times, out=frames), 5, flush)
        times["K5a_score_frames"] = timed(torch, lambda: dc.score(Ss, k=1, quality_scale=qs, frames=frames), 5, flush)
        del coeffs, frames
    out = {}
    for k, ms in times.items():
        wk = work[k]
        gbs = wk["bytes"] / (ms * 1e-3) * 1e-9
        tfs = wk["flop"] / (ms * 1e-3) * 1e-12
        out[k] = {"ms": ms, "n": wk.get("n", n), "bound": wk["bound"], "algorithmic_bytes": wk["bytes"], "algorithmic_flop": wk["flop"],
                  "achieved_gbs": gbs, "achieved_tflops": tfs, "hbm_frac": gbs / hbm_peak, "fp32_frac": tfs / fp32_peak if fp32_peak else None}
    del S
    return out


def roofline_of(gpu, workload, dm, kr):
    """`roofline` of the step's only kernel (K1K6 fused sample + score + arg-min)."""
    dom = "K1K6_sample_score"
    d = kr[dom]
    peaks = gpu.peaks
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    if dm.K > 16:
        # large latent spaces: the step kernel is mgb_score_tc_kernel - sampling operand, quad contraction and GMM log-density
        # all run on tcgen05 as 3xTF32 (three TF32 MMAs per fp32-accurate product).  Denominator: dense TF32 rate = half the
        # MEASURED dense bf16 BURST rate (the kernel runs unthrottled at the max SM clock, see `clocks`), divided by 3.
        bf16 = float(peaks.get("bf16_tflops", 1600.0))
        peak = bf16 / 2.0 / 3.0
        roof = {"kernel": dom + " (mgb_score_tc_kernel)", "bound": "tensor", "achieved": d["achieved_tflops"], "peak": peak,
                "unit": "TFLOP/s", "frac": d["achieved_tflops"] / peak, "traffic": None,
                "peak_source": "fp32-equivalent 3xTF32 rate = bf16_tflops burst (%s) / 2 (TF32) / 3 (split); achieved = algorithmic flop of "
                               "SURVEY.md 8(d) (sampling + quad contraction + FK + GMM log-density) per launch / CUDA-event time"
                               % ("MEASURED_PEAKS.json" if "bf16_tflops" in peaks else "fallback 1600"),
                "frac_vs_sustained_bf16": d["achieved_tflops"] / (float(peaks.get("bf16_tflops_sustained", bf16)) / 6.0)}
        if workload in TENSOR_PIPE_NCU:
            roof["tensor_pipe_active_pct_ncu"] = TENSOR_PIPE_NCU[workload]
    else:
        roof = {"kernel": dom + " (mgb_score_kernel)", "bound": "fp32", "achieved": d["achieved_tflops"], "peak": gpu.fma_peak(), "unit": "TFLOP/s",
                "frac": d["fp32_frac"], "traffic": None,
                "peak_source": "fp32 FMA rate measured live (mgb_fp32_fma_peak); SURVEY.md 8(d): K6 is compute-bound, "
                               "its HBM fraction (%.3f of %s) is not the limiter" % (d["hbm_frac"], peak_src)}
    traffic = TRAFFIC.get(workload)
    if traffic:
        roof["traffic"] = traffic["bytes"]
        roof["traffic_source"] = traffic["source"]
    return roof


def measure_workload(gpu, workload, args, steps, warmup, headline):
    """Everything the line reports for one workload: device step (`value`), strong-scaling variant (N > 1), kernels,
    roofline, e2e."""
    from morphablegraphs_b200 import engine, sharding
    w = WORKLOADS[workload]
    qs = w["quality_scale"]
    case = build_case(workload)
    node, skeleton, mpc, prev = case
    dm = engine.DeviceModel(node, skeleton, gpu.local)
    dc = engine.DeviceConstraints(dm, mpc, prev)
    res = {"description": w["desc"], "n_quads": dc.n_quads, "n_ops": dc.n_ops}

    def run(scaling, nsteps):
        if scaling == "strong":
            n_total = w["n_per_gpu"]
            base, n = sharding.shard_range(n_total, gpu.rank, gpu.world)
        else:
            n, n_total, base = w["n_per_gpu"], w["n_per_gpu"] * gpu.world, gpu.rank * w["n_per_gpu"]
        counts = dm.component_counts(n_total, seed=2024)          # one multinomial split for the whole job
        total_ms, launches = measure_steps(gpu, dm, dc, n, n_total, base, qs, counts, nsteps, warmup)
        return dict(n=n, n_total=n_total, base=base, counts=counts, total_ms=total_ms, launches=launches,
                    value=n_total * nsteps / (total_ms * 1e-3), ms_per_step=total_ms / nsteps)

    main = run(args.scaling, steps)
    res.update(value=main["value"], ms_per_step=main["ms_per_step"], steps=steps, scaling=args.scaling,
               candidates_per_step_all_gpus=main["n_total"], gpu_launches=main["launches"])
    if gpu.world > 1:
        other = "strong" if args.scaling == "weak" else "weak"
        o = run(other, max(3, min(steps, 10)))
        res[other + "_scaling"] = {"value": o["value"], "ms_per_step": o["ms_per_step"], "candidates_per_step_all_gpus": o["n_total"],
                                   "candidates_per_gpu_per_step": o["n"]}
    n, n_total, base, counts = main["n"], main["n_total"], main["base"], main["counts"]
    kr = kernel_rooflines(gpu, dm, dc, n, n_total, qs, counts, staged=headline and not args.no_staged)
    res["kernels"] = kr
    res["roofline"] = roofline_of(gpu, workload, dm, kr)
    res["e2e"] = measure_e2e(gpu, case, dm, dc, n, n_total, base, qs, counts, steps)
    del dc, dm
    gpu.torch.cuda.empty_cache()
    return res


def measure_config3(gpu, n_per_step=1_000_000):
    """BASELINE configs[2]: the action list of the reference's examples/example_input.json (walk, pickRight, carryRight,
    placeRight = 17 motion-primitive steps) over the synthetic 14-primitive graph, loaded from the reference's archive format
    into resident device models; 1M candidates per step in total, sharded over the ranks (125 k per GPU on 8), every step one
    planner-facing call (constraints built on the host from the walk so far -> compile + upload -> fused sample / score /
    arg-min -> exchange of the per-rank best -> winner back-projected and appended).  Wall clock, steps are sequential."""
    import io
    torch = gpu.torch
    from morphablegraphs_b200 import synthetic as syn, MotionPrimitiveGenerator, graph_synthesis as GS, engine
    from morphablegraphs_b200.motion_state_graph import MotionStateGraphLoader, graph_to_bytes
    blob = graph_to_bytes(syn.make_graph_data())
    t0 = time.perf_counter()
    ld = MotionStateGraphLoader()
    ld.set_data_source(io.BytesIO(blob))
    graph = ld.build(resident=True)
    torch.cuda.synchronize()
    load_s = time.perf_counter() - t0
    gen = MotionPrimitiveGenerator(None, {"n_random_samples": n_per_step, "use_local_coordinates": False}, skeleton=graph.skeleton, seed=21)
    synth = GS.GraphWalkSynthesizer(graph, gen)
    synth.run(syn.EXAMPLE_INPUT, max_steps=3)
    torch.cuda.synchronize()
    gpu.barrier()
    l0 = engine.launch_count()
    t0 = time.perf_counter()
    out = synth.run(syn.EXAMPLE_INPUT)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    gpu.windows.append((t0, t1))
    dt = gpu.max_over_ranks(t1 - t0)
    return {"value": len(out) * n_per_step / dt, "unit": UNIT, "motion_primitive_steps": len(out), "candidates_per_step_all_gpus": n_per_step,
            "candidates_per_gpu_per_step": n_per_step // gpu.world, "ms_per_motion_primitive_step": 1e3 * dt / len(out), "total_ms": 1e3 * dt,
            "launches_per_step": (engine.launch_count() - l0) / float(len(out)), "graph_load_s": load_s, "archive_bytes": len(blob),
            "primitives_resident": len(graph.nodes), "scaling": "strong", "timing": "wall clock around the whole synthesis, max over ranks",
            "mean_min_error": float(np.mean([r["min_error"] for r in out])),
            "description": "configs[2]: examples/example_input.json action list over the synthetic walk / pick / carry / place graph "
                           "(graph_synthesis.GraphWalkSynthesizer), non-local path (candidates aligned to the walk so far)"}


def main():
    args = parse_args()
    # exactly ONE line on stdout: libraries (NCCL prints its version banner there) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        os.write(real_stdout, (line + "\n").encode())

    if args.impl == "reference":
        return run_reference(args, emit)
    gpu = Gpu()
    clocks = ClockSampler(gpu.local)
    if gpu.rank == 0:
        clocks.start()
    head = measure_workload(gpu, args.workload, args, args.steps, args.warmup, headline=True)
    others = {}
    if not args.only:
        for name in SECONDARY[args.workload]:
            others[name] = measure_workload(gpu, name, args, max(3, min(args.steps, 10)), args.warmup, headline=False)
    c3 = None
    if not args.only:
        c3 = measure_config3(gpu)
    if gpu.rank == 0:
        clocks.stop()
    gpu.barrier()
    if gpu.rank != 0:
        if gpu.world > 1:
            gpu.dist.destroy_process_group()
        return 0
    out = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": gpu.world, "steps": args.steps, "warmup": max(args.warmup, 3),
           "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": config_of(args.workload, gpu.world),
           "exchange_impl": gpu.exchange_kind, "roofline": head["roofline"], "kernels": head["kernels"], "e2e": head["e2e"],
           "gpu_launches": head["gpu_launches"], "clocks": clocks.summary(gpu.windows)}
    for k in ("strong_scaling", "weak_scaling", "n_quads", "n_ops"):
        if k in head:
            out[k] = head[k]
    if gpu.world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_leg(args.workload, args.cpu_seconds, extras=True)
        for name in others:
            others[name]["cpu_baseline"] = cpu_baseline_leg(name, max(3.0, args.cpu_seconds * 0.5))
    if c3 is not None:
        others["C3"] = c3
    if others:
        out["workloads"] = others
    if gpu.world == 1 and not args.only and not args.no_tree:
        # BASELINE.json configs[4] (cluster-tree build and search over 1M latent samples of the S primitive): tools/tree_bench.py
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("mgb_tree_bench", os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "tree_bench.py"))
            tb = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(tb)
            out["tree"] = tb.run(argparse.Namespace(m=1_000_000, levels=8, no_features=True))
        except Exception as e:          # a side measurement must not cost the headline line
            out["tree"] = {"error": "%s: %s" % (type(e).__name__, e)}
    emit(json.dumps(out))
    if gpu.world > 1:
        gpu.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
