#!/usr/bin/env python
"""BASELINE.json config 5 on one GPU: cluster-tree build (k-means levels + kd forest) and search over M
latent samples of the S primitive, plus the euclidean-PCA feature chain.  Not a bench.py line (the
metric there is candidates/s of the sample-and-evaluate step); prints one JSON object for profiles/.

usage: python tools/tree_bench.py [--m 1000000] [--levels 8] [--no-features]
(the CPU port of the build is timed by bench.py's cpu_baseline leg, key tree_build_port)"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from morphablegraphs_b200 import engine, features as feat, space_partitioning as sp  # noqa: E402
from morphablegraphs_b200.objective_functions import obj_spatial_error_sum  # noqa: E402


def wall(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    return r, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=1_000_000)
    ap.add_argument("--levels", type=int, default=8)
    ap.add_argument("--cpu-m", type=int, default=0, help="unused (the CPU build is timed by bench.py's cpu_baseline leg: tree_build_port)")
    ap.add_argument("--no-features", action="store_true")
    print(json.dumps(run(ap.parse_args())))


def run(a):
    """a: namespace with m, levels, no_features.  Returns the result dictionary (bench.py stores it under its `tree` key)."""
    node, skeleton, mpc, prev = bench.build_case("S1M")
    dm = engine.DeviceModel(node, skeleton, 0)
    dc = engine.DeviceConstraints(dm, mpc, prev)
    M, K, D = a.m, dm.K, dm.D
    out = {"workload": "config 5: S primitive, M=%d latent samples, 4 subdivisions, max_level=%d" % (M, a.levels)}
    S = dm.sample(M, seed=2024)
    data = S.cpu().numpy().astype(np.float64)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", 6650.0))

    # ---- K7 alone: one Lloyd iteration (E-step + sums) over all M samples, root level and a deep level ----
    flush = bench.L2Flusher(torch, S.device)
    k7 = {}
    for name, n_nodes in (("root_level_1_node", 1), ("deep_level_16384_nodes", 16384)):
        per = M // n_nodes
        nd = torch.repeat_interleave(torch.arange(n_nodes, device=S.device, dtype=torch.int32), per)
        x = S[:per * n_nodes].contiguous()
        cent = x[torch.arange(n_nodes, device=S.device)[:, None] * per + torch.arange(4, device=S.device)[None, :] * (per // 4), :K].contiguous()
        fn = lambda: engine.kmeans_level(x, K, nd, n_nodes, 4, cent.clone(), max_iter=0, dim_acc=D)
        ms = bench.timed(torch, fn, 10, flush)
        # max_iter=0 = 3 memsets + ONE E-step launch; algorithmic bytes per sample: row (4*ld) + node id (4) in, label (4) out
        byts = (4.0 * D + 8.0) * per * n_nodes
        # the same pass inside the Lloyd loop of ONE mgb_kmeans_level call (what a build runs): 12 iterations that never converge
        # (tol = 0) + the final pass = 13 E-steps and 12 centroid kernels.  The 56 MB of rows fit the 126 MB L2, so this is the
        # L2-warm figure; the single pass above (L2 flushed, memsets and the call's host work included) is the cold one
        iters = 12
        fn_loop = lambda: engine.kmeans_level(x, K, nd, n_nodes, 4, cent.clone(), max_iter=iters, dim_acc=D)
        ms_loop = bench.timed(torch, fn_loop, 5, flush) / (iters + 1)
        k7[name] = {"ms": ms, "algorithmic_bytes": byts, "achieved_gbs": byts / (ms * 1e-3) * 1e-9, "hbm_frac": byts / (ms * 1e-3) * 1e-9 / hbm,
                    "ms_per_pass_in_loop": ms_loop, "loop_gbs": byts / (ms_loop * 1e-3) * 1e-9}
    out["K7_lloyd_iteration"] = k7

    # ---- build --------------------------------------------------------------------------------------------
    l0 = engine.launch_count()
    tree, t_first = wall(lambda: sp.ClusterTree(4, a.levels, K).construct(data, seed=5))
    l0 = engine.launch_count()
    tree, t_build = wall(lambda: sp.ClusterTree(4, a.levels, K).construct(data, seed=5))      # allocator and torch kernels warm
    out["cluster_tree_build"] = {"seconds": t_build, "first_build_seconds": t_first, "samples_per_s": M / t_build,
                                 "launches": engine.launch_count() - l0, "timing": tree.build_info.get("timing"),
                                 "n_nodes": tree.build_info["n_nodes"], "n_kd_nodes": tree.build_info["n_kd_nodes"],
                                 "levels": tree.build_info["levels"]}
    # ---- search -------------------------------------------------------------------------------------------
    srch = {}
    data_t = (node, mpc, None)
    tree.find_best_example_excluding_search_candidates(obj_spatial_error_sum, data_t, 1)      # uploads the point table
    for nc in (1, 2, 5):
        (v, s), t = wall(lambda: tree.find_best_example_excluding_search_candidates(obj_spatial_error_sum, data_t, nc))
        srch["n_candidates=%d" % nc] = {"seconds": t, "value": v, "lookups": tree.last_search_lookups}
    pts = tree.device_points(obj_spatial_error_sum, data_t)
    ms = bench.timed(torch, lambda: dc.node_costs(pts), 10, flush)
    srch["K8_node_costs"] = {"ms": ms, "points": int(pts.shape[0]), "points_per_s": pts.shape[0] / (ms * 1e-3)}
    exhaustive = dc.score(S, k=1)
    srch["exhaustive_min"] = float(exhaustive["best_cost"][0])
    out["cluster_tree_search"] = srch
    # ---- euclidean PCA features + feature tree ------------------------------------------------------------
    if not a.no_features:
        Mf = min(M, 1_000_000)
        (clouds), t_cloud = wall(lambda: feat.map_samples_to_euclidean_space(dm, S[:Mf]))
        (fp, ratio), t_pca_first = wall(lambda: feat.pca_by_variance(clouds))
        (fp, ratio), t_pca = wall(lambda: feat.pca_by_variance(clouds))
        # K11 alone: the covariance contraction on the tensor cores
        import ctypes as C
        from morphablegraphs_b200 import _lib
        dcl = int(clouds.shape[1])
        mean = clouds.mean(dim=0).contiguous()
        up = torch.zeros((dcl, dcl), dtype=torch.float64, device=clouds.device)
        ms_syrk = bench.timed(torch, lambda: _lib.check(_lib.load().mgb_syrk_upper(
            C.c_void_p(clouds.data_ptr()), Mf, dcl, dcl, C.c_void_p(mean.data_ptr()), C.c_void_p(up.data_ptr()),
            C.c_void_p(torch.cuda.current_stream().cuda_stream))), 3, flush)
        out["euclidean_features"] = {"samples": Mf, "point_cloud_seconds": t_cloud, "pca_seconds": t_pca, "pca_first_call_seconds": t_pca_first,
                                     "n_components": int(fp.shape[1]), "cloud_bytes": int(clouds.numel() * 4),
                                     "K11_syrk": {"ms": ms_syrk, "d": dcl, "algorithmic_tflops": 2.0 * Mf * dcl * dcl / (ms_syrk * 1e-3) * 1e-12,
                                                  "note": "2 n d^2 flop of the full covariance / time; 3xTF32 on tcgen05, upper triangle computed"}}
        del clouds
        opts = {"n_subdivisions": 4, "clustering_method": 0, "use_feature_mean": False}
        f_host = fp.cpu().numpy()
        ftree, t_f = wall(lambda: sp.FeatureClusterTree(f_host, data[:Mf], None, opts, [], seed=3))
        out["feature_tree_build"] = {"seconds": t_f, "samples_per_s": Mf / t_f, "n_nodes": ftree.build_info["n_nodes"],
                                     "levels": len(ftree.build_info["levels"])}
        (v, s), t = wall(lambda: ftree.find_best_example_excluding_search_candidates(obj_spatial_error_sum, data_t, 5))
        (v, s), t = wall(lambda: ftree.find_best_example_excluding_search_candidates(obj_spatial_error_sum, data_t, 5))
        out["feature_tree_search"] = {"seconds": t, "value": v}
    return out


if __name__ == "__main__":
    main()
