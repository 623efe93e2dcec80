import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["MGB_TREE_TIMING"] = "1"
import numpy as np, torch
import bench
from morphablegraphs_b200 import engine, features as feat, space_partitioning as sp
node, skeleton, mpc, prev = bench.build_case("S1M")
dm = engine.DeviceModel(node, skeleton, 0)
M = 1000000
S = dm.sample(M, seed=2024)
data = S.cpu().numpy().astype(np.float64)
clouds = feat.map_samples_to_euclidean_space(dm, S)
fp, ratio = feat.pca_by_variance(clouds)
del clouds
f_host = fp.cpu().numpy()
opts = {"n_subdivisions": 4, "clustering_method": 0, "use_feature_mean": False}
for rep in range(2):
    sp._SUB_TIMING.clear()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ft = sp.FeatureClusterTree(f_host, data, None, opts, [], seed=3)
    torch.cuda.synchronize(); t = time.perf_counter() - t0
    print("feature tree build", round(t, 4), "s", {k: round(v * 1e3, 1) for k, v in sp._SUB_TIMING.items()}, "levels", [(l["nodes"], l["lloyd_iterations"]) for l in ft.build_info["levels"]])
