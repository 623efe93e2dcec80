"""Cluster trees on the CPU: (1) pins the oracle's tree restatements (search, build with injected
seeds, kd-trees) against tests/golden/s_tree.npz, which holds trees BUILT AND SEARCHED BY THE
REFERENCE'S OWN MODULES (oracle/make_golden.py: run_tree_case); (2) checks the product's host-side
tree logic (flat layout, walk, level-synchronous bookkeeping, batched kd build) against the oracle
with the K7 kernel call replaced by the oracle's Lloyd (no GPU here; the kernel itself is checked
in tests/test_gpu_parity.py)."""
import json

import numpy as np
import pytest

from morphablegraphs_b200 import synthetic as syn
from morphablegraphs_b200 import space_partitioning as sp
from morphablegraphs_b200 import Skeleton, MotionStateGraphNode, MotionPrimitiveConstraints, CONSTRAINT_FACTORY
from oracle import np_oracle as O
from tests.common import load_golden

torch = pytest.importorskip("torch")


class TreeCase(object):
    def __init__(self):
        g = load_golden("s_tree")
        self.g = g
        cfg = syn.SHAPES["S"]
        self.skeleton = Skeleton.from_desc(syn.make_skeleton_desc(cfg["n_joints"]))
        self.node = MotionStateGraphNode()
        self.node._initialize_from_json(None, syn.make_model_json("S"))
        self.node.set_skeleton(self.skeleton)
        self.mpc = MotionPrimitiveConstraints()
        self.mpc.skeleton = self.skeleton
        self.mpc.is_local = True
        self.mpc.constraints = syn.build_constraints(CONSTRAINT_FACTORY, self.skeleton, json.loads(str(g["specs_json"])))
        self.data = g["data"]
        self.K = self.node.get_n_spatial_components()
        self.ma = O.ModelArrays(self.node)
        self.n_calls = 0

    def cost(self, point):
        self.n_calls += 1
        return float(O.candidate_costs(self.ma, self.skeleton, self.mpc, np.asarray(point, dtype=np.float64)[None, :])[0])

    def obj(self, s, data):      # reference signature obj(s, data) -> float; no .batch: the lazy per-point path
        return self.cost(s)


@pytest.fixture(scope="module")
def tc():
    return TreeCase()


# ---------------------------------------------------------------------------------------------
# (1) oracle pinned against the reference's own trees
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("key", ["cluster_tree", "cluster_tree_seeded6"])
def test_oracle_cluster_tree_search_matches_reference(tc, key):
    root = json.loads(str(tc.g[key + "_json"]))
    for nc, ref in zip(tc.g["n_candidates"], tc.g[key + "_results"]):
        v, s = O.cluster_tree_search_dict(root, tc.cost, int(nc))
        assert abs(v - ref[0]) <= 1e-9 * max(1.0, abs(ref[0]))
        assert np.array_equal(np.asarray(s), ref[1:])


@pytest.mark.parametrize("key", ["feature_tree", "feature_tree_seeded"])
def test_oracle_feature_tree_search_matches_reference(tc, key):
    root = json.loads(str(tc.g[key + "_json"]))
    for nc, ref in zip(tc.g["n_candidates"], tc.g[key + "_results"]):
        v, s = O.feature_tree_search_dict(root, tc.data, tc.cost, int(nc))
        assert abs(v - ref[0]) <= 1e-9 * max(1.0, abs(ref[0]))
        assert np.array_equal(np.asarray(s), ref[1:])


def _same_cluster_tree(a, b, tol=1e-9):
    assert a["leaf"] == b["leaf"]
    assert np.allclose(a["mean"], b["mean"], rtol=0, atol=tol)
    assert len(a["children"]) == len(b["children"])
    for ca, cb in zip(a["children"], b["children"]):
        assert ("kdtree" in ca) == ("kdtree" in cb)
        if "kdtree" in ca:
            _same_kd(ca["kdtree"], cb["kdtree"])
        else:
            _same_cluster_tree(ca, cb, tol)


def _same_kd(a, b):
    if a is None or b is None:
        assert a is None and b is None
        return
    assert a["type"] == b["type"] and a["point"] == b["point"]
    _same_kd(a["left"], b["left"])
    _same_kd(a["right"], b["right"])


def _same_feature_tree(a, b, tol=1e-9):
    assert np.allclose(a["mean"], b["mean"], rtol=0, atol=tol)
    assert a["indices"] == b["indices"]
    assert len(a["children"]) == len(b["children"])
    for ca, cb in zip(a["children"], b["children"]):
        _same_feature_tree(ca, cb, tol)


@pytest.mark.parametrize("max_level", [6, 2])
def test_oracle_build_matches_reference_builder_with_injected_seeds(tc, max_level):
    ref = json.loads(str(tc.g["cluster_tree_seeded%d_json" % max_level]))
    mine = O.build_cluster_tree_dict(tc.data, 4, max_level, tc.K)
    _same_cluster_tree(mine, ref)


def test_oracle_feature_build_matches_reference_builder_with_injected_seeds(tc):
    ref = json.loads(str(tc.g["feature_tree_seeded_json"]))
    mine = O.build_feature_tree_dict(tc.data[:, :tc.K], tc.data, 4)
    _same_feature_tree(mine, ref)


# ---------------------------------------------------------------------------------------------
# (2) product host logic
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("key", ["cluster_tree", "cluster_tree_seeded6"])
def test_flat_cluster_tree_walk_matches_reference(tc, key):
    tree = sp.ClusterTree.from_dict(json.loads(str(tc.g[key + "_json"])), 4, 6, tc.K)
    for nc, ref in zip(tc.g["n_candidates"], tc.g[key + "_results"]):
        tc.n_calls = 0
        v, s = tree.find_best_example_excluding_search_candidates(tc.obj, (tc.node, tc.mpc, None), int(nc))
        assert abs(v - ref[0]) <= 1e-9 * max(1.0, abs(ref[0]))
        assert np.array_equal(np.asarray(s), ref[1:])
        assert 0 < tc.n_calls <= tree.last_search_lookups        # lazy path: only visited points are evaluated


@pytest.mark.parametrize("key", ["feature_tree", "feature_tree_seeded"])
def test_flat_feature_tree_walk_matches_reference(tc, key):
    tree = sp.FeatureClusterTree.from_dict(json.loads(str(tc.g[key + "_json"])), tc.data)
    for nc, ref in zip(tc.g["n_candidates"], tc.g[key + "_results"]):
        v, s = tree.find_best_example_excluding_search_candidates(tc.obj, (tc.node, tc.mpc, None), int(nc))
        assert abs(v - ref[0]) <= 1e-9 * max(1.0, abs(ref[0]))
        assert np.array_equal(np.asarray(s), ref[1:])


def test_walk_with_batch_objective_scores_the_table_once(tc):
    class BatchObj(object):
        calls = 0

        def batch(self, S, data):
            BatchObj.calls += 1
            return O.candidate_costs(tc.ma, tc.skeleton, tc.mpc, S)

    tree = sp.ClusterTree.from_dict(json.loads(str(tc.g["cluster_tree_json"])), 4, 6, tc.K)
    ref = tc.g["cluster_tree_results"][2]
    v, s = tree.find_best_example_excluding_search_candidates(BatchObj(), (tc.node, tc.mpc, None), 5)
    assert BatchObj.calls == 1
    assert abs(v - ref[0]) <= 1e-9 * max(1.0, abs(ref[0])) and np.array_equal(np.asarray(s), ref[1:])


def _flat_to_dict(t, node=0):
    idx, is_kd = t.children(node)
    d = {"mean": t.points[t.node_row[node]].tolist(), "leaf": bool(t.node_leaf[node]), "children": []}

    def kd(i):
        if i < 0:
            return None
        return {"point": t.points[t.kd_row[i]].tolist(), "type": "inner" if t.kd_inner[i] else "leaf",
                "left": kd(t.kd_left[i]), "right": kd(t.kd_right[i])}
    for c, k in zip(idx, is_kd):
        d["children"].append({"kdtree": kd(int(c))} if k else _flat_to_dict(t, int(c)))
    return d


def test_kd_forest_matches_oracle_kd_trees():
    rng = np.random.RandomState(5)
    data = rng.randn(300, 5)
    data[10:20, 0] = data[10, 0]          # ties on an axis: stability matters
    lens = [1, 2, 3, 7, 40, 1, 100, 16]
    rows = rng.permutation(300)[:sum(lens)]
    rows = np.concatenate([np.sort(r) for r in np.split(rows, np.cumsum(lens)[:-1])])
    X = torch.as_tensor(data)
    pt, left, right, inner, roots = sp._build_kd_forest(X, torch.as_tensor(rows), torch.as_tensor(lens), 3)
    pt, left, right, inner, roots = [a.numpy() for a in (pt, left, right, inner, roots)]

    def to_dict(i):
        if i < 0:
            return None
        return {"point": data[pt[i]].tolist(), "type": "inner" if inner[i] else "leaf", "left": to_dict(left[i]), "right": to_dict(right[i])}
    start = 0
    for s, n in enumerate(lens):
        _same_kd(to_dict(roots[s]), O._kd_dict(data[rows[start:start + n]].tolist(), 3))
        start += n
    assert len(pt) == sum(lens)           # every sample is exactly one kd node


def _oracle_lloyd_level(x, dim, node, n_nodes, n_sub, centroids, tol, dim_acc, max_iter=300):
    """CPU stand-in for the K7 call with the same contract (tests only)."""
    xn = x.double().numpy()
    nd = node.numpy()
    cent = centroids.double().numpy()
    label = np.zeros(len(xn), dtype=np.int32)
    sums = np.zeros((n_nodes, n_sub, dim_acc))
    counts = np.zeros((n_nodes, n_sub), dtype=np.int32)
    inertia = np.zeros(n_nodes)
    for i in range(n_nodes):
        m = nd == i
        X = xn[m][:, :dim]
        if max_iter == 0:
            d = ((X[:, None, :] - cent[i][None, :, :]) ** 2).sum(axis=2)
            lab = d.argmin(axis=1)
            inertia[i] = d[np.arange(len(X)), lab].sum()
        else:
            lab, C, inertia[i] = O.kmeans_lloyd(X, cent[i], max_iter=max_iter)
            cent[i] = C
        label[m] = lab
        for k in range(n_sub):
            sums[i, k] = xn[m][lab == k][:, :dim_acc].sum(axis=0)
            counts[i, k] = (lab == k).sum()
    centroids.copy_(torch.as_tensor(cent, dtype=torch.float32))
    return {"label": torch.as_tensor(label), "sums": torch.as_tensor(sums, dtype=torch.float32), "counts": torch.as_tensor(counts),
            "inertia": torch.as_tensor(inertia, dtype=torch.float32), "iters": 1}


def _injected(depth, gids, starts, counts, x, dim, n_sub=4):
    rows = torch.stack([starts + (counts * j) // n_sub for j in range(n_sub)], dim=1)
    return x[rows][:, :, :dim].contiguous()


@pytest.mark.parametrize("max_level", [6, 2])
def test_level_synchronous_build_bookkeeping(tc, monkeypatch, max_level):
    monkeypatch.setattr(sp, "_lloyd_level", _oracle_lloyd_level)
    seeds = lambda depth, gids, starts, counts, x: _injected(depth, gids, starts, counts, x, tc.K)
    flat, info = sp.build_cluster_tree(tc.data, 4, max_level, tc.K, seeds=seeds, device="cpu")
    ref = json.loads(str(tc.g["cluster_tree_seeded%d_json" % max_level]))
    _same_cluster_tree(_flat_to_dict(flat), ref, tol=2e-5)       # node means: float32 sums on the device path
    assert info["n_kd_nodes"] == len(tc.data)


def test_level_synchronous_feature_build_bookkeeping(tc, monkeypatch):
    monkeypatch.setattr(sp, "_lloyd_level", _oracle_lloyd_level)
    seeds = lambda depth, gids, starts, counts, x: _injected(depth, gids, starts, counts, x, tc.K)
    (points, node_row, child_ptr, child_idx, node_sample), info = sp.build_feature_tree(tc.data[:, :tc.K], tc.data, 4, seeds=seeds, device="cpu")
    ref = sp.FeatureClusterTree.from_dict(json.loads(str(tc.g["feature_tree_seeded_json"])), tc.data)

    # same tree up to node numbering: compare by walking both from the root
    def walk(a_node, b_node):
        assert np.allclose(points[node_row[a_node]], ref.points[ref.node_row[b_node]], rtol=0, atol=2e-5)
        ca = child_idx[child_ptr[a_node]:child_ptr[a_node + 1]]
        cb = ref._children(b_node)
        assert len(ca) == len(cb)
        if len(ca) == 0:
            assert node_sample[a_node] == ref.node_sample[b_node]
        for x, y in zip(ca, cb):
            walk(int(x), int(y))
    walk(0, 0)


def test_oracle_euclidean_pca_features_match_reference(tc):
    """a19: point clouds and PCA(0.95) features of features.map_motions_to_euclidean_pca (component
    signs are sklearn's svd_flip convention: compared up to sign)."""
    g = tc.g
    pc = O.euclidean_point_clouds(tc.ma, tc.skeleton, tc.data[:4])
    assert np.abs(pc - g["euclid_point_clouds"]).max() < 1e-9
    n = int(g["euclid_n_samples"])
    f, ratio = O.pca_by_variance(O.euclidean_point_clouds(tc.ma, tc.skeleton, tc.data[:n]))
    assert f.shape == g["euclid_features"].shape
    assert np.abs(ratio - g["euclid_explained_variance_ratio"]).max() < 1e-12
    assert np.abs(np.abs(f) - np.abs(g["euclid_features"])).max() < 1e-8


@pytest.mark.parametrize("n,max_level", [(1, 3), (3, 3), (4, 3), (5, 1), (37, 2), (200, 4)])
def test_level_synchronous_build_edge_sizes(monkeypatch, n, max_level):
    """Tiny and ragged sample sets, duplicated rows (k-means cannot split them: one child keeps every member
    until max_level), root-is-a-leaf: the level-synchronous build against the oracle's recursion."""
    monkeypatch.setattr(sp, "_lloyd_level", _oracle_lloyd_level)
    rng = np.random.RandomState(n)
    data = rng.randn(n, 6)
    if n >= 37:
        data[5:17] = data[5]                 # a dozen identical samples
    dim = 4
    seeds = lambda depth, gids, starts, counts, x: _injected(depth, gids, starts, counts, x, dim)
    flat, info = sp.build_cluster_tree(data, 4, max_level, dim, seeds=seeds, device="cpu")
    # float32 rows on the product side: build the oracle tree from the same rounded data
    ref = O.build_cluster_tree_dict(data.astype(np.float32).astype(np.float64), 4, max_level, dim)
    got = _flat_to_dict(flat)

    def strip(d):        # kd points are rows of the float64 data on the product side: compare in float32
        if "kdtree" in d:
            def k(x):
                return None if x is None else {"point": np.float32(x["point"]).tolist(), "type": x["type"], "left": k(x["left"]), "right": k(x["right"])}
            return {"kdtree": k(d["kdtree"])}
        return {"mean": d["mean"], "leaf": d["leaf"], "children": [strip(c) for c in d["children"]]}
    _same_cluster_tree(strip(got), strip(ref), tol=2e-5)
    assert info["n_kd_nodes"] == n


def test_feature_tree_edge_sizes(monkeypatch):
    monkeypatch.setattr(sp, "_lloyd_level", _oracle_lloyd_level)
    for n in (1, 2, 4, 5, 23):
        rng = np.random.RandomState(100 + n)
        data = rng.randn(n, 5)
        feats = data[:, :3].copy()
        if n == 23:
            feats[3:9] = feats[3]            # identical feature rows: one cluster swallows them -> single-sample children
        seeds = lambda depth, gids, starts, counts, x: _injected(depth, gids, starts, counts, x, 3)
        (points, node_row, child_ptr, child_idx, node_sample), info = sp.build_feature_tree(feats, data, 4, seeds=seeds, device="cpu")
        leaves = np.where(np.diff(child_ptr) == 0)[0]
        assert np.array_equal(np.sort(node_sample[leaves]), np.arange(n))        # every sample is exactly one leaf
        ref = sp.FeatureClusterTree.from_dict(O.build_feature_tree_dict(feats.astype(np.float32).astype(np.float64),
                                                                        data.astype(np.float32).astype(np.float64), 4), data)

        def walk(a, b):
            assert np.allclose(points[node_row[a]], ref.points[ref.node_row[b]], rtol=0, atol=2e-5)
            ca, cb = child_idx[child_ptr[a]:child_ptr[a + 1]], ref._children(b)
            assert len(ca) == len(cb), (n, a, b)
            for x, y in zip(ca, cb):
                walk(int(x), int(y))
        walk(0, 0)


def test_unsupported_tree_options_are_loud():
    from morphablegraphs_b200 import engine
    with pytest.raises(engine.MgbError):
        sp.build_cluster_tree(np.zeros((10, 3)), 4, 2, 3, use_kd_tree=False, device="cpu")
    with pytest.raises(engine.MgbError):
        sp.FeatureClusterTree(np.zeros((10, 3)), np.zeros((10, 3)), None, {"n_subdivisions": 4, "use_feature_mean": True}, [], device="cpu")
    if not torch.cuda.is_available():
        with pytest.raises(engine.MgbError):          # no CPU fallback for the build itself
            sp.ClusterTree(4, 2, 3).construct(np.zeros((10, 3)))
