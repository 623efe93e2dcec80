"""Cluster trees of the reference (space_partitioning/) behind the reference's names:
`ClusterTree` (k-means hierarchy with kd-tree leaves) and `FeatureClusterTree` (k-means
hierarchy down to single samples), each with
`find_best_example_excluding_search_candidates(obj, data, n_candidates)` as called by
`MotionStateGraphNode.search_best_sample` (reference motion_state_graph_node.py:119-142).

Two things run on the GPU:

* search (a20): the reference calls `obj(child.mean, data)` / `obj(kd_node.point, data)` one
  point at a time while it walks the tree (cluster_tree.py:117-149, cluster_tree_node.py:113-138,
  kdtree.py:138-161,233-250, feature_cluster_tree.py:129-187).  Here every point of the tree
  (node means, kd-tree points) is one row of a device-resident table that is scored in ONE launch
  per search (K8, `mgb_node_costs`); the walk itself then only looks costs up and restates the
  reference's bookkeeping exactly, including its heap-array-prefix candidate selection (quirk Q4).
  Objectives without a `.batch_device` / `.batch` method are called per point, lazily, like the
  reference does.
* build (a18): the reference recurses node by node with sklearn KMeans (cluster_tree_node_builder.py:
  83-137, feature_cluster_tree.py:61-98).  Here a whole tree level is subdivided at once by the K7
  kernels (`mgb_kmeans_level`), samples kept sorted by node between levels; the kd-trees of all
  leaves are built together, one kd level per pass (kdtree.py:44-73: sort by axis `depth % dim`,
  median `len // 2`).

The flat layout: cluster nodes and kd nodes are rows of integer arrays; `points[M, D]` holds every
evaluable point, `node_row` / `kd_row` index into it.
"""
import ctypes as C
import heapq

import numpy as np

from . import engine

KDTREE_INNER = 1
KDTREE_LEAF = 0


class FlatClusterTree(object):
    """Arrays of a ClusterTree.  Children of node i: child_idx[child_ptr[i]:child_ptr[i+1]],
    child_is_kd tells whether an entry is a kd-tree root (index into the kd arrays) or a node."""

    def __init__(self):
        self._points = None         # [M, D] float64 (host); device-built trees keep (node means, data) apart
        self._parts = None          # and only concatenate them if somebody asks for the whole table
        self.device_table = None    # float32 [M, D] torch tensor left on the GPU by the device build
        self.node_row = None        # [n_nodes] row of the node mean
        self.node_leaf = None       # [n_nodes] bool
        self.child_ptr = None       # [n_nodes + 1]
        self.child_idx = None
        self.child_is_kd = None
        self.kd_row = None          # [n_kd]
        self.kd_left = None         # [n_kd] kd index or -1
        self.kd_right = None
        self.kd_inner = None        # [n_kd] bool
        self.root = 0

    @property
    def points(self):
        if self._points is None and self._parts is not None:
            self._points = np.concatenate(self._parts, axis=0)
        return self._points

    @points.setter
    def points(self, value):
        self._points = value

    @property
    def n_points(self):
        if self._points is not None:
            return len(self._points)
        return sum(len(p) for p in self._parts)

    def point(self, row):
        """One row of the point table without materialising it."""
        if self._points is not None:
            return self._points[row]
        for part in self._parts:
            if row < len(part):
                return part[row]
            row -= len(part)
        raise IndexError(row)

    @property
    def n_nodes(self):
        return len(self.node_row)

    def children(self, i):
        a, b = self.child_ptr[i], self.child_ptr[i + 1]
        return self.child_idx[a:b], self.child_is_kd[a:b]


def _tree_dict_from_reference(node):
    """Reference ClusterTreeNode / KDTreeWrapper objects -> nested dicts (duck-typed:
    `.clusters`, `.mean`, `.leaf`, `.type == "kdtree"`, `.kdtree.root` with `.point/.left/.right/.type`)."""
    def kd(n):
        if n is None:
            return None
        return {"point": [float(v) for v in n.point], "type": n.type, "left": kd(n.left), "right": kd(n.right)}

    d = {"mean": np.asarray(node.mean, dtype=np.float64).tolist(), "leaf": bool(node.leaf), "children": []}
    for c in node.clusters:
        if getattr(c, "type", None) == "kdtree" or hasattr(c, "kdtree"):
            d["children"].append({"kdtree": kd(c.kdtree.root)})
        else:
            d["children"].append(_tree_dict_from_reference(c))
    return d


def flatten_cluster_tree(root):
    """Nested dict form ({"mean", "leaf", "children": [node | {"kdtree": kd}]}, kd = {"point", "type",
    "left", "right"}) -> FlatClusterTree.  Node order: breadth-first, children in list order."""
    points = []
    node_row, node_leaf, child_lists = [], [], []
    kd_row, kd_left, kd_right, kd_inner = [], [], [], []

    def add_kd(k):
        # iterative pre-order; returns kd index
        idx0 = len(kd_row)
        stack = [(k, -1, 0)]
        while stack:
            n, parent, side = stack.pop()
            i = len(kd_row)
            kd_row.append(len(points))
            points.append(n["point"])
            kd_left.append(-1)
            kd_right.append(-1)
            kd_inner.append(n["type"] == "inner")
            if parent >= 0:
                (kd_left if side == 0 else kd_right)[parent] = i
            if n.get("right") is not None:
                stack.append((n["right"], i, 1))
            if n.get("left") is not None:
                stack.append((n["left"], i, 0))
        return idx0

    queue = [root]
    node_row.append(0)
    points.append(root["mean"])
    node_leaf.append(bool(root["leaf"]))
    child_lists.append([])
    qi = 0
    ids = {id(root): 0}
    while qi < len(queue):
        n = queue[qi]
        me = ids[id(n)]
        qi += 1
        for c in n["children"]:
            if "kdtree" in c:
                child_lists[me].append((add_kd(c["kdtree"]), True))
            else:
                i = len(node_row)
                ids[id(c)] = i
                node_row.append(len(points))
                points.append(c["mean"])
                node_leaf.append(bool(c["leaf"]))
                child_lists.append([])
                child_lists[me].append((i, False))
                queue.append(c)
    t = FlatClusterTree()
    t.points = np.asarray(points, dtype=np.float64)
    t.node_row = np.asarray(node_row, dtype=np.int64)
    t.node_leaf = np.asarray(node_leaf, dtype=bool)
    t.child_ptr = np.zeros(len(node_row) + 1, dtype=np.int64)
    t.child_ptr[1:] = np.cumsum([len(c) for c in child_lists])
    flat = [e for c in child_lists for e in c]
    t.child_idx = np.asarray([e[0] for e in flat], dtype=np.int64)
    t.child_is_kd = np.asarray([e[1] for e in flat], dtype=bool)
    t.kd_row = np.asarray(kd_row, dtype=np.int64)
    t.kd_left = np.asarray(kd_left, dtype=np.int64)
    t.kd_right = np.asarray(kd_right, dtype=np.int64)
    t.kd_inner = np.asarray(kd_inner, dtype=bool)
    return t


# ---------------------------------------------------------------------------
# cost table (K8)
# ---------------------------------------------------------------------------
class _PointCosts(object):
    """cost(row) for the rows of a tree's point table.  With a batch objective every row is scored
    by one launch when the first cost is asked for; otherwise obj is called per row, lazily (the
    reference's own behaviour)."""

    def __init__(self, tree_owner, obj, data):
        self.owner = tree_owner
        self.obj = obj
        self.data = data
        self.table = None
        self.lazy = {}
        self.n_lookups = 0

    def _fill(self):
        pts = self.owner.points
        if hasattr(self.obj, "batch_device"):
            dev_pts = self.owner.device_points(self.obj, self.data)
            from . import sharding
            if sharding.is_distributed() and getattr(self.owner, "shard_search", False):
                # K8 sharded by point id: every rank scores its slice, the slices are all-gathered
                import torch.distributed as dist
                rank, world = dist.get_rank(), dist.get_world_size()
                torch = engine._torch()
                M = int(dev_pts.shape[0])
                per = (M + world - 1) // world
                lo, hi = min(M, rank * per), min(M, (rank + 1) * per)
                mine = torch.zeros(per, dtype=torch.float32, device=dev_pts.device)
                if hi > lo:
                    mine[:hi - lo] = self.obj.batch_device(dev_pts[lo:hi].contiguous(), self.data)
                full = torch.empty(per * world, dtype=torch.float32, device=dev_pts.device)
                dist.all_gather_into_tensor(full, mine)
                self.table = full[:M].cpu().numpy().astype(np.float64)
                return
            # one D2H of the float32 cost table into a pinned buffer the tree keeps (4.3 MB for 1.09 M points: ~0.1 ms; a pageable
            # copy + float64 widening of the whole table was 1.2 ms of a 1.3 ms query, the kernel is 0.06 ms); lookups widen
            # the few entries the walk reads
            torch = engine._torch()
            costs = self.obj.batch_device(dev_pts, self.data)
            M = int(costs.shape[0])
            host = self.owner.__dict__.get("_cost_pinned")
            if host is None or host.numel() < M:
                host = self.owner.__dict__["_cost_pinned"] = torch.empty(M, dtype=torch.float32, pin_memory=True)
            host[:M].copy_(costs, non_blocking=True)
            torch.cuda.current_stream(costs.device).synchronize()
            self.table = host.numpy()[:M]
        else:
            self.table = np.asarray(self.obj.batch(pts, self.data), dtype=np.float64)

    def __call__(self, row):
        self.n_lookups += 1
        if hasattr(self.obj, "batch_device") or hasattr(self.obj, "batch"):
            if self.table is None:
                self._fill()
            return float(self.table[row])
        if row not in self.lazy:
            self.lazy[row] = float(self.obj(self.owner.point(row), self.data))
        return self.lazy[row]


def _heap_prefix(items, n):
    """Quirk Q4: everything is pushed on a heap and a PREFIX OF THE HEAP ARRAY is taken
    (cluster_tree_node.py:133-138) - not the n smallest."""
    q = []
    for it in items:
        heapq.heappush(q, it)
    return q[:n]


def _kd_find_best_example(t, cost, kd_root):
    """kdtree.py:233-250 with k = 1 (kdtree_wrapper_node.py:53-54): descend towards the cheaper
    child, return (value, row) of the cheapest point on the path (ties: the shallower one)."""
    rows = [t.kd_row[kd_root]]
    q = []
    heapq.heappush(q, (cost(rows[0]), 0))
    node = kd_root
    depth = 0
    while node >= 0 and t.kd_inner[node]:
        depth += 1
        left, right = t.kd_left[node], t.kd_right[node]
        # kdtree.py:138-161: both evaluated, right first; strict '<' prefers right on ties
        if left >= 0 and right >= 0:
            r_d = cost(t.kd_row[right])
            l_d = cost(t.kd_row[left])
            node, least = (left, l_d) if l_d < r_d else (right, r_d)
        elif right >= 0:
            node, least = right, cost(t.kd_row[right])
        elif left >= 0:
            node, least = left, cost(t.kd_row[left])
        else:
            node, least = -1, np.inf
        if node >= 0:
            heapq.heappush(q, (least, depth))
            rows.append(t.kd_row[node])
    value, d = q[0]
    return value, rows[d]


def _leaf_find_best_example(t, cost, node):
    """cluster_tree_node.py:63-76 for a leaf: best over its kd-trees; ties on the value fall back to
    comparing the point lists, like the reference's (value, list) tuples do."""
    idx, is_kd = t.children(node)
    q = []
    if len(idx) > 0:
        for c, kd in zip(idx, is_kd):
            if kd:
                v, row = _kd_find_best_example(t, cost, c)
            else:                                   # leaf of a tree built without kd-trees: a single-sample node
                row = t.node_row[c]
                v = cost(row)
            heapq.heappush(q, (v, tuple(t.point(row).tolist()), row))
    else:
        row = t.node_row[node]
        heapq.heappush(q, (cost(row), tuple(t.point(row).tolist()), row))
    v, _, row = heapq.heappop(q)
    return v, row


def cluster_tree_search(t, cost, n_candidates=1):
    """cluster_tree.py:117-149 on a FlatClusterTree; returns (value, row of the sample in t.points)
    or (inf, row of the root mean) when nothing was found."""
    results = []
    candidates = [(np.inf, 0, t.root)]
    while len(candidates) > 0:
        new_candidates = []
        for c_idx, (value, _tie, node) in enumerate(candidates):
            idx, is_kd = t.children(node)
            # a node that kept > n_subdivisions samples at depth == max_level has kd-tree children but
            # leaf == False (cluster_tree_node_builder.py:139-154); the reference's walk dies there on
            # KDTreeWrapper.mean - here it is searched like a leaf
            if not t.node_leaf[node] and not (len(idx) > 0 and is_kd.all()):
                items = [(cost(t.node_row[c]), ci, int(c)) for ci, c in enumerate(idx)]
                good = _heap_prefix(items, n_candidates)
                for i, c in enumerate(good):
                    heapq.heappush(new_candidates, (c[0], i, c[2]))
            else:
                v, row = _leaf_find_best_example(t, cost, node)
                heapq.heappush(results, (v, c_idx, tuple(t.point(row).tolist()), row))
        candidates = new_candidates[:n_candidates]
    if len(results) > 0:
        r = heapq.heappop(results)
        return r[0], r[3]
    return np.inf, t.node_row[t.root]


class _DevicePointsMixin(object):
    """Point table resident on the GPU of the objective's device model (uploaded once per device; a tree built
    on the device keeps the float32 table it already has there - the score kernels compute in fp32 either way)."""

    def n_points(self):
        flat = getattr(self, "flat", None)
        return flat.n_points if flat is not None else len(self.points)

    def point(self, row):
        flat = getattr(self, "flat", None)
        return flat.point(row) if flat is not None else self.points[row]

    def device_points(self, obj, data):
        torch = engine._torch()
        from .objective_functions import _device_constraints
        dm = _device_constraints(data[0], data[1], data[2]).dmodel
        cache = self.__dict__.setdefault("_dev_points", {})
        key = dm.device
        if key not in cache or cache[key].shape[0] != self.n_points():
            flat = getattr(self, "flat", None)
            table = getattr(flat, "device_table", None) if flat is not None else None
            if table is not None and table.device == dm.torch_device:
                cache[key] = table
            else:
                cache[key] = torch.as_tensor(np.ascontiguousarray(self.points, dtype=np.float64), device=dm.torch_device)
        return cache[key]


class ClusterTree(_DevicePointsMixin):
    """reference space_partitioning/cluster_tree.py:43-149 (constructor arguments, `construct`,
    `find_best_example_excluding_search_candidates`)."""

    def __init__(self, n_subdivisions=4, max_level=4, dim=10, store_indices=False, use_kd_tree=True):
        self.n_subdivisions = max(n_subdivisions, 1)
        self.max_level = max(max_level, 1)
        self.dim = dim
        self.store_indices = store_indices
        self.use_kd_tree = use_kd_tree
        self.data = None
        self.flat = None
        self.build_info = None

    # -- trees built elsewhere -------------------------------------------------------------
    @classmethod
    def from_reference(cls, ref_tree):
        """A tree built (or unpickled) by the reference: ClusterTree with `.root`, `.data`."""
        t = cls(ref_tree.n_subdivisions, ref_tree.max_level, ref_tree.dim)
        t.data = getattr(ref_tree, "data", None)
        t.flat = flatten_cluster_tree(_tree_dict_from_reference(ref_tree.root))
        return t

    @classmethod
    def from_dict(cls, root, n_subdivisions=4, max_level=4, dim=10):
        t = cls(n_subdivisions, max_level, dim)
        t.flat = flatten_cluster_tree(root)
        return t

    @property
    def points(self):
        return self.flat.points

    # -- build (a18) -----------------------------------------------------------------------
    def construct(self, data, seeds=None, seed=0, device=None):
        """cluster_tree.py:98-102 -> cluster_tree_node_builder.py:105-170, level-synchronous on the GPU.
        seeds: optional callable (depth, node_ids, starts, counts, x_sorted) -> float32 [n_live, n_sub, dim]
        initial centroids (parity runs inject the k-means++ picks of the CPU run); default: k-means++
        drawn on the device from `seed`."""
        self.data = np.asarray(data)
        self.dim = min(self.data.shape[1], self.dim)
        from . import sharding
        if sharding.is_distributed():
            # one process per GPU: the Lloyd sums are all-reduced, every rank ends with the same tree.  `seeds` is
            # then a callable (depth, node_ids, global_counts) -> global member positions [n_live, n_sub]
            if not self.use_kd_tree:
                raise engine.MgbError("device build implements use_kd_tree=True")
            self.flat, self.build_info = build_cluster_tree_sharded(self.data, self.n_subdivisions, self.max_level, self.dim,
                                                                    seeds=seeds, seed=seed, device=device)
        else:
            self.flat, self.build_info = build_cluster_tree(self.data, self.n_subdivisions, self.max_level, self.dim,
                                                            self.use_kd_tree, seeds=seeds, seed=seed, device=device)
        self.__dict__.pop("_dev_points", None)
        return self

    # -- search (a20) ----------------------------------------------------------------------
    def find_best_example_excluding_search_candidates(self, obj, data, n_candidates=1):
        cost = _PointCosts(self, obj, data)
        value, row = cluster_tree_search(self.flat, cost, n_candidates)
        self.last_search_lookups = cost.n_lookups
        return value, np.asarray(self.flat.point(row)).tolist()

    def find_best_example_excluding_search(self, obj, data):
        """cluster_tree.py:106-115: greedy descent (first minimum among the children), then the leaf."""
        t = self.flat
        cost = _PointCosts(self, obj, data)
        node, level = t.root, 0
        while level < self.max_level and not t.node_leaf[node]:
            idx, is_kd = t.children(node)
            if len(idx) == 0 or is_kd.all():
                break
            best_v, best_c = np.inf, idx[0]
            for c in idx:
                v = cost(t.node_row[c])
                if v < best_v:
                    best_v, best_c = v, c
            node = int(best_c)
            level += 1
        v, row = _leaf_find_best_example(t, cost, node)
        return v, np.asarray(t.point(row)).tolist()


# ---------------------------------------------------------------------------
# FeatureClusterTree (feature_cluster_tree.py)
# ---------------------------------------------------------------------------
class FeatureClusterTree(_DevicePointsMixin):
    """reference space_partitioning/feature_cluster_tree.py:55-187: k-means hierarchy over a feature
    matrix down to single samples; a node's evaluable point is the mean of its samples' LATENT
    vectors (options["use_feature_mean"] == False, :70-73).  Flat form: node_row (mean rows of
    `points`), child_ptr/child_idx, node_sample (index into `data` of a leaf's first sample)."""

    def __init__(self, features=None, data=None, indices=None, options=None, feature_args=None, seeds=None, seed=0,
                 device=None):
        self.data = None
        self.points = None
        self.node_row = self.child_ptr = self.child_idx = self.node_sample = None
        self.args = feature_args
        self._options = options
        self.build_info = None
        if features is not None:
            self._construct(features, data, indices, options, seeds, seed, device)

    def _construct(self, features, data, indices, options, seeds, seed, device):
        assert indices is None, "the device build starts from the whole sample set"
        if options.get("use_feature_mean", False):
            raise engine.UnsupportedConstraintError("use_feature_mean: node means in feature space cannot be scored by the "
                                                    "latent-space objective (reference feature_cluster_tree.py:70-71)")
        self.data = np.asarray(data)
        flat, info = build_feature_tree(np.asarray(features), self.data, int(options["n_subdivisions"]), seeds=seeds, seed=seed,
                                        device=device)
        self.points, self.node_row, self.child_ptr, self.child_idx, self.node_sample = flat
        self.build_info = info

    @classmethod
    def from_dict(cls, root, data):
        """{"mean", "indices", "children"} nested dicts (a serialised reference tree) -> flat form."""
        t = cls()
        t.data = np.asarray(data)
        means, child_lists, samples = [], [], []
        queue = [root]
        ids = {id(root): 0}
        means.append(root["mean"])
        child_lists.append([])
        samples.append(-1 if not root.get("indices") else int(root["indices"][0]))
        qi = 0
        while qi < len(queue):
            n = queue[qi]
            me = ids[id(n)]
            qi += 1
            for c in n["children"]:
                i = len(means)
                ids[id(c)] = i
                means.append(c["mean"])
                child_lists.append([])
                samples.append(-1 if not c.get("indices") else int(c["indices"][0]))
                child_lists[me].append(i)
                queue.append(c)
        t.points = np.asarray(means, dtype=np.float64)
        t.node_row = np.arange(len(means), dtype=np.int64)
        t.child_ptr = np.zeros(len(means) + 1, dtype=np.int64)
        t.child_ptr[1:] = np.cumsum([len(c) for c in child_lists])
        t.child_idx = np.asarray([e for c in child_lists for e in c], dtype=np.int64)
        t.node_sample = np.asarray(samples, dtype=np.int64)
        return t

    @classmethod
    def from_reference(cls, ref_tree):
        def conv(n):
            return {"mean": np.asarray(n._mean, dtype=np.float64).tolist(),
                    "indices": None if n._indices is None else [int(i) for i in n._indices],
                    "children": [conv(c) for c in n._children]}
        return cls.from_dict(conv(ref_tree), ref_tree.data)

    def _children(self, i):
        return self.child_idx[self.child_ptr[i]:self.child_ptr[i + 1]]

    def find_best_example_excluding_search_candidates(self, obj, args, n_candidates=1):
        """feature_cluster_tree.py:154-187.  The reference appends `self.args` to the objective's data
        tuple (:129); objectives here read data[:3] only."""
        cost = _PointCosts(self, obj, tuple(args))
        results = []
        candidates = [(np.inf, 0)]
        while len(candidates) > 0:
            new_candidates = []
            for value, node in candidates:
                ch = self._children(node)
                if len(ch) > 0:
                    good = _heap_prefix([(cost(self.node_row[c]), int(c)) for c in ch], n_candidates)   # :150-152
                    for c in good:
                        heapq.heappush(new_candidates, c)
                else:
                    heapq.heappush(results, (value, node))
            candidates = new_candidates[:n_candidates]
        self.last_search_lookups = cost.n_lookups
        if len(results) > 0:
            value, node = heapq.heappop(results)
            return value, self.data[self.node_sample[node]]
        return np.inf, self.data[0]

    def find_best_example_excluding_search(self, obj, args):
        """:100-109: greedy descent, returns the leaf's sample."""
        cost = _PointCosts(self, obj, tuple(args))
        node = 0
        while len(self._children(node)) > 0:
            best_v, best_c = np.inf, self._children(node)[0]
            for c in self._children(node):
                v = cost(self.node_row[c])
                if v < best_v:
                    best_v, best_c = v, c
            node = int(best_c)
        return self.data[self.node_sample[node]]


# ---------------------------------------------------------------------------
# device build (a18): level-synchronous k-means + batched kd-trees
# ---------------------------------------------------------------------------
KMEANS_TOL = 1e-4        # sklearn KMeans default tol, relative to the mean per-column variance
KMEANS_MAX_ITER = 300    # sklearn KMeans default


def _lloyd_level(x, dim, node, n_nodes, n_sub, centroids, tol, dim_acc, max_iter=KMEANS_MAX_ITER):
    """The only arithmetic of the build that is not index bookkeeping: K7 (no CPU fallback)."""
    return engine.kmeans_level(x, dim, node, n_nodes, n_sub, centroids, tol=tol, max_iter=max_iter, dim_acc=dim_acc)


def _segment_stats(x, dim, node, n_nodes, dim_acc):
    """Per node: mean over the first dim_acc columns and sklearn's absolute tolerance
    tol * mean_j var_j(x[:, :dim]) (sklearn.cluster._kmeans._tolerance), from two E-step-only passes
    with a single 'cluster' per node: centroid 0 -> sums -> mean; centroid = mean -> inertia."""
    torch = engine._torch()
    zero = torch.zeros((n_nodes, 1, dim), dtype=torch.float32, device=x.device)
    r = _lloyd_level(x, dim, node, n_nodes, 1, zero, None, dim_acc, max_iter=0)
    cnt = r["counts"][:, 0].clamp(min=1).to(torch.float32)
    mean = r["sums"][:, 0, :] / cnt[:, None]
    cen = mean[:, None, :dim].contiguous()
    r2 = _lloyd_level(x, dim, node, n_nodes, 1, cen, None, dim_acc, max_iter=0)
    tol = (KMEANS_TOL * r2["inertia"] / (cnt * dim)).contiguous()
    return mean, tol, r["counts"][:, 0]


def _kmeanspp_seeds(x, dim, starts, counts, n_sub, gen):
    """k-means++ picks for every node of the level at once (one trial per centre; sklearn tries
    2 + log k candidates per centre, which changes the picks, not the algorithm - label parity with
    a CPU run is only defined for injected seeds).  Index bookkeeping in
    torch; x is sorted by node, node i owns rows [starts[i], starts[i] + counts[i])."""
    torch = engine._torch()
    n_nodes = starts.shape[0]
    dev = x.device
    node = torch.repeat_interleave(torch.arange(n_nodes, device=dev), counts)
    cent = torch.empty((n_nodes, n_sub, dim), dtype=torch.float32, device=dev)
    u = torch.rand((n_nodes, n_sub), generator=gen, device=dev, dtype=torch.float64)
    if dev.type == "cuda":
        # per centre: ONE pick kernel for all nodes (mgb_kmeanspp_pick), ONE closest-distance pass (mgb_min_sqdist) and the
        # library prefix sum in between - instead of ~14 element-wise / index kernels
        lib = engine._lib.load()
        p = lambda t: C.c_void_p(t.data_ptr())
        node32 = node.to(torch.int32)
        xc = x if x.is_contiguous() else x.contiguous()
        st, ct = starts.contiguous(), counts.contiguous()
        d2 = torch.empty(xc.shape[0], dtype=torch.float64, device=dev)
        cj = torch.empty((n_nodes, dim), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            for j in range(n_sub):
                cs = torch.cumsum(d2, dim=0) if j > 0 else None
                engine._lib.check(lib.mgb_kmeanspp_pick(p(xc), int(dim), int(xc.shape[1]), p(st), p(ct), int(n_nodes), C.c_void_p(u.data_ptr() + 8 * j),
                                                        int(n_sub), p(cs) if cs is not None else C.c_void_p(0), p(cj), int(dim), stream))
                cent[:, j] = cj
                if j + 1 < n_sub:          # the last centre's distances are never read
                    engine._lib.check(lib.mgb_min_sqdist(p(xc), int(xc.shape[0]), int(dim), int(xc.shape[1]), p(node32), p(cj), p(d2),
                                                         1 if j == 0 else 0, stream))
        return cent.contiguous()
    # CPU tensors: only the world-size-2 gloo tests of the sharded host logic get here (they stand the oracle in for K7; a build on
    # a CPU device without that stand-in fails in mgb_kmeans_level, which needs the CUDA library) - the same picks in torch index ops
    first = starts + (u[:, 0] * counts).long().minimum(counts - 1)
    cent[:, 0] = x[first, :dim]
    d2 = ((x[:, :dim] - cent[node, 0]) ** 2).sum(dim=1).double()
    for j in range(1, n_sub):
        cs = torch.cumsum(d2, dim=0)
        base = torch.where(starts > 0, cs[(starts - 1).clamp(min=0)], torch.zeros_like(cs[:1]).expand(n_nodes))
        total = cs[starts + counts - 1] - base
        target = base + u[:, j] * total
        pick = torch.searchsorted(cs, target, right=True)
        pick = torch.maximum(torch.minimum(pick, starts + counts - 1), starts)
        cent[:, j] = x[pick, :dim]
        d2 = torch.minimum(d2, ((x[:, :dim] - cent[node, j]) ** 2).sum(dim=1).double())
    return cent.contiguous()


def _build_kd_forest(X, rows, seg_len, dim):
    """kd-trees of many point sets at once (kdtree.py:44-73 per set): at kd depth k every set is
    sorted by column k % dim (stable, on top of the order the previous level left, like
    list.sort on the parent's sub-list), the element at len // 2 becomes the node, the halves become
    the next level's sets.  rows: sample ids, concatenated set after set; seg_len [S].
    Returns (kd_point [n] sample ids, kd_left, kd_right, kd_inner, roots [S]) - torch tensors on X's device."""
    torch = engine._torch()
    dev = X.device
    S = int(seg_len.shape[0])
    cur_rows = rows
    cur_seg = torch.repeat_interleave(torch.arange(S, device=dev), seg_len)
    pts, lefts, rights, inners = [], [], [], []
    roots = torch.arange(S, device=dev)
    next_base = S
    k = 0
    n_seg = S
    while cur_rows.numel() > 0:
        vals = X[cur_rows, k % dim]
        p1 = torch.argsort(vals, stable=True)
        p2 = torch.argsort(cur_seg[p1], stable=True)
        perm = p1[p2]
        cur_rows, cur_seg = cur_rows[perm], cur_seg[perm]
        lens = torch.bincount(cur_seg, minlength=n_seg)
        starts = torch.cumsum(lens, 0) - lens
        pos = torch.arange(cur_rows.numel(), device=dev) - starts[cur_seg]
        med = lens // 2
        inner = lens > 1
        left_len = torch.where(inner, med, torch.zeros_like(med))
        right_len = torch.where(inner, lens - med - 1, torch.zeros_like(med))
        exists = torch.stack([left_len > 0, right_len > 0], dim=1).reshape(-1)
        new_id = torch.cumsum(exists.long(), 0) - 1
        pts.append(cur_rows[starts + med])
        lefts.append(torch.where(exists[0::2], next_base + new_id[0::2], torch.full_like(med, -1)))
        rights.append(torch.where(exists[1::2], next_base + new_id[1::2], torch.full_like(med, -1)))
        inners.append(inner)
        seg_med = med[cur_seg]
        keep = pos != seg_med
        side = (pos > seg_med).long()
        cur_seg = new_id[2 * cur_seg + side][keep]
        cur_rows = cur_rows[keep]
        n_seg = int(exists.sum().item())
        next_base += n_seg
        k += 1
    if not pts:
        e = torch.empty(0, dtype=torch.long, device=dev)
        return e, e, e, e.bool(), roots
    return torch.cat(pts), torch.cat(lefts), torch.cat(rights), torch.cat(inners), roots


_STAGE = {}


def _to_host(*tensors):
    """Device tensors -> numpy arrays through ONE kept pinned staging buffer: asynchronous copies, one synchronisation, then a
    plain memcpy out of the staging area (`tensor.cpu()` goes through pageable memory: 11 ms for the 32 MB of kd arrays of a
    1M-sample tree, 4 ms this way)."""
    torch = engine._torch()
    tensors = [t.contiguous() for t in tensors]
    if not tensors or tensors[0].device.type != "cuda":
        return [t.cpu().numpy() for t in tensors]
    nbytes = [t.numel() * t.element_size() for t in tensors]
    offs = np.concatenate([[0], np.cumsum([(b + 63) & ~63 for b in nbytes])]).astype(np.int64)
    total = int(offs[-1])
    key = tensors[0].device.index
    stage = _STAGE.get(key)
    if stage is None or stage.numel() < total:
        torch.cuda.current_stream(tensors[0].device).synchronize()
        stage = _STAGE[key] = torch.empty(max(total, 1 << 20), dtype=torch.uint8, pin_memory=True)
    for t, o, b in zip(tensors, offs[:-1], nbytes):
        if b:
            stage[int(o):int(o) + b].view(t.dtype).view(t.shape).copy_(t, non_blocking=True)
    torch.cuda.current_stream(tensors[0].device).synchronize()
    out = []
    for t, o, b in zip(tensors, offs[:-1], nbytes):
        out.append(np.array(stage[int(o):int(o) + b].view(t.dtype).view(t.shape).numpy(), copy=True))
    return out


def _upload_f32(data, dev):
    """numpy [N, D] (any float type) -> float32 tensor on dev: converted straight into the kept pinned staging buffer and copied
    from there (a pageable source costs a second host copy inside the driver)."""
    torch = engine._torch()
    data = np.asarray(data)
    if dev.type != "cuda" or data.size == 0:
        return torch.as_tensor(np.ascontiguousarray(data, dtype=np.float32), device=dev)
    nbytes = data.size * 4
    stage = _STAGE.get(dev.index)
    if stage is None or stage.numel() < nbytes:
        torch.cuda.current_stream(dev).synchronize()   # the buffer being replaced may still feed a copy
        stage = _STAGE[dev.index] = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, pin_memory=True)
    view = stage[:nbytes].view(torch.float32).view(data.shape)
    ev = _STAGE.get(("uploaded", dev.index))
    if ev is not None:
        ev.synchronize()                      # an earlier upload may still be reading the staging buffer
    np.copyto(view.numpy(), data, casting="same_kind")
    X = torch.empty(data.shape, dtype=torch.float32, device=dev)
    X.copy_(view, non_blocking=True)
    ev = _STAGE[("uploaded", dev.index)] = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(dev))
    return X


def _device_of(device):
    torch = engine._torch()
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise engine.MgbError("no CUDA device: the cluster-tree build has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


_SUB_TIMING = {}


def _subdivide_level(Xl, dim, n_sub, starts, counts, depth, gids, seeds, gen, dim_acc):
    """One tree level: k-means of every live node, samples re-sorted by (node, label).
    Returns (perm, child_node [n_children] index of the parent within the level, child_counts,
    child_means [n_children, dim_acc], parent_mean [n_live, dim_acc], iterations)."""
    torch = engine._torch()
    import os
    import time
    timing = _SUB_TIMING if os.environ.get("MGB_TREE_TIMING") else None

    def tick(name, t0):
        if timing is not None:
            torch.cuda.synchronize() if Xl.device.type == "cuda" else None
            timing[name] = timing.get(name, 0.0) + time.perf_counter() - t0
        return time.perf_counter()

    t0 = time.perf_counter()
    n_live = int(starts.shape[0])
    node64 = torch.repeat_interleave(torch.arange(n_live, device=Xl.device), counts)
    node = node64.to(torch.int32).contiguous()
    mean, tol, _ = _segment_stats(Xl, dim, node, n_live, dim_acc)
    t0 = tick("  node ids + stats", t0)
    if seeds is not None:
        cent = seeds(depth, gids, starts, counts, Xl)
        cent = torch.as_tensor(cent, dtype=torch.float32, device=Xl.device).contiguous()
    else:
        cent = _kmeanspp_seeds(Xl, dim, starts, counts, n_sub, gen)
    t0 = tick("  seeding", t0)
    r = _lloyd_level(Xl, dim, node, n_live, n_sub, cent, tol, dim_acc)
    t0 = tick("  Lloyd", t0)
    key = node64 * n_sub + r["label"].long()
    perm = torch.argsort(key, stable=True)
    flat_counts = r["counts"].reshape(-1).long()
    ck = torch.nonzero(flat_counts > 0).reshape(-1)          # ascending = (node, label) order
    cc = flat_counts[ck]
    means = r["sums"].reshape(-1, dim_acc)[ck] / cc[:, None].to(torch.float32)
    tick("  sort by (node, label) + child means", t0)
    return perm, ck // n_sub, cc, means, mean, r["iters"]


def build_cluster_tree(data, n_sub, max_level, dim, use_kd_tree=True, seeds=None, seed=0, device=None):
    """cluster_tree_node_builder.py:105-170 level by level.  Node rule per depth d and member count c:
    c > n_sub -> k-means into <= n_sub children (nodes while d < max_level, kd-trees at d == max_level);
    otherwise a leaf holding one kd-tree of its members.  Returns (FlatClusterTree, info)."""
    torch = engine._torch()
    if not use_kd_tree:
        raise engine.MgbError("device build implements use_kd_tree=True (the reference builder's setting, "
                              "construction/cluster_tree_builder.py:235-247)")
    dev = _device_of(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(seed))
    data = np.asarray(data)
    N, D = data.shape
    X = _upload_f32(data, dev)
    node_means = [None]                       # list of [*, D] tensors, root first
    node_leaf = [torch.zeros(1, dtype=torch.bool, device=dev)]
    edges_parent, edges_child = [], []        # node -> node edges, in (parent, label) order per level
    kd_owner, kd_len, kd_rows = [], [], []    # kd sets in creation order
    n_nodes = 1
    info = {"levels": [], "n_samples": int(N)}
    Xl = X
    order_l = torch.arange(N, device=dev)
    gids = torch.zeros(1, dtype=torch.long, device=dev)
    starts = torch.zeros(1, dtype=torch.long, device=dev)
    counts = torch.full((1,), N, dtype=torch.long, device=dev)
    if not (N > n_sub and n_sub > 1):
        # the root itself is a leaf (cluster_tree_node_builder.py:131-133)
        node_means[0] = X.double().mean(dim=0, keepdim=True).float()
        node_leaf[0][0] = True
        kd_owner.append(gids)
        kd_len.append(counts)
        kd_rows.append(order_l)
        gids = gids[:0]
    depth = 0
    import os
    import time
    timing = {} if os.environ.get("MGB_TREE_TIMING") else None

    def tick(name, t0):
        if timing is not None:
            torch.cuda.synchronize() if dev.type == "cuda" else None
            timing[name] = timing.get(name, 0.0) + time.perf_counter() - t0
        return time.perf_counter()

    t0 = time.perf_counter()
    while gids.numel() > 0:
        perm, child_node, cc, cmeans, pmean, iters = _subdivide_level(Xl, dim, n_sub, starts, counts, depth, gids, seeds, gen, D)
        t0 = tick("subdivide (stats, seeds, Lloyd, sort key)", t0)
        if depth == 0:
            node_means[0] = pmean
        Xl, order_l = Xl[perm], order_l[perm]
        cstarts = torch.cumsum(cc, 0) - cc
        parents = gids[child_node]
        info["levels"].append({"depth": depth, "nodes": int(gids.numel()), "samples": int(Xl.shape[0]), "lloyd_iterations": int(iters)})
        if depth < max_level:
            n_children = int(cc.numel())
            cg = n_nodes + torch.arange(n_children, device=dev)
            n_nodes += n_children
            live = cc > n_sub
            node_means.append(cmeans)
            node_leaf.append(~live)
            edges_parent.append(parents)
            edges_child.append(cg)
            # leaves: one kd-tree each over their members
            kd_owner.append(cg[~live])
            kd_len.append(cc[~live])
            sample_live = torch.repeat_interleave(live, cc)
            kd_rows.append(order_l[~sample_live])
            Xl, order_l = Xl[sample_live].contiguous(), order_l[sample_live]
            gids, counts = cg[live], cc[live]
            starts = torch.cumsum(counts, 0) - counts
        else:
            # depth == max_level: every cluster becomes a kd-tree of its parent (which keeps leaf == False)
            kd_owner.append(parents)
            kd_len.append(cc)
            kd_rows.append(order_l)
            gids = gids[:0]
        depth += 1
        t0 = tick("level bookkeeping", t0)
    owner = torch.cat(kd_owner) if kd_owner else torch.empty(0, dtype=torch.long, device=dev)
    seg_len = torch.cat(kd_len) if kd_len else owner
    rows = torch.cat(kd_rows) if kd_rows else owner
    kd_pt, kd_l, kd_r, kd_in, roots = _build_kd_forest(X, rows, seg_len, dim)
    t0 = tick("kd forest", t0)
    if timing is not None:
        timing.update(_SUB_TIMING)
        _SUB_TIMING.clear()
        info["timing"] = timing
    # ---- flat arrays on the host ------------------------------------------------------------
    t = FlatClusterTree()
    means_dev = torch.cat(node_means)
    t._parts = [means_dev.double().cpu().numpy(), np.asarray(data, dtype=np.float64)]
    if dev.type == "cuda":
        t.device_table = torch.cat([means_dev.float(), X], dim=0).contiguous()
    t0 = tick("assembly: points", t0)
    t.node_row = np.arange(n_nodes, dtype=np.int64)
    # child lists: node -> node edges first, then the kd roots, stable by parent (sorted on the device, copied once)
    e = torch.empty(0, dtype=torch.long, device=dev)
    ep = torch.cat(edges_parent) if edges_parent else e
    ec = torch.cat(edges_child) if edges_child else e
    par = torch.cat([ep, owner])
    chi = torch.cat([ec, roots])
    iskd = torch.cat([torch.zeros(ep.numel(), dtype=torch.bool, device=dev), torch.ones(owner.numel(), dtype=torch.bool, device=dev)])
    o = torch.argsort(par, stable=True)
    cptr = torch.zeros(n_nodes + 1, dtype=torch.long, device=dev)
    cptr[1:] = torch.cumsum(torch.bincount(par, minlength=n_nodes), 0)
    t.node_leaf, t.child_idx, t.child_is_kd, t.child_ptr = _to_host(torch.cat(node_leaf), chi[o], iskd[o], cptr)
    t0 = tick("assembly: child lists", t0)
    kd_pt_h, t.kd_left, t.kd_right, t.kd_inner = _to_host(kd_pt, kd_l, kd_r, kd_in)
    t.kd_row = n_nodes + kd_pt_h
    t0 = tick("assembly: kd arrays to host", t0)
    info["n_nodes"] = n_nodes
    info["n_kd_nodes"] = int(len(t.kd_row))
    if timing is not None:
        tick("host assembly", t0)
    return t, info


def build_feature_tree(features, data, n_sub, seeds=None, seed=0, device=None):
    """feature_cluster_tree.py:61-98 + clustering.py:105-129 level by level.  Node rule for member count c:
    c > n_sub -> k-means on the feature rows into <= n_sub children (if one cluster swallows every
    member: one single-sample child per member, :90-93); 1 < c <= n_sub -> one single-sample child per
    member; c == 1 -> leaf.  A node's point is the mean of its members' rows of `data` (:72-73).
    Returns ((points, node_row, child_ptr, child_idx, node_sample), info)."""
    torch = engine._torch()
    dev = _device_of(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(seed))
    features = np.asarray(features)
    data = np.asarray(data)
    N, Fd = features.shape
    D = data.shape[1]
    # [features | data] rows in float32 on the device, each part converted into and copied from the pinned staging buffer
    X = torch.empty((N, Fd + D), dtype=torch.float32, device=dev)
    X[:, :Fd] = _upload_f32(features, dev)
    X[:, Fd:] = _upload_f32(data, dev)
    means, exact, first_sample = [None], [torch.full((1,), -1, dtype=torch.long, device=dev)], [torch.zeros(1, dtype=torch.long, device=dev)]
    edges_parent, edges_child = [], []
    n_nodes = 1
    info = {"levels": [], "n_samples": int(N)}
    Xl, order_l = X, torch.arange(N, device=dev)
    gids = torch.zeros(1, dtype=torch.long, device=dev)
    starts = torch.zeros(1, dtype=torch.long, device=dev)
    counts = torch.full((1,), N, dtype=torch.long, device=dev)

    def add_singletons(parent_gid_per_sample, sample_ids):
        nonlocal n_nodes
        m = int(sample_ids.numel())
        if m == 0:
            return
        g = n_nodes + torch.arange(m, device=dev)
        n_nodes += m
        means.append(torch.zeros((m, D), dtype=torch.float32, device=dev))
        exact.append(sample_ids)
        first_sample.append(sample_ids)
        edges_parent.append(parent_gid_per_sample)
        edges_child.append(g)

    if N <= n_sub:
        means[0] = X[:, Fd:].double().mean(dim=0, keepdim=True).float()
        if N > 1:
            add_singletons(torch.zeros(N, dtype=torch.long, device=dev), order_l)
        else:
            exact[0][0] = 0
        gids = gids[:0]
    depth = 0
    while gids.numel() > 0:
        perm, child_node, cc, cmeans, pmean, iters = _subdivide_level(Xl, Fd, n_sub, starts, counts, depth, gids, seeds, gen, Fd + D)
        if depth == 0:
            means[0] = pmean[:, Fd:]
        Xl, order_l = Xl[perm], order_l[perm]
        info["levels"].append({"depth": depth, "nodes": int(gids.numel()), "samples": int(Xl.shape[0]), "lloyd_iterations": int(iters)})
        cstarts = torch.cumsum(cc, 0) - cc
        degenerate = cc == counts[child_node]
        normal = ~degenerate
        n_children = int(normal.sum().item())
        cg = torch.full_like(cc, -1)
        cg[normal] = n_nodes + torch.arange(n_children, device=dev)
        n_nodes += n_children
        single = normal & (cc == 1)
        means.append(cmeans[normal][:, Fd:])
        exact.append(torch.where(single, order_l[cstarts], torch.full_like(cc, -1))[normal])
        first_sample.append(order_l[cstarts][normal])
        edges_parent.append(gids[child_node][normal])
        edges_child.append(cg[normal])
        small = normal & (cc > 1) & (cc <= n_sub)
        live = normal & (cc > n_sub)
        child_of_sample = torch.repeat_interleave(torch.arange(cc.numel(), device=dev), cc)
        expand = (degenerate | small)[child_of_sample]
        parent_of_sample = torch.where(degenerate, gids[child_node], cg)[child_of_sample]
        add_singletons(parent_of_sample[expand], order_l[expand])
        sample_live = live[child_of_sample]
        Xl, order_l = Xl[sample_live].contiguous(), order_l[sample_live]
        gids, counts = cg[live], cc[live]
        starts = torch.cumsum(counts, 0) - counts
        depth += 1
    # child lists sorted by parent and prefix-summed on the device; everything comes back through the pinned staging buffer
    e = torch.empty(0, dtype=torch.long, device=dev)
    par = torch.cat(edges_parent) if edges_parent else e
    chi = torch.cat(edges_child) if edges_child else e
    o = torch.argsort(par, stable=True)
    cptr = torch.zeros(n_nodes + 1, dtype=torch.long, device=dev)
    cptr[1:] = torch.cumsum(torch.bincount(par, minlength=n_nodes), 0)
    pts, ex, child_ptr, child_idx, first = _to_host(torch.cat(means).double(), torch.cat(exact), cptr, chi[o], torch.cat(first_sample))
    sel = ex >= 0
    pts[sel] = np.asarray(data, dtype=np.float64)[ex[sel]]          # single-sample nodes carry the sample itself, in float64
    info["n_nodes"] = n_nodes
    flat = (pts, np.arange(n_nodes, dtype=np.int64), child_ptr, child_idx, first)
    return flat, info


# ---------------------------------------------------------------------------
# sharded build (SURVEY.md section 8e, config 5 on several GPUs): the samples of every tree node are split
# between the ranks by contiguous id ranges; per Lloyd iteration ONE all-reduce of the per-node sums and
# counts is the path's real exchange step (NCCL on the GPUs, gloo in the CPU tests).  Every decision
# (children, live / leaf, centroids) is taken from the all-reduced numbers, so all ranks build the same
# tree; the kd-tree members are all-gathered as sample ids at the end (the data itself is replicated).
# ---------------------------------------------------------------------------
def _dist():
    import torch.distributed as dist
    return dist


def _allreduce(t, group=None):
    _dist().all_reduce(t, group=group)
    return t


def _allgather_var(t, group=None):
    """Concatenation over ranks (rank order) of 1-D tensors of different lengths."""
    torch = engine._torch()
    dist = _dist()
    world = dist.get_world_size(group)
    n = torch.tensor([t.numel()], dtype=torch.long, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(max(sizes), 1)
    buf = torch.zeros(m, dtype=t.dtype, device=t.device)
    buf[:t.numel()] = t
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)])


def _pick_global_rows(Xl, starts, lcounts, pos, group=None):
    """Rows at GLOBAL member positions pos [n_live, k] of every node (a node's global member list is the
    concatenation over ranks of its local members): the owner contributes the row, one all-reduce."""
    torch = engine._torch()
    dist = _dist()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    allc = [torch.zeros_like(lcounts) for _ in range(world)]
    dist.all_gather(allc, lcounts.contiguous(), group=group)
    allc = torch.stack(allc)                                  # [world, n_live]
    offset = torch.cumsum(allc, 0)[rank] - allc[rank]         # members held by lower ranks
    local = pos - offset[:, None]
    mine = (local >= 0) & (local < lcounts[:, None])
    idx = (starts[:, None] + local).clamp(min=0, max=max(Xl.shape[0] - 1, 0))
    rows = torch.zeros(pos.shape + (Xl.shape[1],), dtype=torch.float32, device=Xl.device)
    if Xl.shape[0] > 0:
        rows = torch.where(mine[..., None], Xl[idx], rows)
    return _allreduce(rows, group)


def _sharded_seeds(Xl, dim, node64, starts, lcounts, gcounts, n_sub, seeds, depth, gids, gen, group):
    """Initial centroids, identical on every rank.  seeds: optional callable (depth, gids, global_counts) ->
    global member positions [n_live, n_sub]; default k-means++ (one trial per centre) with the D^2 sampling
    split in two stages: owner rank from the all-gathered per-rank totals, member from the local cumsum."""
    torch = engine._torch()
    dist = _dist()
    dev = Xl.device
    n_live = int(starts.shape[0])
    if seeds is not None:
        pos = torch.as_tensor(seeds(depth, gids, gcounts), dtype=torch.long, device=dev)
        return _pick_global_rows(Xl, starts, lcounts, pos, group)[..., :dim].contiguous()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    u = torch.rand((n_live, n_sub), generator=gen, device=dev, dtype=torch.float64)
    cent = torch.empty((n_live, n_sub, dim), dtype=torch.float32, device=dev)
    pos0 = (u[:, :1] * gcounts[:, None]).long().minimum(gcounts[:, None] - 1)
    cent[:, 0] = _pick_global_rows(Xl, starts, lcounts, pos0, group)[:, 0, :dim]
    d2 = ((Xl[:, :dim] - cent[node64, 0]) ** 2).sum(dim=1).double()
    for j in range(1, n_sub):
        ltot = torch.zeros(n_live, dtype=torch.float64, device=dev).index_add_(0, node64, d2)
        allt = [torch.zeros_like(ltot) for _ in range(world)]
        dist.all_gather(allt, ltot, group=group)
        allt = torch.stack(allt)
        cum = torch.cumsum(allt, 0)
        target = u[:, j] * cum[-1]
        owner = (cum <= target[None, :]).sum(0).clamp(max=world - 1)          # first rank whose cumulative total exceeds it
        ltarget = target - (cum[rank] - allt[rank])
        cs = torch.cumsum(d2, 0)
        base = torch.where(starts > 0, cs[(starts - 1).clamp(min=0, max=max(cs.numel() - 1, 0))] if cs.numel() else torch.zeros(n_live, dtype=torch.float64, device=dev),
                           torch.zeros(n_live, dtype=torch.float64, device=dev))
        pick = torch.searchsorted(cs, base + ltarget, right=True) if cs.numel() else torch.zeros(n_live, dtype=torch.long, device=dev)
        pick = torch.maximum(torch.minimum(pick, starts + lcounts - 1), starts).clamp(min=0, max=max(Xl.shape[0] - 1, 0))
        mine = (owner == rank) & (lcounts > 0)
        row = torch.zeros((n_live, dim), dtype=torch.float32, device=dev)
        if Xl.shape[0] > 0:
            row = torch.where(mine[:, None], Xl[pick, :dim], row)
        cent[:, j] = _allreduce(row, group)
        d2 = torch.minimum(d2, ((Xl[:, :dim] - cent[node64, j]) ** 2).sum(dim=1).double())
    return cent.contiguous()


def _subdivide_level_sharded(Xl, dim, n_sub, starts, lcounts, depth, gids, seeds, gen, dim_acc, group):
    """One level of the sharded build: tolerance, seeds and every Lloyd M-step from all-reduced sums.
    Returns (perm, child_node, global child counts, local child counts, child means, parent means, iterations)."""
    torch = engine._torch()
    dev = Xl.device
    n_live = int(starts.shape[0])
    node64 = torch.repeat_interleave(torch.arange(n_live, device=dev), lcounts)
    node = node64.to(torch.int32).contiguous()

    def e_step(centroids, k):
        r = _lloyd_level(Xl, dim, node, n_live, k, centroids, None, dim_acc, max_iter=0)
        return r, _allreduce(r["sums"].clone(), group), _allreduce(r["counts"].clone(), group), _allreduce(r["inertia"].clone(), group)

    # per-node mean and sklearn's tolerance from the global moments
    zero = torch.zeros((n_live, 1, dim), dtype=torch.float32, device=dev)
    _, gs, gc, _ = e_step(zero, 1)
    gcounts = gc[:, 0].long()
    cnt = gcounts.clamp(min=1).to(torch.float32)
    mean = gs[:, 0, :] / cnt[:, None]
    _, _, _, gin = e_step(mean[:, None, :dim].contiguous(), 1)
    tol = KMEANS_TOL * gin / (cnt * dim)
    cent = _sharded_seeds(Xl, dim, node64, starts, lcounts, gcounts, n_sub, seeds, depth, gids, gen, group)
    conv = torch.zeros(n_live, dtype=torch.bool, device=dev)
    iters = 0
    for it in range(KMEANS_MAX_ITER):
        _, gs, gc, _ = e_step(cent, n_sub)
        has = (gc > 0)[..., None]
        new = torch.where(has, gs[..., :dim] / gc.clamp(min=1)[..., None].to(torch.float32), cent)
        new = torch.where(conv[:, None, None], cent, new)
        shift = ((new - cent) ** 2).sum(dim=(1, 2))
        conv = conv | (shift <= tol)
        cent = new.contiguous()
        iters = it + 1
        if iters % 4 == 0 and bool(conv.all().item()):
            break
    r, gs, gc, _ = e_step(cent, n_sub)                     # labels and sums that belong to the final centroids
    key = node64 * n_sub + r["label"].long()
    perm = torch.argsort(key, stable=True)
    gflat = gc.reshape(-1).long()
    ck = torch.nonzero(gflat > 0).reshape(-1)
    gcc = gflat[ck]
    lcc = r["counts"].reshape(-1).long()[ck]
    means = gs.reshape(-1, dim_acc)[ck] / gcc[:, None].to(torch.float32)
    return perm, ck // n_sub, gcc, lcc, means, mean, iters


def build_cluster_tree_sharded(data, n_sub, max_level, dim, seeds=None, seed=0, device=None, group=None):
    """build_cluster_tree with the samples split between the ranks of `group` (rank r works on the ids
    shard_range(N, r, world)); `data` is the full array on every rank.  Collective: every rank returns the
    same (FlatClusterTree, info)."""
    torch = engine._torch()
    from . import sharding
    dist = _dist()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = _device_of(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(seed))
    data = np.asarray(data)
    N, D = data.shape
    X = _upload_f32(data, dev)
    base, n_mine = sharding.shard_range(N, rank, world)
    if not (N > n_sub and n_sub > 1):
        return build_cluster_tree(data, n_sub, max_level, dim, True, None, seed, device)     # nothing to share
    node_means, node_leaf = [None], [torch.zeros(1, dtype=torch.bool, device=dev)]
    edges_parent, edges_child = [], []
    kd_owner, kd_rows_local, kd_seg_local = [], [], []
    n_nodes, n_seg = 1, 0
    info = {"levels": [], "n_samples": int(N), "world": world}
    Xl = X[base:base + n_mine].contiguous()
    order_l = torch.arange(base, base + n_mine, device=dev)
    gids = torch.zeros(1, dtype=torch.long, device=dev)
    starts = torch.zeros(1, dtype=torch.long, device=dev)
    lcounts = torch.full((1,), n_mine, dtype=torch.long, device=dev)
    depth = 0
    while gids.numel() > 0:
        perm, child_node, gcc, lcc, cmeans, pmean, iters = _subdivide_level_sharded(Xl, dim, n_sub, starts, lcounts, depth, gids, seeds,
                                                                                     gen, D, group)
        if depth == 0:
            node_means[0] = pmean
        Xl, order_l = Xl[perm], order_l[perm]
        parents = gids[child_node]
        info["levels"].append({"depth": depth, "nodes": int(gids.numel()), "local_samples": int(Xl.shape[0]), "lloyd_iterations": int(iters)})
        if depth < max_level:
            n_children = int(gcc.numel())
            cg = n_nodes + torch.arange(n_children, device=dev)
            n_nodes += n_children
            live = gcc > n_sub
            node_means.append(cmeans)
            node_leaf.append(~live)
            edges_parent.append(parents)
            edges_child.append(cg)
            owners, seg_lens = cg[~live], lcc[~live]
            sample_live = torch.repeat_interleave(live, lcc)
            retired = order_l[~sample_live]
            Xl, order_l = Xl[sample_live].contiguous(), order_l[sample_live]
            gids, lcounts = cg[live], lcc[live]
            starts = torch.cumsum(lcounts, 0) - lcounts
        else:
            owners, seg_lens, retired = parents, lcc, order_l
            gids = gids[:0]
        kd_owner.append(owners)
        kd_rows_local.append(retired)
        kd_seg_local.append(n_seg + torch.repeat_interleave(torch.arange(owners.numel(), device=dev), seg_lens))
        n_seg += int(owners.numel())
        depth += 1
    # kd-tree members: ids of every rank, grouped by segment, ascending id inside a segment
    rows = _allgather_var(torch.cat(kd_rows_local), group)
    segs = _allgather_var(torch.cat(kd_seg_local), group)
    o = torch.argsort(rows, stable=True)
    o = o[torch.argsort(segs[o], stable=True)]
    rows, segs = rows[o], segs[o]
    seg_len = torch.bincount(segs, minlength=n_seg)
    owner = torch.cat(kd_owner)
    kd_pt, kd_l, kd_r, kd_in, roots = _build_kd_forest(X, rows, seg_len, dim)
    return _assemble_flat_tree(data, X, n_nodes, node_means, node_leaf, edges_parent, edges_child, owner, roots, kd_pt, kd_l, kd_r, kd_in, info)


def _assemble_flat_tree(data, X, n_nodes, node_means, node_leaf, edges_parent, edges_child, owner, roots, kd_pt, kd_l, kd_r, kd_in, info):
    torch = engine._torch()
    dev = X.device
    t = FlatClusterTree()
    means_dev = torch.cat(node_means)
    t._parts = [means_dev.double().cpu().numpy(), np.asarray(data, dtype=np.float64)]
    if dev.type == "cuda":
        t.device_table = torch.cat([means_dev.float(), X], dim=0).contiguous()
    t.node_row = np.arange(n_nodes, dtype=np.int64)
    t.node_leaf = torch.cat(node_leaf).cpu().numpy()
    ep = torch.cat(edges_parent).cpu().numpy() if edges_parent else np.empty(0, dtype=np.int64)
    ec = torch.cat(edges_child).cpu().numpy() if edges_child else np.empty(0, dtype=np.int64)
    kp, kc = owner.cpu().numpy(), roots.cpu().numpy()
    par = np.concatenate([ep, kp])
    chi = np.concatenate([ec, kc])
    iskd = np.concatenate([np.zeros(len(ep), dtype=bool), np.ones(len(kp), dtype=bool)])
    o = np.argsort(par, kind="stable")
    t.child_idx, t.child_is_kd = chi[o], iskd[o]
    t.child_ptr = np.zeros(n_nodes + 1, dtype=np.int64)
    t.child_ptr[1:] = np.cumsum(np.bincount(par, minlength=n_nodes))
    kd_pt_h, t.kd_left, t.kd_right, t.kd_inner = _to_host(kd_pt, kd_l, kd_r, kd_in)
    t.kd_row = n_nodes + kd_pt_h
    info["n_nodes"] = n_nodes
    info["n_kd_nodes"] = int(len(t.kd_row))
    return t, info
