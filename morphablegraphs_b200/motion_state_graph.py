"""On-disk motion model -> resident device models.

Reads the reference's graph archive (`utilities/zip_io.py:63-233`: `graph_definition.json`,
`skeleton.json`, `elementary_action_models/elementary_action_<action>/<action>_<mp>_quaternion_mm.json`,
`meta_information.json`) into the same nested dictionary `ZipReader.get_graph_data` returns, builds
the graph the way `MotionStateGraphLoader._build_from_zip_file` /
`build_node_group_from_dict` do (`motion_model/motion_state_graph_loader.py:184-307`: nodes keyed
`(action, primitive)`, node groups with `start_states` / `end_states` / `idle_states` / `cycle_states`
from the meta information, `label_to_motion_primitive_map`, `labeled_frames`, transitions keyed
`action_mp` or `action:mp`, `startNode`), and makes every primitive RESIDENT on the GPU at load:
one `mgb_model` handle per node, created once, so that a synthesis step only switches handles
(`MotionStateGraph.make_resident`).

Not read: space-partitioning files (`*_cluster_tree.{json,pck}`: the engine builds its own flat trees,
`space_partitioning.py`), transition models, hand poses, PFNN weights, `.stats` files.  The skeleton entry
is this package's skeleton descriptor (`Skeleton.from_desc`); the reference parses it with anim_utils'
SkeletonBuilder, which is not part of the reference tree.
"""
import collections
import io
import json
import zipfile

from .motion_model import MotionStateGraphNode
from .skeleton import Skeleton

MORPHABLE_MODEL_FILE_ENDING = "mm.json"
MM_TYPE = "quaternion"
ELEMENTARY_ACTION_DIRECTORY = "elementary_action_models"
GRAPH_DEFINITION_FILE = "graph_definition.json"
SKELETON_JSON_FILE = "skeleton.json"
SKELETON_JSON_KEY = "skeleton"
META_INFORMATION_FILE_NAME = "meta_information.json"

NODE_TYPE_START = "start"
NODE_TYPE_STANDARD = "standard"
NODE_TYPE_END = "end"
NODE_TYPE_SINGLE = "single_primitive"
NODE_TYPE_IDLE = "idle"
NODE_TYPE_CYCLE_END = "cycle_end"


# ---------------------------------------------------------------------------
# archive <-> dictionary
# ---------------------------------------------------------------------------
class ZipReader(object):
    """reference utilities/zip_io.py:52-233, format version >= 2 (directory `elementary_action_models`)."""

    def __init__(self, zip_file_path, pickle_objects=False, verbose=False):
        self.zip_file_path = zip_file_path
        self.type_offset = len(MM_TYPE) + 1
        self.format_version = 1.0
        self.verbose = verbose

    def get_graph_data(self):
        with zipfile.ZipFile(self.zip_file_path, "r") as zf:
            data = json.loads(zf.read(GRAPH_DEFINITION_FILE).decode("utf-8"))
            self.format_version = float(data.get("formatVersion", 1.0))
            if self.format_version < 4.0:
                raise ValueError("format version %s stores the skeleton as BVH text; only the JSON skeleton (>= 4.0) is read" % self.format_version)
            data[SKELETON_JSON_KEY] = json.loads(zf.read(SKELETON_JSON_FILE).decode("utf-8"))
            names = zf.namelist()
            actions = collections.OrderedDict()
            for name in names:                                   # zip_io.py:130-147
                parts = name.split("/")
                if len(parts) > 2 and parts[0] == ELEMENTARY_ACTION_DIRECTORY and parts[2].endswith(MORPHABLE_MODEL_FILE_ENDING):
                    actions.setdefault(parts[1], []).append(parts[2][:-8])
            graph = collections.OrderedDict()
            for structure_key, mps in actions.items():           # zip_io.py:168-199
                action = structure_key.split("_")[2]
                entry = {"name": action, "nodes": collections.OrderedDict()}
                meta = ELEMENTARY_ACTION_DIRECTORY + "/" + structure_key + "/" + META_INFORMATION_FILE_NAME
                if meta in names:
                    entry["info"] = json.loads(zf.read(meta).decode("utf-8"))
                for mp_file in mps:
                    action_offset = len(mp_file[:-self.type_offset].split("_")[0]) + 1
                    mp_key = mp_file[action_offset:-self.type_offset]
                    path = ELEMENTARY_ACTION_DIRECTORY + "/" + structure_key + "/" + mp_file + "_mm.json"
                    entry["nodes"][mp_key] = {"name": mp_file[:-self.type_offset], "mm": json.loads(zf.read(path).decode("utf-8"))}
                graph[action] = entry
            data["subgraphs"] = graph
        return data


def write_graph_zip(path_or_file, graph_data):
    """Writer of the archive `ZipReader` reads (the reference writes it in its construction pipeline,
    construction/motion_model_constructor.py:459-523 per primitive): `graph_data` is the dictionary
    `get_graph_data` returns."""
    with zipfile.ZipFile(path_or_file, "w", zipfile.ZIP_DEFLATED) as zf:
        head = {k: v for k, v in graph_data.items() if k not in ("subgraphs", SKELETON_JSON_KEY)}
        head.setdefault("formatVersion", "4.0")
        zf.writestr(GRAPH_DEFINITION_FILE, json.dumps(head))
        zf.writestr(SKELETON_JSON_FILE, json.dumps(graph_data[SKELETON_JSON_KEY]))
        for action, entry in graph_data["subgraphs"].items():
            base = "%s/elementary_action_%s/" % (ELEMENTARY_ACTION_DIRECTORY, action)
            if "info" in entry:
                zf.writestr(base + META_INFORMATION_FILE_NAME, json.dumps(entry["info"]))
            for mp_key, node in entry["nodes"].items():
                zf.writestr(base + "%s_%s_%s_mm.json" % (action, mp_key, MM_TYPE), json.dumps(node["mm"]))


# ---------------------------------------------------------------------------
# graph
# ---------------------------------------------------------------------------
class MotionStateTransition(object):
    def __init__(self, from_node_key, to_node_key, transition_type, transition_model=None):
        self.from_node_key, self.to_node_key = from_node_key, to_node_key
        self.transition_type, self.transition_model = transition_type, transition_model


class MotionStateGroup(object):
    """reference motion_state_group.py:42-103 (the fields the generator path reads)."""

    def __init__(self, ea_name, graph):
        self.ea_name = ea_name
        self.motion_state_graph = graph
        self.nodes = collections.OrderedDict()
        self.start_states, self.end_states, self.idle_states, self.cycle_states = [], [], [], []
        self.label_to_motion_primitive_map = dict()
        self.labeled_frames = dict()
        self.meta_information = None
        self.has_cycle_states = False

    def set_meta_information(self, meta_information=None):
        """action_meta_info.py:58-73 + motion_state_group.py:87-103: node types from the start / end lists."""
        self.meta_information = meta_information
        if meta_information is None:
            keys = list(self.nodes.keys())
            self.start_states = [keys[0][1]] if keys else []
            self.end_states = list(self.start_states)
            for k in keys:
                self.nodes[k].node_type = NODE_TYPE_SINGLE
            return
        for key in ("start_states", "end_states"):
            assert key in meta_information, "meta information of %s lacks %s" % (self.ea_name, key)
        self.start_states = list(meta_information["start_states"])
        self.end_states = list(meta_information["end_states"])
        self.idle_states = list(meta_information.get("idle_states", []))
        self.cycle_states = list(meta_information.get("cycle_states", []))
        self.has_cycle_states = len(self.cycle_states) > 0
        for (action, mp), node in self.nodes.items():
            if mp in self.start_states and mp in self.end_states:
                node.node_type = NODE_TYPE_SINGLE
            elif mp in self.start_states:
                node.node_type = NODE_TYPE_START
            elif mp in self.end_states:
                node.node_type = NODE_TYPE_END
            elif mp in self.idle_states:
                node.node_type = NODE_TYPE_IDLE
            elif mp in self.cycle_states:
                node.node_type = NODE_TYPE_CYCLE_END
            else:
                node.node_type = NODE_TYPE_STANDARD


class MotionStateGraph(object):
    """reference motion_state_graph.py (container part): nodes, node_groups, skeleton, start_node."""

    def __init__(self):
        self.skeleton = None
        self.nodes = collections.OrderedDict()
        self.node_groups = collections.OrderedDict()
        self.start_node = None
        self.action_definitions = None
        self.animated_joints = None
        self._resident = {}

    def get_outgoing(self, node_key, transition_type=None):
        edges = self.nodes[node_key].outgoing_edges
        return [k for k, e in edges.items() if transition_type is None or e.transition_type == transition_type]

    def make_resident(self, device=None):
        """Create every primitive's device model now (one mgb_model handle per node: bases, GMM tables, packed tensor-core
        operands and the skeleton are uploaded once).  Returns {node_key: DeviceModel}; later steps look their handle up."""
        import torch
        dev = torch.cuda.current_device() if device is None else int(device)
        if dev not in self._resident:
            self._resident[dev] = collections.OrderedDict(
                (key, node.motion_primitive.device_model(self.skeleton, dev)) for key, node in self.nodes.items())
        return self._resident[dev]

    def device_model(self, node_key, device=None):
        return self.make_resident(device)[node_key]


class MotionStateGraphLoader(object):
    """reference motion_state_graph_loader.py:60-307 for the archive / dictionary data sources."""

    def __init__(self):
        self.motion_state_graph_path = None
        self.graph_data = None

    def set_data_source(self, motion_state_graph_path):
        """Path WITHOUT the `.zip` suffix, as the reference takes it (:184-185), or a file object / path of the archive."""
        self.motion_state_graph_path = motion_state_graph_path

    def set_graph_data(self, graph_data):
        self.graph_data = graph_data

    def build(self, resident=False, device=None):
        graph = MotionStateGraph()
        data = self.graph_data
        if data is None:
            src = self.motion_state_graph_path
            if isinstance(src, str) and not src.endswith(".zip"):
                src = src + ".zip"
            data = ZipReader(src).get_graph_data()
        self._build_from_dict(graph, data)
        if resident:
            graph.make_resident(device)
        return graph

    def _build_from_dict(self, graph, data):
        if SKELETON_JSON_KEY not in data:
            raise Exception("There is no skeleton defined in the graph file")
        graph.skeleton = Skeleton.from_desc(data[SKELETON_JSON_KEY])
        graph.animated_joints = graph.skeleton.animated_joints
        for action_name, action_data in data["subgraphs"].items():
            group = self.build_node_group_from_dict(action_data, graph)
            graph.nodes.update(group.nodes)
            graph.node_groups[group.ea_name] = group
            if action_name == "walk" and len(group.idle_states) > 0:
                graph.start_node = (action_name, group.idle_states[0])
        self._set_transitions_from_dict(graph, data.get("transitions", {}))
        if "actionDefinitions" in data:
            graph.action_definitions = data["actionDefinitions"]
        if "startNode" in data:
            start = list(data["startNode"])
            if start[1].startswith("walk"):
                start[1] = start[1][5:]
            graph.start_node = tuple(start)

    def build_node_group_from_dict(self, action_data, graph):
        group = MotionStateGroup(action_data["name"], graph)
        for mp_name, nd in action_data["nodes"].items():
            key = (action_data["name"], mp_name)
            node = MotionStateGraphNode(group)
            node._initialize_from_json(None, nd["mm"])
            node.name = mp_name
            node.action_name = action_data["name"]
            node.set_skeleton(graph.skeleton)
            group.nodes[key] = node
        group.set_meta_information(action_data.get("info"))
        for mp_name, nd in action_data["nodes"].items():
            keyframes = nd["mm"].get("keyframes")
            if keyframes:
                for label in keyframes:
                    group.label_to_motion_primitive_map.setdefault(label, []).append(mp_name)
                group.labeled_frames.setdefault(mp_name, dict()).update(keyframes)
        return group

    def _set_transitions_from_dict(self, graph, transition_dict):
        if len(transition_dict) == 0:
            return
        split_key = ":" if ":" in list(transition_dict.keys())[0] else "_"
        for node_key, targets in transition_dict.items():
            from_key = tuple(node_key.split(split_key)[:2])
            if from_key not in graph.nodes:
                continue
            for to in targets:
                to_key = tuple(to.split(split_key)[:2])
                if to_key in graph.nodes:
                    graph.nodes[from_key].outgoing_edges[to_key] = MotionStateTransition(
                        from_key, to_key, self._get_transition_type(graph, from_key, to_key))

    @staticmethod
    def _get_transition_type(graph, from_key, to_key):
        """:255-277"""
        if to_key[0] != from_key[0]:
            return "action_transition"
        ft, tt = graph.nodes[from_key].node_type, graph.nodes[to_key].node_type
        if ft == NODE_TYPE_IDLE:
            return tt if tt in (NODE_TYPE_START, NODE_TYPE_IDLE, NODE_TYPE_END) else "action_transition"
        if tt in (NODE_TYPE_STANDARD, NODE_TYPE_START, NODE_TYPE_IDLE):
            return tt
        if tt == NODE_TYPE_CYCLE_END:
            return "cycle_end"
        return NODE_TYPE_END


def graph_to_bytes(graph_data):
    buf = io.BytesIO()
    write_graph_zip(buf, graph_data)
    return buf.getvalue()
