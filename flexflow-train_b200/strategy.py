"""Reading a parallelisation strategy for the step driver.

Two inputs are understood:

* this repository's own one-line form, `{"tensor_parallel": T}`;
* a parallel computation graph in the reference's v1 file format — the JSON that
  `to_v1(ParallelComputationGraph)` writes (`lib/pcg/include/pcg/file_format/v1/
  v1_parallel_computation_graph.struct.toml`: `raw_graph.node_labels` maps node ids to
  `ParallelLayerAttrs{op_attrs, name}`, `raw_graph.graph` holds `nodes` and `edges`
  `{srcNode, srcIdx, dstNode, dstIdx}`, `v1_graph_edge.struct.toml`). The reference only writes
  this format (`from_v1` does not exist); this module is the reader.

The graph is rebuilt from its EDGES, because the degrees alone mislead: the reference's
`ParallelComputationGraphBuilder::add_weight` hangs the `generate_weight_transform` chain behind
every Weight node (`lib/pcg/src/pcg/parallel_computation_graph/parallel_computation_graph_builder.cc:
588-634`), so a purely data-parallel graph of degree 8 carries a `Replicate{8}` behind each weight.
Those chains describe how a WEIGHT is laid out, not a split of the model:

* a parallel op reached from a Weight node through parallel ops only belongs to that weight's
  layout: Replicate there is the weight replication of data parallelism; Repartition there shards
  the weight (tensor parallelism: an out-channel shard makes the consumer column-parallel, an
  in-channel shard row-parallel);
* a parallel op on an ACTIVATION places a collective in the layer list: Repartition on dim 0 is
  the batch split (data parallelism), Repartition / Combine on another dim slice / gather features,
  Replicate hands an activation to a tensor-parallel group, Reduction sums its partial results.

`parse_strategy` returns the degrees (validated against each other and, when given, the world size)
and `placements`: for every compute layer in topological order, the parallel ops in front of each
of its activation inputs, behind its output, and on each of its weights.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple

# keys of PCGOperatorAttrs in declaration order (lib/op-attrs/include/op-attrs/
# pcg_operator_attrs.variant.toml): a variant is written either with its key or with its index
PCG_OPERATOR_KEYS = [
    "batch_matmul", "batch_norm", "broadcast", "cast", "combine_distributed", "concat", "conv2d",
    "dropout", "element_binary", "element_unary", "embedding", "flat", "gather", "input",
    "layer_norm", "linear", "multi_head_attention", "noop", "pool2d", "reduce", "reduce_distributed",
    "partition_distributed", "replicate_distributed", "reverse", "reshape", "split", "softmax",
    "topk", "transpose", "weight",
]
PARALLEL_OPS = {"replicate_distributed": "replicate", "reduce_distributed": "reduction",
                "partition_distributed": "repartition", "combine_distributed": "combine"}
_TYPE_NAMES = {"ReplicateAttrs": "replicate_distributed", "ReductionAttrs": "reduce_distributed",
               "RepartitionAttrs": "partition_distributed", "CombineAttrs": "combine_distributed",
               "LinearAttrs": "linear", "MultiHeadAttentionAttrs": "multi_head_attention",
               "LayerNormAttrs": "layer_norm", "DropoutAttrs": "dropout", "InputAttrs": "input",
               "WeightAttrs": "weight", "ElementBinaryAttrs": "element_binary", "SoftmaxAttrs": "softmax",
               "ElementUnaryAttrs": "element_unary", "EmbeddingAttrs": "embedding", "ConcatAttrs": "concat"}
_FIELD_HINTS = [("replicate_degree", "replicate_distributed"), ("reduction_degree", "reduce_distributed"),
                ("repartition_degree", "partition_distributed"), ("combine_degree", "combine_distributed"),
                ("out_channels", "linear"), ("num_heads", "multi_head_attention"),
                ("elementwise_affine", "layer_norm"), ("rate", "dropout")]


@dataclass
class ParallelOp:
    op: str            # replicate | reduction | repartition | combine
    degree: int
    dim: int = 0       # ff_dim of repartition / combine
    node: int = -1


@dataclass
class PCGNode:
    id: int
    kind: str                      # a PCG_OPERATOR_KEYS entry, or "unknown"
    attrs: Dict[str, Any]
    name: Optional[str] = None
    inputs: List[Tuple[int, int]] = field(default_factory=list)  # (srcNode, srcIdx) by dstIdx


@dataclass
class LayerPlacement:
    node: int
    kind: str
    name: Optional[str]
    attrs: Dict[str, Any]
    inputs: List[List[ParallelOp]]   # per activation input: parallel ops in front of it
    weights: List[List[ParallelOp]]  # per weight input: the weight's layout chain
    output: List[ParallelOp]         # parallel ops directly behind output 0
    producers: List[int] = field(default_factory=list)  # compute / input node feeding each activation input


@dataclass
class Strategy:
    tensor_parallel: int = 1
    data_parallel: int = 0                 # 0 = "whatever is left of the world size"
    parallel_ops: List[Dict[str, Any]] = field(default_factory=list)  # activation-side ops, for reports
    placements: List[LayerPlacement] = field(default_factory=list)
    weight_replication: int = 1            # Replicate degree found on weight chains (= data-parallel degree)


def _dim_value(v) -> int:
    # ff_dim_t serialises as a bare int or as {"value": n}
    if isinstance(v, dict):
        return int(v.get("value", 0))
    return int(v)


def _pairs(m) -> List[Tuple[int, Any]]:
    """std::unordered_map<int, X>: nlohmann writes a non-string-keyed map as [[k, v], ...]; a
    string-keyed object is accepted too."""
    if isinstance(m, dict):
        return [(int(k), v) for k, v in m.items()]
    return [(int(k), v) for k, v in m]


def _classify(op_attrs) -> Tuple[str, Dict[str, Any]]:
    """(kind, attribute record) of a serialised PCGOperatorAttrs."""
    inner = op_attrs
    kind = None
    if isinstance(op_attrs, dict):
        if "type" in op_attrs and "value" in op_attrs:
            kind, inner = str(op_attrs["type"]), op_attrs["value"]
        elif "index" in op_attrs and "value" in op_attrs:
            i = int(op_attrs["index"])
            kind = PCG_OPERATOR_KEYS[i] if 0 <= i < len(PCG_OPERATOR_KEYS) else None
            inner = op_attrs["value"]
        elif len(op_attrs) == 1 and next(iter(op_attrs)) in PCG_OPERATOR_KEYS:
            kind = next(iter(op_attrs))
            inner = op_attrs[kind]
    if not isinstance(inner, dict):
        inner = {}
    kind = _TYPE_NAMES.get(kind, kind)
    if kind not in PCG_OPERATOR_KEYS:
        kind = None
        for fld, k in _FIELD_HINTS:
            if fld in inner:
                kind = k
                break
    return kind or "unknown", inner


def read_v1_graph(doc: Dict[str, Any]) -> Dict[int, PCGNode]:
    raw = doc["raw_graph"]
    nodes: Dict[int, PCGNode] = {}
    for nid, label in _pairs(raw.get("node_labels", {})):
        kind, attrs = _classify(label.get("op_attrs", label) if isinstance(label, dict) else {})
        name = label.get("name") if isinstance(label, dict) else None
        nodes[nid] = PCGNode(nid, kind, attrs, name)
    graph = raw.get("graph", {})
    for nid in graph.get("nodes", []):
        nodes.setdefault(int(nid), PCGNode(int(nid), "unknown", {}))
    incoming: Dict[int, List[Tuple[int, int, int]]] = {}
    for e in graph.get("edges", []):
        incoming.setdefault(int(e["dstNode"]), []).append((int(e["dstIdx"]), int(e["srcNode"]), int(e["srcIdx"])))
    for dst, lst in incoming.items():
        if dst not in nodes:
            raise ValueError(f"v1 graph: edge into unknown node {dst}")
        for _, src, _ in lst:
            if src not in nodes:
                raise ValueError(f"v1 graph: edge from unknown node {src}")
        nodes[dst].inputs = [(s, i) for _, s, i in sorted(lst)]
    return nodes


def _as_parallel_op(n: PCGNode) -> ParallelOp:
    op = PARALLEL_OPS[n.kind]
    a = n.attrs
    if op == "replicate":
        return ParallelOp(op, int(a.get("replicate_degree", 1)), 0, n.id)
    if op == "reduction":
        return ParallelOp(op, int(a.get("reduction_degree", 1)), 0, n.id)
    if op == "repartition":
        return ParallelOp(op, int(a.get("repartition_degree", 1)), _dim_value(a.get("repartition_dim", 0)), n.id)
    return ParallelOp(op, int(a.get("combine_degree", 1)), _dim_value(a.get("combine_dim", 0)), n.id)


def topological_order(nodes: Dict[int, PCGNode]) -> List[int]:
    indeg = {i: len(n.inputs) for i, n in nodes.items()}
    users: Dict[int, List[int]] = {}
    for i, n in nodes.items():
        for s, _ in n.inputs:
            users.setdefault(s, []).append(i)
    ready = sorted(i for i, d in indeg.items() if d == 0)
    order = []
    while ready:
        i = ready.pop(0)
        order.append(i)
        for u in sorted(users.get(i, [])):
            indeg[u] -= 1
            if indeg[u] == 0:
                ready.append(u)
    if len(order) != len(nodes):
        raise ValueError("v1 graph: the edges contain a cycle")
    return order


def _chain_back(nodes: Dict[int, PCGNode], src: int) -> Tuple[List[ParallelOp], int]:
    """Walks from `src` towards its producer through parallel ops only. Returns the ops in
    dataflow order and the first node that is not a parallel op."""
    ops: List[ParallelOp] = []
    cur = src
    while nodes[cur].kind in PARALLEL_OPS:
        ops.append(_as_parallel_op(nodes[cur]))
        if not nodes[cur].inputs:
            break
        cur = nodes[cur].inputs[0][0]
    ops.reverse()
    return ops, cur


def placements_of(nodes: Dict[int, PCGNode]) -> List[LayerPlacement]:
    users: Dict[int, List[int]] = {}
    for i, n in nodes.items():
        for s, _ in n.inputs:
            users.setdefault(s, []).append(i)
    out: List[LayerPlacement] = []
    for nid in topological_order(nodes):
        n = nodes[nid]
        if n.kind in PARALLEL_OPS or n.kind == "weight":
            continue
        acts, weights, producers = [], [], []
        for s, _ in n.inputs:
            chain, origin = _chain_back(nodes, s)
            if nodes[origin].kind == "weight":
                weights.append(chain)
            else:
                acts.append(chain)
                producers.append(origin)
        behind: List[ParallelOp] = []
        cur = nid
        while True:  # a chain of parallel ops that is the only user of this output
            us = users.get(cur, [])
            if len(us) == 1 and nodes[us[0]].kind in PARALLEL_OPS:
                behind.append(_as_parallel_op(nodes[us[0]]))
                cur = us[0]
            else:
                break
        out.append(LayerPlacement(nid, n.kind, n.name, n.attrs, acts, weights, behind, producers))
    return out


def parse_strategy(doc: Dict[str, Any], world_size: Optional[int] = None) -> Strategy:
    if "raw_graph" not in doc:
        st = Strategy(int(doc.get("tensor_parallel", 1)), int(doc.get("data_parallel", 0)))
        _validate(st, world_size)
        return st
    nodes = read_v1_graph(doc)
    places = placements_of(nodes)
    act_ops: List[ParallelOp] = []
    weight_rep, weight_shard = 1, 1
    seen = set()
    for pl in places:
        for chain in pl.inputs + [pl.output]:
            for op in chain:
                if op.node not in seen:
                    seen.add(op.node)
                    act_ops.append(op)
        for chain in pl.weights:
            for op in chain:
                if op.op == "replicate":
                    weight_rep = max(weight_rep, op.degree)   # data-parallel weight replication
                elif op.op == "repartition":
                    weight_shard = max(weight_shard, op.degree)  # a sharded weight = tensor parallelism
    # data parallelism: the batch dimension (ff dim 0) of an activation is split
    dp = max([o.degree for o in act_ops if o.op in ("repartition", "combine") and o.dim == 0] or [0])
    # tensor parallelism: Replicate on an activation whose partial results a Reduction sums again,
    # feature-dimension Repartition / Combine on activations, or sharded weights
    rep = [o.degree for o in act_ops if o.op == "replicate"]
    red = [o.degree for o in act_ops if o.op == "reduction"]
    tp = 1
    if rep and red:
        tp = max(tp, min(max(rep), max(red)))
    elif red:
        tp = max(tp, max(red))
    tp = max([tp, weight_shard] + [o.degree for o in act_ops if o.op in ("repartition", "combine") and o.dim != 0])
    if dp == 0 and weight_rep > 1 and tp == 1:
        dp = weight_rep
    st = Strategy(tp, dp, [vars(o).copy() for o in act_ops], places, weight_rep)
    _validate(st, world_size)
    return st


def _validate(st: Strategy, world_size: Optional[int]) -> None:
    if st.tensor_parallel < 1 or st.data_parallel < 0:
        raise ValueError(f"strategy: degrees must be positive (tp {st.tensor_parallel}, dp {st.data_parallel})")
    if world_size is None:
        return
    if world_size % st.tensor_parallel != 0:
        raise ValueError(f"strategy: tensor-parallel degree {st.tensor_parallel} does not divide {world_size} ranks")
    if st.data_parallel and st.tensor_parallel * st.data_parallel != world_size:
        raise ValueError(f"strategy: tp {st.tensor_parallel} x dp {st.data_parallel} != world size {world_size}")


def load_strategy(path, world_size: Optional[int] = None) -> Strategy:
    return parse_strategy(json.loads(Path(path).read_text()), world_size)
