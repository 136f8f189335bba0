"""Writes parallel computation graphs in the reference's v1 file format
(`to_v1(ParallelComputationGraph)`, lib/pcg/include/pcg/file_format/v1/*: `raw_graph.node_labels` as
[id, ParallelLayerAttrs] pairs, `raw_graph.graph.{nodes, edges}`), for the models of lib/models with
a Megatron-style placement expressed in the reference's own parallel operators. The reference's
search does not emit a strategy for these models, so this is where the "imported search
strategy" of BASELINE configs[3] comes from; `strategy.load_strategy` reads the files back and
`training.build_from_strategy` turns them into a layer list.

Placement of one BERT layer for tensor-parallel degree t inside data-parallel degree d:

    x --Replicate(t)--> MHA (heads sharded: weight Repartition along its head dimension) --Reduction(t)--> + x
      --> LayerNorm --Replicate(t)--> dense(dim_ff, GELU; weight Repartition along out-channels) --> dropout
      --> dense(hidden; weight Repartition along in-channels) --Reduction(t)--> dropout --> + --> LayerNorm

and the head is vocabulary-parallel: dense(vocab, GELU; out-channel shard) --Combine(feature dim, t)-->
softmax. The input is Repartition'ed along the batch (degree d), the output Combine'd along it, and
every weight carries the Replicate(d) that ParallelComputationGraphBuilder::add_weight would hang
behind it (parallel_computation_graph_builder.cc:588-634)."""
from __future__ import annotations

import json
from typing import Any, Dict, List, Optional


class _Graph:
    def __init__(self):
        self.labels: List[Dict[str, Any]] = []
        self.edges: List[Dict[str, int]] = []

    def node(self, kind: str, value: Dict[str, Any], inputs=(), name: Optional[str] = None) -> int:
        nid = len(self.labels)
        self.labels.append({"op_attrs": {"type": kind, "value": value}, "name": name})
        for i, src in enumerate(inputs):
            self.edges.append({"srcNode": src, "srcIdx": 0, "dstNode": nid, "dstIdx": i})
        return nid

    def doc(self) -> Dict[str, Any]:
        return {"raw_graph": {"node_labels": [[i, l] for i, l in enumerate(self.labels)],
                              "output_labels": [],
                              "graph": {"nodes": list(range(len(self.labels))), "edges": self.edges}}}


def _shape(dims, dtype="FLOAT"):
    return {"dims": {"ff_ordered": list(dims)}, "data_type": dtype}


def bert_pcg(num_layers=12, hidden=768, heads=12, dim_ff=3072, seq=512, batch=64, vocab=30522, dropout=0.1,
             layer_norm_eps=1e-12, tensor_parallel=1, data_parallel=1) -> Dict[str, Any]:
    """lib/models' BERT graph (lib/models/src/models/bert/bert.cc:52-155) with the placement above.
    `batch` is the GLOBAL batch (the graph's input shape); the vocabulary must divide by the
    tensor-parallel degree (pad it, as every vocabulary-parallel implementation does)."""
    t, d = tensor_parallel, data_parallel
    assert heads % t == 0 and dim_ff % t == 0 and vocab % t == 0 and batch % d == 0
    g = _Graph()
    kd = dim_ff // heads

    def weight(shape, shard_dim=None, name=None):
        n = g.node("weight", {"tensor_shape": _shape(shape)}, name=name)
        if shard_dim is not None and t > 1:
            n = g.node("partition_distributed", {"repartition_dim": {"value": shard_dim}, "repartition_degree": t}, [n])
            if d > 1:
                n = g.node("replicate_distributed", {"replicate_degree": d}, [n])
        elif d * t > 1:
            n = g.node("replicate_distributed", {"replicate_degree": d * t}, [n])
        return n

    def replicate(x):
        return g.node("replicate_distributed", {"replicate_degree": t}, [x]) if t > 1 else x

    def reduction(x):
        return g.node("reduce_distributed", {"reduction_degree": t}, [x]) if t > 1 else x

    cur = g.node("input", {"tensor_shape": _shape([batch, seq, hidden])}, name="embedded_sequence")
    if d > 1:
        cur = g.node("partition_distributed", {"repartition_dim": {"value": 0}, "repartition_degree": d}, [cur])
    drop_seed = 22345
    for i in range(num_layers):
        a_in = replicate(cur)
        w_att = weight([3 * hidden * kd + kd * hidden, heads], shard_dim=1, name=f"layer{i}.attention.weight")
        att = g.node("multi_head_attention",
                     {"embed_dim": hidden, "num_heads": heads, "kdim": kd, "vdim": kd, "dropout": dropout, "bias": False,
                      "add_bias_kv": False, "add_zero_attn": False}, [a_in, a_in, a_in, w_att], name=f"layer{i}.attention")
        a_out = reduction(att)
        add1 = g.node("element_binary", {"type": "EW_ADD", "compute_type": "FLOAT", "should_broadcast_lhs": False,
                                         "should_broadcast_rhs": False}, [a_out, cur], name=f"layer{i}.add1")
        ln1 = g.node("layer_norm", {"axes": [{"value": 2}], "elementwise_affine": True, "eps": layer_norm_eps},
                     [add1, weight([hidden]), weight([hidden])], name=f"layer{i}.ln1")
        f_in = replicate(ln1)
        ff1 = g.node("linear", {"out_channels": dim_ff, "use_bias": True, "data_type": "FLOAT", "activation": "GELU",
                                "regularizer": None, "seed": 100 + i},
                     [f_in, weight([hidden, dim_ff], shard_dim=1), weight([dim_ff], shard_dim=0)], name=f"layer{i}.ff1")
        dp1 = g.node("dropout", {"rate": dropout, "seed": drop_seed + 2 * i}, [ff1], name=f"layer{i}.dropout1")
        ff2 = g.node("linear", {"out_channels": hidden, "use_bias": True, "data_type": "FLOAT", "activation": None,
                                "regularizer": None, "seed": 200 + i},
                     [dp1, weight([dim_ff, hidden], shard_dim=0), weight([hidden])], name=f"layer{i}.ff2")
        f_out = reduction(ff2)
        dp2 = g.node("dropout", {"rate": dropout, "seed": drop_seed + 2 * i + 1}, [f_out], name=f"layer{i}.dropout2")
        add2 = g.node("element_binary", {"type": "EW_ADD", "compute_type": "FLOAT", "should_broadcast_lhs": False,
                                         "should_broadcast_rhs": False}, [ln1, dp2], name=f"layer{i}.add2")
        cur = g.node("layer_norm", {"axes": [{"value": 2}], "elementwise_affine": True, "eps": layer_norm_eps},
                     [add2, weight([hidden]), weight([hidden])], name=f"layer{i}.ln2")
    h_in = replicate(cur)
    head = g.node("linear", {"out_channels": vocab, "use_bias": True, "data_type": "FLOAT", "activation": "GELU",
                             "regularizer": None, "seed": 300},
                  [h_in, weight([hidden, vocab], shard_dim=1), weight([vocab], shard_dim=0)], name="head")
    if t > 1:
        head = g.node("combine_distributed", {"combine_dim": {"value": 2}, "combine_degree": t}, [head])
    out = g.node("softmax", {"dim": {"value": 2}}, [head], name="probabilities")
    if d > 1:
        g.node("combine_distributed", {"combine_dim": {"value": 0}, "combine_degree": d}, [out])
    return g.doc()


def write(path, doc) -> None:
    with open(path, "w") as f:
        json.dump(doc, f, separators=(",", ":"))
