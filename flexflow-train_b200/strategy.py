"""Reading a parallelisation strategy for the step driver.

Two inputs are understood:

* this repository's own one-line form, `{"tensor_parallel": T}`;
* a parallel computation graph in the reference's v1 file format — the JSON that
  `to_v1(ParallelComputationGraph)` writes (`lib/pcg/include/pcg/file_format/v1/
  v1_parallel_computation_graph.struct.toml`: `raw_graph.node_labels` maps node ids to
  `ParallelLayerAttrs`, `raw_graph.graph` holds nodes and edges). The reference has no reader for
  it (`from_v1` does not exist). What the driver needs from such a graph is where
  the parallel operators sit and with which degree: Replicate / Reduction pairs around a block
  mean "split this block `degree` ways over a tensor-parallel group"; Repartition / Combine on the
  sample dimension mean data parallelism. The attribute records are found by their field names
  (`replicate_degree`, `reduction_degree`, `repartition_degree` + `repartition_dim`,
  `combine_degree` + `combine_dim`), whatever envelope the variant serialiser put around them.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List


@dataclass
class Strategy:
    tensor_parallel: int = 1
    data_parallel: int = 0                 # 0 = "whatever is left of the world size"
    parallel_ops: List[Dict[str, Any]] = field(default_factory=list)  # what was found, for reports


def _dim_value(v) -> int:
    # ff_dim_t serialises as a bare int or as {"value": n}
    if isinstance(v, dict):
        return int(v.get("value", 0))
    return int(v)


def _walk(node, found: List[Dict[str, Any]]):
    if isinstance(node, dict):
        if "replicate_degree" in node:
            found.append({"op": "replicate", "degree": int(node["replicate_degree"])})
        if "reduction_degree" in node:
            found.append({"op": "reduction", "degree": int(node["reduction_degree"])})
        if "repartition_degree" in node:
            found.append({"op": "repartition", "degree": int(node["repartition_degree"]),
                          "dim": _dim_value(node.get("repartition_dim", 0))})
        if "combine_degree" in node:
            found.append({"op": "combine", "degree": int(node["combine_degree"]),
                          "dim": _dim_value(node.get("combine_dim", 0))})
        for v in node.values():
            _walk(v, found)
    elif isinstance(node, list):
        for v in node:
            _walk(v, found)


def parse_strategy(doc: Dict[str, Any]) -> Strategy:
    if "raw_graph" not in doc:
        return Strategy(int(doc.get("tensor_parallel", 1)), int(doc.get("data_parallel", 0)))
    found: List[Dict[str, Any]] = []
    _walk(doc["raw_graph"].get("node_labels", {}), found)
    tp = max([f["degree"] for f in found if f["op"] in ("replicate", "reduction")] or [1])
    dp = max([f["degree"] for f in found if f["op"] in ("repartition", "combine") and f.get("dim", 0) == 0] or [0])
    # feature-dimension Repartition / Combine pairs are the other way of writing a split block
    tp = max([tp] + [f["degree"] for f in found if f["op"] in ("repartition", "combine") and f.get("dim", 0) != 0])
    return Strategy(tp, dp, found)


def load_strategy(path) -> Strategy:
    return parse_strategy(json.loads(Path(path).read_text()))
