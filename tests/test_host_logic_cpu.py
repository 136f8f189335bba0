"""Host-side helpers that need no GPU: pinned-staging NUMA parsing, the strategy file the BERT
workload imports, and the headers' new entry points being exported by the library."""
import importlib
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_cpulist_parsing(ffk):
    h = importlib.import_module("flexflow_train_b200.hostmem")
    assert h._parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert h._parse_cpulist("") == []
    info = h.bind_host_thread_to_gpu_node(0)   # no GPU here: must report, not raise
    assert info["bound"] in (False, True) and "gpu_numa_node" in info


def test_strategy_file_names_a_tensor_parallel_degree():
    s = json.loads((ROOT / "strategies" / "bert_tp2.json").read_text())
    assert int(s["tensor_parallel"]) == 2


def test_peer_memory_and_static_entry_points_are_exported(ffk):
    lib = ffk.lib()
    for name in ("ffk_p2p_create", "ffk_p2p_connect", "ffk_p2p_local_buffer", "ffk_p2p_all_reduce_sum",
                 "ffk_p2p_status", "ffk_p2p_destroy", "ffk_layer_norm_backward_static"):
        assert name in lib.protos and hasattr(lib.cdll, name)


def test_movement_model_reproduces_measured_collectives(ffk):
    """The NVLink model of the cost estimator against the all-reduce times measured on B200
    (profiles/r01_summary.md), within 25 %."""
    ce = importlib.import_module("flexflow_train_b200.cost_estimator")
    for nbytes, ranks, measured_ms in ((32768, 8, 0.0149), (33.563e6, 8, 0.1387), (33.563e6, 2, 0.0778)):
        got = ce.estimate_movement_ms(int(nbytes), "all_reduce", ranks)
        assert abs(got - measured_ms) <= 0.25 * measured_ms, (nbytes, ranks, got, measured_ms)
    assert ce.estimate_movement_ms(0, "all_reduce", 8) == 0.0
    assert ce.estimate_movement_ms(1 << 20, "all_reduce", 1) == 0.0
    assert ce.estimate_movement_ms(1 << 30, "copy", 2) > ce.estimate_movement_ms(1 << 20, "copy", 2)


def test_strategy_reader_understands_both_forms(ffk, tmp_path):
    S = importlib.import_module("flexflow_train_b200.strategy")
    assert S.load_strategy(ROOT / "strategies" / "bert_tp2.json").tensor_parallel == 2
    # a parallel computation graph in the reference's v1 layout: input -> Repartition(sample dim, 4)
    # -> Replicate(2) -> Linear -> Reduction(2) -> Combine(sample dim, 4)
    def layer(kind, value):
        return {"op_attrs": {"type": kind, "value": value}, "name": None}
    pcg = {"raw_graph": {
        "node_labels": {"0": layer("InputAttrs", {}),
                        "1": layer("RepartitionAttrs", {"repartition_dim": {"value": 0}, "repartition_degree": 4}),
                        "2": layer("ReplicateAttrs", {"replicate_degree": 2}),
                        "3": layer("LinearAttrs", {"out_channels": 3072, "use_bias": True}),
                        "4": layer("ReductionAttrs", {"reduction_degree": 2}),
                        "5": layer("CombineAttrs", {"combine_dim": 0, "combine_degree": 4})},
        "output_labels": {},
        "graph": {"nodes": [0, 1, 2, 3, 4, 5],
                  "edges": [{"srcNode": i, "srcIdx": 0, "dstNode": i + 1, "dstIdx": 0} for i in range(5)]}}}
    p = tmp_path / "pcg.json"
    p.write_text(json.dumps(pcg))
    s = S.load_strategy(p)
    assert s.tensor_parallel == 2 and s.data_parallel == 4
    assert sorted(f["op"] for f in s.parallel_ops) == ["combine", "reduction", "repartition", "replicate"]
