"""GPU: the local cost estimator returns what the reference's (commented-out) test would assert —
elapsed > 0 and memory > 0 for a real op (lib/local-execution/test/src/test_local_cost_estimator.cc) —
zero for parallel ops, and times that scale with the work."""
import importlib

import pytest
import torch


@pytest.mark.gpu
def test_local_cost_estimator(ffk):
    C = importlib.import_module("flexflow_train_b200.cost_estimator")
    est = C.LocalCostEstimator(C.ProfilingSettings(1, 3))
    # the reference test's operator: multi-head attention, embed 32, 8 heads on a 10-batch 12-seq tensor
    c = est.estimate_cost("multihead_attention", (10, 12, 32), num_heads=8, kdim=32, vdim=32)
    assert c.total_elapsed_time > 0 and c.total_mem_usage > 0
    assert est.estimate_cost("repartition", (8, 8)) == C.CostDetails(0.0, 0)
    small = est.estimate_cost("linear", (512, 1024), dtype=torch.bfloat16, out_channels=1024)
    big = est.estimate_cost("linear", (8192, 4096), dtype=torch.bfloat16, out_channels=4096)
    assert big.total_elapsed_time > 4 * small.total_elapsed_time
    assert big.total_mem_usage > small.total_mem_usage


def test_movement_model():
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
    import ffk_import
    ffk_import.load()
    C = importlib.import_module("flexflow_train_b200.cost_estimator")
    assert C.estimate_movement_ms(0) == 0.0 and C.estimate_movement_ms(1 << 30, ranks=1) == 0.0
    one_gib = C.estimate_movement_ms(1 << 30, "all_reduce", 8)
    assert 2.0 < one_gib < 3.5  # 2*(7/8) GiB at ~725 GB/s
    assert C.estimate_movement_ms(1 << 30, "all_gather", 8) < one_gib
