"""Local cost estimator over the sm_100a kernels.

The reference's `LocalCostEstimator::estimate_cost` (lib/local-execution/src/local_cost_estimator.cc:
29-90) builds a one-layer graph, allocates its tensors through a tracking allocator, runs
init / forward / backward under `profiling_wrapper` and returns
`CostDetails{sum of forward + backward milliseconds, tracked bytes}`; parallel ops, Input,
Weight and Noop cost {0, 0} (:36-39) and tensor movement is the constant 0 (:92-100, a TODO).
This module does the same measurement with the step driver's layers and the C-ABI event timer,
and replaces the constant 0 of `estimate_cost(TensorSetMovement)` by an NVLink-5 model so the
machine-mapping search (lib/compiler/.../get_optimal_machine_mapping.cc:159,243) can be fed real
B200 numbers.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from . import kernels as K
from . import training as T
from ._lib import lib


@dataclass
class ProfilingSettings:  # lib/kernels/include/kernels/profiling_settings.struct.toml
    warmup_iters: int = 2
    measure_iters: int = 5


@dataclass
class CostDetails:  # lib/local-execution/include/local-execution/cost_details.struct.toml
    total_elapsed_time: float  # milliseconds, forward + backward, summed over measure_iters
    total_mem_usage: int       # bytes of device memory the op's tensors and state occupy


PARALLEL_OPS = {"repartition", "combine", "replicate", "reduction", "input", "weight", "noop"}


class _Timer:
    def __init__(self):
        self.t = ctypes.c_void_p()
        lib().call("ffk_timer_create", ctypes.byref(self.t))

    def measure(self, fn, settings: ProfilingSettings) -> float:
        """profiling_wrapper (include/kernels/profiling.h:24-49): total ms of the measured iterations."""
        stream = torch.cuda.current_stream().cuda_stream
        ms = ctypes.c_float()
        for i in range(settings.warmup_iters + settings.measure_iters):
            if i == settings.warmup_iters:
                lib().call("ffk_timer_begin", self.t, stream)
            fn()
        lib().call("ffk_timer_end", self.t, stream, ctypes.byref(ms))
        return float(ms.value)

    def __del__(self):
        try:
            lib().call("ffk_timer_destroy", self.t)
        except Exception:
            pass


def _build(op: str, input_shape: Sequence[int], dtype, device, **attrs):
    x = T.Tensor(tuple(input_shape), dtype, device)
    if op == "linear":
        layer = T.Linear(x, attrs["out_channels"], attrs.get("activation", K.ACT_NONE), attrs.get("use_bias", True))
    elif op == "layer_norm":
        layer = T.LayerNorm(x, attrs.get("eps", 1e-5))
    elif op == "softmax":
        layer = T.Softmax(x)
    elif op == "dropout":
        layer = T.Dropout(x, attrs.get("rate", 0.1), attrs.get("seed", 0))
    elif op == "multihead_attention":
        layer = T.SelfAttention(x, attrs["num_heads"], attrs["kdim"], attrs["vdim"])
    elif op == "element_binary":
        layer = T.Add(x, T.Tensor(tuple(input_shape), dtype, device))
    else:
        raise ValueError(f"estimate_cost: no layer wrapper for op {op!r}")
    return x, layer


class LocalCostEstimator:
    def __init__(self, settings: Optional[ProfilingSettings] = None, device="cuda"):
        self.settings = settings or ProfilingSettings()
        self.device = torch.device(device)
        self.timer = _Timer()

    def estimate_cost(self, op: str, input_shape: Sequence[int], dtype=torch.float32, **attrs) -> CostDetails:
        if op in PARALLEL_OPS:
            return CostDetails(0.0, 0)  # local_cost_estimator.cc:36-39
        before = torch.cuda.memory_allocated(self.device)
        x, layer = _build(op, input_shape, dtype, self.device, **attrs)
        mem = torch.cuda.memory_allocated(self.device) - before
        fwd = self.timer.measure(layer.forward, self.settings)
        bwd = self.timer.measure(layer.backward, self.settings)
        return CostDetails(fwd + bwd, int(mem))


# ---- tensor movement (the reference returns 0.0): NVLink 5 through NVSwitch
NVLINK_PEER_GBS = 770.0        # measured peer copy, per direction per GPU (B200_PROFILING.md)
NVLINK_ALLREDUCE_BUS_GBS = 725.0  # measured 8-rank all-reduce bus bandwidth at 1 GiB
HALF_RATE_BYTES = 17e6         # message size at which a collective reaches half of that rate
LAUNCH_LATENCY_MS = 0.015      # one collective: launch + one switch round trip
ONE_SHOT_BYTES = 512 << 10     # up to here every rank talks to all peers directly
ONE_SHOT_GBS = 100.0           # per-rank rate of those small exchanges


def estimate_movement_ms(num_bytes: int, kind: str = "copy", ranks: int = 2) -> float:
    """Milliseconds to move one tensor between GPUs of one NVSwitch domain.
    kind: copy (point to point), all_reduce, all_gather, reduce_scatter, all_to_all.
    Calibrated on this repository's own collectives (profiles/r01_summary.md): 32 KB all-reduce
    at 8 ranks 14.9 us, 33.6 MB 139 us at 8 ranks and 78 us at 2."""
    if num_bytes <= 0 or ranks <= 1:
        return 0.0
    if kind == "copy":
        return LAUNCH_LATENCY_MS + num_bytes / (NVLINK_PEER_GBS * 1e6)
    factor = 2.0 * (ranks - 1) / ranks if kind == "all_reduce" else (ranks - 1) / ranks
    if num_bytes <= ONE_SHOT_BYTES:  # latency-bound protocols (push / one-shot reads)
        return LAUNCH_LATENCY_MS + factor * num_bytes / (ONE_SHOT_GBS * 1e6)
    rate = NVLINK_ALLREDUCE_BUS_GBS * num_bytes / (num_bytes + HALF_RATE_BYTES)  # GB/s at this size
    return LAUNCH_LATENCY_MS + factor * num_bytes / (rate * 1e6)
