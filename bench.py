#!/usr/bin/env python
"""bench.py — the measurement contract of this repository.

Default workload = BASELINE.json configs[2], the configuration the metric "training samples/sec at
1/2/4/8 B200" is quoted on: lib/models' transformer encoder (12 layers, d 1024, 16 heads, seq 512,
batch 64 per GPU, bf16), one STEP = zero grads, forward, MSE gradient, backward, data-parallel
gradient sum, SGD — every arithmetic step a C-ABI call into libffk.so (include/*.h). With
--gpus N > 1 each rank trains on its own batch (weak scaling) and the weight gradients are summed
over the ranks (Replicate backward; SUM, no division) by the peer-memory all-reduce kernels of
libffk.so, issued bucket by bucket under the backward pass (FFK_BENCH_GRAD_SYNC=nccl selects NCCL
all-reduces after backward instead). Before anything is timed at N > 1 one step's summed gradients
are checked: bit-identical on all ranks and within 1 bf16 ulp of an NCCL all-reduce of the same
buckets ("grad_sync_check").

The JSON line carries, besides the contract's keys:
  roofline   the dominant kernel (the tcgen05 GEMM of a feed-forward layer, 275 GFLOP per launch)
             event-timed live inside the timed steps, against the SUSTAINED measured bf16 peak,
             plus the whole step's model FLOPs;
  per_op     (N = 1) BASELINE configs[1], the per-op microbench: every op event-timed on rotating
             buffer sets larger than 2x L2, with SURVEY §8(d)'s algorithmic bytes / FLOP and the
             fraction of the HBM / tensor roofline — measured outside the timed region;
  e2e        the same step fed from pinned HOST memory every step, the step's metric read back;
  cpu_baseline (N = 1) the CPU oracle on a bounded sample of the same workload;
  gpu_launches  kernel nodes of one step captured into a CUDA graph, times the timed steps.

Other workloads: --workload microbench (configs[1] as one step), mlp (configs[0]), bert
[--tp T | --strategy file.json] (configs[3]), dlrm (configs[4]).

`--impl reference` times the reference arm: the CPU restatement of the reference kernels (oracle/,
the reference itself has no CPU path and cannot be built — DESIGN.md) on all host threads, on a
bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

M, KDIM, N = 8192, 4096, 4096          # Linear: batch, in, out
ROWS, COLS = 8192, 4096                # Softmax / LayerNorm / ElementBinary
ENTRIES, EDIM, EBATCH = 1_000_000, 256, 65536
SAMPLES_PER_STEP = M
GEMM_FLOP = 2.0 * M * N * KDIM         # algorithmic FLOP of one Linear-forward launch

# kernels launched by one step, per C-ABI call (bench's own count, reported as gpu_launches)
LAUNCHES_PER_STEP = {
    "copy_dy": 1, "linear_forward": 1, "linear_backward": 5, "softmax_forward": 1, "layer_norm_forward": 1,
    "layer_norm_backward": 2, "element_binary_forward": 1, "element_binary_backward": 1,
    "embedding_forward": 1, "embedding_backward": 1,
}


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return d.get("bf16_tflops", 1590.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 6650.0, "fallback"


def load_peaks_full():
    """(burst bf16 TF/s, sustained bf16 TF/s, HBM GB/s, source) — MEASURED_PEAKS.json, else the
    fallback B200_PROFILING.md states."""
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        burst = d.get("bf16_tflops", 1590.0)
        return burst, d.get("bf16_tflops_sustained", burst), d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 1590.0, 1590.0, 6650.0, "fallback (B200_PROFILING.md)"


def count_step_kernels(torch, step_fn):
    """Kernel launches of one step, counted: the step is captured into a CUDA graph on a side
    stream (nothing executes) and the graph's kernel nodes are counted. Returns
    (kernel nodes, other nodes) or None when the step cannot be captured."""
    try:
        from cuda.bindings import runtime as rt
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph(keep_graph=True)
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                step_fn()
        torch.cuda.current_stream().wait_stream(side)
        raw = g.raw_cuda_graph()
        err, _, n = rt.cudaGraphGetNodes(raw, 0)
        err, nodes, n = rt.cudaGraphGetNodes(raw, n)
        kernels = other = 0
        for nd in nodes[:n]:
            err, ty = rt.cudaGraphNodeGetType(nd)
            if ty == rt.cudaGraphNodeType.cudaGraphNodeTypeKernel:
                kernels += 1
            else:
                other += 1
        del g
        return kernels, other
    except Exception as e:  # noqa: BLE001 - reported in the JSON line
        print(f"bench: kernel count by graph capture failed: {e}", file=sys.stderr)
        try:
            torch.cuda.synchronize()
        except Exception:
            pass
        return None


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- GPU arm
class Workload:
    """Device buffers of one step and the C-ABI calls that make it up."""

    def __init__(self, torch, Kmod, device):
        self.torch, self.K = torch, Kmod
        g = torch.Generator(device=device).manual_seed(1234)
        bf = torch.bfloat16
        r = lambda *s: torch.rand(*s, device=device, generator=g) * 2 - 1
        # Linear
        self.x = r(M, KDIM).to(bf)
        self.w = (r(N, KDIM) / 64).to(bf)
        self.b = r(N).to(bf)
        self.y = torch.empty(M, N, dtype=bf, device=device)
        self.dy_src = r(M, N).to(bf)
        self.dy = torch.empty_like(self.dy_src)
        self.dx = torch.zeros(M, KDIM, dtype=bf, device=device)
        # weight gradients live in flat buckets so that data-parallel sync is one all-reduce each
        self.flat16 = torch.zeros(N * KDIM + N, dtype=bf, device=device)
        self.dw = self.flat16[: N * KDIM].view(N, KDIM)
        self.db = self.flat16[N * KDIM:]
        self.lin = Kmod.Linear.init_kernel(Kmod.ACT_RELU, 0.0, True, Kmod.GEMM_TENSOR, out_dim=N, device=device)
        # Softmax / LayerNorm / ElementBinary (fp32)
        self.a = r(ROWS, COLS)
        self.c = r(ROWS, COLS)
        self.o = torch.empty(ROWS, COLS, device=device)
        self.ga = torch.zeros(ROWS, COLS, device=device)
        self.gc = torch.zeros(ROWS, COLS, device=device)
        self.gamma = r(COLS) + 1
        self.beta = r(COLS)
        self.flat32 = torch.zeros(2 * COLS, device=device)
        self.dgamma = self.flat32[:COLS]
        self.dbeta = self.flat32[COLS:]
        self.ln = Kmod.LayerNorm.init_kernel(True, ROWS, COLS, 1e-5, device=device)
        # Embedding
        self.table = r(ENTRIES, EDIM)
        self.dtable = torch.zeros(ENTRIES, EDIM, device=device)
        self.idx = torch.randint(0, ENTRIES, (EBATCH,), device=device, generator=g, dtype=torch.int64)
        self.eout = torch.empty(EBATCH, EDIM, device=device)
        self.ev_a = torch.cuda.Event(enable_timing=True)
        self.ev_b = torch.cuda.Event(enable_timing=True)
        self.gemm_events = []

    def grads_for_sync(self):
        return [self.flat16, self.flat32]

    def step(self, time_gemm=False, after_linear_backward=None):
        K, t = self.K, self.torch
        K.copy(None, self.dy, self.dy_src)  # backward rewrites dY in place (ReLU gradient)
        if time_gemm:
            a, b = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
            a.record()
        K.Linear.forward_kernel(None, self.lin, self.x, self.y, self.w, self.b, KDIM, N, M)
        if time_gemm:
            b.record()
            self.gemm_events.append((a, b))
        K.Linear.backward_kernel(None, self.lin, self.x, self.dx, self.y, self.dy, self.w, self.dw,
                                 self.db, KDIM, N, M)
        if after_linear_backward is not None:
            after_linear_backward()  # data parallel: start summing dW|db while the rest of the step runs
        K.Softmax.forward_kernel(None, self.a, self.o, ROWS, COLS, 1)
        K.LayerNorm.forward_kernel(None, self.ln, self.a, self.o, self.gamma, self.beta)
        K.LayerNorm.backward_kernel(None, self.ln, self.c, self.a, self.ga, self.gamma, self.dgamma,
                                    self.dbeta)
        K.ElementBinary.forward_kernel(None, K.OperatorType.EW_ADD, self.a, self.c, self.o)
        K.ElementBinary.backward_kernel(None, K.OperatorType.EW_ADD, self.o, self.a, self.c,
                                        self.ga, self.gc)
        K.Embedding.forward_kernel(None, self.idx, self.eout, self.table, K.AGGR_NONE, 1, EDIM, EBATCH)
        K.Embedding.backward_kernel(None, self.idx, self.eout, self.dtable, K.AGGR_NONE, 1, EDIM, EBATCH)


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import ffk_import

    ffk_import.load()
    import importlib
    Kmod = importlib.import_module("flexflow_train_b200.kernels")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    n_gpus = world

    wl = Workload(torch, Kmod, device)
    lib = ffk_import.load().lib()

    pops = None
    peer = None          # gradient sums over NVLink peer memory (default for N > 1)
    grad_sync = "none"
    db_result = wl.db
    if world > 1:
        P = importlib.import_module("flexflow_train_b200.parallel_ops")
        pops = P.ParallelOps()
        peer_error = None
        if os.environ.get("FFK_BENCH_GRAD_SYNC", "peer") == "peer":
            # both gradient buckets and their sums live in one IPC-mapped region per rank
            b16 = (wl.flat16.numel() * 2 + 255) // 256 * 256
            b32 = (wl.flat32.numel() * 4 + 255) // 256 * 256
            try:
                peer = P.PeerMemoryGroup(2 * b16 + 2 * b32)  # raises the same error on every rank
            except RuntimeError as e:
                peer_error = str(e)[:300]
        if peer is not None:
            off = {"g16": 0, "s16": b16, "g32": 2 * b16, "s32": 2 * b16 + b32}
            wl.flat16 = peer.tensor(off["g16"], (wl.flat16.numel(),), torch.bfloat16).zero_()
            wl.dw = wl.flat16[: N * KDIM].view(N, KDIM)
            wl.db = wl.flat16[N * KDIM:]
            wl.flat32 = peer.tensor(off["g32"], (wl.flat32.numel(),), torch.float32).zero_()
            wl.dgamma, wl.dbeta = wl.flat32[:COLS], wl.flat32[COLS:]
            db_result = peer.tensor(off["s16"], (N * KDIM + N,), torch.bfloat16)[N * KDIM:]
            grad_sync = ("peer-memory all-reduce (ffk_p2p_all_reduce_sum, sum over ranks, no division): "
                         "dW|db bf16 32 MiB two-shot on a side stream under the rest of the step, "
                         "dgamma|dbeta fp32 pushed into the peers (one switch traversal)")
        else:
            grad_sync = ("NCCL all-reduce (sum) of two flat gradient buckets per step: dW|db bf16 32 MiB "
                         "(async), dgamma|dbeta fp32")
            if peer_error:
                grad_sync += " [peer-memory path unavailable on this box: " + peer_error + "]"
    side = torch.cuda.Stream(device=device) if peer is not None else None
    lin_done, side_done = torch.cuda.Event(), torch.cuda.Event()

    pending = []

    def start_weight_sync():
        # Replicate-backward / *_nccl_update gradient sync (SUM over ranks, no division) of the
        # 32 MiB dW|db bucket, under the memory-bound rest of the step
        if peer is not None:
            lin_done.record()
            side.wait_event(lin_done)
            peer.all_reduce_sum(off["g16"], off["s16"], wl.flat16.numel(), torch.bfloat16, stream=side)
            side_done.record(side)
        else:
            pending.append(dist.all_reduce(wl.flat16, op=dist.ReduceOp.SUM, async_op=True))

    def sync_grads():
        if peer is not None:
            # same order of collectives on every rank: the small bucket follows the large one
            side.wait_stream(torch.cuda.current_stream())
            peer.all_reduce_sum(off["g32"], off["s32"], wl.flat32.numel(), torch.float32, stream=side)
            torch.cuda.current_stream().wait_stream(side)
        elif pops is not None:
            pops.allreduce_gradients([wl.flat32])
            for w in pending:
                w.wait()
            pending.clear()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def full_step(time_gemm=False):
        wl.step(time_gemm, start_weight_sync if pops is not None else None)
        sync_grads()

    for _ in range(args.warmup):
        full_step()
    barrier()

    check = None
    if world > 1:
        # the summed gradients of one step: bit-identical on all ranks, within 1 ulp of NCCL's sums
        full_step()
        torch.cuda.synchronize()
        for flat, name, key in ((wl.flat16, "dW|db", "s16"), (wl.flat32, "dgamma|dbeta", "s32")):
            want = flat.clone()
            dist.all_reduce(want, op=dist.ReduceOp.SUM)
            got = peer.tensor(off[key], (flat.numel(),), flat.dtype) if peer is not None else want
            if peer is None:
                continue   # NCCL path: the bucket itself was all-reduced in place by the step
            ref0 = got.clone()
            dist.broadcast(ref0, src=0)
            if not torch.equal(ref0, got):
                raise SystemExit(f"grad_sync_check: {name}: rank {rank} holds a different sum than rank 0")
            sc = flat.float().abs()
            dist.all_reduce(sc, op=dist.ReduceOp.SUM)
            a, b = got.float(), want.float()
            ulp = sc * (2.0 ** -7 if flat.dtype == torch.bfloat16 else 2.0 ** -22)   # 1 ulp of the partial sums
            bad = ((a - b).abs() > ulp + 1e-30).sum().item()
            if bad or not torch.isfinite(a).all():
                raise SystemExit(f"grad_sync_check: {name}: {bad} elements differ from NCCL by more than 1 ulp")
        check = "ok"
        barrier()

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    launches0 = lib.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        full_step(time_gemm=True)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    abi_calls = lib.launches - launches0
    clock_info = clocks.stop() if rank == 0 else None
    gemm_ms = [a.elapsed_time(b) for a, b in wl.gemm_events]

    # ---- e2e: the same step fed from pinned HOST memory every step, result read back.
    # The loader is double-buffered, as any input pipeline in front of this API would be:
    # while step i computes from device buffer i%2, the copy of step i+1's batch into buffer
    # (i+1)%2 runs on a second stream. Every step still moves its own h2d bytes and its own
    # result inside the timed region; only the overlap is new.
    hostmem = importlib.import_module("flexflow_train_b200.hostmem")
    numa = hostmem.bind_host_thread_to_gpu_node(local_rank)  # staging pages on the GPU's own node
    hx = hostmem.pinned_empty((M, KDIM), torch.bfloat16)
    hx.copy_(wl.x.cpu())
    hidx = hostmem.pinned_empty((EBATCH,), torch.int64)
    hidx.copy_(wl.idx.cpu())
    hres = [hostmem.pinned_empty((N,), torch.bfloat16) for _ in range(2)]
    h2d = hx.numel() * hx.element_size() + hidx.numel() * hidx.element_size()
    d2h = hres[0].numel() * hres[0].element_size()
    xbuf = [wl.x, torch.empty_like(wl.x)]
    ibuf = [wl.idx, torch.empty_like(wl.idx)]
    copy_stream = torch.cuda.Stream(device=device)
    ready = [torch.cuda.Event() for _ in range(2)]     # batch landed in buffer b
    consumed = [torch.cuda.Event() for _ in range(2)]  # compute finished reading buffer b
    main = torch.cuda.current_stream()

    def prefetch(b):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[b])
            if trace is not None:
                c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                c0.record(copy_stream)
                copies.append((c0, c1))
            xbuf[b].copy_(hx, non_blocking=True)
            if trace is not None:
                c1.record(copy_stream)
            ibuf[b].copy_(hidx, non_blocking=True)
            ready[b].record(copy_stream)

    trace = [] if os.environ.get("FFK_BENCH_TRACE") else None
    copies = []

    def e2e_loop(steps):
        for b in range(2):
            consumed[b].record(main)
        prefetch(0)
        for i in range(steps):
            b = i % 2
            if i + 1 < steps:
                prefetch(1 - b)
            main.wait_event(ready[b])
            wl.x, wl.idx = xbuf[b], ibuf[b]
            full_step()
            consumed[b].record(main)
            hres[b].copy_(db_result, non_blocking=True)
            if trace is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(main)
                trace.append(ev)
        main.synchronize()                              # the caller consumes the last result
        wl.x, wl.idx = xbuf[0], ibuf[0]

    # The pinned-host -> device path itself warms up on these boxes (the same 64 MiB copy goes from
    # ~28 to ~55 GB/s over its first few dozen uses), so the untimed part runs at least 12 steps.
    e2e_warmup = max(args.warmup, 12)
    e2e_loop(e2e_warmup)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    e2e_loop(args.steps)
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    if trace:
        tail = trace[-args.steps:]
        print("e2e step ends (ms since first):", [round(tail[0].elapsed_time(t), 2) for t in tail], file=sys.stderr)
        print("h2d copy ms:", [round(a.elapsed_time(b), 2) for a, b in copies[-args.steps:]], file=sys.stderr)

    # max over ranks, as the contract asks
    if world > 1:
        t = torch.tensor([ms_total, ms_e2e], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e = t.tolist()

    if peer is not None:
        peer.check()  # raises if any bounded wait for a peer rank expired
    if rank != 0:
        if world > 1:
            dist.barrier()
            if peer is not None:
                peer.close()
            dist.destroy_process_group()
        return

    tf_peak, hbm_peak, peak_src = load_peaks()
    ms_per_step = ms_total / args.steps
    value = SAMPLES_PER_STEP * n_gpus / (ms_per_step * 1e-3)
    gemm_avg_ms = sum(gemm_ms) / len(gemm_ms)
    achieved = GEMM_FLOP / (gemm_avg_ms * 1e-3) / 1e12
    traffic = None
    tr = ROOT / "profiles" / "gemm_traffic_bytes.json"
    if tr.exists():
        try:
            traffic = json.loads(tr.read_text()).get("dram_bytes_per_launch")
        except Exception:
            traffic = None

    out = {
        "metric": "training samples/sec (per-op microbench step, fwd+bwd)",
        "value": value,
        "unit": "samples/s",
        "n_gpus": n_gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": ms_per_step,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "bf16 (Linear, fp32 accumulate) + fp32 (Softmax/LayerNorm/ElementBinary/Embedding)",
        "data": "synthetic",
        "config": {
            "workload": "BASELINE configs[1]: Linear 8192x4096x4096 bf16 fwd+bwd; Softmax fwd, LayerNorm "
                        "fwd+bwd, ElementBinary add fwd+bwd on 8192x4096 fp32; Embedding 1Mx256 fp32 "
                        "fwd+bwd with 65536 int64 indices",
            "samples_per_step": SAMPLES_PER_STEP,
            "l2": "inputs larger than L2 (step touches >2 GB; L2 is 126 MB), no explicit flush",
            "parallelism": f"dp{n_gpus}" if n_gpus > 1 else "single",
            "grad_sync": grad_sync,
        },
        "roofline": {
            "kernel": "gemm_tc_pair_kernel (tcgen05 cta_group::2, Linear forward 8192x4096x4096 bf16, fused bias+ReLU)",
            "bound": "tensor",
            "achieved": achieved,
            "peak": tf_peak,
            "peak_source": f"{peak_src} MEASURED_PEAKS.json bf16_tflops (burst: the timed region is tens of ms)",
            "unit": "TFLOP/s",
            "frac": achieved / tf_peak,
            "traffic": traffic,
            "avg_launch_ms": gemm_avg_ms,
            "launches_timed": len(gemm_ms),
        },
        "e2e": {
            "value": SAMPLES_PER_STEP * n_gpus / (ms_e2e / args.steps * 1e-3),
            "unit": "samples/s",
            "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h,
            "ms_per_step": ms_e2e / args.steps,
            "warmup_steps": e2e_warmup,
            "host_staging": numa,
        },
        "gpu_launches": args.steps * (sum(LAUNCHES_PER_STEP.values()) + (2 if peer is not None else 0)),
        "abi_calls_timed": abi_calls,
        "clocks": clock_info,
    }
    if check is not None:
        out["grad_sync_check"] = check
    if n_gpus == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(threads=os.cpu_count() or 1, budget_s=20.0, kind_note="port")
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        if peer is not None:
            peer.close()
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- per-op table
L2_BYTES = 126 * 1024 * 1024


def per_op_table(torch, Kmod, device, iters=10):
    """BASELINE configs[1] op by op: event-timed over `iters` launches that rotate through buffer
    sets whose total size exceeds 2x L2 (so no launch finds its operands in the 126 MB L2), with
    SURVEY §8(d)'s algorithmic bytes / FLOP per launch. HBM-bound ops are judged against the
    measured copy bandwidth, GEMMs against the measured burst bf16 peak (each op is timed alone)."""
    K = Kmod
    burst, _, hbm, src = load_peaks_full()
    bf, f32 = torch.bfloat16, torch.float32
    g = torch.Generator(device=device).manual_seed(4321)
    r = lambda *shape: torch.rand(*shape, device=device, generator=g) * 2 - 1
    rows = []

    def timed(name, bound, work, set_bytes, make, run, note=None):
        nsets = max(2, -(-2 * L2_BYTES // max(set_bytes, 1)) + 1)
        sets = [make() for _ in range(nsets)]
        for i in range(3):
            run(sets[i % nsets])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            run(sets[i % nsets])
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        if bound == "tensor":
            ach, peak, unit = work / (us * 1e-6) / 1e12, burst, "TFLOP/s"
        else:
            ach, peak, unit = work / (us * 1e-6) / 1e9, hbm, "GB/s"
        row = {"op": name, "us": round(us, 2), "bound": bound, "algorithmic": work, "achieved": round(ach, 1),
               "peak": peak, "unit": unit, "frac": round(ach / peak, 3), "buffer_sets": nsets}
        if note:
            row["note"] = note
        rows.append(row)
        del sets
        torch.cuda.empty_cache()

    # ---- Linear 8192 x 4096 x 4096 bf16
    flop = 2.0 * M * N * KDIM
    lin = K.Linear.init_kernel(K.ACT_RELU, 0.0, True, K.GEMM_TENSOR, out_dim=N, device=device)
    w = (r(N, KDIM) / 64).to(bf)
    b = r(N).to(bf)
    mk_fwd = lambda: (r(M, KDIM).to(bf), torch.empty(M, N, dtype=bf, device=device))
    timed("Linear fwd 8192x4096x4096 bf16 (+bias+ReLU)", "tensor", flop, 2 * M * KDIM * 2,
          mk_fwd, lambda s: K.Linear.forward_kernel(None, lin, s[0], s[1], w, b, KDIM, N, M))
    mk_d = lambda: (r(M, N).to(bf), torch.zeros(M, KDIM, dtype=bf, device=device))
    timed("Linear dgrad (NN, beta=1)", "tensor", flop, 2 * M * N * 2, mk_d,
          lambda s: K.Gemm.run(s[0], w, s[1], M, KDIM, N, b_mn=True, accumulate=True, path=K.GEMM_TENSOR))
    mk_w = lambda: (r(M, N).to(bf), r(M, KDIM).to(bf), torch.zeros(N, KDIM, dtype=bf, device=device))
    timed("Linear wgrad (TN, beta=1)", "tensor", flop, 2 * M * N * 2 + N * KDIM * 2, mk_w,
          lambda s: K.Gemm.run(s[0], s[1], s[2], N, KDIM, M, a_mn=True, b_mn=True, accumulate=True,
                               path=K.GEMM_TENSOR),
          note="256 tiles = 3.46 waves of 74 CTA pairs: with scratch registered the 34 tiles of the last wave are cut "
               "along K in two (DESIGN 3.1); the rest of the gap to the forward product is the power cap — filling "
               "idle pairs lowers the clock of the busy ones (profiles/r02_kcut_timing.txt)")
    mk_b = lambda: (r(M, KDIM).to(bf), torch.zeros(M, KDIM, dtype=bf, device=device), r(M, N).relu().to(bf),
                    r(M, N).to(bf), torch.zeros(N, KDIM, dtype=bf, device=device), torch.zeros(N, dtype=bf, device=device))
    timed("Linear bwd (ReLU grad + wgrad + bias grad + dgrad)", "tensor", 2 * flop, 4 * M * N * 2 + N * KDIM * 2, mk_b,
          lambda s: K.Linear.backward_kernel(None, lin, s[0], s[1], s[2], s[3], w, s[4], s[5], KDIM, N, M),
          note="4 launches + 2 reduction launches; the elementwise and column-sum passes are HBM-bound")
    del w, b

    # ---- row ops on 8192 x 4096 fp32
    es = 4
    t = ROWS * COLS * es
    mk2 = lambda: (r(ROWS, COLS), torch.empty(ROWS, COLS, device=device))
    timed("Softmax fwd 8192x4096 fp32", "hbm", 2 * t, 2 * t, mk2,
          lambda s: K.Softmax.forward_kernel(None, s[0], s[1], ROWS, COLS, 1))
    gamma, beta = r(COLS) + 1, r(COLS)
    ln = K.LayerNorm.init_kernel(True, ROWS, COLS, 1e-5, device=device)
    timed("LayerNorm fwd 8192x4096 fp32", "hbm", 2 * t + 2 * COLS * 4 + 2 * ROWS * 4, 2 * t, mk2,
          lambda s: K.LayerNorm.forward_kernel(None, ln, s[0], s[1], gamma, beta))
    dg, db = torch.zeros(COLS, device=device), torch.zeros(COLS, device=device)
    x_ln = r(ROWS, COLS)
    K.LayerNorm.forward_kernel(None, ln, x_ln, torch.empty_like(x_ln), gamma, beta)  # mean / rstd of x_ln
    mk_lb = lambda: (r(ROWS, COLS), x_ln.clone(), torch.zeros(ROWS, COLS, device=device))
    timed("LayerNorm bwd (dX, dgamma, dbeta) 8192x4096 fp32", "hbm", 3 * t, 3 * t, mk_lb,
          lambda s: K.LayerNorm.backward_kernel(None, ln, s[0], s[1], s[2], gamma, dg, db, accumulate_input_grad=False),
          note="first-writer form (ffk_layer_norm_backward_ex): read dY, X, write dX = SURVEY a9's 3 passes")
    timed("LayerNorm bwd, accumulating dX (reference convention)", "hbm", 4 * t, 3 * t, mk_lb,
          lambda s: K.LayerNorm.backward_kernel(None, ln, s[0], s[1], s[2], gamma, dg, db),
          note="dX += ...: the old dX is read as well, 4 passes")
    mk3 = lambda: (r(ROWS, COLS), r(ROWS, COLS), torch.empty(ROWS, COLS, device=device))
    timed("ElementBinary add fwd 8192x4096 fp32", "hbm", 3 * t, 3 * t, mk3,
          lambda s: K.ElementBinary.forward_kernel(None, K.OperatorType.EW_ADD, s[0], s[1], s[2]))
    mk5 = lambda: (r(ROWS, COLS), r(ROWS, COLS), r(ROWS, COLS), torch.zeros(ROWS, COLS, device=device),
                   torch.zeros(ROWS, COLS, device=device))
    timed("ElementBinary add bwd 8192x4096 fp32", "hbm", 5 * t, 5 * t, mk5,
          lambda s: K.ElementBinary.backward_kernel(None, K.OperatorType.EW_ADD, s[0], s[1], s[2], s[3], s[4]))
    del x_ln

    # ---- Embedding 1M x 256 fp32, 65536 int64 indices: the table itself is 1 GB, the gathered
    # rows are scattered over it, so only index / output sets rotate
    table = r(ENTRIES, EDIM)
    mk_e = lambda: (torch.randint(0, ENTRIES, (EBATCH,), device=device, generator=g, dtype=torch.int64),
                    torch.empty(EBATCH, EDIM, device=device))
    timed("Embedding fwd 1Mx256 fp32, 65536 i64 indices", "hbm", EBATCH * (8 + 2 * EDIM * 4),
          EBATCH * (8 + 2 * EDIM * 4), mk_e,
          lambda s: K.Embedding.forward_kernel(None, s[0], s[1], table, K.AGGR_NONE, 1, EDIM, EBATCH),
          note="random 1 KB rows of a 1 GB table: every row is a DRAM page miss; 4 rows in flight per thread "
               "(ncu: profiles/r02_perop_full_subset.csv)")
    dtable = table.zero_()
    mk_eb = lambda: (torch.randint(0, ENTRIES, (EBATCH,), device=device, generator=g, dtype=torch.int64),
                     r(EBATCH, EDIM))
    timed("Embedding bwd 1Mx256 fp32 (scatter-add)", "hbm", EBATCH * (8 + 3 * EDIM * 4),
          EBATCH * (8 + 3 * EDIM * 4), mk_eb,
          lambda s: K.Embedding.backward_kernel(None, s[0], s[1], dtable, K.AGGR_NONE, 1, EDIM, EBATCH),
          note="red.global.add does not return the old value: dW is written, not read back over HBM")
    del table, dtable

    # ---- the encoder's own shapes (configs[2]): bf16 rows of 1024 and the attention block
    TOK, D = 64 * 512, 1024
    mk2b = lambda: (r(TOK, D).to(bf), torch.empty(TOK, D, dtype=bf, device=device))
    gb, bb = (r(D) + 1).to(bf), r(D).to(bf)
    lnb = K.LayerNorm.init_kernel(True, TOK, D, 1e-5, device=device)
    timed("LayerNorm fwd 32768x1024 bf16", "hbm", 2 * TOK * D * 2, 2 * TOK * D * 2, mk2b,
          lambda s: K.LayerNorm.forward_kernel(None, lnb, s[0], s[1], gb, bb),
          note="2 KB rows, one warp per row: 30 us launches are latency-bound (two dependent reductions per row, "
               "next row prefetched); ncu in profiles/r02_perop_full_subset.csv")
    dpo = K.Dropout.init_kernel(0.1, 12345, TOK * 4096, device=device)
    mk_dp = lambda: (r(TOK, 4096).to(bf), torch.empty(TOK, 4096, dtype=bf, device=device))
    timed("Dropout fwd 32768x4096 bf16", "hbm", 2 * TOK * 4096 * 2 + TOK * 4096 // 8, 2 * TOK * 4096 * 2, mk_dp,
          lambda s: K.Dropout.forward_kernel(None, dpo, s[0], s[1]),
          note="ALU-bound, not HBM-bound: Philox4x32-10 per 8 elements, 61 % SM throughput under ncu "
               "(profiles/r02_perop_full_subset.csv); the fused encoder step runs dropout in the GEMM epilogue instead")
    torch.cuda.synchronize()
    return {"peaks": {"bf16_tflops_burst": burst, "hbm_gbs": hbm, "source": src},
            "l2": "each op rotates through buffer sets of more than 2 x 126 MB in total", "iters": iters, "rows": rows}


# ----------------------------------------------------------------------------- encoder workload
def grad_sync_check(torch, dist, model, sync, dp_group, rank):
    """N > 1, before timing: one step's gradients summed by the product's own path must be
    bit-identical on every rank and agree with an NCCL all-reduce of the same buckets to 1 bf16
    (fp32 for fp32 buckets) ulp. Returns "ok" or raises."""
    flats = model.grad_buckets
    for b in flats:   # rank-dependent, reproducible contents
        gen = torch.Generator(device=b.device).manual_seed(4242 + rank)
        b.copy_(((torch.rand(b.numel(), device=b.device, generator=gen) - 0.5) * 2e-2).to(b.dtype))
    want = [b.clone() for b in flats]
    scale = [b.float().abs() for b in flats]        # sum over ranks of |g|: the size of the partial sums
    for w, sc in zip(want, scale):
        dist.all_reduce(w, op=dist.ReduceOp.SUM, group=dp_group)
        dist.all_reduce(sc, op=dist.ReduceOp.SUM, group=dp_group)
    if hasattr(sync, "after_layer"):
        for l in reversed(model.layers):
            if l.grads:
                sync.after_layer(l)
        sync.finish()
        got = sync.sums
    else:
        sync.allreduce_gradients(flats)
        got = flats
    torch.cuda.synchronize()
    for gsum, w, sc in zip(got, want, scale):
        if not torch.isfinite(gsum.float()).all():
            raise SystemExit("grad_sync_check: non-finite values in the summed gradients")
        ref0 = gsum.clone()
        dist.broadcast(ref0, src=dist.get_global_rank(dp_group, 0) if dp_group is not None else 0, group=dp_group)
        if not torch.equal(ref0, gsum):
            raise SystemExit(f"grad_sync_check: rank {rank} holds a different sum than rank 0")
        # The product sums in fp32 in rank order and rounds once; NCCL rounds every partial sum of
        # its ring / tree to the storage type. Both are within one ulp (of the storage type) of
        # the LARGEST partial sum, i.e. of sum_r |g_r| — where the summands cancel that is more
        # than an ulp of the result, so the bound is taken on that scale.
        a, b = gsum.float(), w.float()
        ulp = sc * (2.0 ** -7 if gsum.dtype == torch.bfloat16 else 2.0 ** -10 if gsum.dtype == torch.float16 else 2.0 ** -22)
        bad = ((a - b).abs() > ulp + 1e-30).sum().item()
        if bad:
            raise SystemExit(f"grad_sync_check: {bad} elements differ from the NCCL all-reduce by more than 1 ulp "
                             f"of the partial sums")
    if hasattr(sync, "check"):
        sync.check()
    return "ok"


def run_encoder(args):
    """BASELINE configs[2] (default) / configs[3] (--workload bert): lib/models' transformer encoder
    (12 layers, d 1024, 16 heads, seq 512, batch 64 per GPU, bf16, dim_ff 4096 assumed) trained with
    MSE + SGD — or the BERT-base graph with sparse CCE — through the step driver
    (flexflow-train_b200/training.py), data-parallel over --gpus ranks."""
    import importlib

    import torch
    import torch.distributed as dist

    import ffk_import

    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    Kmod = importlib.import_module("flexflow_train_b200.kernels")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    pops = None
    bert = args.workload == "bert"
    tp = 1
    strategy_note = None
    st = None
    if args.strategy:
        # "imported search strategy": {"tensor_parallel": T}, or a parallel computation graph in the
        # reference's v1 JSON layout (flexflow-train_b200/strategy.py reads its edges and places the
        # parallel ops per layer); the rest of the world size is data parallel. "auto:tpT" writes the
        # BERT graph with a tensor-parallel placement of degree T for this world size and reads it back.
        S = importlib.import_module("flexflow_train_b200.strategy")
        if args.strategy.startswith("auto:tp"):
            X = importlib.import_module("flexflow_train_b200.pcg_export")
            t_ = int(args.strategy[len("auto:tp"):])
            st = S.parse_strategy(X.bert_pcg(batch=args.batch * (world // t_), vocab=30528, tensor_parallel=t_,
                                             data_parallel=world // t_), world)
        else:
            st = S.load_strategy(args.strategy, world)
        tp = st.tensor_parallel
        strategy_note = {"file": os.path.basename(args.strategy), "tensor_parallel": st.tensor_parallel,
                         "data_parallel": st.data_parallel or world // st.tensor_parallel,
                         "placed_layers": len(st.placements),
                         "parallel_ops_on_activations": len(st.parallel_ops)}
    if args.tp:
        tp = args.tp
    if not bert or world % tp != 0:
        tp = 1
    dp = world // tp
    tp_group, tp_rank, dp_group = None, 0, None
    grad_sync = "none"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
        if tp > 1:
            # consecutive ranks form a tensor-parallel group; ranks with the same index in their
            # group form a data-parallel group (every rank must create every group)
            for b in range(dp):
                grp = dist.new_group(list(range(b * tp, (b + 1) * tp)))
                if rank // tp == b:
                    tp_group = grp
            for i in range(tp):
                grp = dist.new_group(list(range(i, world, tp)))
                if rank % tp == i:
                    dp_group = grp
            tp_rank = rank % tp
    B, S, L = args.batch, 512, args.layers
    D, H, FF, V = (768, 12, 3072, 30522) if bert else (1024, 16, 4096, 0)
    hd = args.head_dim if args.head_dim > 0 else FF // H
    g = torch.Generator(device=device).manual_seed(1234 + rank // tp)  # one batch per data-parallel replica
    # fused-operator execution is the default; FFK_BENCH_NO_FUSE=1 runs one kernel per operator
    fuse = os.environ.get("FFK_BENCH_NO_FUSE") is None
    from_pcg = bert and st is not None and len(st.placements) > 0
    if from_pcg:
        # the layer list comes from the graph: Replicate / Reduction / Combine placed where the file says
        V = max(int(pl.attrs.get("out_channels", 0)) for pl in st.placements if pl.kind == "linear")
        L = len([pl for pl in st.placements if pl.kind == "multi_head_attention"])
        Kmod.set_gemm_workspace(int(B * S * ((V + 7) // 8 * 8) * 2 * 1.05))
        model = T.build_from_strategy(st, batch=B, dtype=torch.bfloat16, device=device, tp_rank=tp_rank,
                                      tp_group=tp_group, fuse=fuse)
        labels = torch.randint(0, V, (B * S,), generator=g, device=device, dtype=torch.int32)
    elif bert:
        # the vocabulary head's gradients go through repacked operands: that needs scratch
        Kmod.set_gemm_workspace(int(B * S * ((V + 7) // 8 * 8) * 2 * 1.05))
        model = T.build_bert(L, D, H, FF, S, B, V, dropout=0.1, dtype=torch.bfloat16, device=device,
                             tp_rank=tp_rank, tp_world=tp, tp_group=tp_group, fuse=fuse)
        labels = torch.randint(0, V, (B * S,), generator=g, device=device, dtype=torch.int32)
    else:
        # scratch for the K-cut GEMM schedule (configs[1]'s weight-gradient product in the per-op table: 256 tiles
        # = 3.46 waves of 74 CTA pairs; the encoder's own 64-tile weight gradients measured faster whole)
        Kmod.set_gemm_workspace(32 << 20)
        model = T.build_transformer_encoder(L, D, H, S, B, FF, hd, dropout=0.1, dtype=torch.bfloat16, device=device,
                                            fuse=fuse)
    sync_launches_per_step = 0
    if dp > 1:
        P = importlib.import_module("flexflow_train_b200.parallel_ops")
        # tensor-parallel runs keep ONE communication library: their activation all-reduces are NCCL's, and
        # a spinning peer-memory kernel of the data-parallel group on a side stream could wait on a rank whose
        # NCCL kernel waits on a third (a cycle across two libraries)
        if os.environ.get("FFK_BENCH_GRAD_SYNC", "peer") == "peer" and tp == 1:
            try:
                min_mb = int(os.environ.get("FFK_BENCH_SYNC_MIN_MB", "150"))
                blocks = int(os.environ.get("FFK_BENCH_SYNC_BLOCKS", "0"))
                pops = P.PeerGradSync(model, group=dp_group,   # raises the same error on every rank
                                      min_bytes=min_mb << 20, background_blocks=blocks)
                grad_sync = (f"peer-memory all-reduce over NVLink ({'ffk_p2p_all_reduce_sum_background, ' + str(blocks) + ' CTAs' if blocks else 'ffk_p2p_all_reduce_sum'}: "
                             f"SUM over ranks in fp32 rank order, no division), one launch per >= {min_mb} MiB of "
                             f"finished gradients on a side stream while backward continues; SGD reads the summed buckets")
            except RuntimeError as e:
                grad_sync = "NCCL all-reduce [peer-memory path unavailable on this box: " + str(e)[:200] + "]"
        if pops is None:
            pops = P.ParallelOps(group=dp_group)
            if grad_sync == "none":
                grad_sync = "NCCL all-reduce (SUM) of the flat gradient buckets after backward"
    check = None
    if dp > 1:
        check = grad_sync_check(torch, dist, model, pops, dp_group, rank)
    opt = T.SGD(lr=1e-3, weight_decay=1e-4)
    hx = (torch.rand(B, S, D, generator=g, device=device) * 2 - 1).to(torch.bfloat16).cpu().pin_memory()
    target = None if bert else (torch.rand(B, S, D, generator=g, device=device) * 2 - 1).to(torch.bfloat16)
    model.input.data.copy_(hx)
    lib = ffk_import.load().lib()
    perf = Kmod.PerfMetrics(device=device)
    out32 = None if bert else torch.empty(B * S, D, device=device)
    tgt32 = None if bert else target.float().view(B * S, D).contiguous()
    hres = torch.empty(7, dtype=torch.int32).pin_memory()

    def step(sync=pops):
        if bert:
            T.sparse_cce_step(model, labels, opt, sync, global_batch=B * S * dp)
        else:
            T.mse_step(model, target, opt, sync, replicas=dp)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the dominant kernel, timed live: the plain CTA-pair tcgen05 GEMM (40 of a layer's 46 GEMM launches).
    # ffk_gemm_time_next brackets the NEXT such launch with an event pair: set right before the first
    # attention layer's forward, that is its query projection — tokens x (heads*head_dim) x d_model.
    import ctypes
    att0 = next(l for l in model.layers if isinstance(l, T.SelfAttention))
    tok = B * S
    proj_n, proj_k = (H // tp) * hd, D
    gemm_flop = 2.0 * tok * proj_n * proj_k
    gemm_timers = []
    plain_att_forward = att0.forward

    def timed_forward():
        t = ctypes.c_void_p()
        lib.call("ffk_timer_create", ctypes.byref(t))
        lib.call("ffk_gemm_time_next", t)
        plain_att_forward()
        gemm_timers.append(t)

    for _ in range(args.warmup):
        step()
    barrier()
    counted = None
    if tp == 1:
        counted = count_step_kernels(torch, lambda: step(None))   # the compute part of one step
        if hasattr(pops, "launches"):
            l0 = pops.launches
            step()
            sync_launches_per_step = pops.launches - l0
        elif pops is not None:
            sync_launches_per_step = len(model.grad_buckets)
        barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    calls0 = lib.launches
    att0.forward = timed_forward
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    att0.forward = plain_att_forward
    ms_total = e0.elapsed_time(e1)
    abi_calls = lib.launches - calls0 - 2 * len(gemm_timers)
    clock_info = clocks.stop() if rank == 0 else None
    gemm_ms = []
    for t in gemm_timers:
        ms = ctypes.c_float()
        lib.call("ffk_timer_read", t, ctypes.byref(ms))
        lib.call("ffk_timer_destroy", t)
        gemm_ms.append(ms.value)

    def e2e_step():
        model.input.data.copy_(hx, non_blocking=True)          # the batch arrives from the host
        step()
        Kmod._call("ffk_fill", Kmod._stream(None), perf.buf.data_ptr(), 7, 0.0, Kmod.INT32)
        if not bert:  # the metrics kernels take fp32 (reference signature); BERT reads back train_all only
            Kmod.Cast.forward_kernel(None, model.output.data.view(B * S, D), out32)
            Kmod.update_metrics_label_kernel(None, Kmod.METRIC_MSE, out32, tgt32, B * S, D, perf)
        hres.copy_(perf.buf, non_blocking=True)                # the step's metric goes back
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        e2e_step()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(args.steps):
        e2e_step()
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    if world > 1:
        t = torch.tensor([ms_total, ms_e2e], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e = t.tolist()
    if hasattr(pops, "check"):
        pops.check()  # raises if a bounded wait for a peer rank expired
    if rank != 0:
        if world > 1:
            dist.barrier()
            if hasattr(pops, "close"):
                pops.close()
            dist.destroy_process_group()
        return
    ms_per_step = ms_total / args.steps
    model_fused = bool(getattr(model, "fuse", False))
    params = sum(p.numel() for p in model.params)
    # fwd+bwd FLOPs per sample: 3 x (projections + scores + FFN), SURVEY §8d
    per_tok = 2 * (3 * D * H * hd + H * hd * D + 2 * S * H * hd + 2 * D * FF)
    flops_per_sample = 3 * per_tok * S * L + 3 * 2 * D * V * S
    burst, sustained, _, peak_src = load_peaks_full()
    step_tflops = flops_per_sample * B / tp / (ms_per_step * 1e-3) / 1e12  # per GPU
    gemm_avg_ms = sum(gemm_ms) / len(gemm_ms)
    achieved = gemm_flop / (gemm_avg_ms * 1e-3) / 1e12
    traffic = None
    tr = ROOT / "profiles" / "gemm_traffic_bytes.json"
    if tr.exists():
        try:
            traffic = json.loads(tr.read_text()).get(f"{tok}x{proj_n}x{proj_k}")
        except Exception:
            traffic = None
    launches = None
    if counted is not None:
        launches = args.steps * (counted[0] + sync_launches_per_step)
    out = {
        "metric": "training samples/sec (BERT-base graph of lib/models, fwd+bwd+SGD)" if bert else
                  "training samples/sec (transformer encoder, fwd+bwd+SGD)",
        "value": B * dp / (ms_per_step * 1e-3), "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": (f"BASELINE configs[3], data-parallel form: BERT-base graph, {L} layers, hidden {D}, "
                                f"{H} heads, head dim {hd}, dim_ff {FF} GELU, vocabulary head {V} + softmax, seq {S}, "
                                f"batch {B} per GPU, sparse CCE + SGD") if bert else
                               (f"BASELINE configs[2]: transformer encoder {L} layers, d {D}, {H} heads, head dim {hd}, "
                                f"dim_ff {FF}, seq {S}, batch {B} per GPU, dropout 0.1, MSE + SGD"),
                   "global_batch": B * dp, "seq_len": S, "parameters": params,
                   "parallelism": (f"dp{dp}" + (f" x tp{tp}" if tp > 1 else "")) if world > 1 else "single",
                   "grad_sync": grad_sync,
                   "execution": ("fused operators: MHA+residual, dense+activation+dropout, dense+dropout+residual as "
                                 "single GEMM launches; first-writer gradients (no zero fills), shared residual "
                                 "gradient buffers, gated data gradients") if model_fused else "one kernel per operator",
                   "l2": "inputs larger than L2: one layer's activations are >= 0.5 GB against 126 MB of L2, no explicit flush"},
        "roofline": {"kernel": f"gemm_tc_pair_kernel (tcgen05 cta_group::2, plain epilogue): query projection of the first "
                               f"attention layer, {tok}x{proj_n}x{proj_k} bf16, one launch per timed step bracketed by "
                               f"ffk_gemm_time_next; this instantiation is 40 of a layer's 46 GEMM launches and the GEMM "
                               f"family is 88 % of the step (profiles/r02_summary.md)",
                     "bound": "tensor", "achieved": achieved, "peak": sustained,
                     "peak_source": f"{peak_src} bf16_tflops_sustained (kernel timed inside a long step)",
                     "unit": "TFLOP/s", "frac": achieved / sustained, "traffic": traffic,
                     "avg_launch_ms": gemm_avg_ms, "launches_timed": len(gemm_ms),
                     "flop_per_launch": gemm_flop,
                     "whole_step": {"model_tflops_per_gpu": step_tflops, "frac_of_sustained": step_tflops / sustained,
                                    "frac_of_burst": step_tflops / burst, "flop_per_sample": flops_per_sample}},
        "e2e": {"value": B * dp / (ms_e2e / args.steps * 1e-3), "unit": "samples/s",
                "h2d_bytes_per_step": hx.numel() * 2, "d2h_bytes_per_step": 28,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "gpu_launches_how": (f"{counted[0]} kernel nodes (+{counted[1]} other nodes) in one step captured as a CUDA "
                             f"graph, + {sync_launches_per_step} gradient-sum launches, x {args.steps} steps")
                            if counted is not None else "not counted (capture failed or tensor-parallel run)",
        "abi_calls_timed": abi_calls, "clocks": clock_info,
    }
    if check is not None:
        out["grad_sync_check"] = check
    if strategy_note is not None:
        out["config"]["strategy"] = strategy_note
    if world == 1 and not args.no_per_op:
        del model, opt
        torch.cuda.empty_cache()
        out["per_op"] = per_op_table(torch, Kmod, device)
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_encoder_baseline(os.cpu_count() or 1, repeats=2, bert=bert)
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        if hasattr(pops, "close"):
            pops.close()
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- DLRM workload
def run_dlrm(args):
    """BASELINE configs[4]: the DLRM of examples/cpp/DLRM at the shapes of its run script
    (run_random.sh:9): 8 tables of 1 M x 64 fp32, bag size 1, bottom MLP 64-512-512-64, "cat"
    interaction, top MLP 576-1024-1024-1024-1, MSE + SGD; tables sharded over the ranks, lookups
    repartitioned with an all-to-all, MLPs data parallel; the tables are updated row-sparsely
    (ffk_embedding_sgd_update). `--workload dlrm [--batch B]` (B per GPU, default 8192)."""
    import importlib

    import torch
    import torch.distributed as dist

    import ffk_import

    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    pops = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
        pops = importlib.import_module("flexflow_train_b200.parallel_ops").ParallelOps()
    B = args.batch if args.batch != 64 else 8192
    tables = 8
    # what the veneer's ManagedPerDeviceFFHandle does for an fp32 caller of the reference API: GEMM
    # scratch registered and allowTensorOpMathConversion on, so the fp32 MLPs run on the tensor cores
    Kd = importlib.import_module("flexflow_train_b200.kernels")
    Kd.set_gemm_workspace(256 << 20, fp32_tensor_math=True)
    net = T.DLRM(batch=B, embedding_size=(1000000,) * tables, rank=rank, world=world, device=device)
    opt = T.SGD(lr=0.01, weight_decay=0.0)
    g = torch.Generator(device=device).manual_seed(77 + rank)
    hdense = torch.rand(B, 64, generator=g, device=device).cpu().pin_memory()
    target = (torch.rand(B, 1, generator=g, device=device) > 0.5).float()
    hout = torch.empty(B, 1).pin_memory()
    net.dense.data.copy_(hdense)
    lib = ffk_import.load().lib()

    def step():
        T.dlrm_step(net, target, opt, pops)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        return e0.elapsed_time(e1)

    for _ in range(max(args.warmup, 3)):
        step()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    calls0 = lib.launches
    ms_total = timed(step, args.steps)
    abi_calls = lib.launches - calls0
    clock_info = clocks.stop() if rank == 0 else None

    def e2e_step():
        net.dense.data.copy_(hdense, non_blocking=True)
        step()
        hout.copy_(net.model.output.data, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    if world > 1:
        t = torch.tensor([ms_total, ms_e2e], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e = t.tolist()
    if rank == 0:
        ms = ms_total / args.steps
        # bytes one step must move per GPU for the tables: gather of the global batch's rows, and the
        # row-sparse update (dOut read, row read-modify-write); the MLPs (2.6 M parameters) add
        # 3 x 2 x B x params FLOP on the tensor-core side, small next to the table traffic
        tl = tables // world
        row = 64 * 4
        lookups = tl * B * world
        algo = lookups * (8 + 2 * row) + lookups * (8 + 3 * row)
        _, hbm_peak, peak_src = load_peaks()
        print(json.dumps({
            "metric": "training samples/sec (DLRM, fwd+bwd+SGD)", "value": B * world / (ms * 1e-3),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"BASELINE configs[4]: DLRM, {tables} tables of 1M x 64 fp32 sharded over {world} "
                                   f"rank(s), bag 1, bottom 64-512-512-64, cat, top 576-1024-1024-1024-1, batch {B} per GPU, "
                                   f"MSE + SGD, row-sparse table update",
                       "global_batch": B * world, "parallelism": f"tables/{world} x dp{world}" if world > 1 else "single",
                       "l2": "tables exceed L2 (256 MB each)"},
            "roofline": {"kernel": "whole step against the embedding traffic alone (gather + row-sparse update)", "bound": "hbm",
                         "achieved": algo / (ms * 1e-3) / 1e9, "peak": hbm_peak, "peak_source": peak_src,
                         "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / hbm_peak, "traffic": None},
            "e2e": {"value": B * world / (ms_e2e / args.steps * 1e-3), "unit": "samples/s",
                    "h2d_bytes_per_step": B * 64 * 4, "d2h_bytes_per_step": B * 1 * 4,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": None, "abi_calls_timed": abi_calls, "clocks": clock_info}), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- MLP workload
def run_mlp(args):
    """BASELINE configs[0]: the reference's own CPU-runnable case — MLP 784-512-512-10, batch 64,
    fp32, sparse categorical cross-entropy + SGD through the step driver. The step is launch-bound
    (18 small kernels), so it is timed twice: eager C-ABI calls and one CUDA-graph replay per step.
    Single GPU only (`--workload mlp`); a parity-test configuration, not the headline line."""
    import importlib

    import torch

    import ffk_import

    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    torch.cuda.set_device(0)
    device = torch.device("cuda", 0)
    B = 64
    model = T.build_mlp(batch=B)
    opt = T.SGD(lr=0.01, weight_decay=1e-4)
    g = torch.Generator(device=device).manual_seed(7)
    model.input.data.copy_(torch.rand(B, 784, device=device, generator=g))
    labels = torch.randint(0, 10, (B,), device=device, generator=g, dtype=torch.int32)
    lib = ffk_import.load().lib()

    def step():
        T.sparse_cce_step(model, labels, opt)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    calls0 = lib.launches
    step()
    calls_per_step = lib.launches - calls0
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            step()
    torch.cuda.current_stream().wait_stream(side)

    def timed(fn, n):
        for _ in range(max(args.warmup, 3)):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    steps = max(args.steps, 200)
    clocks = ClockSampler(0)
    clocks.start()
    ms_eager = timed(step, steps)
    ms_graph = timed(graph.replay, steps)
    clock_info = clocks.stop()
    print(json.dumps({
        "metric": "training samples/sec (MLP 784-512-512-10, fwd+bwd+SGD)",
        "value": B / (ms_graph * 1e-3), "unit": "samples/s", "n_gpus": 1, "steps": steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_graph, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (exact CUDA-core GEMM, K split over a thread-block cluster)",
        "data": "synthetic",
        "config": {"workload": "BASELINE configs[0]: MLP 784-512-512-10, batch 64, sparse CCE + SGD, one CUDA-graph "
                               "replay per step", "global_batch": B, "l2": "L2-resident by nature (2.7 MB of weights)"},
        "eager": {"ms_per_step": ms_eager, "samples_per_s": B / (ms_eager * 1e-3)},
        "gpu_launches": None, "abi_calls_per_step": calls_per_step, "clocks": clock_info,
    }), flush=True)


# ----------------------------------------------------------------------------- CPU arm
def oracle_step(o, np, rows, erows, bufs=None):
    """The same step on the CPU oracle at `rows` of the 8192 samples (erows embedding indices)."""
    if bufs is None:
        rng = np.random.default_rng(1234)
        u = lambda *s: rng.uniform(-1, 1, size=s).astype(np.float32)
        bufs = dict(x=u(rows, KDIM), w=u(N, KDIM) / 64, b=u(N), y=np.zeros((rows, N), np.float32),
                    dy=u(rows, N), dx=np.zeros((rows, KDIM), np.float32), dw=np.zeros((N, KDIM), np.float32),
                    db=np.zeros(N, np.float32), a=u(rows, COLS), c=u(rows, COLS),
                    o=np.zeros((rows, COLS), np.float32), ga=np.zeros((rows, COLS), np.float32),
                    gc=np.zeros((rows, COLS), np.float32), gamma=u(COLS) + 1, beta=u(COLS),
                    dgamma=np.zeros(COLS, np.float32), dbeta=np.zeros(COLS, np.float32),
                    mean=np.zeros(rows, np.float32), rstd=np.zeros(rows, np.float32),
                    table=u(20000, EDIM), dtable=np.zeros((20000, EDIM), np.float32),
                    idx=rng.integers(0, 20000, erows, dtype=np.int64), eout=np.zeros((erows, EDIM), np.float32))
    B = bufs
    dy = B["dy"].copy()
    o.linear_forward(B["x"], B["w"], B["b"], B["y"], KDIM, N, rows, 0)
    o.linear_backward(B["x"], B["dx"], B["y"], dy, B["w"], B["dw"], B["db"], KDIM, N, rows, 0, 0.0, None)
    o.softmax_forward(B["a"], B["o"], rows, COLS, 1)
    o.layer_norm_forward(B["a"], B["o"], B["gamma"], B["beta"], B["mean"], B["rstd"], rows, COLS, 1e-5)
    o.layer_norm_backward(B["c"], B["a"], B["ga"], B["gamma"], B["dgamma"], B["dbeta"], B["mean"], B["rstd"], rows, COLS, 0)
    shp = [rows, COLS]
    o.element_binary_forward(28, B["a"], B["c"], B["o"], 2, shp, shp, shp)
    o.element_binary_backward(28, B["o"], B["a"], B["c"], B["ga"], B["gc"], 2, shp, shp, shp)
    o.embedding_forward(B["idx"], B["eout"], B["table"], -1, 1, EDIM, erows)
    o.embedding_backward(B["idx"], B["eout"], B["dtable"], -1, 1, EDIM, erows)
    return bufs


def cpu_baseline(threads, budget_s, kind_note):
    import numpy as np
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle as oracle_mod

    os.environ["OMP_NUM_THREADS"] = str(threads)
    o = oracle_mod.load()
    rows = 128
    bufs = oracle_step(o, np, rows, rows * 8)  # warm-up + allocation
    t0 = time.perf_counter()
    n = 0
    while True:
        oracle_step(o, np, rows, rows * 8, bufs)
        n += 1
        if time.perf_counter() - t0 > budget_s / 2 or n >= 5:
            break
    dt = (time.perf_counter() - t0) / n
    return {"value": rows / dt, "unit": "samples/s", "cores": threads, "kind": kind_note,
            "sample": f"{rows} of {M} rows of the same step ({rows * 8} of {EBATCH} embedding indices, "
                      f"20000-row table) on the CPU oracle (OpenMP, {threads} threads); {n} repeats, "
                      f"{dt:.2f} s each"}


ENC = dict(D=1024, H=16, hd=256, FF=4096, S=512, L=12, B=64)   # configs[2]


def cpu_encoder_sample(o, np, layers, enc=None, state=None):
    """One sample (512 tokens) through `layers` encoder layers on the CPU oracle: forward, MSE
    gradient, backward, SGD. Attention, LayerNorm and the Linears treat samples independently, so
    one sample of the batch is a faithful 1/64 of a step; layers are identical, so `layers` of 12
    is a faithful fraction of a sample."""
    sys.path.insert(0, str(ROOT / "tests"))
    from oracle_encoder import OracleEncoder
    if state is None:
        W = OracleEncoder.random_weights(layers, ENC["D"], ENC["H"], ENC["hd"], ENC["FF"], seed=4321)
        rng = np.random.default_rng(1234)
        x = rng.uniform(-1, 1, (1, ENC["S"], ENC["D"])).astype(np.float32)
        t = rng.uniform(-1, 1, (1, ENC["S"], ENC["D"])).astype(np.float32)
        words = lambda n: rng.integers(0, 2 ** 32, (n + 31) // 32, dtype=np.uint32)
        masks = [(words(ENC["S"] * ENC["FF"]), words(ENC["S"] * ENC["D"])) for _ in range(layers)]
        state = (OracleEncoder(o, W, 1, ENC["S"], ENC["D"], ENC["H"], ENC["hd"], ENC["FF"], rate=0.1), x, t, masks)
    encm, x, t, masks = state
    encm.step(x, t, 1e-3, 1e-4, masks=masks, global_volume=ENC["B"] * ENC["S"] * ENC["D"])
    return state


def cpu_encoder_baseline(threads, repeats, bert=False, layers=1):
    import numpy as np
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle as oracle_mod

    os.environ["OMP_NUM_THREADS"] = str(threads)
    o = oracle_mod.load()
    state = cpu_encoder_sample(o, np, layers)          # warm-up + allocation
    t0 = time.perf_counter()
    for _ in range(repeats):
        cpu_encoder_sample(o, np, layers, state=state)
    dt = (time.perf_counter() - t0) / repeats
    value = 1.0 / (dt * ENC["L"] / layers)
    return {"value": value, "unit": "samples/s", "cores": threads, "kind": "port",
            "sample": f"1 of 64 samples (512 tokens) through {layers} of 12 encoder layers (forward, MSE gradient, "
                      f"backward, SGD; d 1024, 16 heads of 256, dim_ff 4096) on the CPU oracle (OpenMP, {threads} "
                      f"threads), {repeats} repeats of {dt:.1f} s; samples/s = 1 / (time x 12 / {layers})"
                      + (" [BERT run: the encoder sample stands in]" if bert else "")}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle as oracle_mod

    threads = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(threads)
    o = oracle_mod.load()
    note = ("the reference has no CPU op kernels and cannot be compiled here (SURVEY F1/F2): this arm times "
            "oracle/cpu_ref.cc, the CPU restatement of its .cu bodies")
    if args.workload != "microbench":
        # configs[2]: each step = one sample through one encoder layer (about 10 s of CPU work)
        layers = 1
        state = cpu_encoder_sample(o, np, layers)
        warm = min(args.warmup, 1)
        for _ in range(warm):
            cpu_encoder_sample(o, np, layers, state=state)
        steps = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(steps):
            cpu_encoder_sample(o, np, layers, state=state)
        dt = (time.perf_counter() - t0) / steps
        value = 1.0 / (dt * ENC["L"] / layers)
        sample = (f"each step = 1 of 64 samples (512 tokens) through {layers} of 12 encoder layers, forward + MSE "
                  f"gradient + backward + SGD; {steps} timed steps of {dt:.1f} s; samples/s = 1 / (time x 12 / {layers})")
        out = {
            "impl": "reference",
            "metric": "training samples/sec (transformer encoder, fwd+bwd+SGD)",
            "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (double accumulation)", "data": "synthetic",
            "config": {"workload": "BASELINE configs[2]: transformer encoder 12 layers, d 1024, 16 heads, head dim 256, "
                                   "dim_ff 4096, seq 512, dropout 0.1, MSE + SGD — bounded sample on host cores",
                       "global_batch": 64, "seq_len": 512},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": note,
        }
        print(json.dumps(out), flush=True)
        return
    rows = 128
    bufs = oracle_step(o, np, rows, rows * 8)
    for _ in range(min(args.warmup, 1)):
        oracle_step(o, np, rows, rows * 8, bufs)
    steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle_step(o, np, rows, rows * 8, bufs)
    dt = (time.perf_counter() - t0) / steps
    value = rows / dt
    sample = (f"each step = {rows} of {M} rows of the workload step ({rows * 8} embedding indices, 20000-row "
              f"table); {steps} timed steps of {dt:.2f} s")
    out = {
        "impl": "reference",
        "metric": "training samples/sec (per-op microbench step, fwd+bwd)",
        "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": steps, "warmup": min(args.warmup, 1),
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (double accumulation)", "data": "synthetic",
        "config": {"workload": "BASELINE configs[1] step, bounded sample on host cores", "samples_per_step": rows},
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": note,
    }
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-per-op", action="store_true", help="skip the per-op table of the default workload")
    ap.add_argument("--workload", default="encoder", choices=["encoder", "microbench", "bert", "dlrm", "mlp"])
    ap.add_argument("--layers", type=int, default=12)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--head-dim", type=int, default=0, help="0 = the reference's dim_ff/heads (256)")
    ap.add_argument("--tp", type=int, default=0, help="bert: tensor-parallel degree (divides --gpus)")
    ap.add_argument("--strategy", default="", help='bert: {"tensor_parallel": T} file, a parallel computation graph '
                    'in the reference v1 JSON layout (strategies/*.v1.json), or auto:tpT')
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference(args)
    elif args.workload in ("encoder", "bert"):
        run_encoder(args)
    elif args.workload == "mlp":
        run_mlp(args)
    elif args.workload == "dlrm":
        run_dlrm(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
