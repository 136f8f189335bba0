"""CPU, world_size 2 over gloo: the rank arithmetic of the multi-GPU parallel operators.
The local copy/accumulate/strided-copy steps are served by an oracle-backed stand-in (the
product's local ops are GPU-only); the collectives are the real torch.distributed calls."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


class OracleLocalOps:
    def __init__(self):
        sys.path.insert(0, str(ROOT / "tests"))
        import oracle

        self.o = oracle.load()

    def copy(self, dst, src):
        self.o.copy(dst.numpy(), src.numpy(), src.numel() * src.element_size())

    def accumulate(self, dst, src):
        self.o.accumulate(dst.numpy(), src.numpy(), src.numel())

    def copy_blocks(self, out, out_off, inp, in_off, num_blocks, out_blk, in_blk, add=False):
        o = out.view(-1).numpy()[out_off:]
        i = inp.contiguous().view(-1).numpy()[in_off:]
        self.o.copy_with_stride(o, i, num_blocks, out_blk, in_blk, int(add))


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, str(ROOT))
    import ffk_import

    ffk_import.load()
    from flexflow_train_b200.parallel_ops import ParallelOps

    P = ParallelOps(local=OracleLocalOps())
    ok = {}
    g = torch.Generator().manual_seed(1234)
    full = torch.rand(4, 6, 5, generator=g)  # identical on every rank (same seed)

    # Replicate: forward broadcast is bit-exact; backward sums the replica gradients
    x = full.clone() if rank == 0 else torch.zeros_like(full)
    P.replicate_forward(x, src=0)
    ok["replicate_fwd"] = torch.equal(x, full)
    dy = full * (rank + 1)
    dx = torch.ones_like(full)
    P.replicate_backward(dx, dy.clone())
    ok["replicate_bwd"] = torch.allclose(dx, 1 + full * sum(r + 1 for r in range(world)))

    # Reduction: forward sums partials; backward copies
    out = torch.zeros_like(full)
    P.reduction_forward(full * (rank + 1), out)
    ok["reduction_fwd"] = torch.allclose(out, full * 3)
    gi = torch.zeros_like(full)
    P.reduction_backward(gi, full)
    ok["reduction_bwd"] = torch.equal(gi, full)

    for dim in (0, 1):
        slab_shape = P.slab_shape(full.shape, dim)
        slab = torch.zeros(slab_shape)
        P.repartition_forward(full, slab, dim)
        expect = torch.chunk(full, world, dim=dim)[rank]
        ok[f"repartition_fwd_{dim}"] = torch.equal(slab, expect)
        # Combine forward is the inverse: gathering the slabs restores the tensor bit for bit
        back = torch.zeros_like(full)
        P.combine_forward(slab.contiguous(), back, dim)
        ok[f"combine_fwd_{dim}"] = torch.equal(back, full)
        # backward forms accumulate
        gfull = torch.ones_like(full)
        P.repartition_backward(gfull, slab * (rank + 1), dim)
        parts = [c * (r + 1) for r, c in enumerate(torch.chunk(full, world, dim=dim))]
        ok[f"repartition_bwd_{dim}"] = torch.allclose(gfull, 1 + torch.cat(parts, dim=dim))
        gslab = torch.ones(slab_shape)
        P.combine_backward(full, gslab, dim)
        ok[f"combine_bwd_{dim}"] = torch.allclose(gslab, 1 + expect)

    flat = torch.full((10,), float(rank + 1))
    P.allreduce_gradients([flat])
    ok["allreduce_no_division"] = torch.equal(flat, torch.full((10,), 3.0))
    results[rank] = ok
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_parallel_ops_world_size_2_gloo():
    world = 2
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), results), nprocs=world, join=True)
    assert len(results) == world
    for rank, ok in results.items():
        bad = [k for k, v in ok.items() if not v]
        assert not bad, f"rank {rank}: {bad}"


def test_gpu_local_ops_fail_loudly_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("checks the no-GPU behaviour")
    sys.path.insert(0, str(ROOT))
    import ffk_import

    ffk_import.load()
    from flexflow_train_b200.parallel_ops import GpuLocalOps

    ops = GpuLocalOps()
    a, b = torch.zeros(4), torch.ones(4)
    with pytest.raises((ValueError, RuntimeError)):
        ops.copy(a, b)  # CPU tensors / no driver are refused: there is no CPU fallback
