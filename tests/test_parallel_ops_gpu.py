"""GPU, 2 / 4 / 8 ranks over NCCL (each skipped when the box has fewer GPUs): the multi-GPU
parallel operators with the real local kernels, checked against the single-process statement of
the same math — bit-exact for Replicate forward and every Repartition / Combine direction along
every dimension (the reference's exact-equality tests: lib/kernels/test/src/
test_partition_kernel.cc:30-58, test_replicate_kernel.cc:28-34, test_combine_kernel.cc)."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.gpu


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sys.path.insert(0, str(ROOT))
    import ffk_import

    ffk_import.load()
    from flexflow_train_b200.parallel_ops import ParallelOps

    P = ParallelOps()
    ok = {}
    g = torch.Generator(device="cuda").manual_seed(1234)
    tri = world * (world + 1) // 2
    for dtype in (torch.float32, torch.bfloat16):
        full = torch.rand(8, 16, 64, generator=g, device="cuda").to(dtype)
        x = full.clone() if rank == 0 else torch.zeros_like(full)
        P.replicate_forward(x)
        ok[f"replicate_fwd_{dtype}"] = torch.equal(x, full)
        dx = torch.ones_like(full)
        P.replicate_backward(dx, (full * (rank + 1)).contiguous())
        ok[f"replicate_bwd_{dtype}"] = torch.allclose(dx.float(), 1 + full.float() * tri, rtol=2e-2)
        out = torch.zeros_like(full)
        P.reduction_forward((full * (rank + 1)).contiguous(), out)
        ok[f"reduction_fwd_{dtype}"] = torch.allclose(out.float(), full.float() * tri, rtol=2e-2)
        for dim in (0, 1, 2):
            slab = torch.zeros(P.slab_shape(full.shape, dim), dtype=dtype, device="cuda")
            P.repartition_forward(full, slab, dim)
            expect = torch.chunk(full, world, dim=dim)[rank].contiguous()
            ok[f"repartition_fwd_{dim}_{dtype}"] = torch.equal(slab, expect)
            back = torch.zeros_like(full)
            P.combine_forward(slab, back, dim)
            ok[f"combine_fwd_{dim}_{dtype}"] = torch.equal(back, full)
            gfull = torch.ones_like(full)
            P.repartition_backward(gfull, slab, dim)
            ok[f"repartition_bwd_{dim}_{dtype}"] = torch.equal(gfull, 1 + full)
            gslab = torch.ones_like(slab)
            P.combine_backward(full, gslab, dim)
            ok[f"combine_bwd_{dim}_{dtype}"] = torch.equal(gslab, 1 + expect)
    torch.cuda.synchronize()
    results[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_parallel_ops_nccl(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run with gpurun --gpus {world})")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
    assert len(results) == world
    for rank, ok in results.items():
        bad = [k for k, v in ok.items() if not v]
        assert not bad, f"rank {rank}: {bad}"


def _nccl_abi_worker(rank, world, id_path, results):
    """The NCCL entry points of the C ABI themselves (no torch.distributed): rank 0 creates the
    unique id and publishes it through a file, every rank builds a communicator and runs the
    nccl-SGD update; the result must equal the single-process update on the summed gradient."""
    import ctypes
    import time as _t

    torch.cuda.set_device(rank)
    sys.path.insert(0, str(ROOT))
    import ffk_import

    ffk = ffk_import.load()
    L = ffk.lib()
    idbuf = (ctypes.c_char * 128)()
    if rank == 0:
        L.call("ffk_nccl_get_unique_id", ctypes.cast(idbuf, ctypes.c_void_p))
        Path(id_path + ".tmp").write_bytes(bytes(idbuf))
        os.replace(id_path + ".tmp", id_path)
    else:
        while not os.path.exists(id_path):
            _t.sleep(0.05)
        ctypes.memmove(idbuf, Path(id_path).read_bytes(), 128)
    comm = ctypes.c_void_p()
    L.call("ffk_nccl_comm_create", ctypes.byref(comm), ctypes.cast(idbuf, ctypes.c_void_p), world, rank)
    n = 100_003
    g = torch.Generator(device="cuda").manual_seed(7)
    w0 = torch.rand(n, device="cuda", generator=g)
    base = torch.rand(n, device="cuda", generator=g)
    grad = base * (rank + 1)
    w = w0.clone()
    v = torch.zeros(n, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    L.call("ffk_sgd_nccl_update", stream, comm, 0.1, 0.9, 0, 1e-4, grad.data_ptr(), n, w.data_ptr(),
           v.data_ptr(), 4, None)
    torch.cuda.synchronize()
    gsum = base * sum(r + 1 for r in range(world))       # SUM over ranks, not the mean
    gt = gsum + 1e-4 * w0
    expect = w0 - 0.1 * gt                                  # first step: V = g
    results[rank] = bool(torch.allclose(w, expect, rtol=1e-5, atol=1e-6) and torch.allclose(grad, gsum, rtol=1e-6))
    L.call("ffk_nccl_comm_destroy", comm)


def test_nccl_c_abi_two_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_nccl_abi_worker, args=(2, str(tmp_path / "nccl_id"), results), nprocs=2, join=True)
    assert dict(results) == {0: True, 1: True}
