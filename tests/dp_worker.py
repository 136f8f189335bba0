"""Worker of tests/test_p2p_gpu.py::test_data_parallel_encoder_peer_sums_match_nccl: one rank of a
data-parallel toy encoder. Trains the same model twice from the same weights — gradient sums by
PeerGradSync (peer memory, issued under backward) and by NCCL all-reduces after backward — and
compares the weights. Exits 0 when every check passed."""
import importlib
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

import torch
import torch.distributed as dist

import ffk_import


def exact_sum_case(T, P, rank, world, B, S, D, H, FF, L):
    """The peer-memory sum itself, without a training step around it: every rank fills its gradient
    buckets with its own data, the slices go out the way a backward pass would send them, and the
    result must equal, BIT FOR BIT, the fp32 sum in rank order of all ranks' buckets rounded once."""
    model = T.build_transformer_encoder(num_layers=L, d_model=D, heads=H, seq=S, batch=B, dim_ff=FF, dropout=0.0,
                                        dtype=torch.bfloat16)
    sync = P.PeerGradSync(model, min_bytes=64 << 10, background_blocks=8 if rank >= 0 else 0)
    ok = True
    for trial, blocks in enumerate((8, 0)):     # the background kernel and the full-size one
        sync.blocks = blocks
        for b in model.grad_buckets:
            gen = torch.Generator(device=b.device).manual_seed(1000 * trial + 17 + rank)
            b.copy_(((torch.rand(b.numel(), device=b.device, generator=gen) - 0.5) * 0.3).to(b.dtype))
        torch.cuda.synchronize()
        dist.barrier()
        for l in reversed(model.layers):
            if l.grads:
                sync.after_layer(l)
        sync.finish()
        torch.cuda.synchronize()
        for b, got in zip(model.grad_buckets, sync.sums):
            every = [torch.empty_like(b) for _ in range(world)]
            dist.all_gather(every, b)
            acc = torch.zeros(b.numel(), dtype=torch.float32, device=b.device)
            for e in every:                      # rank order, fp32
                acc += e.float()
            want = acc.to(b.dtype)
            if not torch.equal(got, want):
                n = (got != want).sum().item()
                print(f"rank {rank}: peer sum (blocks={blocks}) differs from the fp32 rank-order sum in {n} elements", flush=True)
                ok = False
    sync.check()
    dist.barrier()
    sync.close()
    return ok


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    P = importlib.import_module("flexflow_train_b200.parallel_ops")
    B, S, D, H, FF, L = 2, 32, 128, 4, 512, 3
    g = torch.Generator(device="cuda").manual_seed(100 + rank)          # every rank its own batch
    x = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
    target = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)

    def train(kind):
        model = T.build_transformer_encoder(num_layers=L, d_model=D, heads=H, seq=S, batch=B, dim_ff=FF,
                                            dropout=0.0, dtype=torch.bfloat16)
        # small slices so that several background launches go out per backward pass
        sync = P.PeerGradSync(model, min_bytes=64 << 10, background_blocks=8) if kind == "peer" else P.ParallelOps()
        opt = T.SGD(lr=0.05, weight_decay=1e-3)
        for _ in range(3):
            model.input.data.copy_(x)
            T.mse_step(model, target, opt, sync)
        torch.cuda.synchronize()
        launches = getattr(sync, "launches", 0)
        if kind == "peer":
            sync.check()
            dist.barrier()
            sync.close()
        return [p.clone() for p in model.params], launches

    ok = exact_sum_case(T, P, rank, world, B, S, D, H, FF, L)
    p0 = [p.clone().float() for p in T.build_transformer_encoder(num_layers=L, d_model=D, heads=H, seq=S, batch=B,
                                                                 dim_ff=FF, dropout=0.0, dtype=torch.bfloat16).params]
    a, launches = train("peer")
    b, _ = train("nccl")
    if launches < 6:            # more than one slice per step went out
        print(f"rank {rank}: only {launches} background launches", flush=True)
        ok = False
    for i, (p, q) in enumerate(zip(a, b)):
        pf, qf = p.float(), q.float()
        if not torch.isfinite(pf).all():
            print(f"rank {rank}: parameter {i} is not finite", flush=True)
            ok = False
        # NCCL rounds every partial sum of its ring to bf16, the peer-memory kernel sums in fp32 and
        # rounds once (exact_sum_case pins that bit for bit). Where the ranks' gradients cancel, the two
        # sums differ by a sizeable fraction of the RESULT, so the trained weights are compared on the
        # scale of what three steps changed: a bf16 ulp of the weight plus 15 % of the largest update.
        tol = torch.maximum(pf.abs(), qf.abs()) * 2.0 ** -6 + (qf - p0[i]).abs().max() * 0.15 + 1e-7
        if ((pf - qf).abs() > tol).any():
            print(f"rank {rank}: parameter {i} differs from the NCCL run by {(pf - qf).abs().max().item()}", flush=True)
            ok = False
        ref = p.clone()
        dist.broadcast(ref, src=0)
        if not torch.equal(ref, p):
            print(f"rank {rank}: parameter {i} differs from rank 0", flush=True)
            ok = False
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
