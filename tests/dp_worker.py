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

    a, launches = train("peer")
    b, _ = train("nccl")
    ok = launches >= 6          # more than one slice per step went out
    if not ok:
        print(f"rank {rank}: only {launches} background launches", flush=True)
    for i, (p, q) in enumerate(zip(a, b)):
        pf, qf = p.float(), q.float()
        if not torch.isfinite(pf).all():
            print(f"rank {rank}: parameter {i} is not finite", flush=True)
            ok = False
        # NCCL rounds every partial sum of its ring to bf16, the peer-memory kernel sums in fp32 and
        # rounds once: after three steps the weights agree to a bf16 ulp of the tensor's scale
        tol = torch.maximum(pf.abs(), qf.abs()) * 2.0 ** -6 + qf.abs().max() * 2.0 ** -6 + 1e-6
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
