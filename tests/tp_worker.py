"""Worker of tests/test_tensor_parallel_gpu.py: one rank of a tensor-parallel BERT (toy size). Every
rank also trains the unsharded model on the same data and compares. Exits 0 when they agree."""
import importlib
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parent))

import numpy as np
import torch
import torch.distributed as dist

import ffk_import
from helpers import assert_close_l2


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    K = importlib.import_module("flexflow_train_b200.kernels")
    K.set_gemm_workspace(32 << 20)
    cfg = dict(num_layers=2, hidden=64, heads=4, dim_ff=128, seq=16, batch=4, vocab=203, dropout=0.0)
    g = torch.Generator(device="cuda").manual_seed(1)
    x = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
    labels = torch.randint(0, 203, (64,), device="cuda", generator=g, dtype=torch.int32)

    def train(model, steps=3):
        opt = T.SGD(lr=0.2, weight_decay=0.0)
        for _ in range(steps):
            model.input.data.copy_(x)
            T.sparse_cce_step(model, labels, opt)
        torch.cuda.synchronize()
        return model

    full = train(T.build_bert(**cfg))
    part = train(T.build_bert(**cfg, tp_rank=rank, tp_world=world))
    ok = True
    # the same model built from a parallel computation graph in the reference's v1 file format:
    # Replicate / Reduction around attention and feed-forward, a vocabulary-parallel head whose
    # logits are Combine'd along the feature dimension (vocabulary padded to divide by the degree)
    S = importlib.import_module("flexflow_train_b200.strategy")
    X = importlib.import_module("flexflow_train_b200.pcg_export")
    vocab = 208
    pcfg = dict(cfg, vocab=vocab)
    labels_p = labels % vocab
    doc = X.bert_pcg(num_layers=2, hidden=64, heads=4, dim_ff=128, seq=16, batch=4, vocab=vocab, dropout=0.0,
                     tensor_parallel=world, data_parallel=1)
    st = S.parse_strategy(doc, world_size=world)

    def train_p(model, steps=3):
        opt = T.SGD(lr=0.2, weight_decay=0.0)
        for _ in range(steps):
            model.input.data.copy_(x)
            T.sparse_cce_step(model, labels_p, opt)
        torch.cuda.synchronize()
        return model

    full_p = train_p(T.build_bert(**pcfg))
    from_pcg = train_p(T.build_from_strategy(st, batch=4, tp_rank=rank, fuse=True))
    try:
        assert st.tensor_parallel == world
        kinds = [type(l).__name__ for l in from_pcg.layers]
        assert "CombineGather" in kinds and "ReduceSum" in kinds and "ReplicateIn" in kinds, kinds
        assert_close_l2(from_pcg.output.data, full_p.output.data.float().cpu().numpy(), what="probabilities (PCG build)")
        head_full = [l for l in full_p.layers if isinstance(l, T.Linear)][-1]
        head_part = [l for l in from_pcg.layers if isinstance(l, T.Linear)][-1]
        n = head_part.out_dim
        assert n == vocab // world
        assert_close_l2(head_part.w, head_full.w[rank * n:(rank + 1) * n].float().cpu().numpy(), what="vocabulary shard")
    except AssertionError as e:
        print(f"rank {rank}: {e}", flush=True)
        ok = False
    try:
        assert_close_l2(part.output.data, full.output.data.float().cpu().numpy(), what="probabilities")
        ff1_full = [l for l in full.layers if isinstance(l, T.Linear)][0]
        ff1_part = [l for l in part.layers if isinstance(l, T.Linear)][0]
        n = ff1_part.out_dim
        assert_close_l2(ff1_part.w, ff1_full.w[rank * n:(rank + 1) * n].float().cpu().numpy(), what="ff1 shard")
        head_full = [l for l in full.layers if isinstance(l, T.Linear)][-1]
        head_part = [l for l in part.layers if isinstance(l, T.Linear)][-1]
        assert_close_l2(head_part.w, head_full.w.float().cpu().numpy(), what="replicated head")
    except AssertionError as e:
        print(f"rank {rank}: {e}", flush=True)
        ok = False
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
