"""Worker of tests/test_dlrm_sharded_gpu.py: one rank of a table-sharded DLRM (toy size); every
rank also trains the unsharded model on the global batch and compares its rows and its tables."""
import importlib
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parent))

import torch
import torch.distributed as dist

import ffk_import
from helpers import assert_close


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ffk_import.load()
    T = importlib.import_module("flexflow_train_b200.training")
    P = importlib.import_module("flexflow_train_b200.parallel_ops")
    B = 128
    cfg = dict(embedding_size=(1000,) * 4, sparse_feature_size=16, bag=2, mlp_bot=(4, 32, 16), mlp_top=(32, 16, 2))
    g = torch.Generator(device="cuda").manual_seed(2)
    dense = torch.rand(world * B, 4, device="cuda", generator=g)
    target = (torch.rand(world * B, 2, device="cuda", generator=g) > 0.5).float()

    full = T.DLRM(batch=world * B, **cfg)
    part = T.DLRM(batch=B, rank=rank, world=world, **cfg)
    full.dense.data.copy_(dense)
    part.dense.data.copy_(dense[rank * B:(rank + 1) * B])
    opt_f, opt_p = T.SGD(lr=0.5, weight_decay=0.0), T.SGD(lr=0.5, weight_decay=0.0)
    pops = P.ParallelOps()
    for _ in range(3):
        T.dlrm_step(full, target, opt_f)
        T.dlrm_step(part, target[rank * B:(rank + 1) * B].contiguous(), opt_p, pops)
    torch.cuda.synchronize()
    ok = True
    try:
        assert_close(part.model.output.data, full.model.output.data[rank * B:(rank + 1) * B].cpu().numpy(),
                     rtol=1e-4, what="outputs of this rank's rows")
        tl = len(part.emb.tables)
        for j in range(tl):
            assert_close(part.emb.tables[j], full.emb.tables[rank * tl + j].cpu().numpy(), rtol=1e-4,
                         what=f"table {rank * tl + j}")
        lf = [l for l in full.model.layers if isinstance(l, T.Linear)][-1]
        lp = [l for l in part.model.layers if isinstance(l, T.Linear)][-1]
        assert_close(lp.w, lf.w.cpu().numpy(), rtol=1e-4, what="top MLP weight")
    except AssertionError as e:
        print(f"rank {rank}: {e}", flush=True)
        ok = False
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
