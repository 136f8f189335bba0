"""Worker of tests/test_p2p_gpu.py: one rank of a peer-memory group. Exits 0 when every check
passed. Launched with RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT in the environment."""
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
    dist.init_process_group("gloo")  # the handle exchange only; no NCCL anywhere in this test
    ffk_import.load()
    P = importlib.import_module("flexflow_train_b200.parallel_ops")
    big = 6 << 20                       # elements of the two-shot case
    region = 2 * (big * 4) + (1 << 20)
    g = P.PeerMemoryGroup(region)
    ok = True
    # push protocol (<= 64 KB), one-shot reads (<= 512 KB), two-shot above
    for dtype, count in ((torch.float32, 1000 * 4), (torch.bfloat16, 24 * 8), (torch.float32, 30000 * 4), (torch.float32, big),
                         (torch.bfloat16, big), (torch.float16, 70000 * 8)):
        es = torch.empty(0, dtype=dtype).element_size()
        nb = count * es
        in_off, out_off = 0, (nb + 255) // 256 * 256
        src = g.tensor(in_off, (count,), dtype)
        dst = g.tensor(out_off, (count,), dtype)
        for rep in range(3):  # repeated launches exercise the epoch / flag protocol
            gen = torch.Generator(device="cuda").manual_seed(100 * rep + 7)
            base = torch.rand(count, device="cuda", generator=gen) - 0.5   # same on every rank
            src.copy_(((rank + 1) * base).to(dtype))
            dst.fill_(-7)
            torch.cuda.synchronize()
            g.all_reduce_sum(in_off, out_off, count, dtype)
            torch.cuda.synchronize()
            g.check()
            # fp32 sum in rank order of the values each rank stored
            want = torch.zeros(count, device="cuda")
            for r in range(world):
                want += ((r + 1) * base).to(dtype).float()
            want = want.to(dtype)
            if not torch.equal(dst, want):
                bad = (dst.float() - want.float()).abs().max().item()
                print(f"rank {rank}: {dtype} count {count} rep {rep}: max diff {bad}", flush=True)
                ok = False
            dist.barrier()  # keep the ranks' fill / check phases apart from the next launch
    g.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
