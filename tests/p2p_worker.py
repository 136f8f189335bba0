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


def timeout_case(g, rank, world):
    """Rank 1 skips the first collective; see test_p2p_missing_rank_poisons_the_group."""
    import time
    count = 40000 * 4      # one-shot protocol
    src = g.tensor(0, (count,), torch.float32)
    dst = g.tensor(count * 4, (count,), torch.float32)
    src.fill_(1.0)
    dst.fill_(-7.0)
    torch.cuda.synchronize()
    dist.barrier()
    ok = True
    if rank != 1:
        g.all_reduce_sum(0, count * 4, count, torch.float32)
        torch.cuda.synchronize()
        ok &= bool(torch.isnan(dst).all())          # no partial sum was published
    else:
        time.sleep(1.5)                              # well past the other ranks' timeout
    dist.barrier()
    # the next collective: everybody takes part, the group is poisoned on every rank
    dst.fill_(-7.0)
    torch.cuda.synchronize()
    g.all_reduce_sum(0, count * 4, count, torch.float32)
    torch.cuda.synchronize()
    ok &= bool(torch.isnan(dst).all())
    try:
        g.check()
        ok = False
        print(f"rank {rank}: status did not report the expired wait", flush=True)
    except Exception:
        pass
    if not ok:
        print(f"rank {rank}: timeout case failed (dst[0] = {dst[0].item()})", flush=True)
    return ok


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
    if len(sys.argv) > 1 and sys.argv[1] == "timeout":
        ok = timeout_case(g, rank, world)
        dist.barrier()
        g.close()
        dist.destroy_process_group()
        sys.exit(0 if ok else 1)
    # push protocol (<= 64 KB), one-shot reads (<= 512 KB), two-shot above
    for dtype, count in ((torch.float32, 1000 * 4), (torch.bfloat16, 24 * 8), (torch.float32, 30000 * 4), (torch.float32, big),
                         (torch.bfloat16, big), (torch.float16, 70000 * 8)):
        es = torch.empty(0, dtype=dtype).element_size()
        nb = count * es
        in_off, out_off = 0, (nb + 255) // 256 * 256
        src = g.tensor(in_off, (count,), dtype)
        dst = g.tensor(out_off, (count,), dtype)
        for rep in range(4):  # repeated launches exercise the epoch / flag protocol; the last one
            # uses the background kernel (128 threads, fits beside a GEMM CTA)
            gen = torch.Generator(device="cuda").manual_seed(100 * rep + 7)
            base = torch.rand(count, device="cuda", generator=gen) - 0.5   # same on every rank
            src.copy_(((rank + 1) * base).to(dtype))
            dst.fill_(-7)
            torch.cuda.synchronize()
            g.all_reduce_sum(in_off, out_off, count, dtype, background_blocks=24 if rep == 3 else 0)
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
