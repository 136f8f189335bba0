"""Probe (torchrun): peer-memory all-reduce vs NCCL for the two gradient buckets of bench.py."""
import importlib, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch, torch.distributed as dist
import ffk_import
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
ffk_import.load()
P = importlib.import_module("flexflow_train_b200.parallel_ops")
def timed(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
for dtype, count in ((torch.float32, 8192), (torch.bfloat16, 16 * 1024 * 1024 + 4096)):
    es = torch.empty(0, dtype=dtype).element_size()
    nb = (count * es + 255) // 256 * 256
    g = P.PeerMemoryGroup(2 * nb)
    x = g.tensor(0, (count,), dtype); x.fill_(1)
    ref = torch.ones(count, dtype=dtype, device="cuda")
    t_p2p = timed(lambda: g.all_reduce_sum(0, nb, count, dtype))
    g.check()
    t_nccl = timed(lambda: dist.all_reduce(ref.fill_(1)))
    t_fill = timed(lambda: ref.fill_(1))
    if rank == 0:
        print(f"{count * es / 1e6:8.3f} MB {dtype}: p2p {t_p2p:7.1f} us | nccl {t_nccl - t_fill:7.1f} us (fill {t_fill:.1f} subtracted) | world {world}", flush=True)
    assert torch.all(g.tensor(nb, (count,), dtype) == world)
    g.close()
dist.destroy_process_group()
