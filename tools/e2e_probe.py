"""Probe: H2D copy duration while different kernels run concurrently."""
import importlib, sys
sys.path.insert(0, ".")
import torch
import bench, ffk_import
ffk_import.load()
Kmod = importlib.import_module("flexflow_train_b200.kernels")
dev = torch.device("cuda", 0)
wl = bench.Workload(torch, Kmod, dev)
hx = torch.empty(bench.M, bench.KDIM, dtype=torch.bfloat16, pin_memory=True)
x2 = torch.empty_like(wl.x)
cs = torch.cuda.Stream()
main = torch.cuda.current_stream()

def gemms():
    for _ in range(6):
        Kmod.Linear.forward_kernel(None, wl.lin, wl.x, wl.y, wl.w, wl.b, bench.KDIM, bench.N, bench.M)
def streams():
    for _ in range(8):
        Kmod.ElementBinary.forward_kernel(None, Kmod.OperatorType.EW_ADD, wl.a, wl.c, wl.o)
        Kmod.Softmax.forward_kernel(None, wl.a, wl.o, bench.ROWS, bench.COLS, 1)
def embed():
    for _ in range(20):
        Kmod.Embedding.forward_kernel(None, wl.idx, wl.eout, wl.table, Kmod.AGGR_NONE, 1, bench.EDIM, bench.EBATCH)
        Kmod.Embedding.backward_kernel(None, wl.idx, wl.eout, wl.dtable, Kmod.AGGR_NONE, 1, bench.EDIM, bench.EBATCH)
def lnb():
    for _ in range(8):
        Kmod.LayerNorm.backward_kernel(None, wl.ln, wl.c, wl.a, wl.ga, wl.gamma, wl.dgamma, wl.dbeta)
def nothing():
    pass

for name, fn in (("idle", nothing), ("gemms", gemms), ("streaming", streams), ("embedding", embed), ("ln bwd", lnb), ("whole step", wl.step), ("idle", nothing)):
    ts = []
    for rep in range(6):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn()                      # keep the GPU busy on the main stream
        with torch.cuda.stream(cs):
            a.record(cs)
            x2.copy_(hx, non_blocking=True)
            b.record(cs)
        fn()
        torch.cuda.synchronize()
        ts.append(round(a.elapsed_time(b), 3))
    print(f"{name:12s} copy ms: {ts}")
