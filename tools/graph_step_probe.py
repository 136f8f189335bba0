"""Probe: the configs[1] microbench step, eager C-ABI calls vs one CUDA-graph replay."""
import importlib, sys
sys.path.insert(0, ".")
import torch
import bench, ffk_import
ffk_import.load()
Kmod = importlib.import_module("flexflow_train_b200.kernels")
wl = bench.Workload(torch, Kmod, torch.device("cuda", 0))
def timed(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("eager ms/step", timed(wl.step))
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        wl.step()
torch.cuda.current_stream().wait_stream(s)
print("graph ms/step", timed(g.replay))
print("eager ms/step", timed(wl.step))
