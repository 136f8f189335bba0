"""Probe: Linear forward/backward with an out_dim that breaks TMA's pitch rule (BERT's vocabulary),
with and without registered scratch (repack onto the tensor cores vs CUDA-core fallback)."""
import importlib, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import ffk_import
ffk_import.load()
K = importlib.import_module("flexflow_train_b200.kernels")
dt = torch.bfloat16
M, I, O = 8192, 768, 30522
r = lambda *s: (torch.rand(*s, device="cuda") - 0.5).to(dt)
x, w, b = r(M, I), r(O, I) * 0.05, r(O)
y, dy = torch.zeros(M, O, dtype=dt, device="cuda"), r(M, O)
dx, dw, db = torch.zeros_like(x), torch.zeros_like(w), torch.zeros_like(b)
m = K.Linear.init_kernel(K.ACT_NONE, 0.0, True, K.GEMM_AUTO, out_dim=O)
def timed(fn, iters=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
fl = 2.0 * M * I * O
for ws in (None, 1 << 30):
    K.set_gemm_workspace(ws)
    tf = timed(lambda: K.Linear.forward_kernel(None, m, x, y, w, b, I, O, M))
    tb = timed(lambda: K.Linear.backward_kernel(None, m, x, dx, y, dy, w, dw, db, I, O, M))
    print(f"scratch {ws}: fwd {tf:.2f} ms ({fl/tf/1e9:.0f} TF/s)  bwd {tb:.2f} ms ({2*fl/tb/1e9:.0f} TF/s)", flush=True)
K.set_gemm_workspace(None)
