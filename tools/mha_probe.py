"""Developer probe: MultiHeadAttention forward / backward at the encoder's shape (BASELINE configs[2])."""
import importlib, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import ffk_import
ffk_import.load()
K = importlib.import_module("flexflow_train_b200.kernels")
dt = torch.bfloat16
hd = int(sys.argv[1]) if len(sys.argv) > 1 else 256
N, H, E, S = 64, 16, 1024, 512
dims = [N, H, E, E, E, hd, hd, hd, E, S, S]
m = K.MultiHeadAttention.init_kernel(*dims, dtype=dt)
r = lambda *s: (torch.rand(*s, device="cuda") - 0.5).to(dt)
x, out, dout = r(N, S, E), r(N, S, E), r(N, S, E)
w = r(m.weightSize // 2) * 0.05
dw = torch.zeros_like(w)
dq, dk, dv = torch.zeros_like(x), torch.zeros_like(x), torch.zeros_like(x)
def timed(fn, iters=5):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
print(f"mha fwd head dim {hd}: {timed(lambda: K.MultiHeadAttention.forward_kernel(None, m, x, x, x, w, out)):.1f} us")
print(f"mha bwd head dim {hd}: {timed(lambda: K.MultiHeadAttention.backward_kernel(None, m, x, dq, x, dk, x, dv, w, dw, dout)):.1f} us")
