#!/usr/bin/env python
"""Launches the memory-bound kernels of the per-op table a few times each on rotating buffers
(for `ncu --set full -k regex:...` captures; see profiles/). Not a benchmark."""
import importlib
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import ffk_import  # noqa: E402

ffk_import.load()
K = importlib.import_module("flexflow_train_b200.kernels")
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(1)
r = lambda *s: torch.rand(*s, device=dev, generator=g) * 2 - 1
bf = torch.bfloat16
REPS = 3

# LayerNorm fwd / bwd fp32 8192 x 4096
M, N = 8192, 4096
xs = [r(M, N) for _ in range(3)]
ys = [torch.empty(M, N, device=dev) for _ in range(3)]
gamma, beta = r(N) + 1, r(N)
ln = K.LayerNorm.init_kernel(True, M, N, 1e-5)
dg, db = torch.zeros(N, device=dev), torch.zeros(N, device=dev)
for i in range(REPS):
    K.LayerNorm.forward_kernel(None, ln, xs[i % 3], ys[i % 3], gamma, beta)
for i in range(REPS):
    K.LayerNorm.backward_kernel(None, ln, ys[i % 3], xs[i % 3], ys[(i + 1) % 3], gamma, dg, db, accumulate_input_grad=False)
# Softmax fwd
for i in range(REPS):
    K.Softmax.forward_kernel(None, xs[i % 3], ys[i % 3], M, N, 1)
del xs, ys
# LayerNorm fwd bf16 32768 x 1024 and Dropout fwd bf16 32768 x 4096
T_, D = 32768, 1024
xb = [r(T_, D).to(bf) for _ in range(4)]
yb = [torch.empty(T_, D, dtype=bf, device=dev) for _ in range(4)]
gb, bb = (r(D) + 1).to(bf), r(D).to(bf)
lnb = K.LayerNorm.init_kernel(True, T_, D, 1e-5)
for i in range(REPS):
    K.LayerNorm.forward_kernel(None, lnb, xb[i % 4], yb[i % 4], gb, bb)
xd = [r(T_, 4096).to(bf) for _ in range(2)]
yd = [torch.empty(T_, 4096, dtype=bf, device=dev) for _ in range(2)]
dp = K.Dropout.init_kernel(0.1, 1, T_ * 4096)
for i in range(REPS):
    K.Dropout.forward_kernel(None, dp, xd[i % 2], yd[i % 2])
del xd, yd
# Embedding fwd / bwd 1M x 256 fp32
E, ED, EB = 1_000_000, 256, 65536
table = r(E, ED)
idx = [torch.randint(0, E, (EB,), device=dev, generator=g, dtype=torch.int64) for _ in range(3)]
out = [torch.empty(EB, ED, device=dev) for _ in range(3)]
for i in range(REPS):
    K.Embedding.forward_kernel(None, idx[i % 3], out[i % 3], table, K.AGGR_NONE, 1, ED, EB)
# the fused feed-forward GEMM of the encoder: dense + ReLU + dropout, and a plain dense of the same shape
x = r(T_, D).to(bf)
w = (r(4096, D) / 32).to(bf)
b = r(4096).to(bf)
y = torch.empty(T_, 4096, dtype=bf, device=dev)
m = K.Linear.init_kernel(K.ACT_RELU, 0.0, True, K.GEMM_TENSOR, out_dim=4096)
dpo = K.Dropout.init_kernel(0.1, 2, T_ * 4096)
for i in range(REPS):
    K.Linear.forward_kernel(None, m, x, y, w, b, D, 4096, T_)
for i in range(REPS):
    K.FusedLinear.forward_kernel(None, m, x, y, w, b, D, 4096, T_, dropout=dpo)
torch.cuda.synchronize()
print("probe: ok")
