"""Developer probe: the memory-bound row kernels at the encoder's bf16 shapes (BASELINE configs[2]),
timed with CUDA events and reported against the bytes they must move."""
import importlib, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import ffk_import
ffk_import.load()
K = importlib.import_module("flexflow_train_b200.kernels")
dt = torch.bfloat16
r = lambda *s: (torch.rand(*s, device="cuda") - 0.5).to(dt)

def timed(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3

def report(name, us, nbytes):
    print(f"{name:34s} {us:8.1f} us  {nbytes/us/1e6:6.2f} TB/s", flush=True)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
M, N = 32768, 1024
x, y, dy, dx = r(M, N), r(M, N), r(M, N), r(M, N)
gamma, beta = r(N), r(N)
dgamma, dbeta = torch.zeros_like(gamma), torch.zeros_like(beta)
ln = K.LayerNorm.init_kernel(True, M, N, 1e-5)
B = M * N * 2
if which in ("all", "ln"):
    report("layer_norm fwd 32768x1024", timed(lambda: K.LayerNorm.forward_kernel(None, ln, x, y, gamma, beta)), 2 * B)
    report("layer_norm bwd 32768x1024", timed(lambda: K.LayerNorm.backward_kernel(None, ln, dy, x, dx, gamma, dgamma, dbeta, True)), 4 * B)
if which == "lnsweep":
    for dtt in (torch.bfloat16, torch.float32):
        for NN in (128, 256, 512, 768, 1024):
            MM = (32 << 20) // NN
            xs = (torch.rand(MM, NN, device="cuda") - 0.5).to(dtt)
            dys, dxs, ys = xs.clone(), xs.clone(), xs.clone()
            gm = torch.ones(NN, device="cuda", dtype=dtt)
            dg, dbb = torch.zeros_like(gm), torch.zeros_like(gm)
            st = K.LayerNorm.init_kernel(True, MM, NN, 1e-5)
            eb = MM * NN * xs.element_size()
            report(f"ln fwd {str(dtt)[6:]} {MM}x{NN}", timed(lambda: K.LayerNorm.forward_kernel(None, st, xs, ys, gm, gm)), 2 * eb)
            report(f"ln bwd {str(dtt)[6:]} {MM}x{NN}", timed(lambda: K.LayerNorm.backward_kernel(None, st, dys, xs, dxs, gm, dg, dbb, True)), 4 * eb)
    sys.exit(0)
if which in ("all", "softmax"):
    R, C = 64 * 16 * 512, 512
    p, q, g = r(R, C), r(R, C), r(R, C)
    report("softmax fwd 524288x512 in place", timed(lambda: K.Softmax.forward_kernel(None, p, p, R, C, 1)), 2 * R * C * 2)
    report("softmax fwd 524288x512", timed(lambda: K.Softmax.forward_kernel(None, p, q, R, C, 1)), 2 * R * C * 2)
    report("softmax jacobian 524288x512", timed(lambda: K.Softmax.backward_kernel(None, g, p, q, R, C, 1, mode=1)), 3 * R * C * 2)
    del p, q, g
if which in ("all", "dropout"):
    big, bo = r(M, 4096), r(M, 4096)
    d = K.Dropout.init_kernel(0.1, 7, M * 4096)
    report("dropout fwd 32768x4096", timed(lambda: K.Dropout.forward_kernel(None, d, big, bo)), 2 * M * 4096 * 2 + M * 4096 // 8)
    report("dropout bwd 32768x4096", timed(lambda: K.Dropout.backward_kernel(None, d, bo, big)), 2 * M * 4096 * 2 + M * 4096 // 8)
    d2 = K.Dropout.init_kernel(0.1, 7, M * N)
    report("dropout fwd 32768x1024", timed(lambda: K.Dropout.forward_kernel(None, d2, x, y)), 2 * B + M * N // 8)
    del big, bo
if which in ("all", "add"):
    OT = K.OperatorType
    ga, gb = torch.zeros_like(x), torch.zeros_like(x)
    report("add fwd 32768x1024", timed(lambda: K.ElementBinary.forward_kernel(None, OT.EW_ADD, x, y, dx)), 3 * B)
    report("add bwd 32768x1024", timed(lambda: K.ElementBinary.backward_kernel(None, OT.EW_ADD, dy, x, y, ga, gb)), 5 * B)
if which in ("all", "bias"):
    for out_dim in (1024, 4096):
        xx, w, b = r(M, 64), r(out_dim, 64), r(out_dim)
        dyy, dxx, dw, db = r(M, out_dim), torch.zeros(M, 64, dtype=dt, device="cuda"), torch.zeros_like(w), torch.zeros_like(b)
        m = K.Linear.init_kernel(K.ACT_NONE, 0.0, True, K.GEMM_AUTO, out_dim=out_dim)
        yy = torch.zeros(M, out_dim, dtype=dt, device="cuda")
        t_all = timed(lambda: K.Linear.backward_kernel(None, m, xx, dxx, yy, dyy, w, dw, db, 64, out_dim, M))
        t_nob = timed(lambda: K.Linear.backward_kernel(None, m, xx, dxx, yy, dyy, w, dw, None, 64, out_dim, M))
        report(f"bias grad column sum 32768x{out_dim}", t_all - t_nob, M * out_dim * 2)
if which in ("all", "metrics"):
    perf = K.PerfMetrics()
    a32, b32 = torch.rand(M, N, device="cuda"), torch.rand(M, N, device="cuda")
    report("metrics (mse) 32768x1024 fp32", timed(lambda: K.update_metrics_label_kernel(None, K.METRIC_MSE, a32, b32, M, N, perf)), 2 * M * N * 4)
