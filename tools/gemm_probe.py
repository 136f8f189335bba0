"""Developer probe: checks the tcgen05 GEMM against torch on every operand layout and
times the headline shapes. Not part of the test-suite (tests/ compares with the oracle)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import ffk_import
ffk = ffk_import.load()
L = ffk.lib()
DT = {torch.float32: 4, torch.bfloat16: 6, torch.float16: 3}

def gemm(a, b, M, N, K, batch, a_mn, b_mn, out_dtype, path, accumulate=False, bias=None, act=-1, d=None):
    if d is None:
        d = torch.zeros(batch, M, N, dtype=out_dtype, device="cuda")
    lda = a.shape[-1]; ldb = b.shape[-1]
    L.call("ffk_gemm", 0, a.data_ptr(), b.data_ptr(), d.data_ptr(), M, N, K, batch, lda, ldb, N,
           a.shape[-2] * lda, b.shape[-2] * ldb, M * N, DT[a.dtype], DT[out_dtype], int(a_mn), int(b_mn),
           1.0, int(accumulate), bias.data_ptr() if bias is not None else None, act, path)
    return d

def check(M, N, K, batch, a_mn, b_mn, dt, path=1):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    A = (torch.rand(batch, M, K, device="cuda", generator=g) * 2 - 1).to(dt)
    B = (torch.rand(batch, N, K, device="cuda", generator=g) * 2 - 1).to(dt)
    a = A.transpose(1, 2).contiguous() if a_mn else A
    b = B.transpose(1, 2).contiguous() if b_mn else B
    out_dt = torch.float32
    d = gemm(a, b, M, N, K, batch, a_mn, b_mn, out_dt, path)
    torch.cuda.synchronize()
    ref = torch.matmul(A.double(), B.double().transpose(1, 2))
    err = (d.double() - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = (2e-2 if dt == torch.float32 and path == 1 else 1e-3) * max(scale, 1.0)
    ok = err <= tol
    print(f"{'OK ' if ok else 'BAD'} M={M} N={N} K={K} b={batch} a_mn={int(a_mn)} b_mn={int(b_mn)} {dt} path={path} err={err:.3e} scale={scale:.2e}", flush=True)
    return ok

def bench_epilogues(M=8192, N=4096, K=4096, dt=torch.bfloat16, iters=20):
    A = (torch.rand(1, M, K, device="cuda") - 0.5).to(dt)
    B = (torch.rand(1, N, K, device="cuda") - 0.5).to(dt)
    bias = torch.rand(N, device="cuda").to(dt)
    d = torch.zeros(1, M, N, dtype=dt, device="cuda")
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    for name, kw in (("none", {}), ("bias", dict(bias=bias)), ("relu", dict(act=0)), ("bias+relu", dict(bias=bias, act=0)),
                     ("bias+gelu", dict(bias=bias, act=3)), ("accumulate", dict(accumulate=True))):
        for _ in range(3):
            gemm(A, B, M, N, K, 1, False, False, dt, 1, d=d, **kw)
        torch.cuda.synchronize(); e0.record()
        for _ in range(iters):
            gemm(A, B, M, N, K, 1, False, False, dt, 1, d=d, **kw)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        print(f"epilogue {name:12s}: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)


def bench(M, N, K, a_mn, b_mn, dt=torch.bfloat16, iters=20):
    A = (torch.rand(1, K if a_mn else M, M if a_mn else K, device="cuda") - 0.5).to(dt)
    B = (torch.rand(1, K if b_mn else N, N if b_mn else K, device="cuda") - 0.5).to(dt)
    d = torch.zeros(1, M, N, dtype=dt, device="cuda")
    for _ in range(3):
        gemm(A, B, M, N, K, 1, a_mn, b_mn, dt, 1, d=d)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters):
        gemm(A, B, M, N, K, 1, a_mn, b_mn, dt, 1, d=d)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"bench M={M} N={N} K={K} a_mn={int(a_mn)} b_mn={int(b_mn)} {dt}: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
    # torch/cuBLAS for context
    At = A[0].t() if a_mn else A[0]; Bt = B[0].t() if b_mn else B[0]
    for _ in range(3): torch.matmul(At, Bt.t())
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): torch.matmul(At, Bt.t())
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"   cuBLAS same shape: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)

def timed(fn, iters=10):
    for _ in range(3): fn()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench_encoder_shapes(dt=torch.bfloat16):
    """The GEMM shapes of one encoder layer (BASELINE configs[2]) with their epilogues."""
    r = lambda *s: (torch.rand(*s, device="cuda") - 0.5).to(dt)
    for (name, M, N, K, batch, a_mn, b_mn, kw) in (
            ("proj  NT plain       ", 32768, 4096, 1024, 1, False, False, {}),
            ("ffn1  NT bias+relu   ", 32768, 4096, 1024, 1, False, False, dict(bias=True, act=0)),
            ("ffn2  NT bias        ", 32768, 1024, 4096, 1, False, False, dict(bias=True)),
            ("dgrad NN accumulate  ", 32768, 4096, 1024, 1, False, True, dict(accumulate=True)),
            ("dgrad NN accumulate k", 32768, 1024, 4096, 1, False, True, dict(accumulate=True)),
            ("wgrad TN accumulate  ", 4096, 1024, 32768, 1, True, True, dict(accumulate=True)),
            ("S=QK^T batched NT    ", 512, 512, 256, 1024, False, False, {}),
            ("O=PV   batched NN    ", 512, 256, 512, 1024, False, True, {}),
            ("dV=P^T dO batched TN ", 512, 256, 512, 1024, True, True, {})):
        A = r(batch, K if a_mn else M, M if a_mn else K)
        B = r(batch, K if b_mn else N, N if b_mn else K)
        d = torch.zeros(batch, M, N, dtype=dt, device="cuda")
        kw = dict(kw)
        if kw.get("bias"):
            kw["bias"] = r(N)
        ms = timed(lambda: gemm(A, B, M, N, K, batch, a_mn, b_mn, dt, 1, d=d, **kw))
        At = A.transpose(1, 2) if a_mn else A
        Bt = B if b_mn else B.transpose(1, 2)
        ms_ref = timed(lambda: torch.bmm(At, Bt))
        fl = 2.0 * M * N * K * batch
        print(f"{name}: {ms*1e3:7.1f} us {fl/ms/1e9:7.1f} TF/s | cuBLAS plain {ms_ref*1e3:7.1f} us {fl/ms_ref/1e9:7.1f} TF/s", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "enc":
        bench_encoder_shapes()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "fp32":
        ws = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
        L.call("ffk_set_gemm_workspace", ws.data_ptr(), ws.numel())
        M = N = Kd = 4096
        A = torch.rand(1, M, Kd, device="cuda") - 0.5
        B = torch.rand(1, N, Kd, device="cuda") - 0.5
        d = torch.zeros(1, M, N, device="cuda")
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        for path, name in ((3, "split-bf16 tensor"), (2, "CUDA-core fp32"), (1, "tf32 tensor")):
            for _ in range(2): gemm(A, B, M, N, Kd, 1, False, False, torch.float32, path, d=d)
            torch.cuda.synchronize(); e0.record()
            for _ in range(5): gemm(A, B, M, N, Kd, 1, False, False, torch.float32, path, d=d)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            ref = torch.matmul(A[0].double(), B[0].double().t())
            err = ((d[0].double() - ref).abs().max() / ref.abs().max()).item()
            print(f"fp32 4096^3 {name:18s}: {ms*1e3:8.1f} us  {2*M*N*Kd/ms/1e9:7.1f} TFLOP/s  max err/scale {err:.2e}", flush=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "bench":
        bench(8192, 4096, 4096, False, False, iters=5)
        sys.exit(0)
    ok = True
    for dt in (torch.bfloat16, torch.float16, torch.float32):
        for a_mn in (False, True):
            for b_mn in (False, True):
                if dt == torch.float32 and (a_mn or b_mn):
                    continue  # TF32 tiles are K-major only
                ok &= check(128, 256, 64, 1, a_mn, b_mn, dt)
                ok &= check(256, 512, 256, 1, a_mn, b_mn, dt)
                ok &= check(200, 136, 104, 3, a_mn, b_mn, dt)
                ok &= check(1024, 1024, 1024, 1, a_mn, b_mn, dt)
    ok &= check(100, 10, 784, 1, False, False, torch.float32, path=2)
    ok &= check(64, 512, 10, 1, False, True, torch.float32, path=2)
    ok &= check(77, 33, 19, 2, True, True, torch.bfloat16, path=2)
    print("ALL OK" if ok else "SOME BAD")
    if ok:
        bench_epilogues()
        bench(8192, 4096, 4096, False, False)
        bench(8192, 4096, 4096, False, True)
        bench(4096, 4096, 8192, True, True)
