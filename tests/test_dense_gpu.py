"""GPU parity: Linear / BatchMatmul / GEMM through the C ABI vs the CPU oracle.
fp32 tensors take the exact CUDA-core path (rtol 1e-4); bf16/half take the tcgen05 tiles
(rtol 2e-2). Inputs: U(-1,1) activations (seed 1234), weights scaled by 1/sqrt(K) (seed 4321)."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_W, assert_close, to_dev, uniform

pytestmark = pytest.mark.gpu

ACTS = {"none": -1, "relu": 0, "sigmoid": 1, "tanh": 2, "gelu": 3}


def _linear_case(K, oracle, dtype, batch, in_dim, out_dim, act, use_bias, l2=0.0, path=0):
    x, xn = to_dev(uniform((batch, in_dim), SEED_ACT), dtype)
    w, wn = to_dev(uniform((out_dim, in_dim), SEED_W) / np.sqrt(in_dim), dtype)
    b, bn = to_dev(uniform((out_dim,), SEED_W + 1), dtype) if use_bias else (None, None)
    y = torch.zeros(batch, out_dim, dtype=dtype, device="cuda")
    m = K.Linear.init_kernel(ACTS[act], l2, use_bias, path)
    K.Linear.forward_kernel(None, m, x, y, w, b, in_dim, out_dim, batch)
    yref = np.zeros((batch, out_dim), np.float32)
    oracle.linear_forward(xn, wn, bn, yref, in_dim, out_dim, batch, ACTS[act])
    assert_close(y, yref, dtype, what=f"linear fwd {dtype} {batch}x{in_dim}x{out_dim} {act}")

    # backward: grads start non-zero to check the accumulate convention
    dy, dyn = to_dev(uniform((batch, out_dim), SEED_ACT + 7), dtype)
    dx, dxn = to_dev(uniform((batch, in_dim), SEED_ACT + 8) * 0.1, dtype)
    dw, dwn = to_dev(uniform((out_dim, in_dim), SEED_W + 9) * 0.1, dtype)
    db, dbn = to_dev(uniform((out_dim,), SEED_W + 10) * 0.1, dtype) if use_bias else (None, None)
    preact = pre_n = None
    if act == "gelu":
        pre_n = np.zeros((batch, out_dim), np.float32)
        oracle.linear_forward(xn, wn, bn, pre_n, in_dim, out_dim, batch, -1)
        preact, pre_n = to_dev(pre_n, dtype)
    yq, yn = to_dev(yref, dtype)  # both sides see the same stored activations
    K.Linear.backward_kernel(None, m, x, dx, yq, dy, w, dw, db, in_dim, out_dim, batch, preact)
    dyn2 = dyn.copy()
    oracle.linear_backward(xn, dxn, yn, dyn2, wn, dwn, dbn, in_dim, out_dim, batch, ACTS[act],
                           float(l2), pre_n)
    tag = f"linear bwd {dtype} {batch}x{in_dim}x{out_dim} {act}"
    assert_close(dy, dyn2, dtype, what=tag + " dY(in place)")
    assert_close(dw, dwn, dtype, what=tag + " dW")
    assert_close(dx, dxn, dtype, what=tag + " dX")
    if use_bias:
        assert_close(db, dbn, dtype, what=tag + " db")


@pytest.mark.parametrize("act", ["none", "relu", "sigmoid", "tanh", "gelu"])
def test_linear_fp32_mlp_shapes(K, oracle, act):
    # config[0]: MLP 784-512-512-10, batch 64 (ragged 784 and 10 included)
    _linear_case(K, oracle, torch.float32, 64, 784, 512, act, True)
    _linear_case(K, oracle, torch.float32, 64, 512, 10, act, True, l2=1e-4)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("act", ["none", "relu", "gelu"])
def test_linear_16bit_tensor_core(K, oracle, dtype, act):
    _linear_case(K, oracle, dtype, 256, 512, 384, act, True, path=1)
    _linear_case(K, oracle, dtype, 200, 136, 104, act, False, l2=1e-3, path=1)  # ragged tiles


def test_linear_bf16_unaligned_falls_back_to_simt(K, oracle):
    _linear_case(K, oracle, torch.bfloat16, 33, 70, 10, "relu", True)


def test_linear_empty_batch_is_a_no_op(K):
    m = K.Linear.init_kernel()
    x = torch.zeros(0, 16, device="cuda")
    y = torch.zeros(0, 8, device="cuda")
    w = torch.zeros(8, 16, device="cuda")
    K.Linear.forward_kernel(None, m, x, y, w, None, 16, 8, 0)


def test_linear_full_size_bf16_linearity(K):
    """BASELINE config[1] size (8192x4096x4096 bf16), checked through size-independent
    properties: exact row-selection (a one-hot activation row returns a weight column),
    zero input -> bias, and homogeneity f(2x) = 2 f(x) (exact in binary floating point)."""
    M, Kd, N = 8192, 4096, 4096
    g = torch.Generator(device="cuda").manual_seed(SEED_W)
    w = ((torch.rand(N, Kd, device="cuda", generator=g) * 2 - 1) / 64).to(torch.bfloat16)
    x = torch.zeros(M, Kd, device="cuda", dtype=torch.bfloat16)
    rows = torch.arange(M, device="cuda")
    x[rows, (rows * 7) % Kd] = 1.0
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    m = K.Linear.init_kernel(path=1)
    K.Linear.forward_kernel(None, m, x, y, w, None, Kd, N, M)
    expect = w[:, (rows * 7) % Kd].t().contiguous()
    assert torch.equal(y, expect), "one-hot rows must return weight columns exactly"
    xr = ((torch.rand(M, Kd, device="cuda", generator=g) * 2 - 1)).to(torch.bfloat16)
    y1 = torch.empty_like(y)
    y2 = torch.empty_like(y)
    K.Linear.forward_kernel(None, m, xr, y1, w, None, Kd, N, M)
    K.Linear.forward_kernel(None, m, (xr * 2).contiguous(), y2, w, None, Kd, N, M)
    assert torch.equal(y2.float(), y1.float() * 2)
    # spot-check 64 random outputs against fp64 dot products
    idx = torch.randint(0, M, (64,), device="cuda", generator=g)
    jdx = torch.randint(0, N, (64,), device="cuda", generator=g)
    ref = (xr[idx].double() * w[jdx].double()).sum(1)
    assert_close(y1[idx, jdx], ref.float().cpu().numpy(), torch.bfloat16, what="full-size spot check")


@pytest.mark.parametrize("dtype,path", [(torch.float32, 0), (torch.bfloat16, 1)])
def test_batch_matmul(K, oracle, dtype, path):
    # reference fixture shape: m=10 n=10 k=10 batch=5 (lib/kernels/test/src/test_batch_matmul_kernel.cc:9-13)
    for (m, n, k, batch) in [(10, 10, 10, 5), (64, 48, 40, 3), (128, 256, 64, 2)]:
        if path == 1 and (m % 8 or k % 8):
            continue
        a, an = to_dev(uniform((batch, n, k), SEED_ACT), dtype)
        b, bn = to_dev(uniform((batch, k, m), SEED_W) / np.sqrt(k), dtype)
        o = torch.zeros(batch, n, m, dtype=dtype, device="cuda")
        K.BatchMatmul.forward_kernel(None, o, a, b, m, n, k, batch, path=path)
        oref = np.zeros((batch, n, m), np.float32)
        oracle.batch_matmul_forward(oref, an, bn, m, n, k, batch, -1, -1, -1)
        assert_close(o, oref, dtype, what=f"bmm fwd {m},{n},{k},{batch}")
        do, don = to_dev(uniform((batch, n, m), SEED_ACT + 3), dtype)
        da, dan = to_dev(uniform((batch, n, k), SEED_ACT + 4) * 0.1, dtype)
        db, dbn = to_dev(uniform((batch, k, m), SEED_ACT + 5) * 0.1, dtype)
        K.BatchMatmul.backward_kernel(None, o, do, a, da, b, db, m, n, k, batch, path=path)
        oracle.batch_matmul_backward(don, an, dan, bn, dbn, m, n, k, batch)
        assert_close(da, dan, dtype, what="bmm dA")
        assert_close(db, dbn, dtype, what="bmm dB")


def test_batch_matmul_seq_length_truncation(K, oracle):
    m, n, k, batch = 12, 10, 8, 2
    a, an = to_dev(uniform((batch, n, k), SEED_ACT))
    b, bn = to_dev(uniform((batch, k, m), SEED_W))
    for (ad, bd, sl) in [(1, -1, 6), (0, 1, 8), (-1, 0, 5)]:
        o = torch.full((batch, n, m), 7.0, device="cuda")
        oref = np.full((batch, n, m), 7.0, np.float32)
        K.BatchMatmul.forward_kernel(None, o, a, b, m, n, k, batch, ad, bd, sl)
        oracle.batch_matmul_forward(oref, an, bn, m, n, k, batch, ad, bd, sl)
        assert_close(o, oref, what=f"bmm truncation {ad},{bd},{sl}")


def test_fp32_on_tensor_cores_through_the_bf16_split(K, oracle):
    """fp32 operands, tcgen05 arithmetic: x = hi + lo in bf16, three products, fp32 accumulation.
    Must meet the fp32 tolerance (rtol 1e-4) against the fp64-accumulating oracle."""
    K.set_gemm_workspace(64 << 20, fp32_tensor_math=True)
    try:
        _linear_case(K, oracle, torch.float32, 512, 1024, 768, "relu", True, l2=1e-4, path=K.GEMM_SPLIT_BF16)
        _linear_case(K, oracle, torch.float32, 384, 520, 264, "none", True, path=K.GEMM_SPLIT_BF16)
        # AUTO picks the split path for large fp32 products once the fp32 policy allows it
        _linear_case(K, oracle, torch.float32, 1024, 1024, 512, "gelu", True, path=K.GEMM_AUTO)
        m, n, k, batch = 256, 192, 128, 6
        a, an = to_dev(uniform((batch, n, k), SEED_ACT))
        b, bn = to_dev(uniform((batch, k, m), SEED_W) / np.sqrt(k))
        o = torch.zeros(batch, n, m, device="cuda")
        K.BatchMatmul.forward_kernel(None, o, a, b, m, n, k, batch, path=K.GEMM_SPLIT_BF16)
        oref = np.zeros((batch, n, m), np.float32)
        oracle.batch_matmul_forward(oref, an, bn, m, n, k, batch, -1, -1, -1)
        assert_close(o, oref, what="bmm split-bf16")
    finally:
        K.set_gemm_workspace(None)


def test_registration_alone_keeps_fp32_exact(K):
    """Registering scratch must not change fp32 numerics: without the explicit policy
    (ffk_set_gemm_fp32_tensor_math) AUTO stays on the exact CUDA-core kernel, bit-identical to the
    result without scratch; with the policy the split path answers within 1e-4 and, on different
    streams of one device, calls that share the scratch are ordered (same result as serial calls)."""
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(1024, 512, device="cuda", generator=g) * 2 - 1
    w = (torch.rand(1024, 512, device="cuda", generator=g) * 2 - 1) / 16
    m = K.Linear.init_kernel(path=K.GEMM_AUTO)

    def run(stream=None):
        y = torch.zeros(1024, 1024, device="cuda")
        K.Linear.forward_kernel(stream, m, x, y, w, None, 512, 1024, 1024)
        return y

    K.set_gemm_workspace(None)
    exact = run()
    K.set_gemm_workspace(64 << 20)
    try:
        assert torch.equal(run(), exact)
        K.set_gemm_workspace(64 << 20, fp32_tensor_math=True)
        split = run()
        assert not torch.equal(split, exact)
        assert_close(split, exact.cpu().numpy(), what="fp32 split vs exact")
        torch.cuda.synchronize()
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        outs = [run(s1 if i % 2 else s2) for i in range(6)]
        torch.cuda.synchronize()
        for o in outs:
            assert torch.equal(o, split)
    finally:
        K.set_gemm_workspace(None)


def test_split_path_without_scratch_is_refused(K, ffk):
    K.set_gemm_workspace(None)
    x = torch.zeros(256, 256, device="cuda")
    y = torch.zeros(256, 256, device="cuda")
    m = K.Linear.init_kernel(path=K.GEMM_SPLIT_BF16)
    with pytest.raises(ffk.FFKError) as e:
        K.Linear.forward_kernel(None, m, x, y, x, None, 256, 256, 256)
    assert e.value.code == -3


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_cuda_core_gemm_cluster_split_k(K, a_mn, b_mn):
    """Few output tiles and a long K: the CUDA-core GEMM splits K over a thread-block cluster and
    reduces the partial tiles through distributed shared memory. Checked against fp64 products for
    every operand layout, ragged edges, batches, the accumulate flag — and twice for determinism."""
    g = torch.Generator(device="cuda").manual_seed(99)
    for (M, N, Kd, batch) in [(64, 512, 784, 1), (64, 10, 512, 1), (7, 5, 1000, 1), (20, 30, 300, 3),
                              (64, 64, 4096, 1), (130, 70, 129, 2)]:
        a = torch.rand((batch, Kd, M) if a_mn else (batch, M, Kd), device="cuda", generator=g) * 2 - 1
        b = torch.rand((batch, Kd, N) if b_mn else (batch, N, Kd), device="cuda", generator=g) * 2 - 1
        d0 = torch.rand(batch, M, N, device="cuda", generator=g)
        A = a.transpose(1, 2) if a_mn else a
        B = b.transpose(1, 2) if b_mn else b
        ref = (0.5 * torch.bmm(A.double(), B.double().transpose(1, 2)) + d0.double()).float()
        outs = []
        for _ in range(2):
            d = d0.clone()
            K.Gemm.run(a, b, d, M, N, Kd, batch=batch, a_mn=a_mn, b_mn=b_mn, alpha=0.5, accumulate=True,
                       path=K.GEMM_SIMT)
            outs.append(d)
        assert torch.equal(outs[0], outs[1])
        assert_close(outs[0], ref.cpu().numpy(), rtol=1e-5, what=f"split-k gemm {M}x{N}x{Kd}x{batch} {a_mn}{b_mn}")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_unaligned_pitch_is_repacked_onto_the_tensor_cores(K, oracle, dtype):
    """A 16-bit Linear whose out_dim is not a multiple of 8 (BERT's 30522-word vocabulary) breaks
    TMA's 16-byte pitch rule for dY in both backward products. With scratch registered the operand
    is repacked once and the products stay on the tensor cores; results must match the oracle like
    any other 16-bit Linear, and the CUDA-core answer of the same call without scratch."""
    K.set_gemm_workspace(96 << 20)
    try:
        _linear_case(K, oracle, dtype, 1024, 256, 1003, "gelu", True)
        _linear_case(K, oracle, dtype, 520, 264, 250, "none", False, l2=1e-3)
    finally:
        K.set_gemm_workspace(None)
