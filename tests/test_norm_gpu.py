"""GPU parity: Softmax / LayerNorm / BatchNorm vs the oracle."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_W, assert_close, assert_equal_bits, to_dev, uniform

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("outer,dim,inner", [(1, 100, 1), (64, 10, 1), (33, 4096, 1), (7, 1000, 1),
                                             (2, 5, 12), (4, 32, 100), (3, 8200, 1), (5, 3001, 1), (9, 30522, 1)])
def test_softmax_forward(K, oracle, dtype, outer, dim, inner):
    x, xn = to_dev(uniform((outer, dim, inner), SEED_ACT) * 4, dtype)
    y = torch.empty_like(x)
    K.Softmax.forward_kernel(None, x, y, outer, dim, inner)
    ref = np.zeros_like(xn)
    oracle.softmax_forward(xn, ref, outer, dim, inner)
    assert_close(y, ref, dtype, what=f"softmax {outer},{dim},{inner}")


def test_softmax_full_size_rows_sum_to_one(K):
    x = torch.randn(8192, 4096, device="cuda", generator=torch.Generator(device="cuda").manual_seed(SEED_ACT))
    y = torch.empty_like(x)
    K.Softmax.forward_kernel(None, x, y, 8192, 4096, 1)
    s = y.double().sum(1)
    assert (s - 1).abs().max().item() < 1e-5
    assert torch.equal(y.argmax(1), x.argmax(1)) and (y >= 0).all()
    # shift invariance: softmax(x + c) == softmax(x) up to rounding of x + c
    y2 = torch.empty_like(x)
    K.Softmax.forward_kernel(None, x + 3.0, y2, 8192, 4096, 1)
    assert (y2 - y).abs().max().item() < 1e-5


def test_softmax_backward_reference_copy_and_jacobian(K, oracle):
    # reference fixture: the backward is a device copy (lib/kernels/test/src/test_softmax_kernel.cc:52-57)
    dy, dyn = to_dev(uniform((100,), SEED_ACT))
    dx = torch.zeros_like(dy)
    K.Softmax.backward_kernel(None, dx, dy)
    assert_equal_bits(dx, dy, "softmax bwd copy")
    outer, dim, inner = 6, 50, 3
    x, xn = to_dev(uniform((outer, dim, inner), SEED_ACT))
    y = torch.empty_like(x)
    K.Softmax.forward_kernel(None, x, y, outer, dim, inner)
    g, gn = to_dev(uniform((outer, dim, inner), SEED_W))
    dx = torch.zeros_like(x)
    K.Softmax.backward_kernel(None, dx, g, y, outer, dim, inner, mode=1)
    ref = np.zeros_like(xn)
    oracle.softmax_backward_jacobian(ref, gn, y.cpu().numpy(), outer, dim, inner)
    assert_close(dx, ref, what="softmax jacobian bwd")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,N,affine", [(10, 10, True), (64, 1024, True), (33, 4096, True),
                                        (17, 1000, False), (5, 16388, True), (300, 256, True)])
@pytest.mark.parametrize("workspace", [True, False])
def test_layer_norm(K, oracle, dtype, M, N, affine, workspace):
    eps = 1e-5  # reference fixture: lib/kernels/test/src/test_layer_norm_kernels.cc:13
    x, xn = to_dev(uniform((M, N), SEED_ACT) + 0.25, dtype)
    g, gn = to_dev(uniform((N,), SEED_W) + 1.0, dtype) if affine else (None, None)
    b, bn = to_dev(uniform((N,), SEED_W + 1), dtype) if affine else (None, None)
    y = torch.empty_like(x)
    m = K.LayerNorm.init_kernel(affine, M, N, eps, workspace=workspace)
    K.LayerNorm.forward_kernel(None, m, x, y, g, b)
    yref = np.zeros_like(xn)
    mean, rstd = np.zeros(M, np.float32), np.zeros(M, np.float32)
    oracle.layer_norm_forward(xn, yref, gn, bn, mean, rstd, M, N, eps)
    assert_close(y, yref, dtype, what=f"layernorm fwd {M}x{N}")
    assert_close(m.mean, mean, rtol=1e-4, what="mean")
    assert_close(m.rstd, rstd, rtol=1e-4, what="rstd")
    dy, dyn = to_dev(uniform((M, N), SEED_ACT + 1), dtype)
    dx, dxn = to_dev(uniform((M, N), SEED_ACT + 2), dtype)
    dg = torch.full((N,), 5.0, dtype=dtype, device="cuda")
    db = torch.full((N,), 5.0, dtype=dtype, device="cuda")
    dgn, dbn = np.full(N, 5.0, np.float32), np.full(N, 5.0, np.float32)
    K.LayerNorm.backward_kernel(None, m, dy, x, dx, g, dg, db)
    # the oracle uses the GPU's saved statistics so both sides differentiate the same function
    oracle.layer_norm_backward(dyn, xn, dxn, gn, dgn, dbn, m.mean.cpu().numpy(), m.rstd.cpu().numpy(), M, N, 0)
    assert_close(dx, dxn, dtype, what=f"layernorm dX {M}x{N} ws={workspace}")
    assert_close(dg, dgn, dtype, what="layernorm dgamma (overwritten like the reference)")
    assert_close(db, dbn, dtype, what="layernorm dbeta")
    # accumulate mode
    K.LayerNorm.backward_kernel(None, m, dy, x, None, g, dg, db, accumulate_param_grads=True)
    oracle.layer_norm_backward(dyn, xn, None, gn, dgn, dbn, m.mean.cpu().numpy(), m.rstd.cpu().numpy(), M, N, 1)
    assert_close(dg, dgn, dtype, what="layernorm dgamma accumulate")


def test_layer_norm_full_size_statistics(K):
    M, N = 8192, 4096
    x = torch.randn(M, N, device="cuda", generator=torch.Generator(device="cuda").manual_seed(SEED_ACT)) * 3 + 1
    y = torch.empty_like(x)
    m = K.LayerNorm.init_kernel(False, M, N, 1e-5)
    K.LayerNorm.forward_kernel(None, m, x, y, None, None)
    assert y.double().mean(1).abs().max().item() < 1e-5
    assert (y.double().var(1, unbiased=False) - 1).abs().max().item() < 1e-3
    # idempotence up to eps: normalising a normalised row changes nothing
    y2 = torch.empty_like(x)
    K.LayerNorm.forward_kernel(None, m, y, y2, None, None)
    assert (y2 - y).abs().max().item() < 1e-4


@pytest.mark.parametrize("relu", [False, True])
@pytest.mark.parametrize("n,c,h,w", [(1, 10, 10, 10), (8, 16, 7, 5), (4, 3, 1, 1)])
def test_batch_norm(K, oracle, n, c, h, w, relu):
    # reference fixture shape 1x10x10x10 relu (lib/kernels/test/src/test_batch_norm_kernel.cc:9-10)
    eps = 1e-5
    x, xn = to_dev(uniform((n, c, h, w), SEED_ACT) + 0.5)
    sc, scn = to_dev(uniform((c,), SEED_W) + 1.0)
    bi, bin_ = to_dev(uniform((c,), SEED_W + 1))
    m = K.BatchNorm.init_kernel(n, c, h, w, relu, eps=eps)
    y = torch.empty_like(x)
    K.BatchNorm.forward_kernel(None, m, x, y, sc, bi)
    yref = np.zeros_like(xn)
    rm, rv, sm, si = (np.zeros(c, np.float32) for _ in range(4))
    oracle.batch_norm_forward(xn, yref, scn, bin_, rm, rv, sm, si, n, c, h, w, eps, 1.0)
    if n * h * w > 1:
        assert_close(y, yref, what="batchnorm fwd")
        assert_close(m.running_var, rv, what="running var")
    assert_close(m.running_mean, rm, what="running mean")
    assert_close(m.save_mean, sm, what="save mean")
    dy, dyn = to_dev(uniform((n, c, h, w), SEED_ACT + 1))
    dx, dxn = to_dev(uniform((n, c, h, w), SEED_ACT + 2))
    dsc, dscn = to_dev(uniform((c,), SEED_W + 2))
    dbi, dbin = to_dev(uniform((c,), SEED_W + 3))
    K.BatchNorm.backward_kernel(None, m, x, dy, y, dx, sc, dsc, dbi)
    oracle.batch_norm_backward(xn, dyn, y.cpu().numpy(), dxn, scn, dscn, dbin,
                               m.save_mean.cpu().numpy(), m.save_invstd.cpu().numpy(), n, c, h, w, int(relu))
    if n * h * w > 1:
        assert_close(dx, dxn, rtol=1e-3, what="batchnorm dX")
        assert_close(dsc, dscn, rtol=1e-3, what="batchnorm dscale")
    assert_close(dbi, dbin, what="batchnorm dbias")
    assert_close(dy, dyn, what="batchnorm dY after relu grad")
