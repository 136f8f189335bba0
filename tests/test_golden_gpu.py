"""GPU: the sm_100a kernels against the golden vectors produced by the cuDNN calls the
reference makes (tests/golden/cudnn_ops.npz; see tests/test_golden_cpu.py for provenance)."""
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def G():
    return np.load(GOLD / "cudnn_ops.npz")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def near(got, ref, rtol=1e-5, atol=1e-6):
    np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=rtol, atol=atol)


def test_softmax(K, G):
    x = dev(G["softmax_x"])
    n, c, h, w = x.shape
    y = torch.empty_like(x)
    K.Softmax.forward_kernel(None, x, y, n, c, h * w)
    near(y, G["softmax_y"], rtol=2e-5)


@pytest.mark.parametrize("name", ["relu", "sigmoid", "tanh", "elu"])
def test_activations(K, G, name):
    op = getattr(K.OperatorType, name.upper())
    x, dy, dx = dev(G["act_x"]), dev(G["act_dy"]), dev(G["act_dx0"])
    y = torch.empty_like(x)
    K.ElementUnary.forward_kernel(None, op, 0.0, x, y)
    near(y, G[f"act_{name}_y"], rtol=2e-5)
    K.ElementUnary.backward_kernel(None, op, 0.0, x, dx, dev(G[f"act_{name}_y"]), dy)
    near(dx, G[f"act_{name}_dx"], rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("name", ["add", "sub", "mul", "max", "min"])
def test_element_binary_with_channel_broadcast(K, G, name):
    op = getattr(K.OperatorType, "EW_" + name.upper())
    a, b = dev(G["bin_lhs"]), dev(G["bin_rhs"])
    o = torch.empty_like(a)
    K.ElementBinary.forward_kernel(None, op, a, b, o)
    near(o, G[f"bin_{name}"])
    if name == "add":
        dr = dev(G["bin_drhs0"])
        K.ElementBinary.backward_kernel(None, op, dev(G["bin_dy"]), a, b, None, dr)
        near(dr, G["bin_add_drhs"], atol=1e-5)


@pytest.mark.parametrize("name,op", [("reduce_sum", "REDUCE_SUM"), ("reduce_mean", "REDUCE_MEAN")])
def test_reduce(K, G, name, op):
    x = dev(G["reduce_x"])
    out = torch.zeros(G[name].shape, device="cuda")
    K.Reduce.forward_kernel(None, getattr(K.OperatorType, op), x, out)
    near(out, G[name], atol=1e-5)


def test_batch_norm(K, G):
    x = dev(G["bn_x"])
    n, c, h, w = x.shape
    m = K.BatchNorm.init_kernel(n, c, h, w, False, eps=float(G["bn_eps"][0]))
    y = torch.empty_like(x)
    sc, bi = dev(G["bn_scale"]), dev(G["bn_bias"])
    K.BatchNorm.forward_kernel(None, m, x, y, sc, bi)
    near(y, G["bn_y"], rtol=1e-4, atol=1e-5)
    near(m.running_mean, G["bn_running_mean"])
    near(m.running_var, G["bn_running_var"], rtol=1e-4)
    near(m.save_invstd, G["bn_save_invvar"], rtol=1e-4)
    dy, dx, ds, db = dev(G["bn_dy"]), dev(G["bn_dx0"]), dev(G["bn_dscale0"]), dev(G["bn_dbias0"])
    K.BatchNorm.backward_kernel(None, m, x, dy, y, dx, sc, ds, db)
    near(dx, G["bn_dx"], rtol=1e-3, atol=1e-4)
    near(ds, G["bn_dscale"], rtol=1e-4, atol=1e-4)
    near(db, G["bn_dbias"], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("name,ptype", [("pool_max", "POOL_MAX"), ("pool_avg", "POOL_AVG")])
def test_pool2d(K, G, name, ptype):
    x = dev(G["pool_x"])
    n, c, ih, iw = x.shape
    _, _, oh, ow = G["pool_dy"].shape
    m = K.Pool2D.init_kernel(iw, ih, c, n, ow, oh, c, n, 1, 1, 3, 3, 2, 2, getattr(K, ptype))
    y = torch.empty(n, c, oh, ow, device="cuda")
    K.Pool2D.forward_kernel(None, m, x, y)
    near(y, G[name + "_y"])
    dx = dev(G["pool_dx0"])
    K.Pool2D.backward_kernel(None, m, x, dx, y, dev(G["pool_dy"]))
    near(dx, G[name + "_dx"], atol=1e-5)


@pytest.mark.parametrize("name", ["conv_a", "conv_b", "conv_c", "conv_d"])
def test_conv2d(K, name):
    """Conv2D against the cuDNN calls of conv_2d_kernels.cu:276-388 (tests/golden/cudnn_conv.npz,
    oracle/cudnn_ref/conv_golden.cu): forward + bias + ReLU, gradients accumulating onto dX / dW / db."""
    GC = np.load(GOLD / "cudnn_conv.npz")
    n, ci, ih, iw, co, oh, ow, k, _, p, _, s, _, g, relu = (int(v) for v in GC[name + "_geo"])
    x, w, b = dev(GC[name + "_x"]), dev(GC[name + "_w"]), dev(GC[name + "_b"])
    m = K.Conv2D.init_kernel(K.ACT_RELU if relu else None, k, k, g, p, p, s, s, (n, ci, ih, iw), (n, co, oh, ow))
    y = torch.zeros(n, co, oh, ow, device="cuda")
    K.Conv2D.forward_kernel(None, m, x, y, w, b)
    near(y, GC[name + "_y"], rtol=1e-4, atol=1e-5)
    dy, dx, dw, db = dev(GC[name + "_dy"]), dev(GC[name + "_dx0"]), dev(GC[name + "_dw0"]), dev(GC[name + "_db0"])
    K.Conv2D.backward_kernel(None, m, x, dx, dev(GC[name + "_y"]), dy, w, dw, db)
    near(dw, GC[name + "_dw"], rtol=1e-4, atol=1e-4)
    near(db, GC[name + "_db"], rtol=1e-4, atol=1e-4)
    near(dx, GC[name + "_dx"], rtol=1e-4, atol=1e-4)
    near(dy, GC[name + "_dy_after"], rtol=0, atol=0)   # the ReLU gradient is applied to dY in place


import json


@pytest.mark.parametrize("case", json.loads((GOLD / "copy_like.json").read_text())["cases"], ids=lambda c: c["name"])
def test_reference_exact_vectors(K, case):
    """The vectors the reference's own kernel tests assert with == (tests/golden/copy_like.json)."""
    n = case["n"]
    src = torch.full((n * case.get("reps", 1),), case["src_fill"], device="cuda")
    dst = torch.full((n,), case.get("dst_fill", 0.0), device="cuda")
    if case["op"] == "copy":
        K.copy(None, dst, src)
    elif case["op"] == "accumulate":
        K.accumulate(None, dst, src)
    else:
        K._call("ffk_sum_replicas", None, dst.data_ptr(), src.data_ptr(), n, case["reps"], case["accumulate"], K.FLOAT)
    assert torch.equal(dst, torch.full((n,), case["expect_fill"], device="cuda"))
