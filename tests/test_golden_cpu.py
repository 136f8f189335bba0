"""CPU: the oracle against the committed golden vectors.

tests/golden/mha_*.npz were produced by the reference's own arithmetic provider: cuDNN's
legacy multi-head attention, configured exactly as lib/kernels/src/cuda/ops/attention_kernels.cu
configures it, run on a B200 (oracle/cudnn_ref/mha_golden.cu, tools/make_mha_golden.py).
cuDNN evaluates those products with TF32 tensor-op math (10-bit mantissa), so the agreement
bound against them is 2e-3 of the tensor's scale; the oracle itself accumulates in fp64.

tests/golden/copy_like.json restates the exact-equality cases of the reference's kernel tests."""
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
CASES = ["fixture", "ragged", "cross"]


def load(case):
    z = np.load(GOLD / f"mha_{case}.npz")
    layout = json.loads(str(z["layout"]))
    c = layout["config"]
    dims = [c[k] for k in ("N", "H", "qSize", "kSize", "vSize", "qProj", "kProj", "vProj", "oProj", "qoSeq", "kvSeq")]
    return z, layout, dims


def close_tf32(got, ref, what):
    scale = np.abs(ref).max()
    err = np.abs(got - ref).max()
    assert err <= 2e-3 * scale, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("case", CASES)
def test_weight_layout_matches_cudnn_report(case):
    _, layout, dims = load(case)
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    kinds = layout["kinds"]
    assert layout["weightSizeBytes"] == 4 * (qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP)
    assert kinds["Q"] == {"offset_elems": 0, "dims": [H, qP, qS], "strides": [qP, 1, H * qP]}
    assert kinds["K"] == {"offset_elems": qS * H * qP, "dims": [H, kP, kS], "strides": [kP, 1, H * kP]}
    assert kinds["V"] == {"offset_elems": qS * H * qP + kS * H * kP, "dims": [H, vP, vS], "strides": [vP, 1, H * vP]}
    assert kinds["O"] == {"offset_elems": qS * H * qP + kS * H * kP + vS * H * vP, "dims": [H, oP, vP],
                          "strides": [vP * oP, 1, oP]}


@pytest.mark.parametrize("case", CASES)
def test_oracle_mha_against_cudnn_vectors(oracle, case):
    z, _, dims = load(case)
    q, k, v, w, dout = (np.ascontiguousarray(z[n]) for n in ("q", "k", "v", "w", "dout"))
    out = np.zeros_like(z["out"])
    oracle.mha_forward(dims, q, k, v, w, out)
    close_tf32(out, z["out"], "forward")
    dq, dk, dv, dw = (np.zeros_like(z[n]) for n in ("dq", "dk", "dv", "dw"))
    oracle.mha_backward(dims, q, dq, k, dk, v, dv, w, dw, dout)
    close_tf32(dq, z["dq"], "dq")
    close_tf32(dk, z["dk"], "dk")
    close_tf32(dv, z["dv"], "dv")
    close_tf32(dw, z["dw"], "dw")


def test_oracle_mha_against_torch(oracle):
    import torch

    dims = [2, 3, 10, 12, 14, 4, 4, 6, 8, 5, 7]
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    rng = np.random.default_rng(0)
    u = lambda *s: rng.uniform(-1, 1, s).astype(np.float32)
    q, k, v, dout = u(N, Sq, qS), u(N, Sk, kS), u(N, Sk, vS), u(N, Sq, oP)
    w = u(qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP) * 0.3
    tq, tk, tv, tw = (torch.from_numpy(a.copy()).double().requires_grad_() for a in (q, k, v, w))
    o = 0
    wq = tw[o:o + qS * H * qP].view(qS, H, qP); o += qS * H * qP
    wk = tw[o:o + kS * H * kP].view(kS, H, kP); o += kS * H * kP
    wv = tw[o:o + vS * H * vP].view(vS, H, vP); o += vS * H * vP
    wo = tw[o:].view(H, vP, oP)
    qp = torch.einsum("nsi,ihp->nhsp", tq, wq)
    kp = torch.einsum("nti,ihp->nhtp", tk, wk)
    vp = torch.einsum("nti,ihp->nhtp", tv, wv)
    p = torch.softmax(torch.einsum("nhsp,nhtp->nhst", qp, kp), -1)
    out_t = torch.einsum("nshp,hpo->nso", torch.einsum("nhst,nhtp->nshp", p, vp), wo)
    out_t.backward(torch.from_numpy(dout).double())
    out = np.zeros((N, Sq, oP), np.float32)
    oracle.mha_forward(dims, q, k, v, w, out)
    np.testing.assert_allclose(out, out_t.detach().numpy(), rtol=1e-5, atol=1e-6)
    dq, dk, dv, dw = np.full_like(q, 3.0), np.zeros_like(k), np.zeros_like(v), np.ones_like(w)
    oracle.mha_backward(dims, q, dq, k, dk, v, dv, w, dw, dout)
    np.testing.assert_allclose(dq, tq.grad.numpy(), rtol=1e-5, atol=1e-6)  # overwritten
    np.testing.assert_allclose(dk, tk.grad.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(dv, tv.grad.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(dw, 1 + tw.grad.numpy(), rtol=1e-5, atol=1e-5)  # accumulated


# ---------------------------------------------------------------------------------------------
# tests/golden/cudnn_ops.npz: results of the cuDNN calls the reference makes for Softmax,
# ElementUnary (RELU/SIGMOID/TANH/ELU), ElementBinary, Reduce, BatchNorm, Pool2D and Dropout,
# issued with the reference's descriptor settings on a B200 (oracle/cudnn_ref/ops_golden.cu,
# tools/make_ops_golden.py). fp32 throughout; cuDNN's own rounding differs from an fp64
# accumulation by a few ulp, hence rtol 1e-5 / atol 1e-6 unless noted.
@pytest.fixture(scope="module")
def G():
    return np.load(GOLD / "cudnn_ops.npz")


def near(got, ref, rtol=1e-5, atol=1e-6):
    np.testing.assert_allclose(got, ref, rtol=rtol, atol=atol)


def test_oracle_softmax_vs_cudnn(oracle, G):
    x = np.ascontiguousarray(G["softmax_x"])
    n, c, h, w = x.shape
    y = np.zeros_like(x)
    oracle.softmax_forward(x, y, n, c, h * w)
    near(y, G["softmax_y"])


@pytest.mark.parametrize("name,op", [("relu", 13), ("sigmoid", 15), ("tanh", 16), ("elu", 17)])
def test_oracle_activations_vs_cudnn(oracle, G, name, op):
    x = np.ascontiguousarray(G["act_x"])
    y = np.zeros_like(x)
    oracle.element_unary_forward(op, 0.0, x, y, x.size)  # ELU coefficient 0.0, as the reference sets
    near(y, G[f"act_{name}_y"])
    dx = np.array(G["act_dx0"], copy=True)
    oracle.element_unary_backward(op, 0.0, x, dx, np.ascontiguousarray(G[f"act_{name}_y"]),
                                  np.ascontiguousarray(G["act_dy"]), x.size)
    near(dx, G[f"act_{name}_dx"])  # beta = 1: accumulates onto dx0


@pytest.mark.parametrize("name,op", [("add", 28), ("sub", 35), ("mul", 29), ("max", 40), ("min", 41)])
def test_oracle_element_binary_vs_cudnn(oracle, G, name, op):
    a, b = np.ascontiguousarray(G["bin_lhs"]), np.ascontiguousarray(G["bin_rhs"])
    o = np.zeros_like(a)
    oracle.element_binary_forward(op, a, b, o, 4, list(a.shape), list(b.shape), list(a.shape))
    near(o, G[f"bin_{name}"])


def test_oracle_broadcast_gradient_vs_cudnn_reduce(oracle, G):
    a, b, dy = (np.ascontiguousarray(G[k]) for k in ("bin_lhs", "bin_rhs", "bin_dy"))
    dr = np.array(G["bin_drhs0"], copy=True)
    oracle.element_binary_backward(28, dy, a, b, None, dr, 4, list(a.shape), list(b.shape), list(a.shape))
    near(dr, G["bin_add_drhs"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name,mean", [("reduce_sum", 0), ("reduce_mean", 1)])
def test_oracle_reduce_vs_cudnn(oracle, G, name, mean):
    x = np.ascontiguousarray(G["reduce_x"])
    out = np.zeros(G[name].shape, np.float32)
    oracle.reduce_forward(x, out, 4, list(x.shape), list(out.shape), mean)
    near(out, G[name], rtol=1e-5, atol=1e-5)


def test_oracle_batch_norm_vs_cudnn(oracle, G):
    x = np.ascontiguousarray(G["bn_x"])
    n, c, h, w = x.shape
    eps = float(G["bn_eps"][0])
    y = np.zeros_like(x)
    rm, rv, sm, si = (np.zeros(c, np.float32) for _ in range(4))
    oracle.batch_norm_forward(x, y, np.ascontiguousarray(G["bn_scale"]), np.ascontiguousarray(G["bn_bias"]),
                              rm, rv, sm, si, n, c, h, w, eps, 1.0)
    near(y, G["bn_y"], rtol=1e-4, atol=1e-5)
    near(rm, G["bn_running_mean"])
    near(rv, G["bn_running_var"], rtol=1e-4)
    near(sm, G["bn_save_mean"])
    near(si, G["bn_save_invvar"], rtol=1e-4)  # cuDNN's "saved inverse variance" is 1/sqrt(var+eps)
    dx, ds, db = (np.array(G[k], copy=True) for k in ("bn_dx0", "bn_dscale0", "bn_dbias0"))
    oracle.batch_norm_backward(x, np.array(G["bn_dy"], copy=True), y, dx, np.ascontiguousarray(G["bn_scale"]),
                               ds, db, sm, si, n, c, h, w, 0)
    near(dx, G["bn_dx"], rtol=1e-3, atol=1e-4)
    near(ds, G["bn_dscale"], rtol=1e-4, atol=1e-4)
    near(db, G["bn_dbias"], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("name,ptype", [("pool_max", 0), ("pool_avg", 1)])
def test_oracle_pool2d_vs_cudnn(oracle, G, name, ptype):
    x = np.ascontiguousarray(G["pool_x"])
    n, c, ih, iw = x.shape
    _, _, oh, ow = G["pool_dy"].shape
    y = np.zeros((n, c, oh, ow), np.float32)
    oracle.pool2d(x, y, None, None, n, c, ih, iw, oh, ow, 3, 3, 2, 2, 1, 1, ptype, 0)
    near(y, G[name + "_y"])
    dx = np.array(G["pool_dx0"], copy=True)
    oracle.pool2d(x, None, dx, np.ascontiguousarray(G["pool_dy"]), n, c, ih, iw, oh, ow, 3, 3, 2, 2, 1, 1, ptype, 1)
    near(dx, G[name + "_dx"], rtol=1e-5, atol=1e-5)


def test_dropout_value_rule_of_cudnn(G):
    """cuDNN's mask cannot be reproduced; what it pins is the value rule our kernels share:
    y is 0 or x/(1-p), the backward pass reuses the forward mask, about 1-p of entries survive."""
    x, y, dy, dx = (G[k] for k in ("dropout_x", "dropout_y", "dropout_dy", "dropout_dx"))
    p = float(G["dropout_rate"][0])
    kept = y != 0
    near(y[kept], x[kept] / (1 - p))
    assert ((dx != 0) == kept).all()
    near(dx[kept], dy[kept] / (1 - p))
    assert abs(kept.mean() - (1 - p)) < 0.02


# ---------------------------------------------------------------------------------------------
def _copy_like_cases():
    return json.loads((GOLD / "copy_like.json").read_text())["cases"]


@pytest.mark.parametrize("case", _copy_like_cases(), ids=lambda c: c["name"])
def test_oracle_against_the_reference_tests_exact_vectors(oracle, case):
    n = case["n"]
    src = np.full(n * case.get("reps", 1), case["src_fill"], np.float32)
    dst = np.full(n, case.get("dst_fill", 0.0), np.float32)
    if case["op"] == "copy":
        oracle.copy(dst, src, n * 4)
    elif case["op"] == "accumulate":
        oracle.accumulate(dst, src, n)
    else:
        oracle.sum_replicas(dst, src, n, case["reps"], case["accumulate"])
    assert (dst == np.float32(case["expect_fill"])).all()


# ---------------------------------------------------------------------------------------------
# tests/golden/cudnn_conv.npz: Conv2D through the cuDNN calls of conv_2d_kernels.cu:276-388
# (forward + bias + ReLU; ReLU gradient in place on dY; filter / bias / data gradients with
# beta = 1) on four geometries incl. groups, recorded on a B200 by oracle/cudnn_ref/conv_golden.cu
# with CUDNN_FMA_MATH and fixed algorithms (see the header of that file).
CONV_CASES = ["conv_a", "conv_b", "conv_c", "conv_d"]


@pytest.fixture(scope="module")
def GC():
    return np.load(GOLD / "cudnn_conv.npz")


@pytest.mark.parametrize("name", CONV_CASES)
def test_oracle_conv2d_vs_cudnn(oracle, GC, name):
    geo = [int(v) for v in GC[name + "_geo"]]
    relu = geo.pop()
    x, w, b = (np.ascontiguousarray(GC[f"{name}_{k}"]) for k in ("x", "w", "b"))
    y = np.zeros_like(GC[name + "_y"])
    oracle.conv2d_forward(geo, x, y, w, b, relu)
    near(y, GC[name + "_y"], rtol=1e-5, atol=1e-5)
    dy, dx, dw, db = (np.array(GC[f"{name}_{k}"], copy=True) for k in ("dy", "dx0", "dw0", "db0"))
    oracle.conv2d_backward(geo, x, dx, np.ascontiguousarray(GC[name + "_y"]), dy, w, dw, db, relu)
    near(dy, GC[name + "_dy_after"], rtol=0, atol=0)          # the ReLU gradient is applied to dY in place
    near(dw, GC[name + "_dw"], rtol=1e-4, atol=1e-4)
    near(db, GC[name + "_db"], rtol=1e-4, atol=1e-4)
    near(dx, GC[name + "_dx"], rtol=1e-4, atol=1e-5)
