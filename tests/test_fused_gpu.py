"""Fused-operator entry points (ffk_linear_forward_fused / ffk_linear_backward_fused,
ffk_layer_norm_backward_ex, ffk_mha_forward_residual / ffk_mha_backward_accumulate) against the
chains of per-operator entry points they stand for, through the C ABI, and against the oracle.
Shapes cover the one-kernel path (16-bit, CTA-pair and single-CTA tiles, ragged row counts) and the
built-in unfused chain (fp32, widths that are not a multiple of 32)."""
import numpy as np
import pytest
import torch

from helpers import assert_close, to_dev, uniform

pytestmark = pytest.mark.gpu

ACTS = {"none": -1, "relu": 0, "gelu": 3}


def _unfused_forward(K, x, w, b, act, drop_state, residual):
    """Linear -> Dropout -> ElementBinary add with the per-operator kernels."""
    B, I = x.shape
    O = w.shape[0]
    y = torch.zeros(B, O, dtype=x.dtype, device="cuda")
    pre = torch.zeros_like(y)
    K.Linear.forward_kernel(None, K.Linear.init_kernel(K.ACT_NONE, 0.0, True, K.GEMM_AUTO), x, pre, w, b, I, O, B)
    K.Linear.forward_kernel(None, K.Linear.init_kernel(act, 0.0, True, K.GEMM_AUTO), x, y, w, b, I, O, B)
    if drop_state is not None:
        d = torch.zeros_like(y)
        K.Dropout.forward_kernel(None, drop_state, y, d)
        y = d
    if residual is not None:
        o = torch.zeros_like(y)
        K.ElementBinary.forward_kernel(None, K.OperatorType.EW_ADD, y, residual, o)
        y = o
    return y, pre


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(512, 256, 512), (500, 192, 384), (96, 64, 128), (300, 128, 200)])
@pytest.mark.parametrize("act", ["none", "relu", "gelu"])
def test_fused_linear_forward_equals_the_chain(K, dtype, shape, act):
    B, I, O = shape
    x, _ = to_dev(uniform((B, I), 1), dtype)
    w, _ = to_dev(uniform((O, I), 2) / np.sqrt(I), dtype)
    b, _ = to_dev(uniform((O,), 3), dtype)
    res, _ = to_dev(uniform((B, O), 4), dtype)
    for use_drop, use_res, use_pre in ((True, True, True), (True, False, False), (False, True, False), (False, False, True)):
        d_ref = K.Dropout.init_kernel(0.25, 777, B * O) if use_drop else None
        d_fus = K.Dropout.init_kernel(0.25, 777, B * O) if use_drop else None
        want, want_pre = _unfused_forward(K, x, w, b, ACTS[act], d_ref, res if use_res else None)
        y = torch.full((B, O), 7.0, dtype=dtype, device="cuda")
        pre = torch.full((B, O), 7.0, dtype=dtype, device="cuda") if use_pre else None
        m = K.Linear.init_kernel(ACTS[act], 0.0, True, K.GEMM_AUTO)
        K.FusedLinear.forward_kernel(None, m, x, y, w, b, I, O, B, preact_out=pre, residual=res if use_res else None,
                                     dropout=d_fus)
        torch.cuda.synchronize()
        tag = f"{dtype} {shape} {act} drop={use_drop} res={use_res}"
        if use_drop:   # the epilogue draws the very bits of the stand-alone kernel
            assert torch.equal(d_fus.mask, d_ref.mask), f"{tag}: dropout masks differ"
            assert d_fus.offset == d_ref.offset == 1
        # the chain rounds every intermediate to the storage type, the fused GEMM rounds once
        assert_close(y, want.float().cpu().numpy(), dtype, what=f"fused forward {tag}")
        if use_pre:
            assert_close(pre, want_pre.float().cpu().numpy(), dtype, what=f"fused pre-activation {tag}")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(512, 256, 512), (300, 128, 192), (64, 40, 72)])
def test_fused_feed_forward_backward_equals_the_chain(K, oracle, dtype, shape):
    """Linear(ReLU) -> Dropout -> Linear -> Dropout (+ residual): the fused backward (gated data
    gradient, pre-activation gradient handed down, first-writer data gradient) against the chain of
    per-operator backward kernels on the same masks."""
    B, D, F = shape
    rate = 0.2
    x, _ = to_dev(uniform((B, D), 1), dtype)
    w1, _ = to_dev(uniform((F, D), 2) / np.sqrt(D), dtype)
    b1, _ = to_dev(uniform((F,), 3), dtype)
    w2, _ = to_dev(uniform((D, F), 4) / np.sqrt(F), dtype)
    b2, _ = to_dev(uniform((D,), 5), dtype)
    gout, _ = to_dev(uniform((B, D), 6), dtype)
    z = lambda *s: torch.zeros(*s, dtype=dtype, device="cuda")
    # ---- fused forward
    m1 = K.Linear.init_kernel(K.ACT_RELU, 0.0, True, K.GEMM_AUTO, out_dim=F)
    m2 = K.Linear.init_kernel(K.ACT_NONE, 0.0, True, K.GEMM_AUTO, out_dim=D)
    d1, d2 = K.Dropout.init_kernel(rate, 11, B * F), K.Dropout.init_kernel(rate, 12, B * D)
    h, y = z(B, F), z(B, D)
    K.FusedLinear.forward_kernel(None, m1, x, h, w1, b1, D, F, B, dropout=d1)
    K.FusedLinear.forward_kernel(None, m2, h, y, w2, b2, F, D, B, dropout=d2, residual=x)
    # ---- fused backward: dz2 = dropout_bwd(gout); ff2 with the gate; ff1 takes the pre-activation gradient
    dz2 = z(B, D)
    K.Dropout.backward_kernel(None, d2, gout, dz2)
    dh = torch.full((B, F), 9.0, dtype=dtype, device="cuda")            # overwritten: first writer
    dw2, db2, dw1, db1 = z(D, F), z(D), z(F, D), z(F)
    dx = gout.clone()                                                     # residual branch: dx starts as gout
    K.FusedLinear.backward_kernel(None, m2, h, dh, y, dz2, w2, dw2, db2, F, D, B, accumulate_input_grad=False,
                                  input_gate=h, input_gate_scale=1.0 / (1.0 - rate))
    K.FusedLinear.backward_kernel(None, m1, x, dx, h, dh, w1, dw1, db1, D, F, B, accumulate_input_grad=True,
                                  output_grad_is_preactivation=True)
    # ---- the chain: same masks, per-operator kernels
    u1 = K.Linear.init_kernel(K.ACT_RELU, 0.0, True, K.GEMM_AUTO, out_dim=F)
    r1 = z(B, F)
    K.Linear.forward_kernel(None, u1, x, r1, w1, b1, D, F, B)
    hd_ = z(B, F)
    K._call("ffk_dropout_backward", K._stream(None), r1.data_ptr(), hd_.data_ptr(), d1.mask.data_ptr(), B * F, rate,
            K.dtype_code(r1))                                              # y = x * keep / (1 - rate): same map as forward
    c_dz2 = z(B, D)
    K.Dropout.backward_kernel(None, d2, gout, c_dz2)
    c_dhd, c_dw2, c_db2 = z(B, F), z(D, F), z(D)
    K.Linear.backward_kernel(None, m2, hd_, c_dhd, y, c_dz2, w2, c_dw2, c_db2, F, D, B)
    c_dh = z(B, F)
    K.Dropout.backward_kernel(None, d1, c_dhd, c_dh)
    c_dx, c_dw1, c_db1 = gout.clone(), z(F, D), z(F)
    K.Linear.backward_kernel(None, u1, x, c_dx, r1, c_dh, w1, c_dw1, c_db1, D, F, B)
    torch.cuda.synchronize()
    assert_close(h, hd_.float().cpu().numpy(), dtype, what="fused ReLU + dropout output")
    for name, got, want in (("dW2", dw2, c_dw2), ("db2", db2, c_db2), ("dW1", dw1, c_dw1), ("db1", db1, c_db1),
                            ("dX", dx, c_dx)):
        what = f"fused feed-forward backward {name} {dtype} {shape}"
        if dtype == torch.float32:
            assert_close(got, want.float().cpu().numpy(), dtype, what=what)
        else:
            # two 16-bit pipelines that round their intermediates at different points (the chain rounds
            # the ReLU output before the dropout scale, the fused GEMM rounds once): sums over hundreds of
            # rows that cancel differ by more than an elementwise 2e-2; bounded in relative L2 instead
            from helpers import assert_close_l2
            assert_close_l2(got, want.float().cpu().numpy(), what=what)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(64, 1024), (300, 96), (40, 4096), (17, 200), (9, 8192)])
def test_layer_norm_backward_overwrite_equals_accumulate_on_zero(K, oracle, dtype, shape):
    """accumulate_input_grad = 0 must give, on a buffer full of garbage, what the accumulating form
    gives on a zero-filled one — for every kernel variant the dispatcher picks."""
    M, N = shape
    x, xn = to_dev(uniform((M, N), 1), dtype)
    g, gn = to_dev(uniform((N,), 2) + 1, dtype)
    be, _ = to_dev(uniform((N,), 3), dtype)
    dy, dyn = to_dev(uniform((M, N), 4), dtype)
    ln = K.LayerNorm.init_kernel(True, M, N, 1e-5)
    y = torch.empty_like(x)
    K.LayerNorm.forward_kernel(None, ln, x, y, g, be)
    outs = []
    for acc in (True, False):
        dx = torch.zeros_like(x) if acc else torch.full_like(x, 123.0)
        dg, db = torch.zeros_like(g), torch.zeros_like(g)
        K.LayerNorm.backward_kernel(None, ln, dy, x, dx, g, dg, db, accumulate_input_grad=acc)
        outs.append((dx, dg, db))
    torch.cuda.synchronize()
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)
    dxr, dgr, dbr = np.zeros((M, N), np.float32), np.zeros(N, np.float32), np.zeros(N, np.float32)
    oracle.layer_norm_backward(dyn, xn, dxr, gn, dgr, dbr, ln.mean[:M].cpu().numpy(), ln.rstd[:M].cpu().numpy(), M, N, 0)
    assert_close(outs[1][0], dxr, dtype, what=f"layer norm dX overwrite {dtype} {shape}")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_mha_residual_and_accumulating_backward(K, dtype):
    """out = MHA(x,x,x) + x in one call equals MHA then add; the accumulating backward on ONE
    gradient buffer equals dq + dk + dv (+ the buffer's previous contents)."""
    N, S, E, H, hd = 3, 24, 64, 4, 16
    x, _ = to_dev(uniform((N, S, E), 1), dtype)
    st = K.MultiHeadAttention.init_kernel(N, H, E, E, E, hd, hd, hd, E, S, S, dtype=dtype)
    nw = st.weightSize // x.element_size()
    w, _ = to_dev(uniform((nw,), 2) / 8, dtype)
    dout, _ = to_dev(uniform((N, S, E), 3), dtype)
    plain, fused = torch.zeros_like(x), torch.zeros_like(x)
    K.MultiHeadAttention.forward_kernel(None, st, x, x, x, w, plain)
    K.MultiHeadAttention.forward_residual_kernel(None, st, x, x, x, w, fused, x)
    torch.cuda.synchronize()
    assert_close(fused, (plain.float() + x.float()).cpu().numpy(), dtype, what="mha + residual")
    dq, dk, dv = torch.zeros_like(x), torch.zeros_like(x), torch.zeros_like(x)
    dw_a, dw_b = torch.zeros_like(w), torch.zeros_like(w)
    K.MultiHeadAttention.backward_kernel(None, st, x, dq, x, dk, x, dv, w, dw_a, dout)
    for first in (True, False):
        buf = torch.full_like(x, 55.0) if first else dout.clone()
        dw_b.zero_()
        K.MultiHeadAttention.backward_accumulate_kernel(None, st, x, buf, x, buf, x, buf, w, dw_b, dout,
                                                        accumulate=(not first, True, True))
        torch.cuda.synchronize()
        want = dq.float() + dk.float() + dv.float() + (0 if first else dout.float())
        assert_close(buf, want.cpu().numpy(), dtype, what=f"mha accumulate backward first={first}")
        assert_close(dw_b, dw_a.float().cpu().numpy(), dtype, what="mha weight gradient")


def test_fused_and_unfused_models_agree(ffk):
    """The same encoder (same seeds, same dropout masks) built for per-operator and for fused-operator
    execution: three training steps must leave the same weights, up to bf16 rounding."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    B, S, D, H, FF, L = 4, 64, 256, 4, 512, 2
    g = torch.Generator(device="cuda").manual_seed(5)
    x = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
    target = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
    finals = []
    for fuse in (False, True):
        model = T.build_transformer_encoder(num_layers=L, d_model=D, heads=H, seq=S, batch=B, dim_ff=FF,
                                            dropout=0.1, dtype=torch.bfloat16, fuse=fuse)
        opt = T.SGD(lr=0.05, weight_decay=1e-3)
        for _ in range(3):
            model.input.data.copy_(x)
            T.mse_step(model, target, opt)
        torch.cuda.synchronize()
        finals.append(([p.float().cpu().numpy() for p in model.params], model.output.data.float().cpu().numpy(),
                       len(model.layers)))
    (pa, oa, la), (pb, ob, lb) = finals
    assert lb < la
    from helpers import assert_close_l2
    assert_close_l2(ob, oa, what="fused vs unfused output")
    # parameters that start at zero (biases, beta) hold nothing but three steps of gradient: two bf16
    # pipelines that round intermediates at different points agree on them to a few per cent; the
    # fp32 run of tests/test_training_gpu.py (fuse=True against the oracle, rtol 1e-4) is the exact check
    for i, (a, b) in enumerate(zip(pa, pb)):
        assert_close_l2(b, a, rel=6e-2, max_rel=0.3, what=f"fused vs unfused parameter {i}")
