"""CPU: the oracle restatement is cross-checked against PyTorch's CPU implementations of the
same math (the convention of the reference's legacy alignment tests, tests/align/align_utils.py:
seeded random inputs, torch.testing.assert_close). The reference pins none of these ops
numerically, so agreement of two independent statements is the pin we have."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import SEED_ACT, SEED_W, uniform

TOL = dict(rtol=1e-5, atol=1e-6)


def t(a):
    return torch.from_numpy(np.array(a, copy=True))


@pytest.mark.parametrize("act,fn", [(-1, lambda v: v), (0, torch.relu), (1, torch.sigmoid), (2, torch.tanh),
                                    (3, lambda v: F.gelu(v, approximate="tanh"))])
def test_linear(oracle, act, fn):
    B, I, O = 16, 40, 24
    x, w, b = uniform((B, I), SEED_ACT), uniform((O, I), SEED_W), uniform((O,), 3)
    y = np.zeros((B, O), np.float32)
    oracle.linear_forward(x, w, b, y, I, O, B, act)
    tx, tw, tb = t(x).requires_grad_(), t(w).requires_grad_(), t(b).requires_grad_()
    pre = F.linear(tx, tw, tb)
    ty = fn(pre)
    torch.testing.assert_close(t(y), ty.detach(), **TOL)
    dy = uniform((B, O), 5)
    ty.backward(t(dy))
    dx, dw, db = np.zeros_like(x), np.zeros_like(w), np.zeros_like(b)
    dy2 = dy.copy()
    oracle.linear_backward(x, dx, y, dy2, w, dw, db, I, O, B, act, 0.0, pre.detach().numpy().copy())
    torch.testing.assert_close(t(dx), tx.grad, **TOL)
    torch.testing.assert_close(t(dw), tw.grad, **TOL)
    torch.testing.assert_close(t(db), tb.grad, **TOL)


def test_linear_backward_accumulates_and_regularises(oracle):
    B, I, O = 4, 6, 5
    x, w = uniform((B, I), 1), uniform((O, I), 2)
    dy = uniform((B, O), 3)
    dw0 = uniform((O, I), 4)
    dw = dw0.copy()
    oracle.linear_backward(x, None, None, dy.copy(), w, dw, None, I, O, B, -1, 0.5, None)
    np.testing.assert_allclose(dw, dw0 + dy.T @ x + 0.5 * w, rtol=1e-5, atol=1e-6)


def test_batch_matmul(oracle):
    b, n, k, m = 3, 7, 5, 6
    A, Bm = uniform((b, n, k), 1), uniform((b, k, m), 2)
    o = np.zeros((b, n, m), np.float32)
    oracle.batch_matmul_forward(o, A, Bm, m, n, k, b, -1, -1, -1)
    ta, tb = t(A).requires_grad_(), t(Bm).requires_grad_()
    to = torch.bmm(ta, tb)
    torch.testing.assert_close(t(o), to.detach(), **TOL)
    do = uniform((b, n, m), 3)
    to.backward(t(do))
    da, db = np.zeros_like(A), np.zeros_like(Bm)
    oracle.batch_matmul_backward(do, A, da, Bm, db, m, n, k, b)
    torch.testing.assert_close(t(da), ta.grad, **TOL)
    torch.testing.assert_close(t(db), tb.grad, **TOL)


UNARY = {
    8: lambda v, s: v * s, 9: lambda v, s: v + s, 12: lambda v, s: v - s, 11: lambda v, s: v / s,
    71: lambda v, s: v ** s, 13: lambda v, s: torch.relu(v), 15: lambda v, s: torch.sigmoid(v),
    16: lambda v, s: torch.tanh(v), 17: lambda v, s: F.elu(v, alpha=s), 56: lambda v, s: torch.exp(v),
    14: lambda v, s: v * 1.0, 67: lambda v, s: F.gelu(v), 70: lambda v, s: torch.rsqrt(v),
    61: lambda v, s: torch.sin(v), 62: lambda v, s: torch.cos(v),
}


@pytest.mark.parametrize("op", sorted(UNARY))
def test_element_unary(oracle, op):
    n = 257
    x = uniform((n,), SEED_ACT, 0.2, 2.0) if op in (70, 71) else uniform((n,), SEED_ACT)
    s = 1.5 if op != 17 else 1.0
    y = np.zeros(n, np.float32)
    assert oracle.element_unary_forward(op, s, x, y, n) == 0
    tx = t(x).requires_grad_()
    ty = UNARY[op](tx, s)
    torch.testing.assert_close(t(y), ty.detach(), **TOL)
    dy = uniform((n,), 9)
    ty.backward(t(dy))
    dx = np.zeros(n, np.float32)
    oracle.element_unary_backward(op, s, x, dx, y, dy, n)
    torch.testing.assert_close(t(dx), tx.grad, rtol=1e-5, atol=1e-5)


BIN = {28: torch.add, 35: torch.sub, 29: torch.mul, 36: torch.div, 40: torch.maximum, 41: torch.minimum}


@pytest.mark.parametrize("op", sorted(BIN))
@pytest.mark.parametrize("ls,rs", [((3, 4, 5), (3, 4, 5)), ((3, 1, 5), (1, 4, 1)), ((1, 1, 5), (3, 4, 5))])
def test_element_binary(oracle, op, ls, rs):
    os_ = tuple(max(a, b) for a, b in zip(ls, rs))
    l, r = uniform(ls, 1), uniform(rs, 2)
    if op == 36:
        r = np.sign(r) * (np.abs(r) + 0.5)
    o = np.zeros(os_, np.float32)
    oracle.element_binary_forward(op, l, r, o, 3, ls, rs, os_)
    tl, tr = t(l).requires_grad_(), t(r).requires_grad_()
    to = BIN[op](tl, tr)
    torch.testing.assert_close(t(o), to.detach(), **TOL)
    dy = uniform(os_, 3)
    to.backward(t(dy))
    dl, dr = np.zeros_like(l), np.zeros_like(r)
    oracle.element_binary_backward(op, dy, l, r, dl, dr, 3, ls, rs, os_)
    # continuous random inputs: no ties, so max/min subgradients agree with torch
    torch.testing.assert_close(t(dl), tl.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(t(dr), tr.grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("outer,dim,inner", [(4, 10, 1), (2, 7, 3)])
def test_softmax(oracle, outer, dim, inner):
    x = uniform((outer, dim, inner), SEED_ACT) * 3
    y = np.zeros_like(x)
    oracle.softmax_forward(x, y, outer, dim, inner)
    tx = t(x).requires_grad_()
    ty = torch.softmax(tx, dim=1)
    torch.testing.assert_close(t(y), ty.detach(), **TOL)
    dy = uniform(x.shape, 2)
    ty.backward(t(dy))
    dx = np.zeros_like(x)
    oracle.softmax_backward_jacobian(dx, dy, y, outer, dim, inner)
    torch.testing.assert_close(t(dx), tx.grad, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("affine", [True, False])
def test_layer_norm(oracle, affine):
    M, N, eps = 10, 10, 1e-5  # the reference's fixture size
    x = uniform((M, N), SEED_ACT)
    g = uniform((N,), 2) + 1 if affine else None
    b = uniform((N,), 3) if affine else None
    y, mean, rstd = np.zeros_like(x), np.zeros(M, np.float32), np.zeros(M, np.float32)
    oracle.layer_norm_forward(x, y, g, b, mean, rstd, M, N, eps)
    tx = t(x).requires_grad_()
    tg = t(g).requires_grad_() if affine else None
    tb = t(b).requires_grad_() if affine else None
    ty = F.layer_norm(tx, (N,), tg, tb, eps)
    torch.testing.assert_close(t(y), ty.detach(), rtol=1e-5, atol=1e-5)
    dy = uniform((M, N), 4)
    ty.backward(t(dy))
    dx, dg, db = np.zeros_like(x), np.zeros(N, np.float32), np.zeros(N, np.float32)
    oracle.layer_norm_backward(dy, x, dx, g, dg, db, mean, rstd, M, N, 0)
    torch.testing.assert_close(t(dx), tx.grad, rtol=1e-4, atol=1e-5)
    if affine:
        torch.testing.assert_close(t(dg), tg.grad, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(t(db), tb.grad, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("relu", [False, True])
def test_batch_norm(oracle, relu):
    n, c, h, w, eps = 4, 6, 5, 3, 1e-5
    x = uniform((n, c, h, w), 1)
    sc, bi = uniform((c,), 2) + 1, uniform((c,), 3)
    y = np.zeros_like(x)
    rm, rv, sm, si = (np.zeros(c, np.float32) for _ in range(4))
    oracle.batch_norm_forward(x, y, sc, bi, rm, rv, sm, si, n, c, h, w, eps, 1.0)
    tx, tsc, tbi = t(x).requires_grad_(), t(sc).requires_grad_(), t(bi).requires_grad_()
    trm, trv = torch.zeros(c), torch.zeros(c)
    ty = F.batch_norm(tx, trm, trv, tsc, tbi, training=True, momentum=1.0, eps=eps)
    torch.testing.assert_close(t(y), ty.detach(), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(t(rm), trm, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(t(rv), trv, rtol=1e-5, atol=1e-6)
    dy = uniform(x.shape, 4)
    out = torch.relu(ty) if relu else ty
    out.backward(t(dy))
    dx, dsc, dbi = np.zeros_like(x), np.zeros(c, np.float32), np.zeros(c, np.float32)
    oracle.batch_norm_backward(x, dy.copy(), y, dx, sc, dsc, dbi, sm, si, n, c, h, w, int(relu))
    torch.testing.assert_close(t(dx), tx.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(t(dsc), tsc.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(t(dbi), tbi.grad, rtol=1e-4, atol=1e-5)


def test_embedding_gather_topk(oracle):
    table = uniform((50, 8), 1)
    idx = np.random.default_rng(2).integers(0, 50, (20, 3)).astype(np.int64)
    out = np.zeros((20, 8), np.float32)
    oracle.embedding_forward(idx, out, table, 0, 3, 8, 20)
    tt = t(table).requires_grad_()
    to = F.embedding_bag(t(idx), tt, mode="sum")
    torch.testing.assert_close(t(out), to.detach(), **TOL)
    oracle.embedding_forward(idx, out, table, 1, 3, 8, 20)
    torch.testing.assert_close(t(out), F.embedding_bag(t(idx), t(table), mode="mean"), **TOL)
    dout = uniform((20, 8), 3)
    to.backward(t(dout))
    dtab = np.zeros_like(table)
    oracle.embedding_backward(idx, dout, dtab, 0, 3, 8, 20)
    torch.testing.assert_close(t(dtab), tt.grad, rtol=1e-5, atol=1e-6)
    x = uniform((4, 6, 5), 4)
    gi = np.random.default_rng(5).integers(0, 6, (4, 3, 5)).astype(np.int64)
    go = np.zeros((4, 3, 5), np.float32)
    oracle.gather_forward(x, gi, go, go.size, 5, 6, 3)
    torch.testing.assert_close(t(go), torch.gather(t(x), 1, t(gi)))
    row = uniform((3, 40), 6)
    vals, inds = np.zeros((3, 5), np.float32), np.zeros((3, 5), np.int32)
    oracle.topk_forward(row, vals, inds, 3, 40, 5)
    tv, ti = torch.topk(t(row), 5, dim=1)
    torch.testing.assert_close(t(vals), tv)
    assert (inds == ti.numpy()).all()


def test_losses_and_optimizers(oracle):
    rows, classes = 8, 5
    logits = uniform((rows, classes), 1)
    tl = t(logits).requires_grad_()
    p = torch.softmax(tl, 1)
    lab = np.random.default_rng(2).integers(0, classes, rows).astype(np.int32)
    loss = F.nll_loss(torch.log(p), t(lab).long(), reduction="mean")
    loss.backward()
    g = np.zeros_like(logits)
    # softmax bwd is a copy in the reference, so dLoss/dlogits == the loss kernel's output
    oracle.sparse_cce_loss_backward(g, p.detach().numpy().copy(), lab, rows, classes, 1, 1.0 / rows)
    torch.testing.assert_close(t(g), tl.grad, rtol=1e-5, atol=1e-6)
    n = 100
    w0, gr = uniform((n,), 3), uniform((n,), 4)
    for mom, nest in ((0.0, False), (0.9, False), (0.9, True)):
        w, v = w0.copy(), np.zeros(n, np.float32)
        tw = t(w0).requires_grad_()
        opt = torch.optim.SGD([tw], lr=0.01, momentum=mom, nesterov=nest, weight_decay=1e-4)
        for _ in range(3):
            tw.grad = t(gr)
            opt.step()
            oracle.sgd_update(0.01, mom, int(nest), 1e-4, gr, n, 1, w, v)
        torch.testing.assert_close(t(w), tw.detach(), rtol=1e-5, atol=1e-6)
    w, m, v = w0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)
    tw = t(w0).requires_grad_()
    opt = torch.optim.Adam([tw], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4)
    for step in range(1, 4):
        tw.grad = t(gr)
        opt.step()
        alpha_t = 1e-3 * np.sqrt(1 - 0.999 ** step) / (1 - 0.9 ** step)
        oracle.adam_update(alpha_t, 0.9, 0.999, 1e-4, 1e-8, gr, n, 1, m, v, w)
    # the reference folds the bias correction into alpha_t (eps is not rescaled): 1e-4 agreement
    torch.testing.assert_close(t(w), tw.detach(), rtol=1e-4, atol=1e-6)


def test_layout_ops(oracle):
    x = uniform((3, 4, 5), 1)
    out = np.zeros((5, 3, 4), np.float32)
    oracle.transpose(x, out, 3, [3, 4, 5], [2, 0, 1], 0)
    assert (out == np.transpose(x, (2, 0, 1))).all()
    r = np.zeros_like(x)
    oracle.reverse(x, r, 3, 4, 5)
    assert (r == x[:, ::-1, :]).all()
    red = np.zeros((3, 1, 5), np.float32)
    oracle.reduce_forward(x, red, 3, [3, 4, 5], [3, 1, 5], 1)
    np.testing.assert_allclose(red, x.mean(1, keepdims=True), rtol=1e-6)
    xi = uniform((2, 3, 9, 11), 2)
    for ptype, fn in ((0, F.max_pool2d), (1, lambda a, k, s, p: F.avg_pool2d(a, k, s, p, count_include_pad=False))):
        y = np.zeros((2, 3, 5, 6), np.float32)
        oracle.pool2d(xi, y, None, None, 2, 3, 9, 11, 5, 6, 3, 3, 2, 2, 1, 1, ptype, 0)
        tx = t(xi).requires_grad_()
        ty = fn(tx, 3, 2, 1)
        torch.testing.assert_close(t(y), ty.detach(), **TOL)
        dy = uniform((2, 3, 5, 6), 3)
        ty.backward(t(dy))
        dx = np.zeros_like(xi)
        oracle.pool2d(xi, None, dx, dy, 2, 3, 9, 11, 5, 6, 3, 3, 2, 2, 1, 1, ptype, 1)
        torch.testing.assert_close(t(dx), tx.grad, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("relu", [False, True])
def test_conv2d(oracle, relu):
    n, ci, ih, iw, co, k, s, p, g = 2, 4, 7, 6, 6, 3, 2, 1, 2
    oh, ow = (ih + 2 * p - k) // s + 1, (iw + 2 * p - k) // s + 1
    geo = [n, ci, ih, iw, co, oh, ow, k, k, p, p, s, s, g]
    x, w, b = uniform((n, ci, ih, iw), 1), uniform((co, ci // g, k, k), 2), uniform((co,), 3)
    y = np.zeros((n, co, oh, ow), np.float32)
    oracle.conv2d_forward(geo, x, y, w, b, int(relu))
    tx, tw, tb = t(x).requires_grad_(), t(w).requires_grad_(), t(b).requires_grad_()
    ty = F.conv2d(tx, tw, tb, stride=s, padding=p, groups=g)
    if relu:
        ty = torch.relu(ty)
    torch.testing.assert_close(t(y), ty.detach(), **TOL)
    dy = uniform((n, co, oh, ow), 4)
    ty.backward(t(dy))
    dx, dw, db = np.zeros_like(x), np.zeros_like(w), np.zeros_like(b)
    oracle.conv2d_backward(geo, x, dx, y, dy.copy(), w, dw, db, int(relu))
    torch.testing.assert_close(t(dx), tx.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(t(dw), tw.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(t(db), tb.grad, rtol=1e-5, atol=1e-5)
