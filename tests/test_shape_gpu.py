"""GPU parity: Concat / Split / Transpose / Reverse / Reduce / Pool2D vs the oracle (and
torch's own layout ops where the result must be bit-identical)."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_W, assert_close, to_dev, uniform

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shapes,axis", [([(4, 3, 8), (4, 5, 8)], 1), ([(2, 6), (3, 6), (1, 6)], 0),
                                         ([(5, 7, 3), (5, 7, 4)], 2), ([(100, 100), (100, 100)], 1)])
def test_concat(K, oracle, shapes, axis):
    ins = [to_dev(uniform(s, SEED_ACT + i)) for i, s in enumerate(shapes)]
    oshape = list(shapes[0])
    oshape[axis] = sum(s[axis] for s in shapes)
    out = torch.zeros(oshape, device="cuda")
    K.Concat.forward_kernel(None, out, [t for t, _ in ins], axis)
    assert torch.equal(out, torch.cat([t for t, _ in ins], dim=axis))  # pure copies: bit-exact
    g, gn = to_dev(uniform(tuple(oshape), SEED_W))
    grads = [to_dev(uniform(s, SEED_W + 1 + i)) for i, s in enumerate(shapes)]
    expect = [gr + part for (gr, _), part in zip(grads, torch.split(g, [s[axis] for s in shapes], dim=axis))]
    K.Concat.backward_kernel(None, g, [t for t, _ in grads], axis)
    for (gr, _), e in zip(grads, expect):
        assert torch.equal(gr, e)
    # oracle statement of the primitive
    nb, oblk = K.Concat._blocks(oshape, axis)
    ref = np.zeros(oshape, np.float32)
    off = 0
    flat = ref.reshape(-1)
    for (_, n), s in zip(ins, shapes):
        _, iblk = K.Concat._blocks(s, axis)
        oracle.copy_with_stride(flat[off:], n, nb, oblk, iblk, 0)
        off += iblk
    assert (out.cpu().numpy() == ref).all()


def test_split(K):
    # reference fixture: 2 outputs, block sizes 50 + 50 of 100 (lib/kernels/test/src/test_split_kernel.cc:11-16)
    x, _ = to_dev(uniform((4, 100), SEED_ACT))
    outs = [torch.zeros(4, 50, device="cuda"), torch.zeros(4, 50, device="cuda")]
    K.Split.forward_kernel(None, outs, x, [50, 50], 100, 4)
    for o, e in zip(outs, torch.split(x, 50, dim=1)):
        assert torch.equal(o, e)
    xg = torch.ones(4, 100, device="cuda")
    K.Split.backward_kernel(None, xg, outs, [50, 50], 100, 4)
    assert torch.equal(xg, 1 + x)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape,perm", [((10, 10), (1, 0)), ((3, 40, 70), (0, 2, 1)), ((4, 5, 6), (2, 0, 1)),
                                        ((2, 3, 4, 5), (3, 1, 0, 2)), ((7,), (0,)), ((2, 2, 3, 2, 4), (4, 3, 2, 1, 0))])
def test_transpose(K, oracle, shape, perm, dtype):
    x, xn = to_dev(uniform(shape, SEED_ACT), dtype)
    oshape = tuple(shape[p] for p in perm)
    out = torch.zeros(oshape, dtype=dtype, device="cuda")
    m = K.Transpose.init_kernel(len(shape), perm)
    K.Transpose.forward_kernel(None, m, x, out)
    assert torch.equal(out, x.permute(perm).contiguous())
    ref = np.zeros(oshape, np.float32)
    oracle.transpose(xn, ref, len(shape), list(shape), list(perm), 0)
    assert (out.float().cpu().numpy() == ref).all()
    g, gn = to_dev(uniform(oshape, SEED_W), dtype)
    xg, xgn = to_dev(uniform(shape, SEED_W + 1), dtype)
    K.Transpose.backward_kernel(None, m, xg, g)
    inv = np.argsort(perm).tolist()
    oracle.transpose(gn, xgn, len(shape), list(oshape), inv, 1)
    assert_close(xg, xgn, dtype, what="transpose bwd")


def test_reverse(K, oracle):
    # reference fixture: 1 x 10 x 10 (lib/kernels/test/src/test_reverse_kernels.cc:9-12)
    for blks, rs, inner in [(1, 10, 10), (3, 7, 1), (5, 2, 33)]:
        x, xn = to_dev(uniform((blks, rs, inner), SEED_ACT))
        out = torch.zeros_like(x)
        K.Reverse.forward_kernel(None, x, out, blks, rs, inner)
        assert torch.equal(out, torch.flip(x, [1]))
        ref = np.zeros_like(xn)
        oracle.reverse(xn, ref, blks, rs, inner)
        assert (out.cpu().numpy() == ref).all()


@pytest.mark.parametrize("op", ["REDUCE_SUM", "REDUCE_MEAN"])
@pytest.mark.parametrize("ishape,oshape", [((8, 4096), (8, 1)), ((10, 10, 10, 10), (10, 1, 1, 10)),
                                           ((6, 5, 4), (1, 5, 1)), ((3, 7), (3, 7)), ((33, 9), (1, 1))])
def test_reduce(K, oracle, op, ishape, oshape):
    code = getattr(K.OperatorType, op)
    x, xn = to_dev(uniform(ishape, SEED_ACT))
    out = torch.zeros(oshape, device="cuda")
    K.Reduce.forward_kernel(None, code, x, out)
    ref = np.zeros(oshape, np.float32)
    oracle.reduce_forward(xn, ref, len(ishape), list(ishape), list(oshape), int(op == "REDUCE_MEAN"))
    assert_close(out, ref, what=f"reduce {op} {ishape}->{oshape}")
    g, gn = to_dev(uniform(oshape, SEED_W))
    xg, xgn = to_dev(uniform(ishape, SEED_W + 1))
    K.Reduce.backward_kernel(None, code, g, xg)
    oracle.reduce_backward(gn, xgn, len(ishape), list(ishape), list(oshape), int(op == "REDUCE_MEAN"))
    assert_close(xg, xgn, what="reduce bwd")


@pytest.mark.parametrize("ptype", ["MAX", "AVG"])
@pytest.mark.parametrize("n,c,ih,iw,k,s,p", [(10, 10, 10, 10, 2, 2, 0), (2, 3, 9, 11, 3, 2, 1), (1, 4, 7, 7, 3, 1, 1)])
def test_pool2d(K, oracle, ptype, n, c, ih, iw, k, s, p):
    code = getattr(K, "POOL_" + ptype)
    oh, ow = (ih + 2 * p - k) // s + 1, (iw + 2 * p - k) // s + 1
    x, xn = to_dev(uniform((n, c, ih, iw), SEED_ACT))
    y = torch.zeros(n, c, oh, ow, device="cuda")
    m = K.Pool2D.init_kernel(iw, ih, c, n, ow, oh, c, n, p, p, k, k, s, s, code)
    K.Pool2D.forward_kernel(None, m, x, y)
    ref = np.zeros((n, c, oh, ow), np.float32)
    oracle.pool2d(xn, ref, None, None, n, c, ih, iw, oh, ow, k, k, s, s, p, p, code, 0)
    assert_close(y, ref, rtol=1e-6, what=f"pool2d {ptype}")
    dy, dyn = to_dev(uniform((n, c, oh, ow), SEED_W))
    dx, dxn = to_dev(uniform((n, c, ih, iw), SEED_W + 1))
    K.Pool2D.backward_kernel(None, m, x, dx, y, dy)
    oracle.pool2d(xn, None, dxn, dyn, n, c, ih, iw, oh, ow, k, k, s, s, p, p, code, 1)
    assert_close(dx, dxn, rtol=1e-5, what=f"pool2d bwd {ptype}")


@pytest.mark.parametrize("relu", [False, True])
@pytest.mark.parametrize("n,ci,ih,iw,co,k,s,p,g", [(2, 3, 9, 11, 8, 3, 1, 1, 1), (3, 4, 8, 8, 6, 3, 2, 0, 2),
                                                   (1, 16, 12, 12, 32, 1, 1, 0, 1), (2, 6, 7, 5, 6, 5, 2, 2, 3)])
def test_conv2d(K, oracle, n, ci, ih, iw, co, k, s, p, g, relu):
    oh, ow = (ih + 2 * p - k) // s + 1, (iw + 2 * p - k) // s + 1
    geo = [n, ci, ih, iw, co, oh, ow, k, k, p, p, s, s, g]
    x, xn = to_dev(uniform((n, ci, ih, iw), SEED_ACT))
    w, wn = to_dev(uniform((co, ci // g, k, k), SEED_W) / np.sqrt(ci // g * k * k))
    b, bn = to_dev(uniform((co,), SEED_W + 1))
    m = K.Conv2D.init_kernel(K.ACT_RELU if relu else None, k, k, g, p, p, s, s, (n, ci, ih, iw), (n, co, oh, ow))
    y = torch.zeros(n, co, oh, ow, device="cuda")
    K.Conv2D.forward_kernel(None, m, x, y, w, b)
    yref = np.zeros((n, co, oh, ow), np.float32)
    oracle.conv2d_forward(geo, xn, yref, wn, bn, int(relu))
    assert_close(y, yref, what="conv2d fwd")
    dy, dyn = to_dev(uniform((n, co, oh, ow), SEED_ACT + 1))
    dx, dxn = to_dev(uniform((n, ci, ih, iw), SEED_ACT + 2))
    dw, dwn = to_dev(uniform((co, ci // g, k, k), SEED_W + 2))
    db, dbn = to_dev(uniform((co,), SEED_W + 3))
    K.Conv2D.backward_kernel(None, m, x, dx, y, dy, w, dw, db)
    oracle.conv2d_backward(geo, xn, dxn, y.cpu().numpy(), dyn, wn, dwn, dbn, int(relu))
    assert_close(dx, dxn, what="conv2d dX")
    assert_close(dw, dwn, what="conv2d dW")
    assert_close(db, dbn, what="conv2d db")
